// Hermite4B200 member functions: one C-ABI call each (see Hermite4B200.hpp).  C++98, compiled with the reference's
// own headers.  No CUDA, no particle arithmetic on this side of the boundary.
#include <cstdlib>
#include <cstring>
#include "Hermite4B200.hpp"

Hermite4B200::Hermite4B200(NbodySystem *ns, Logger *logger, NbodyUtils *nu) : Hermite4(ns, logger, nu)
{
    ctx = NULL;
    const char *h = getenv("GRAVIDY_B200_HOOKS");
    hooks_mode = h != NULL && strcmp(h, "0") != 0;
    // the reference's -g option counts GPUs (src/include/utils/OptionsParser.cpp:75); this binding drives one
    // (device 0): multi-GPU runs go through gdy_create_sharded as gravidy_b200/host/Hermite4B200.cpp shows
    check(gdy_create(&ctx, ns->n, ns->e2, ns->eta, 0), "gdy_create");
    // double4 {x,y,z,w} and Forces {a[3],a1[3]} are plain doubles: the reference's arrays ARE the C ABI's layout
    check(gdy_upload_particles(ctx, &ns->h_r[0].x, &ns->h_v[0].x), "gdy_upload_particles");
}

Hermite4B200::~Hermite4B200()
{
    gdy_destroy(ctx);
}

void Hermite4B200::check(int rc, const char *what)
{
    if (rc == GDY_OK) return;
    // the reference's convention: report and exit (src/include/gpu/Hermite4GPU.cuh:43-50)
    std::cerr << "[ERROR] " << what << ": " << gdy_last_error(ctx) << std::endl;
    exit(EXIT_FAILURE);
}

void Hermite4B200::init_acc_jrk()
{
    check(gdy_init_acc_jrk(ctx), "gdy_init_acc_jrk");
    // Hermite4::init_dt (host, libm log/pow: src/include/Hermite4.cpp:108-140) reads h_f once
    check(gdy_download_forces(ctx, &ns->h_f[0].a[0]), "gdy_download_forces");
}

void Hermite4B200::upload_times()
{
    check(gdy_upload_times(ctx, ns->h_t, ns->h_dt), "gdy_upload_times");
}

unsigned int Hermite4B200::find_particles_to_move(double ITIME)
{
    uint32_t nact = 0;
    check(gdy_find_particles_to_move(ctx, ITIME, &nact), "gdy_find_particles_to_move");
    return nact;
}

void Hermite4B200::save_old_acc_jrk(unsigned int nact)
{
    check(gdy_save_old_acc_jrk(ctx, nact), "gdy_save_old_acc_jrk");
}

void Hermite4B200::predicted_pos_vel(double ITIME)
{
    ns->gtime.prediction_ini = omp_get_wtime();
    check(gdy_predicted_pos_vel(ctx, ITIME), "gdy_predicted_pos_vel");
    ns->gtime.prediction_end += omp_get_wtime() - ns->gtime.prediction_ini;
}

void Hermite4B200::update_acc_jrk(unsigned int nact)
{
    ns->gtime.update_ini = omp_get_wtime();
    check(gdy_update_acc_jrk(ctx, nact), "gdy_update_acc_jrk");
    check(gdy_sync(ctx), "gdy_sync");   // so that the update timer means what the log's GFLOPS column says
    ns->gtime.update_end += omp_get_wtime() - ns->gtime.update_ini;
}

void Hermite4B200::correction_pos_vel(double ITIME, unsigned int nact)
{
    ns->gtime.correction_ini = omp_get_wtime();
    check(gdy_correction_pos_vel(ctx, ITIME, nact), "gdy_correction_pos_vel");
    ns->gtime.correction_end += omp_get_wtime() - ns->gtime.correction_ini;
}

void Hermite4B200::next_integration_time(double &ATIME)
{
    check(gdy_next_integration_time(ctx, &ATIME), "gdy_next_integration_time");
}

double Hermite4B200::get_energy_gpu()
{
    double ek = 0.0, ep = 0.0;
    check(gdy_energy(ctx, &ek, &ep), "gdy_energy");
    ns->en.kinetic = ek;
    ns->en.potential = ep;
    return ek + ep;
}

void Hermite4B200::refresh_mirrors()
{
    check(gdy_download_state(ctx, &ns->h_r[0].x, &ns->h_v[0].x, &ns->h_f[0].a[0], NULL, NULL, NULL, ns->h_t, ns->h_dt,
                             NULL, NULL), "gdy_download_state");
}
