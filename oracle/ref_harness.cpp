// TEST INFRASTRUCTURE ONLY.  Never linked into, imported by or called from the product path.
//
// C-callable harness around the UNMODIFIED reference sources under /root/reference/src, compiled
// where they lie by oracle/Makefile into oracle/_ref/libgravidy_ref.so.  It exposes the hooks of
// the reference's Hermite4CPU back end (reference: src/include/cpu/Hermite4CPU.hpp:56-68 and
// src/include/Hermite4.hpp:60-63) so that tests can drive them one by one on the same inputs as the
// CUDA path, and a driver that calls them in the order of the reference's own loop
// (src/include/cpu/Hermite4CPU_integration.cpp:81-124).
//
// The hooks are private members in the reference; this TU (and only this TU) is compiled with
// -fno-access-control.  All arithmetic executed below is the reference's own object code.
#include <stdint.h>
#include <cstring>
#include <cstdlib>
#include <limits>

#include "cpu/Hermite4CPU.hpp"

struct gref
{
    OptionsParser *op;
    NbodySystem *ns;
    Logger *logger;
    NbodyUtils *nu;
    Hermite4CPU *h4;
};

struct gref_run_stats
{
    uint64_t iterations;   // block steps done
    uint64_t nsteps;       // sum of nact
    uint64_t interactions; // sum of nact * n, 64-bit (the reference wraps at 2^32)
    double itime;          // ITIME at exit
    double atime;          // next block time at exit
    double update_s;       // seconds spent in update_acc_jrk
    double wall_s;         // seconds in the loop
};

extern "C" {

gref *gref_create_from_file(const char *input_path, const char *output_prefix, float softening,
                            float eta, float integration_time, float interval_time)
{
    gref *g = new gref;
    char arg0[] = "gravidy-ref";
    char *argv[] = {arg0, NULL};
    g->op = new OptionsParser(1, argv);
    g->op->input_filename = input_path;
    g->op->output_filename = std::string(output_prefix) + ".out"; // OptionsParser.cpp:402-417
    g->op->softening = softening;
    g->op->eta = eta;
    g->op->integration_time = integration_time;
    g->op->interval_time = interval_time;
    // OptionsParser.cpp:395-401
    g->op->length_output_number = 0;
    unsigned int tmp = (unsigned int)(integration_time / interval_time);
    do { ++g->op->length_output_number; tmp /= 10; } while (tmp);
    g->op->length_output_number += 1;

    // Same sequence as main(): src/gravidy.cpp:63-96
    g->ns = new NbodySystem(*g->op);
    g->ns->read_input_file();
    g->ns->alloc_base_attributes(0);
    g->ns->copy_input_data();
    g->logger = new Logger(g->ns);
    g->nu = new NbodyUtils(g->ns);
    g->h4 = new Hermite4CPU(g->ns, g->logger, g->nu);
    return g;
}

void gref_destroy(gref *g)
{
    if (!g) return;
    delete g->h4;
    delete g->nu;
    delete g->logger;
    delete g->ns;
    delete g->op;
    delete g;
}

uint32_t gref_n(gref *g) { return g->ns->n; }
double gref_e2(gref *g) { return g->ns->e2; }
float gref_eta(gref *g) { return g->ns->eta; }
float gref_total_mass(gref *g) { return g->ns->total_mass; }
void gref_set_threads(int nthreads) { omp_set_num_threads(nthreads); }
int gref_max_threads(void) { return omp_get_max_threads(); }

// ---- the hooks, one call each ------------------------------------------------------------
void gref_init_acc_jrk(gref *g) { g->h4->init_acc_jrk(); }

double gref_init_dt(gref *g, double itime)
{
    double atime = 1.0e+10; // Hermite4CPU_integration.cpp:43
    g->h4->init_dt(atime, ETA_S, itime); // Hermite4CPU_integration.cpp:54
    return atime;
}

uint32_t gref_find_particles_to_move(gref *g, double itime)
{
    return g->h4->find_particles_to_move(itime);
}
void gref_save_old_acc_jrk(gref *g, uint32_t nact) { g->h4->save_old_acc_jrk(nact); }
void gref_predicted_pos_vel(gref *g, double itime) { g->h4->predicted_pos_vel(itime); }
void gref_update_acc_jrk(gref *g, uint32_t nact) { g->h4->update_acc_jrk(nact); }
void gref_correction_pos_vel(gref *g, double itime, uint32_t nact)
{
    g->h4->correction_pos_vel(itime, nact);
}
double gref_next_integration_time(gref *g)
{
    double atime = 0.0;
    g->h4->next_integration_time(atime);
    return atime;
}
double gref_get_energy(gref *g, double *ekin, double *epot)
{
    double e = g->nu->get_energy(0);
    if (ekin) *ekin = g->ns->en.kinetic;
    if (epot) *epot = g->ns->en.potential;
    return e;
}
double gref_get_timestep_normal(gref *g, uint32_t i, float eta)
{
    return g->nu->get_timestep_normal(i, eta);
}
double gref_normalize_dt(gref *g, double new_dt, double old_dt, double t)
{
    return g->nu->normalize_dt(new_dt, old_dt, t, 0);
}

// NbodyUtils::get_center_of_density (reference: src/include/utils/NbodyUtils.cpp:163-220): N sorts of N distances,
// density from the J-th neighbour (J = 6, common.hpp:96), float accumulators.  The reference's OpenMP loop shares
// radius/aa/bb between threads: call it with ONE thread (gref_set_threads(1)) for a defined result.
void gref_center_of_density(gref *g, double *out3)
{
    double3 c = g->nu->get_center_of_density();
    out3[0] = c.x; out3[1] = c.y; out3[2] = c.z;
}
int gref_J(void) { return J; }

// The reference's own driver, unmodified: writes <prefix>.out.{info,log,snapshot_*}.
void gref_integration(gref *g) { g->h4->integration(); }

// -a / -l of the command line (OptionsParser.cpp:372-376): all-particle dumps and Lagrange radii at
// every output point.
void gref_set_print_options(gref *g, int print_all, int print_lagrange)
{
    g->ns->ops.print_all = print_all;
    g->ns->ops.print_lagrange = print_lagrange;
}

// -l in the reference writes past the end of layers_radii: get_layers (NbodyUtils.cpp:300-328) keeps counting layers
// after the sixth, and the vector holds six (NbodyUtils.cpp:46).  The member is public, so the harness gives it room
// for one entry per particle BEFORE the run; the reference's code itself is untouched and the logger prints the first
// six entries only (Logger.cpp:257-261).
void gref_pad_layers(gref *g)
{
    g->nu->layers_radii.resize(g->ns->n + 16, 0.0);
}
// NbodyUtils::lagrange_radii (NbodyUtils.cpp:262-278) on the current positions: density centre, sorted radii, layers.
void gref_lagrange_radii(gref *g, double *out6)
{
    gref_pad_layers(g);
    g->nu->lagrange_radii();
    for (int k = 0; k < 6; ++k) out6[k] = g->nu->layers_radii[k];
}
// get_core_radius (NbodyUtils.cpp:113-144) and get_half_mass_relaxation_time (:149-160, sets ns->r_hm) around the
// density centre the previous call left in nu->cod.  out3 = {r_core, t_rlx, r_hm}.
void gref_core_and_relaxation(gref *g, double *out3)
{
    out3[0] = g->nu->get_core_radius();
    out3[1] = g->nu->get_half_mass_relaxation_time();
    out3[2] = g->ns->r_hm;
}

// ---- state copy-out / copy-in (plain doubles; NULL pointers are skipped) -------------------
// r4: n x {x,y,z,m}   v4: n x {x,y,z,0}   f6/old6: n x {a[3],a1[3]}   a2/a3: n x 3
// p7: n x {r[3],v[3],m} with m widened from float       move: n x uint32
void gref_get_state(gref *g, double *r4, double *v4, double *f6, double *old6, double *a2,
                    double *a3, double *t, double *dt, double *p7, uint32_t *move)
{
    NbodySystem *ns = g->ns;
    for (unsigned int i = 0; i < ns->n; i++)
    {
        if (r4) { r4[4*i] = ns->h_r[i].x; r4[4*i+1] = ns->h_r[i].y; r4[4*i+2] = ns->h_r[i].z; r4[4*i+3] = ns->h_r[i].w; }
        if (v4) { v4[4*i] = ns->h_v[i].x; v4[4*i+1] = ns->h_v[i].y; v4[4*i+2] = ns->h_v[i].z; v4[4*i+3] = 0.0; }
        if (f6) for (int k = 0; k < 3; k++) { f6[6*i+k] = ns->h_f[i].a[k]; f6[6*i+3+k] = ns->h_f[i].a1[k]; }
        if (old6) for (int k = 0; k < 3; k++) { old6[6*i+k] = ns->h_old[i].a[k]; old6[6*i+3+k] = ns->h_old[i].a1[k]; }
        if (a2) { a2[3*i] = ns->h_a2[i].x; a2[3*i+1] = ns->h_a2[i].y; a2[3*i+2] = ns->h_a2[i].z; }
        if (a3) { a3[3*i] = ns->h_a3[i].x; a3[3*i+1] = ns->h_a3[i].y; a3[3*i+2] = ns->h_a3[i].z; }
        if (t) t[i] = ns->h_t[i];
        if (dt) dt[i] = ns->h_dt[i];
        if (p7) { for (int k = 0; k < 3; k++) { p7[7*i+k] = ns->h_p[i].r[k]; p7[7*i+3+k] = ns->h_p[i].v[k]; } p7[7*i+6] = ns->h_p[i].m; }
        if (move) move[i] = ns->h_move[i];
    }
}

void gref_set_state(gref *g, const double *r4, const double *v4, const double *f6,
                    const double *old6, const double *a2, const double *a3, const double *t,
                    const double *dt, const double *p7, const uint32_t *move)
{
    NbodySystem *ns = g->ns;
    for (unsigned int i = 0; i < ns->n; i++)
    {
        if (r4) { ns->h_r[i].x = r4[4*i]; ns->h_r[i].y = r4[4*i+1]; ns->h_r[i].z = r4[4*i+2]; ns->h_r[i].w = r4[4*i+3]; }
        if (v4) { ns->h_v[i].x = v4[4*i]; ns->h_v[i].y = v4[4*i+1]; ns->h_v[i].z = v4[4*i+2]; }
        if (f6) for (int k = 0; k < 3; k++) { ns->h_f[i].a[k] = f6[6*i+k]; ns->h_f[i].a1[k] = f6[6*i+3+k]; }
        if (old6) for (int k = 0; k < 3; k++) { ns->h_old[i].a[k] = old6[6*i+k]; ns->h_old[i].a1[k] = old6[6*i+3+k]; }
        if (a2) { ns->h_a2[i].x = a2[3*i]; ns->h_a2[i].y = a2[3*i+1]; ns->h_a2[i].z = a2[3*i+2]; }
        if (a3) { ns->h_a3[i].x = a3[3*i]; ns->h_a3[i].y = a3[3*i+1]; ns->h_a3[i].z = a3[3*i+2]; }
        if (t) ns->h_t[i] = t[i];
        if (dt) ns->h_dt[i] = dt[i];
        if (p7) { for (int k = 0; k < 3; k++) { ns->h_p[i].r[k] = p7[7*i+k]; ns->h_p[i].v[k] = p7[7*i+3+k]; } ns->h_p[i].m = (float)p7[7*i+6]; }
        if (move) ns->h_move[i] = move[i];
    }
}

// ---- the reference loop without its file output ---------------------------------------------
// Same call order as Hermite4CPU_integration.cpp:81-124; the caller has already run
// gref_init_acc_jrk + gref_init_dt (their ATIME is passed in).  Stops when ITIME >= t_end or after
// max_steps block steps (0 = unlimited).  trace_nact / trace_itime (may be NULL) receive the first
// trace_cap steps' active counts and block times.
void gref_run(gref *g, double atime, double t_end, uint64_t max_steps, gref_run_stats *out,
              uint32_t *trace_nact, double *trace_itime, uint64_t trace_cap)
{
    double ATIME = atime;
    double ITIME = g->ns->snapshot_time;
    uint64_t it = 0, nsteps = 0, inter = 0;
    double upd0 = g->ns->gtime.update_end;
    double t0 = omp_get_wtime();
    double upd = 0.0;
    while (ITIME < t_end)
    {
        if (max_steps && it >= max_steps) break;
        ITIME = ATIME;
        unsigned int nact = g->h4->find_particles_to_move(ITIME);
        g->h4->save_old_acc_jrk(nact);
        g->h4->predicted_pos_vel(ITIME);
        double u0 = omp_get_wtime();
        g->h4->update_acc_jrk(nact);
        upd += omp_get_wtime() - u0;
        g->h4->correction_pos_vel(ITIME, nact);
        inter += (uint64_t)nact * (uint64_t)g->ns->n;
        g->h4->next_integration_time(ATIME);
        if (it < trace_cap)
        {
            if (trace_nact) trace_nact[it] = nact;
            if (trace_itime) trace_itime[it] = ITIME;
        }
        nsteps += nact;
        it++;
        g->ns->iterations++;
    }
    (void)upd0;
    if (out)
    {
        out->iterations = it;
        out->nsteps = nsteps;
        out->interactions = inter;
        out->itime = ITIME;
        out->atime = ATIME;
        out->update_s = upd;
        out->wall_s = omp_get_wtime() - t0;
    }
}

// Timed force sweep over an explicit active list (bench cpu_baseline / --impl reference leg):
// loads `move`, runs the reference's update_acc_jrk(nact) on the current predictors.
double gref_timed_update(gref *g, const uint32_t *move, uint32_t nact)
{
    for (uint32_t k = 0; k < nact; k++) g->ns->h_move[k] = move[k];
    double t0 = omp_get_wtime();
    g->h4->update_acc_jrk(nact);
    return omp_get_wtime() - t0;
}

} // extern "C"
