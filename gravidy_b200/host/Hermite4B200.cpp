#include "Hermite4B200.hpp"

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <thread>

namespace gravidy {

namespace {
double now_s()
{
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
}  // namespace

// Errors keep the reference's convention: report and exit(EXIT_FAILURE)
// (reference: CSC macro, src/include/gpu/Hermite4GPU.cuh:43-50).
void Hermite4B200::check(int rc, int shard, const char *what)
{
    if (rc == GDY_OK) return;
    std::ostringstream m;
    m << what << " failed on shard " << shard << " (code " << rc << "): "
      << gdy_last_error(shard < (int)shards.size() ? shards[shard] : nullptr);
    logger->log_error(m.str());
    std::exit(EXIT_FAILURE);
}

void Hermite4B200::each_shard(const std::function<void(int)> &fn)
{
    if (ngpus == 1) { fn(0); return; }
    std::vector<std::thread> th;
    for (int g = 0; g < ngpus; ++g) th.emplace_back(fn, g);
    for (std::thread &t : th) t.join();
}

// Replaces Hermite4GPU::Hermite4GPU + alloc_arrays_device (reference: Hermite4GPU.cu:43-165).
// -g 0 (default) = every visible GPU, rounded down to 1, 2, 4 or 8 shards.
Hermite4B200::Hermite4B200(NbodySystem *ns_, Logger *logger_, NbodyUtils *nu_) : Hermite4(ns_, logger_, nu_)
{
    int want = ns->gpus;
    int avail = gdy_device_count();
    if (avail <= 0)
    {
        logger->log_error("no CUDA device found; this build has no CPU back end");
        std::exit(EXIT_FAILURE);
    }
    if (want <= 0 || want > avail) want = avail;
    ngpus = 1;
    while (ngpus * 2 <= want && ngpus < 8) ngpus *= 2;
    {
        std::ostringstream m;
        m << "GPUs: " << ngpus << " (j-shards of " << (ns->n + ngpus - 1) / ngpus << " particles)";
        logger->log_info(m.str());
    }
    shards.assign(ngpus, nullptr);
    uint8_t uid[GDY_NCCL_UID_BYTES] = {0};
    if (ngpus > 1) check(gdy_nccl_unique_id(uid), 0, "gdy_nccl_unique_id");
    each_shard([&](int g) {
        check(gdy_create_sharded(&shards[g], ns->n, ns->e2, ns->eta, g, g, ngpus, ngpus > 1 ? uid : nullptr), g,
              "gdy_create_sharded");
        check(gdy_upload_particles(shards[g], &ns->h_r[0].x, &ns->h_v[0].x), g, "gdy_upload_particles");
    });
    if (ngpus > 1)
    {
        // peer memory for the fused NVLink exchange of the persistent small-block kernel
        std::vector<uint8_t> handles((size_t)ngpus * GDY_PEER_HANDLE_BYTES);
        each_shard([&](int g) { check(gdy_peer_export(shards[g], &handles[(size_t)g * GDY_PEER_HANDLE_BYTES]), g, "gdy_peer_export"); });
        each_shard([&](int g) { check(gdy_peer_attach(shards[g], handles.data()), g, "gdy_peer_attach"); });
    }
    nu->neighbour_distance = [this](int k, std::vector<double> &out) {
        out.resize(ns->n);
        check(gdy_neighbour_distance(shards[0], k, out.data()), 0, "gdy_neighbour_distance");
    };
}

Hermite4B200::~Hermite4B200()
{
    for (gdy_ctx *c : shards) gdy_destroy(c);
}

// ---- hooks ---------------------------------------------------------------------------------------------
void Hermite4B200::init_acc_jrk()
{
    each_shard([&](int g) { check(gdy_init_acc_jrk(shards[g]), g, "gdy_init_acc_jrk"); });
    // the host init_dt needs the forces once
    check(gdy_download_forces(shards[0], &ns->h_f[0].a[0]), 0, "gdy_download_forces");
}

void Hermite4B200::upload_times()
{
    each_shard([&](int g) { check(gdy_upload_times(shards[g], ns->h_t, ns->h_dt), g, "gdy_upload_times"); });
}

unsigned int Hermite4B200::find_particles_to_move(double ITIME)
{
    std::vector<uint32_t> nact(ngpus, 0);
    each_shard([&](int g) { check(gdy_find_particles_to_move(shards[g], ITIME, &nact[g]), g, "gdy_find_particles_to_move"); });
    return nact[0];
}

void Hermite4B200::save_old_acc_jrk(unsigned int nact)
{
    each_shard([&](int g) { check(gdy_save_old_acc_jrk(shards[g], nact), g, "gdy_save_old_acc_jrk"); });
}

void Hermite4B200::predicted_pos_vel(double ITIME)
{
    const double t0 = now_s();
    each_shard([&](int g) { check(gdy_predicted_pos_vel(shards[g], ITIME), g, "gdy_predicted_pos_vel"); });
    ns->gtime.prediction_end += now_s() - t0;
}

void Hermite4B200::update_acc_jrk(unsigned int nact)
{
    const double t0 = now_s();
    each_shard([&](int g) {
        check(gdy_update_acc_jrk(shards[g], nact), g, "gdy_update_acc_jrk");
        check(gdy_sync(shards[g]), g, "gdy_sync");   // so that the update timer means what the log says
    });
    ns->gtime.update_end += now_s() - t0;
}

void Hermite4B200::correction_pos_vel(double ITIME, unsigned int nact)
{
    const double t0 = now_s();
    each_shard([&](int g) { check(gdy_correction_pos_vel(shards[g], ITIME, nact), g, "gdy_correction_pos_vel"); });
    ns->gtime.correction_end += now_s() - t0;
}

void Hermite4B200::next_integration_time(double &ATIME)
{
    std::vector<double> at(ngpus, 0.0);
    each_shard([&](int g) { check(gdy_next_integration_time(shards[g], &at[g]), g, "gdy_next_integration_time"); });
    ATIME = at[0];
}

double Hermite4B200::get_energy_gpu()
{
    // collective on a sharded run: every rank sums its i-chunks and the chunk sums are all-reduced (same bits on every rank)
    std::vector<double> ek(ngpus, 0.0), ep(ngpus, 0.0);
    each_shard([&](int g) { check(gdy_energy(shards[g], &ek[g], &ep[g]), g, "gdy_energy"); });
    return ek[0] + ep[0];
}

void Hermite4B200::refresh_mirrors(bool forces)
{
    check(gdy_download_state(shards[0], &ns->h_r[0].x, &ns->h_v[0].x, forces ? &ns->h_f[0].a[0] : nullptr, nullptr,
                             nullptr, nullptr, forces ? ns->h_t : nullptr, forces ? ns->h_dt : nullptr, nullptr,
                             nullptr),
          0, "gdy_download_state");
}

// Fused driver: the loop body of the reference for every block step up to the first one with
// ITIME >= t_stop (at least one step), without leaving the device (one 16-byte readback per block step).
gdy_run_stats Hermite4B200::run_until(double t_stop, double &ATIME, double &ITIME)
{
    std::vector<gdy_run_stats> st(ngpus);
    const double at = ATIME, it = ITIME;
    const bool one_step = !(it < t_stop);   // an output is already due: the reference takes one step per output
    const double stop = one_step ? (double)ns->integration_time : t_stop;
    each_shard([&](int g) {
        check(gdy_run(shards[g], at, it, stop, one_step ? 1 : 0, &st[g], nullptr, nullptr, nullptr, 0), g, "gdy_run");
    });
    ATIME = st[0].atime;
    ITIME = st[0].itime;
    ns->gtime.update_end += st[0].force_ms * 1e-3;
    return st[0];
}

// ---- checkpoint ------------------------------------------------------------------------------------------
// The reference can only resume from a TEXT snapshot (src/include/utils/Logger.cpp:77-111 writes it with 4-6
// digits, src/include/utils/OptionsParser.cpp:122-324 reads it back), i.e. it restarts from rounded
// positions with freshly initialised forces and timesteps: the continued run is a different run.  The
// checkpoint keeps what the loop needs in full precision — r, v, acc, jerk, t, dt, the loop counters and
// the energy references — so that `--restart` continues the block-timestep schedule bit for bit
// (a2, a3 and the old forces are rewritten by every step before they are read, so they are not part of it).
namespace {
struct CkptHeader
{
    char magic[8];
    uint32_t n, snap_number, output_factor, iterations, nsteps, reserved;
    int64_t interactions;
    double e2;
    float eta, reserved2;
    double ITIME, ATIME, e_ini, e_tmp;
    double update_end;   // seconds spent in force evaluations so far: the log's GFLOPS column is interactions / this
};
const char CKPT_MAGIC[8] = {'G', 'D', 'Y', 'C', 'K', 'P', 'T', '2'};
}  // namespace

void Hermite4B200::write_checkpoint(const LoopState &ls)
{
    refresh_mirrors(true);
    CkptHeader h;
    std::memset(&h, 0, sizeof(h));
    std::memcpy(h.magic, CKPT_MAGIC, 8);
    h.n = ns->n; h.snap_number = ls.snap_number; h.output_factor = ls.output_factor;
    h.iterations = ns->iterations; h.nsteps = ls.nsteps; h.interactions = ls.interactions;
    h.e2 = ns->e2; h.eta = ns->eta;
    h.ITIME = ls.ITIME; h.ATIME = ls.ATIME; h.e_ini = ns->en.ini; h.e_tmp = ns->en.tmp;
    h.update_end = ns->gtime.update_end;
    const std::string name = ns->output_filename + ".ckpt", tmp = name + ".tmp";
    {
        std::ofstream out(tmp.c_str(), std::ios::out | std::ios::binary | std::ios::trunc);
        const size_t n = ns->n;
        out.write(reinterpret_cast<const char *>(&h), sizeof(h));
        out.write(reinterpret_cast<const char *>(ns->h_r), n * sizeof(double4));
        out.write(reinterpret_cast<const char *>(ns->h_v), n * sizeof(double4));
        out.write(reinterpret_cast<const char *>(ns->h_f), n * sizeof(Forces));
        out.write(reinterpret_cast<const char *>(ns->h_t), n * sizeof(double));
        out.write(reinterpret_cast<const char *>(ns->h_dt), n * sizeof(double));
        if (!out.good())
        {
            logger->log_error("cannot write " + tmp);
            std::exit(EXIT_FAILURE);
        }
    }
    // the previous checkpoint stays valid until the new one is complete
    if (std::rename(tmp.c_str(), name.c_str()))
    {
        logger->log_error("cannot rename " + tmp + " to " + name);
        std::exit(EXIT_FAILURE);
    }
}

void Hermite4B200::read_checkpoint(LoopState &ls)
{
    std::ifstream in(ns->restart_filename.c_str(), std::ios::in | std::ios::binary);
    CkptHeader h;
    in.read(reinterpret_cast<char *>(&h), sizeof(h));
    if (in.good() && std::memcmp(h.magic, CKPT_MAGIC, 7) == 0 && h.magic[7] != CKPT_MAGIC[7])
    {
        // an older (or newer) layout of the same file family: say so instead of "not a checkpoint"
        logger->log_error(ns->restart_filename + " is a gravidy-b200 checkpoint of format version " + std::string(1, h.magic[7]) +
                          "; this build reads version " + std::string(1, CKPT_MAGIC[7]) + ": write it again with this build");
        std::exit(EXIT_FAILURE);
    }
    if (!in.good() || std::memcmp(h.magic, CKPT_MAGIC, 8) != 0)
    {
        logger->log_error(ns->restart_filename + " is not a gravidy-b200 checkpoint");
        std::exit(EXIT_FAILURE);
    }
    if (h.n != ns->n)
    {
        std::ostringstream m;
        m << ns->restart_filename << " holds " << h.n << " particles, the input file " << ns->n;
        logger->log_error(m.str());
        std::exit(EXIT_FAILURE);
    }
    if (h.e2 != ns->e2 || h.eta != ns->eta)
        logger->log_warning("softening/eta differ from the checkpointed run: the schedule will not be the same");
    if (!(h.ITIME < (double)ns->integration_time))
    {
        logger->log_error("the checkpoint is already at or past the requested integration time (-t)");
        std::exit(EXIT_FAILURE);
    }
    const size_t n = ns->n;
    in.read(reinterpret_cast<char *>(ns->h_r), n * sizeof(double4));
    in.read(reinterpret_cast<char *>(ns->h_v), n * sizeof(double4));
    in.read(reinterpret_cast<char *>(ns->h_f), n * sizeof(Forces));
    in.read(reinterpret_cast<char *>(ns->h_t), n * sizeof(double));
    in.read(reinterpret_cast<char *>(ns->h_dt), n * sizeof(double));
    if (!in.good())
    {
        logger->log_error(ns->restart_filename + " is truncated");
        std::exit(EXIT_FAILURE);
    }
    ls.ITIME = h.ITIME; ls.ATIME = h.ATIME; ls.nsteps = h.nsteps; ls.output_factor = h.output_factor;
    ls.snap_number = h.snap_number; ls.interactions = h.interactions;
    ns->iterations = h.iterations;
    ns->en.ini = h.e_ini; ns->en.tmp = h.e_tmp;
    ns->gtime.update_end = h.update_end;   // GFLOPS of the restarted run's log lines stay interactions / force time
    each_shard([&](int g) {
        check(gdy_upload_particles(shards[g], &ns->h_r[0].x, &ns->h_v[0].x), g, "gdy_upload_particles");
        check(gdy_upload_forces(shards[g], &ns->h_f[0].a[0]), g, "gdy_upload_forces");
        check(gdy_upload_times(shards[g], ns->h_t, ns->h_dt), g, "gdy_upload_times");
    });
}

}  // namespace gravidy
