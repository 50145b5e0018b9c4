// Hermite4B200 — the back end of the host program: the reference's Hermite4GPU member functions
// (reference: src/include/gpu/Hermite4GPU.cuh:131-142) re-implemented as thin calls into the C ABI of
// include/gravidy_b200.h.  No kernel, no CUDA call and no particle arithmetic lives on this side of the
// boundary; the particle set stays on the device between output points.
//
// Multi-GPU: one gdy_ctx (= one j-shard, one NCCL rank) per GPU, all inside this process, each driven by
// its own host thread; state is replicated, so every shard takes the same decisions and only shard 0
// talks to the logger.
#pragma once
#include <cstdint>
#include <functional>
#include <vector>
#include "../../include/gravidy_b200.h"
#include "Hermite4.hpp"

namespace gravidy {

class Hermite4B200 : public Hermite4
{
public:
    Hermite4B200(NbodySystem *ns, Logger *logger, NbodyUtils *nu);
    ~Hermite4B200();

    // hooks, names and argument meaning of the reference
    void init_acc_jrk();
    void update_acc_jrk(unsigned int nact);
    void predicted_pos_vel(double ITIME);
    void correction_pos_vel(double ITIME, unsigned int nact);
    double get_energy_gpu();
    unsigned int find_particles_to_move(double ITIME) override;
    void next_integration_time(double &ATIME) override;
    void save_old_acc_jrk(unsigned int nact) override;

    void integration() override;

private:
    std::vector<gdy_ctx *> shards;
    int ngpus = 1;

    void check(int rc, int shard, const char *what);
    void each_shard(const std::function<void(int)> &fn);
    void upload_times();                 // h_t, h_dt -> device (after the host init_dt)
    void refresh_mirrors(bool forces);   // device -> h_r, h_v (, h_f, h_t, h_dt) at output points
    gdy_run_stats run_until(double t_stop, double &ATIME, double &ITIME);

    // binary full-precision checkpoint: everything the loop needs to go on bit for bit
    struct LoopState
    {
        double ITIME, ATIME;
        unsigned int nsteps, output_factor, snap_number;
        long long interactions;
    };
    void write_checkpoint(const LoopState &ls);
    void read_checkpoint(LoopState &ls);
};

}  // namespace gravidy
