// The integration loop.  Same sequence, same output points and same log/snapshot calls as the
// reference (src/include/gpu/Hermite4GPU_integration.cpp:38-130, identical to the CPU back end's
// src/include/cpu/Hermite4CPU_integration.cpp:38-130).  Two ways through the loop body:
//   --hooks   every block step goes hook by hook, exactly as written in the reference;
//   default   the block steps between two output points run inside gdy_run (same kernels, same
//             arithmetic, no host round trip per hook) — the loop below then iterates once per output.
#include <chrono>
#include <sstream>
#include "Hermite4B200.hpp"

namespace gravidy {

namespace {
double now_s()
{
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
}  // namespace

void Hermite4B200::integration()
{
    ns->gtime.integration_ini = now_s();

    double ATIME = 1.0e+10;                // next block time
    double ITIME = ns->snapshot_time;      // current integration time
    unsigned int nact = 0;
    unsigned int nsteps = 0;
    long long interactions = 0;            // 64-bit (the reference's nact * n wraps at 2^32)
    unsigned int output_factor = 1;

    unsigned int snap_number = ns->snapshot_number;
    const bool restart = !ns->restart_filename.empty();
    if (restart)
    {
        // --restart: r, v, acc, jerk, t, dt and the loop counters come from the checkpoint in full precision;
        // no init_acc_jrk / init_dt, no second snapshot of the restart time
        LoopState ls;
        read_checkpoint(ls);
        ITIME = ls.ITIME; ATIME = ls.ATIME; nsteps = ls.nsteps; interactions = ls.interactions;
        output_factor = ls.output_factor; snap_number = ls.snap_number;
        nu->nbody_attributes();
        logger->print_info();
        logger->write_info();
        const double e_tmp = ns->en.tmp;
        logger->print_energy_log(ITIME, ns->iterations, interactions, nsteps, get_energy_gpu());
        ns->en.tmp = e_tmp;   // RelE of the next line stays relative to the last line of the checkpointed run
    }
    else
    {
        init_acc_jrk();
        init_dt(ATIME, ETA_S, ITIME);
        upload_times();

        ns->en.ini = get_energy_gpu();
        ns->en.tmp = ns->en.ini;

        nu->nbody_attributes();

        logger->print_info();
        logger->write_info();
        logger->print_energy_log(ITIME, ns->iterations, interactions, nsteps, ns->en.ini);

        logger->write_snapshot(snap_number, ITIME);
        snap_number++;

        if (ns->ops.print_all) logger->print_all(ITIME, snap_number);
        if (ns->ops.print_lagrange)
        {
            nu->lagrange_radii();
            logger->print_lagrange_radii(ITIME, nu->layers_radii);
        }
    }

    while (ITIME < ns->integration_time)
    {
        bool wrote_output = false;
        if (ns->hooks_mode)
        {
            ITIME = ATIME;
            nact = find_particles_to_move(ITIME);
            save_old_acc_jrk(nact);
            predicted_pos_vel(ITIME);
            update_acc_jrk(nact);
            correction_pos_vel(ITIME, nact);
            interactions += (long long)nact * (long long)ns->n;
            next_integration_time(ATIME);
        }
        else
        {
            // every block step up to (and including) the first one that reaches the next output time;
            // the counters are left where the reference has them at its output check: the step that
            // triggers the output is added after the check, below
            const double t_out = ns->interval_time * output_factor;
            const double t_stop = t_out < ns->integration_time ? t_out : (double)ns->integration_time;
            const gdy_run_stats st = run_until(t_stop, ATIME, ITIME);
            nact = st.last_nact;
            interactions += (long long)st.interactions;
            nsteps += (unsigned int)(st.nsteps - nact);
            ns->iterations += (unsigned int)(st.iterations - 1);
        }

        if (ITIME >= ns->interval_time * output_factor)
        {
            refresh_mirrors(ns->ops.print_all != 0);
            logger->print_energy_log(ITIME, ns->iterations, interactions, nsteps, get_energy_gpu());
            if (ns->ops.print_all) logger->print_all(ITIME, snap_number);
            if (ns->ops.print_lagrange)
            {
                nu->lagrange_radii();
                logger->print_lagrange_radii(ITIME, nu->layers_radii);
            }
            logger->write_snapshot(snap_number, ITIME);
            snap_number++;
            output_factor += 1;
            wrote_output = true;
        }

        nsteps += nact;
        ns->iterations++;
        if (ns->checkpoint && wrote_output)
            write_checkpoint(LoopState{ITIME, ATIME, nsteps, output_factor, snap_number, interactions});
    }

    ns->gtime.integration_end = now_s() - ns->gtime.integration_ini;
    if (ns->checkpoint) write_checkpoint(LoopState{ITIME, ATIME, nsteps, output_factor, snap_number, interactions});
    refresh_mirrors(false);
    logger->write_snapshot(snap_number, ITIME);
    std::ostringstream s;
    s << snap_number;
    logger->add_info("SnapshotNumber:", s.str());
}

}  // namespace gravidy
