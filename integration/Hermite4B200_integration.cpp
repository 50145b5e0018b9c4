// Hermite4B200::integration — the loop of the reference's back ends (src/include/gpu/Hermite4GPU_integration.cpp:
// 38-130 and src/include/cpu/Hermite4CPU_integration.cpp:38-130 make the same calls in the same order) over the
// B200 hooks.  What differs from those files, and why:
//   * init_dt stays the base class's host function; its t, dt go up once (upload_times);
//   * the particle set lives on the device, so the host arrays the Logger prints from are refreshed at output points
//     (refresh_mirrors) instead of after every hook;
//   * the interaction counter is 64-bit (the reference's `nact * ns->n` wraps at 2^32);
//   * default mode hands all block steps up to the next output time to gdy_run (same kernels, same arithmetic, no host
//     round trip per hook); GRAVIDY_B200_HOOKS=1 walks the loop hook by hook exactly as the reference does.
#include "Hermite4B200.hpp"

void Hermite4B200::integration()
{
    ns->gtime.integration_ini = omp_get_wtime();

    double ATIME = 1.0e+10;
    double ITIME = ns->snapshot_time;
    unsigned int nact = 0;
    unsigned int nsteps = 0;
    long long interactions = 0;
    unsigned int output_factor = 1;

    init_acc_jrk();
    init_dt(ATIME, ETA_S, ITIME);
    upload_times();

    ns->en.ini = get_energy_gpu();
    ns->en.tmp = ns->en.ini;

    nu->nbody_attributes();

    logger->print_info();
    logger->write_info();
    logger->print_energy_log(ITIME, ns->iterations, interactions, nsteps, ns->en.ini);

    unsigned int snap_number = ns->snapshot_number;
    logger->write_snapshot(snap_number, ITIME);
    snap_number++;

    if (ns->ops.print_all) logger->print_all(ITIME, snap_number);
    if (ns->ops.print_lagrange)
    {
        nu->lagrange_radii();
        logger->print_lagrange_radii(ITIME, nu->layers_radii);
    }

    while (ITIME < ns->integration_time)
    {
        if (hooks_mode)
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
            // every block step up to and including the first one that reaches the next output time; the counters are
            // left where the reference has them at its output check (the triggering step is added below, after it)
            const double t_out = (double)ns->interval_time * output_factor;
            const double t_stop = t_out < (double)ns->integration_time ? t_out : (double)ns->integration_time;
            gdy_run_stats st;
            check(gdy_run(ctx, ATIME, ITIME, t_stop, 0, &st, NULL, NULL, NULL, 0), "gdy_run");
            ITIME = st.itime;
            ATIME = st.atime;
            nact = st.last_nact;
            interactions += (long long)st.interactions;
            nsteps += (unsigned int)(st.nsteps - nact);
            ns->iterations += (long long)st.iterations - 1;
            ns->gtime.update_end += st.force_ms * 1e-3;
        }

        if (ITIME >= ns->interval_time * output_factor)
        {
            refresh_mirrors();
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
        }

        nsteps += nact;
        ns->iterations++;
    }

    ns->gtime.integration_end = omp_get_wtime() - ns->gtime.integration_ini;
    refresh_mirrors();
    logger->write_snapshot(snap_number, ITIME);
    logger->add_info(std::string("SnapshotNumber:"), std::string(SSTR(snap_number)));
}
