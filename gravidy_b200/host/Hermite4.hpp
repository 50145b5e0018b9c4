// Hermite4 — integrator base class: owns the pointers to the system, logger and utilities and declares
// the scheduler entry points under the reference's names (reference: src/include/Hermite4.hpp:48-88;
// bodies in src/include/Hermite4.cpp:63-153).  In this build find_particles_to_move,
// next_integration_time and save_old_acc_jrk keep their names and meaning but run on the device, so the
// back end supplies their bodies; init_dt stays host arithmetic with libm log/pow exactly as the
// reference (a device log could round differently and flip a timestep level).
#pragma once
#include "Logger.hpp"
#include "NbodySystem.hpp"
#include "NbodyUtils.hpp"

namespace gravidy {

class Hermite4
{
public:
    Hermite4(NbodySystem *ns, Logger *logger, NbodyUtils *nu);
    virtual ~Hermite4();

    NbodySystem *ns;
    Logger *logger;
    NbodyUtils *nu;

    virtual unsigned int find_particles_to_move(double ITIME) = 0;
    virtual void next_integration_time(double &ATIME) = 0;
    virtual void save_old_acc_jrk(unsigned int nact) = 0;
    void init_dt(double &ATIME, float ETA, double ITIME);

    virtual void integration() = 0;

protected:
    void alloc_arrays_host();
    void free_arrays_host();
};

}  // namespace gravidy
