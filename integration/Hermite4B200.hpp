// Hermite4B200 — the back end a maintainer of the reference adds beside Hermite4CPU / Hermite4GPU / Hermite4MPI:
// a subclass of the reference's OWN `Hermite4` (src/include/Hermite4.hpp:48-88, included unmodified) whose member
// functions are one C-ABI call each into libgravidy_b200.so (include/gravidy_b200.h).  Same member names, argument
// meaning and error behaviour (print + exit(EXIT_FAILURE), src/include/gpu/Hermite4GPU.cuh:43-50) as Hermite4GPU
// (src/include/gpu/Hermite4GPU.cuh:131-142).  This file and its two .cpp files are compiled with the reference's
// headers (NbodySystem.hpp, Logger.hpp, NbodyUtils.hpp) where they lie under /root/reference, in the reference's
// language level (-std=gnu++98): nothing here is C++11.
//
// The reference selects its back end at compile time in src/gravidy.cpp:37-43, 90-96.  The maintainer's change
// there is two `#elif defined(B200)` branches.  integration/Makefile proves the claim
// "host side and API unchanged" more strictly: it compiles the UNMODIFIED src/gravidy.cpp by pre-including this
// header with -DGDY_TAKE_CPU_SLOT, which claims the include guard of cpu/Hermite4CPU.hpp and names this class
// where main() says Hermite4CPU — no line of the reference is edited or copied.
#ifndef HERMITE4B200_HPP
#define HERMITE4B200_HPP
#include "Hermite4.hpp"
#include "gravidy_b200.h"

class Hermite4B200 : public Hermite4 {
    public:
        Hermite4B200(NbodySystem *ns, Logger *logger, NbodyUtils *nu);
        ~Hermite4B200();

        // the hooks of the loop, as Hermite4GPU declares them
        void init_acc_jrk();
        void update_acc_jrk(unsigned int nact);
        void predicted_pos_vel(double ITIME);
        void correction_pos_vel(double ITIME, unsigned int nact);
        double get_energy_gpu();
        // the scheduler entry points of the base class, now device-side
        unsigned int find_particles_to_move(double ITIME);
        void save_old_acc_jrk(unsigned int nact);
        void next_integration_time(double &ATIME);

        void integration();

    private:
        gdy_ctx *ctx;
        bool hooks_mode;                       // GRAVIDY_B200_HOOKS=1: hook by hook, exactly the reference's loop body
        void check(int rc, const char *what);
        void upload_times();                   // h_t, h_dt -> device, after the base class's (host) init_dt
        void refresh_mirrors();                // device -> h_r, h_v, h_f, h_t, h_dt at output points
};

#ifdef GDY_TAKE_CPU_SLOT
#define HERMITE4CPU_HPP
#define Hermite4CPU Hermite4B200
#endif
#endif // HERMITE4B200_HPP
