#include "Hermite4.hpp"

#include <cmath>
#include <cstring>

namespace gravidy {

Hermite4::Hermite4(NbodySystem *ns_, Logger *logger_, NbodyUtils *nu_) : ns(ns_), logger(logger_), nu(nu_)
{
    alloc_arrays_host();
}

Hermite4::~Hermite4() { free_arrays_host(); }

// host mirrors read by Logger and NbodyUtils (reference allocates many more: Hermite4.cpp:158-189; the
// rest of the integrator state — old forces, higher derivatives, predicted set — lives on the device only)
void Hermite4::alloc_arrays_host()
{
    ns->h_f = new Forces[ns->n];
    ns->h_t = new double[ns->n];
    ns->h_dt = new double[ns->n];
    ns->h_move = new unsigned int[ns->n];
    std::memset(ns->h_f, 0, sizeof(Forces) * ns->n);
    std::memset(ns->h_t, 0, sizeof(double) * ns->n);
    std::memset(ns->h_dt, 0, sizeof(double) * ns->n);
    std::memset(ns->h_move, 0, sizeof(unsigned int) * ns->n);
}

void Hermite4::free_arrays_host()
{
    delete[] ns->h_f; delete[] ns->h_t; delete[] ns->h_dt; delete[] ns->h_move;
    ns->h_f = nullptr; ns->h_t = nullptr; ns->h_dt = nullptr; ns->h_move = nullptr;
}

// Aarseth's initial step ETA*|a|/|j| rounded down to a power of two and clamped to
// [D_TIME_MIN, D_TIME_MAX]; every particle starts at ITIME; ATIME becomes the smallest step
// (reference: src/include/Hermite4.cpp:108-140).
void Hermite4::init_dt(double &ATIME, float ETA, double ITIME)
{
    for (unsigned int i = 0; i < ns->n; ++i)
    {
        const Forces &f = ns->h_f[i];
        const double amag = nu->get_magnitude(f.a[0], f.a[1], f.a[2]);
        const double jmag = nu->get_magnitude(f.a1[0], f.a1[1], f.a1[2]);
        double step = ETA * (amag / jmag);
        const int level = (int)(std::ceil(std::log(step) / std::log(2.0)) - 1);
        step = std::pow(2, level);
        if (step < D_TIME_MIN) step = D_TIME_MIN;
        else if (step > D_TIME_MAX) step = D_TIME_MAX;
        ns->h_dt[i] = step;
        ns->h_t[i] = ITIME;
        if (step < ATIME) ATIME = step;
    }
}

}  // namespace gravidy
