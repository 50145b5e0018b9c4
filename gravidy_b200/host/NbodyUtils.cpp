#include "NbodyUtils.hpp"

#include <algorithm>
#include <cmath>

namespace gravidy {

NbodyUtils::NbodyUtils(NbodySystem *ns_) : ns(ns_), cod{0.0, 0.0, 0.0}, radii(ns_->n), layers_radii(N_LAGRANGE, 0.0) {}

// reference: NbodyUtils.cpp:59-78
void NbodyUtils::nbody_attributes()
{
    ns->r_virial = get_virial_radius(ns->en.ini);
    ns->t_cr = get_crossing_time(ns->r_virial);
    ns->r_cl = get_close_encounter_radius(ns->r_virial);
    ns->dt_cl = get_close_encounter_timestep(ns->r_cl);
    ns->r_core = get_core_radius();
    ns->t_rlx = get_half_mass_relaxation_time();
}

double NbodyUtils::get_virial_radius(double energy)
{
    return (-G * ns->total_mass * ns->total_mass) / (4 * energy);
}

double NbodyUtils::get_crossing_time(double r_virial)
{
    const float M = ns->total_mass;
    const float Rv3 = (float)(r_virial * r_virial * r_virial);
    const float Ut = (float)std::sqrt(Rv3 / (G * M));
    const float t_cr = (float)(2 * std::sqrt(2) * Ut);
    return t_cr;
}

double NbodyUtils::get_close_encounter_radius(double r_virial) { return (4.0 * r_virial) / ns->n; }

double NbodyUtils::get_close_encounter_timestep(double r_cl)
{
    const float r_cl3 = (float)(r_cl * r_cl * r_cl);
    return 0.04 * std::sqrt(r_cl3 * ns->n);
}

void NbodyUtils::get_radii()
{
    for (unsigned int i = 0; i < ns->n; ++i)
    {
        const double rx = ns->h_r[i].x - cod.x, ry = ns->h_r[i].y - cod.y, rz = ns->h_r[i].z - cod.z;
        radii[i].index = i;
        radii[i].value = std::sqrt(rx * rx + ry * ry + rz * rz);
    }
}

// radius that encloses CORE_MASS of the total mass around the density centre (reference: :113-144)
double NbodyUtils::get_core_radius()
{
    get_radii();
    std::vector<Distance> d(radii);
    std::sort(d.begin(), d.end());
    double core_mass = 0.0, radius = 0.0;
    for (unsigned int i = 0; i < ns->n; ++i)
    {
        if (core_mass > ns->total_mass * CORE_MASS)
        {
            radius = d[i - 1].value;
            break;
        }
        core_mass += ns->h_r[d[i].index].w;
    }
    return radius;
}

// reference: :149-160 (empirical Spitzer form with a softening-independent Coulomb logarithm)
double NbodyUtils::get_half_mass_relaxation_time()
{
    cod = get_center_of_density();
    const float r_h = (float)get_halfmass_radius();
    ns->r_hm = r_h;
    const float a = (float)std::sqrt((ns->n * r_h * r_h * r_h) / (G * (ns->total_mass / (ns->n))));
    const float b = (float)(1 / std::log(0.11 * ns->n));
    const float t_rh = (float)(0.138 * a * b);
    return t_rh;
}

// Casertano & Hut density centre from the J-th neighbour distance (reference: :163-220).  The reference
// finds that distance by sorting all N distances for every particle on the host; here the search is one
// O(N^2) device kernel and only the weighted average is left to the host.
double3 NbodyUtils::get_center_of_density()
{
    std::vector<double> rj;
    neighbour_distance(J_NEIGHBOUR, rj);
    double3 c = {0.0, 0.0, 0.0};
    float dsum = 0.0f;
    for (unsigned int i = 0; i < ns->n; ++i)
    {
        const float radius = (float)rj[i];
        const float aa = (float)((J_NEIGHBOUR - 1) * ns->h_r[i].w);
        const float bb = (float)((4.0 * M_PI * radius * radius * radius) / 3.0);
        const double p = aa / bb;
        dsum += (float)p;
        c.x += ns->h_r[i].x * p; c.y += ns->h_r[i].y * p; c.z += ns->h_r[i].z * p;
    }
    c.x /= dsum; c.y /= dsum; c.z /= dsum;
    return c;
}

double NbodyUtils::get_halfmass_radius()
{
    std::vector<Distance> d(ns->n);
    for (unsigned int i = 0; i < ns->n; ++i)
    {
        const double rx = cod.x - ns->h_r[i].x, ry = cod.y - ns->h_r[i].y, rz = cod.z - ns->h_r[i].z;
        d[i].index = i;
        d[i].value = std::sqrt(rx * rx + ry * ry + rz * rz);
    }
    std::sort(d.begin(), d.end());
    float half_mass = 0;
    unsigned int j = ns->n;
    for (unsigned int i = 0; i < ns->n; ++i)
    {
        if (half_mass >= ns->total_mass / 2.0) { j = i; break; }
        half_mass += (float)ns->h_r[d[i].index].w;
    }
    return d[j > 0 ? j - 1 : 0].value;
}

void NbodyUtils::lagrange_radii()
{
    cod = get_center_of_density();
    get_radii();
    std::sort(radii.begin(), radii.end());
    get_layers();
}

// radii at which the enclosed mass passes each fraction in LAGRANGE_RADII (reference: :301-328)
void NbodyUtils::get_layers()
{
    float tmp_mass = 0.0f;
    unsigned int layer = 0;
    for (unsigned int i = 0; i < ns->n && layer < (unsigned int)N_LAGRANGE; ++i)
    {
        if (tmp_mass >= ns->total_mass) break;
        if (tmp_mass > LAGRANGE_RADII[layer])
        {
            layers_radii[layer] = radii[i].value;
            layer += 1;
        }
        tmp_mass += (float)ns->h_r[radii[i].index].w;
    }
}

double NbodyUtils::get_magnitude(double x, double y, double z) { return std::sqrt(x * x + y * y + z * z); }

}  // namespace gravidy
