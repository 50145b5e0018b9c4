// NbodyUtils — system diagnostics of the host program (reference: src/include/utils/NbodyUtils.cpp).
//
// Energy and the k-th-neighbour search behind the density centre are O(N^2) and run on the GPU through
// the C ABI (gdy_energy, gdy_neighbour_distance); the O(N log N) bookkeeping around them (sorting radii,
// cumulative masses) stays on the host, with the reference's float/double types so that the printed
// Lagrange radii agree.  The timestep functions of the reference's NbodyUtils (get_timestep_normal,
// normalize_dt) live on the device, inside the corrector kernel (csrc/gdy_step.cuh).
#pragma once
#include <functional>
#include <vector>
#include "NbodySystem.hpp"

namespace gravidy {

class NbodyUtils
{
public:
    explicit NbodyUtils(NbodySystem *ns);

    NbodySystem *ns;
    double3 cod;                         // centre of density
    std::vector<Distance> radii;
    std::vector<double> layers_radii;    // Lagrange radii

    // set by the back end: distance of every particle to its k-th nearest neighbour, from the device
    std::function<void(int k, std::vector<double> &out)> neighbour_distance;

    void nbody_attributes();
    double get_virial_radius(double energy);
    double get_crossing_time(double r_virial);
    double get_close_encounter_radius(double r_virial);
    double get_close_encounter_timestep(double r_cl);
    double get_core_radius();
    double get_half_mass_relaxation_time();
    double3 get_center_of_density();
    double get_halfmass_radius();
    void lagrange_radii();
    double get_magnitude(double x, double y, double z);

private:
    void get_radii();
    void get_layers();
};

}  // namespace gravidy
