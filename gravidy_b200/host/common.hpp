// common.hpp — constants and plain data types of the host program.
//
// Values and meanings follow the reference so that option defaults, timestep bounds, log columns and
// file formats stay the same (reference: src/include/common.hpp:82-129 constants, :171-329 types).
// The device-side copies of the timestep bounds live in csrc/gdy_step.cuh.
#pragma once
#include <string>
#include <vector>

namespace gravidy {

struct double3 { double x, y, z; };
struct double4 { double x, y, z, w; };   // positions carry the mass in w (reference: NbodySystem.hpp:106)

const double G = 1.0;                    // N-body units
const int J_NEIGHBOUR = 6;               // Casertano & Hut neighbour for the density centre (common.hpp:96)
const float CORE_MASS = 0.1f;            // mass fraction that defines the core radius (common.hpp:102)
const float LAGRANGE_RADII[] = {0.01f, 0.05f, 0.1f, 0.2f, 0.5f, 0.75f};   // common.hpp:104
const int N_LAGRANGE = sizeof(LAGRANGE_RADII) / sizeof(float);

const double SOFTENING_DEFAULT = 1e-4;   // "E"      common.hpp:107
const float ETA_S = 0.01f;               // initial-step accuracy parameter, common.hpp:114
const float ETA_N = 0.01f;               // default -e,                      common.hpp:119
const double D_TIME_MIN = 1.1920928955078125e-07;   // 2^-23, common.hpp:122
const double D_TIME_MAX = 0.125;                    // 2^-3,  common.hpp:129
const double KERNEL_GFLOP = 48e-9;       // flop-per-pair convention of the log's GFLOPS column (common.hpp:331)

// {a, a1} = acceleration and jerk (reference "Forces", common.hpp:228-231)
struct Forces { double a[3]; double a1[3]; };

// wall-clock accumulators, in seconds (the reference keeps these as float, common.hpp:287-310; double
// here so that sub-second resolution survives long uptimes — SURVEY.md A.10)
struct Gtime
{
    double integration_ini = 0.0, integration_end = 0.0;
    double prediction_end = 0.0, update_end = 0.0, correction_end = 0.0;
    double gflops = 0.0;
};

struct Energy { double ini = 0.0, end = 0.0, tmp = 0.0; };

struct PrintOptions { int print_screen = 0, print_all = 0, print_lagrange = 0; };

struct InputRow { unsigned int id; float m; double r[3]; double v[3]; };   // reference file_data, common.hpp:321-329

struct Distance
{
    unsigned int index;
    double value;
    bool operator<(const Distance &o) const { return value < o.value; }
};

}  // namespace gravidy
