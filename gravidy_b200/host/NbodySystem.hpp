// NbodySystem — the particle set and run parameters of the host program.
//
// Same role and public member names as the reference's NbodySystem (src/include/NbodySystem.hpp:47-141):
// every other component reads the h_* host arrays directly.  Here they are MIRRORS of the
// device-resident state, refreshed from the GPU at output points only (the reference's GPU build moved
// the whole set over PCIe every block step: src/include/gpu/Hermite4GPU.cu:226-265).
#pragma once
#include <string>
#include <vector>
#include "OptionsParser.hpp"
#include "common.hpp"

namespace gravidy {

class NbodySystem
{
public:
    explicit NbodySystem(const OptionsParser &op);
    ~NbodySystem();
    NbodySystem(const NbodySystem &) = delete;
    NbodySystem &operator=(const NbodySystem &) = delete;

    // run parameters
    std::string input_filename, output_filename, resume_filename, snapshot_filename;
    float integration_time, interval_time, snapshot_time;
    unsigned int snapshot_number, length_snap;
    bool resume;
    PrintOptions ops;
    int gpus;
    bool hooks_mode;
    bool checkpoint;
    std::string restart_filename;
    double e2;          // softening squared: FLOAT product widened (reference: NbodySystem.cpp:50)
    float eta;          // accuracy parameter, float as in the reference (NbodySystem.hpp:67)

    // system
    unsigned int n = 0;
    unsigned int iterations = 0;
    float total_mass = 0.0f;   // float accumulators as in the reference (NbodySystem.hpp:68,81)
    float max_mass = 0.0f;
    Energy en;
    Gtime gtime;
    double r_virial = 0, t_cr = 0, r_cl = 0, dt_cl = 0, r_core = 0, r_hm = 0, t_rlx = 0;

    // host mirrors (AoS, the shapes the C ABI takes: r4/v4 = n x 4, f6 = n x 6)
    unsigned int *h_id = nullptr;
    double4 *h_r = nullptr, *h_v = nullptr;
    Forces *h_f = nullptr;
    double *h_t = nullptr, *h_dt = nullptr;
    unsigned int *h_move = nullptr;

    void read_input_file();
    void alloc_base_attributes(int rank);
    void free_base_attributes();
    void copy_input_data();

private:
    std::vector<InputRow> reader;
};

}  // namespace gravidy
