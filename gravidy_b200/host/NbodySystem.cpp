#include "NbodySystem.hpp"

#include <algorithm>
#include <cctype>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <sstream>

namespace gravidy {

NbodySystem::NbodySystem(const OptionsParser &op)
    : input_filename(op.input_filename), output_filename(op.output_filename),
      resume_filename(op.resume_filename), snapshot_filename(op.snapshot_filename),
      integration_time(op.integration_time), interval_time(op.interval_time),
      snapshot_time(op.snapshot_time), snapshot_number(op.snapshot_number),
      length_snap(op.length_output_number), resume(op.resume), ops(op.ops), gpus(op.gpus),
      hooks_mode(op.hooks_mode), checkpoint(op.checkpoint), restart_filename(op.restart_filename)
{
    const float e2f = op.softening * op.softening;   // float arithmetic first, then widened
    e2 = e2f;
    eta = op.eta;
}

NbodySystem::~NbodySystem() { free_base_attributes(); }

namespace {
std::vector<std::string> split_ws(const std::string &line)
{
    std::vector<std::string> out;
    std::istringstream iss(line);
    std::string tok;
    while (iss >> tok) out.push_back(tok);
    return out;
}
}  // namespace

// Eight whitespace-separated columns per particle: id m x y z vx vy vz; the mass is narrowed to float
// (reference reader: src/include/NbodySystem.cpp:79-207).  A first '#' line may carry "Time: <t>"
// (snapshots written by Logger::write_snapshot); later '#' lines are comments.
void NbodySystem::read_input_file()
{
    total_mass = 0;
    max_mass = 0.0f;
    bool got_time = false;
    const std::string &fname = resume ? snapshot_filename : input_filename;
    std::ifstream file(fname.c_str());
    std::string line;
    if (file.is_open())
    {
        while (std::getline(file, line))
        {
            const size_t first = line.find_first_not_of(' ');
            if (first == std::string::npos) continue;
            if (line[first] == '#')
            {
                if (!got_time)
                {
                    std::string l = line;
                    l.erase(std::remove(l.begin(), l.end(), '#'), l.end());
                    l.erase(std::remove(l.begin(), l.end(), ':'), l.end());
                    std::transform(l.begin(), l.end(), l.begin(), ::tolower);
                    const std::vector<std::string> tok = split_ws(l);
                    for (size_t t = 0; t + 1 < tok.size(); ++t)
                        if (tok[t] == "time")
                        {
                            snapshot_time = (float)strtod(tok[t + 1].c_str(), nullptr);
                            got_time = true;
                            break;
                        }
                }
                continue;
            }
            const std::vector<std::string> tok = split_ws(line);
            if (tok.size() != 8)
            {
                std::cerr << "gravidy: wrong line format in" << input_filename
                          << ": only eight numbers per line are allowed" << std::endl;
                break;
            }
            InputRow row;
            row.id = (unsigned int)std::atoi(tok[0].c_str());
            row.m = (float)strtod(tok[1].c_str(), nullptr);
            for (int k = 0; k < 3; ++k) row.r[k] = strtod(tok[2 + k].c_str(), nullptr);
            for (int k = 0; k < 3; ++k) row.v[k] = strtod(tok[5 + k].c_str(), nullptr);
            if (row.m > max_mass) max_mass = row.m;
            reader.push_back(row);
            total_mass += row.m;
        }
        file.close();
    }
    n = (unsigned int)reader.size();
}

void NbodySystem::alloc_base_attributes(int /*rank*/)
{
    h_id = new unsigned int[n];
    h_r = new double4[n];
    h_v = new double4[n];
}

void NbodySystem::free_base_attributes()
{
    delete[] h_id; delete[] h_r; delete[] h_v;
    h_id = nullptr; h_r = nullptr; h_v = nullptr;
}

void NbodySystem::copy_input_data()
{
    for (unsigned int i = 0; i < n; ++i)
    {
        const InputRow &row = reader[i];
        h_id[i] = row.id;
        h_r[i] = double4{row.r[0], row.r[1], row.r[2], (double)row.m};
        h_v[i] = double4{row.v[0], row.v[1], row.v[2], 0.0};
    }
    std::vector<InputRow>().swap(reader);
}

}  // namespace gravidy
