// OptionsParser — the reference's command line without boost::program_options.
//
// Same options, short/long names, defaults and error behaviour (reference:
// src/include/utils/OptionsParser.cpp:41-89 option table, :111-423 check_options, incl. the resume
// path that reads "<out>.info").  Public fields keep the reference's names
// (src/include/utils/OptionsParser.hpp:55-70) because NbodySystem copies them one by one.
#pragma once
#include <map>
#include <string>
#include "common.hpp"

namespace gravidy {

class OptionsParser
{
public:
    OptionsParser(int argc, char *argv[]);

    std::string input_filename, output_filename, resume_filename, snapshot_filename;
    float integration_time = 1.0f;
    float interval_time = (float)D_TIME_MAX;
    int gpus = 0;
    float snapshot_time = 0.0f;
    float softening = (float)SOFTENING_DEFAULT;
    float eta = ETA_N;
    unsigned int snapshot_number = 0;
    unsigned int length_output_number = 0;
    bool resume = false;
    PrintOptions ops;
    bool hooks_mode = false;   // --hooks: drive the loop hook by hook instead of the fused device driver
    bool checkpoint = false;   // --checkpoint: binary full-precision state beside the text snapshots
    std::string restart_filename;   // --restart: continue from such a checkpoint

    bool file_exists(const std::string &filename) const;
    bool check_options();
    std::string usage() const;

private:
    std::map<std::string, std::string> given;   // long option name -> value ("" for flags)
    std::string parse_error;
    bool has(const char *name) const { return given.count(name) != 0; }
};

}  // namespace gravidy
