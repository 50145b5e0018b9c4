// TEST INFRASTRUCTURE ONLY (oracle build).
//
// Replacement translation unit for the reference's src/include/utils/OptionsParser.cpp, which is
// CLI glue over boost::program_options (absent here) and is not on the hot path.  It defines the
// same class members declared in the UNMODIFIED reference header
// (src/include/utils/OptionsParser.hpp:53-79) with the reference's defaults
// (src/include/utils/OptionsParser.cpp:325-423): softening = E, eta = ETA_N,
// interval_time = D_TIME_MAX, integration_time = 1.  The harness overwrites the public fields.
#include "utils/OptionsParser.hpp"

OptionsParser::OptionsParser(int argc, char *argv[])
{
    (void)argc; (void)argv;
    input_filename = "";
    output_filename = "";
    resume_filename = "";
    snapshot_filename = "";
    integration_time = 1;
    interval_time = D_TIME_MAX;
    gpus = 0;
    snapshot_time = 0;
    softening = E;
    eta = ETA_N;
    snapshot_number = 0;
    length_output_number = 2;
    resume = false;
    ops.print_screen = 0;
    ops.print_all = 0;
    ops.print_lagrange = 0;
}

OptionsParser::~OptionsParser() {}

bool OptionsParser::file_exists(std::string filename)
{
    struct stat buffer;
    return stat(filename.c_str(), &buffer) != -1;
}

bool OptionsParser::check_options() { return true; }
