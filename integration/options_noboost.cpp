// The reference's OptionsParser (declared in its own, unmodified src/include/utils/OptionsParser.hpp) without
// boost::program_options, which is not installed on this image: the command line is parsed by the boost-free parser of
// the product's host program (gravidy_b200/host/OptionsParser.cpp: same options, defaults and messages as
// src/include/utils/OptionsParser.cpp:41-423) and the result is copied field by field into the reference's class.
// Only CLI glue is replaced; every other reference translation unit is compiled from /root/reference as it is.
#include "utils/OptionsParser.hpp"                     // the reference's class (global namespace)
#include "../gravidy_b200/host/OptionsParser.hpp"      // gravidy::OptionsParser

static int g_argc = 0;
static char **g_argv = 0;

OptionsParser::OptionsParser(int argc_, char *argv_[])
{
    g_argc = argc_;
    g_argv = argv_;
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

bool OptionsParser::check_options()
{
    gravidy::OptionsParser p(g_argc, g_argv);
    if (!p.check_options()) return false;
    input_filename = p.input_filename;
    output_filename = p.output_filename;
    resume_filename = p.resume_filename;
    snapshot_filename = p.snapshot_filename;
    integration_time = p.integration_time;
    interval_time = p.interval_time;
    gpus = p.gpus;
    snapshot_time = p.snapshot_time;
    softening = p.softening;
    eta = p.eta;
    snapshot_number = p.snapshot_number;
    length_output_number = p.length_output_number;
    resume = p.resume;
    ops.print_screen = p.ops.print_screen != 0;
    ops.print_all = p.ops.print_all != 0;
    ops.print_lagrange = p.ops.print_lagrange != 0;
    return true;
}
