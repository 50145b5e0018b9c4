#include "OptionsParser.hpp"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <set>
#include <sstream>
#include <sys/stat.h>

namespace gravidy {

namespace {
struct OptSpec { const char *long_name; char short_name; bool takes_value; };
const OptSpec SPECS[] = {
    {"help", 'h', false},     {"input", 'i', true},    {"time", 't', true},   {"resume", 'r', true},
    {"output", 'o', true},    {"softening", 's', true}, {"eta", 'e', true},    {"screen", 'p', false},
    {"interval", 'z', true},  {"lagrange", 'l', false}, {"all", 'a', false},   {"gpu", 'g', true},
    {"hooks", 0, false},      {"checkpoint", 0, false}, {"restart", 0, true},
};
const OptSpec *find_long(const std::string &n)
{
    for (const OptSpec &s : SPECS) if (n == s.long_name) return &s;
    return nullptr;
}
const OptSpec *find_short(char ch)
{
    for (const OptSpec &s : SPECS) if (s.short_name && ch == s.short_name) return &s;
    return nullptr;
}
std::string strip(std::string s, const char *chars)
{
    for (const char *p = chars; *p; ++p) s.erase(std::remove(s.begin(), s.end(), *p), s.end());
    return s;
}
}  // namespace

OptionsParser::OptionsParser(int argc, char *argv[])
{
    for (int k = 1; k < argc; ++k)
    {
        std::string tok = argv[k];
        const OptSpec *spec = nullptr;
        std::string value;
        bool have_value = false;
        if (tok.size() > 2 && tok[0] == '-' && tok[1] == '-')
        {
            std::string name = tok.substr(2);
            const size_t eq = name.find('=');
            if (eq != std::string::npos) { value = name.substr(eq + 1); name = name.substr(0, eq); have_value = true; }
            spec = find_long(name);
        }
        else if (tok.size() >= 2 && tok[0] == '-' && tok[1] != '-')
        {
            spec = find_short(tok[1]);
            if (spec && tok.size() > 2) { value = tok.substr(2); have_value = true; }   // -t1.0
        }
        if (!spec) { parse_error = "unrecognised option '" + tok + "'"; return; }
        if (spec->takes_value)
        {
            if (!have_value)
            {
                if (k + 1 >= argc) { parse_error = std::string("the required argument for option '--") + spec->long_name + "' is missing"; return; }
                value = argv[++k];
            }
            given[spec->long_name] = value;
        }
        else
            given[spec->long_name] = "";
    }
}

std::string OptionsParser::usage() const
{
    std::ostringstream u;
    u << "GraviDy:\n"
      << "  -h [ --help ]                   Display this message\n\n"
      << "Required options:\n"
      << "  -i [ --input ] <filename>       Input data filename\n"
      << "  -t [ --time ] <value> (=1)      Integration time (In N-body units)\n"
      << "  -r [ --resume ] <filename>.info Resume a simulation with an .info file\n\n"
      << "Elective options:\n"
      << "  -o [ --output ] <filename>      Output data filename\n"
      << "  -s [ --softening ] <value>      Softening parameter (default 1e-4)\n"
      << "  -e [ --eta ] <value>            ETA of time-step calculation (default 0.01)\n"
      << "  -p [ --screen ]                 Print summary in the screen instead of a file\n"
      << "  -z [ --interval ] <value>       Output time interval (default 0.125)\n\n"
      << "Extra options:\n"
      << "  -l [ --lagrange ]               Print information of the Lagrange Radii in every\n"
      << "                                  integration time\n"
      << "  -a [ --all ]                    Print all the information of N-particles in every\n"
      << "                                  integration time\n"
      << "  -g [ --gpu ] <value> (=0)       GPUs to use, by default is the maximum available\n"
      << "                                  devices (1, 2, 4 or 8 j-shards)\n"
      << "  --hooks                         Drive every block step through the individual hooks\n"
      << "                                  instead of the fused device driver (same results)\n"
      << "  --checkpoint                    Write <output>.ckpt (binary, full precision: r, v, acc, jerk,\n"
      << "                                  t, dt, counters) at every output time and at the end\n"
      << "  --restart <filename>.ckpt       Continue the run of -i <filename> from a checkpoint: the\n"
      << "                                  block-timestep schedule goes on exactly where it stopped\n";
    return u.str();
}

bool OptionsParser::file_exists(const std::string &filename) const
{
    struct stat st;
    return stat(filename.c_str(), &st) != -1;
}

bool OptionsParser::check_options()
{
    if (!parse_error.empty())
    {
        std::cerr << "gravidy: " << parse_error << std::endl << usage() << std::endl;
        return false;
    }
    if (has("help"))
    {
        std::cerr << usage() << std::endl;
        return false;
    }
    hooks_mode = has("hooks");
    checkpoint = has("checkpoint");
    if (has("restart"))
    {
        restart_filename = given["restart"];
        if (!file_exists(restart_filename))
        {
            std::cerr << "gravidy: cannot access " << restart_filename << ": No such file or directory" << std::endl;
            return false;
        }
        if (has("resume"))
        {
            std::cerr << "gravidy: --restart and --resume exclude each other" << std::endl << usage() << std::endl;
            return false;
        }
    }
    auto as_float = [&](const char *name) { return (float)strtod(given[name].c_str(), nullptr); };
    const bool time_given = true;   // "-t" has a default value of 1 in the reference, so it always counts
    integration_time = has("time") ? as_float("time") : 1.0f;
    snapshot_time = 0.0f;

    if (has("resume"))
    {
        resume_filename = given["resume"];
        resume = true;
        if (!file_exists(resume_filename))
        {
            std::cerr << "gravidy: cannot access " << resume_filename << ": No such file or directory" << std::endl;
            return false;
        }
        std::cerr << "Reading resume_filename: " << resume_filename << std::endl;
        // "# Key:      value" lines written by Logger::write_info / add_info
        std::map<std::string, std::string> cfg;
        std::ifstream rfile(resume_filename.c_str());
        std::string key, value;
        while (rfile.good() && std::getline(rfile, key, ':'))
        {
            key = strip(key, "# ");
            if (std::getline(rfile, value)) cfg[key] = strip(value, " ");
        }
        rfile.close();
        static const char *allowed[] = {"NumberParticles", "Softening", "EtaTimestep", "IntegrationTime",
                                        "PrintScreen", "InputFilename", "OutputFilename", "SnapshotNumber"};
        for (const auto &kv : cfg)
        {
            bool ok = false;
            for (const char *a : allowed) ok = ok || kv.first == a;
            if (ok) std::cerr << kv.first << " -> " << kv.second << std::endl;
            else
            {
                std::cerr << "Invalid param" << std::endl
                          << "gravidy: Invalid option -- " << kv.first << " in the info file" << std::endl
                          << usage() << std::endl;
                return false;
            }
        }
        if (time_given)
        {
            const double config_time = strtod(cfg["IntegrationTime"].c_str(), nullptr);
            if (config_time >= integration_time)
            {
                std::cerr << "gravidy: option requires an argument -- 'time'"
                          << " greater than the one specifid in the info file" << std::endl << usage() << std::endl;
                return false;
            }
            snapshot_time = (float)config_time;
        }
        softening = (float)strtod(cfg["Softening"].c_str(), nullptr);
        if (has("softening"))
        {
            const float tmp = as_float("softening");
            if (tmp > 0) softening = tmp;
            else std::cerr << "gravidy: softening cannot be negative" << std::endl;
        }
        eta = (float)strtod(cfg["EtaTimestep"].c_str(), nullptr);
        if (has("eta"))
        {
            const float tmp = as_float("eta");
            if (tmp > 0) eta = tmp;
            else std::cerr << "gravidy: eta cannot be negative" << std::endl;
        }
        ops.print_screen = std::atoi(cfg["PrintScreen"].c_str());
        if (has("screen")) ops.print_screen = 1;
        if (has("gpu")) gpus = std::atoi(given["gpu"].c_str());
        if (has("interval")) interval_time = as_float("interval");   // the reference leaves this unset on
                                                                     // resume; default kept
        input_filename = cfg["InputFilename"];
        output_filename = input_filename + ".out";
        const std::string out_name = cfg["OutputFilename"];
        snapshot_number = (unsigned int)std::atoi(cfg["SnapshotNumber"].c_str());
        // The reference looks for a 4-digit suffix only (OptionsParser.cpp:303-306) although
        // write_snapshot pads to digits(time/interval)+1 (Logger.cpp:80), so its resume finds the file
        // only for runs with 100..999 outputs.  Same first choice here, then the other widths.
        std::string snap_name;
        static const int widths[] = {4, 2, 3, 5, 6, 7};
        for (int w : widths)
        {
            std::ostringstream s;
            s << std::setw(w) << std::setfill('0') << snapshot_number;
            snap_name = out_name + ".snapshot_" + s.str();
            if (file_exists(snap_name)) break;
        }
        if (!file_exists(snap_name))
        {
            std::cerr << "gravidy: cannot access " << snap_name << ": No such file or directory\n"
                      << "(Check the OutputFilename and SnapshotNumber)" << std::endl;
            return false;
        }
        snapshot_filename = snap_name;
    }
    else
    {
        resume = false;
        snapshot_number = 0;
        if (!has("input"))
        {
            std::cerr << "gravidy: option requires an argument -- 'input'" << std::endl << usage() << std::endl;
            return false;
        }
        input_filename = given["input"];
        if (!file_exists(input_filename))
        {
            std::cerr << "gravidy: cannot access " << input_filename << ": No such file or directory" << std::endl;
            return false;
        }
        ops.print_screen = has("screen") ? 1 : 0;
        ops.print_all = has("all") ? 1 : 0;
        ops.print_lagrange = has("lagrange") ? 1 : 0;
        gpus = has("gpu") ? std::atoi(given["gpu"].c_str()) : 0;
        softening = has("softening") ? as_float("softening") : (float)SOFTENING_DEFAULT;
        eta = has("eta") ? as_float("eta") : ETA_N;
        interval_time = has("interval") ? as_float("interval") : (float)D_TIME_MAX;
        output_filename = (has("output") ? given["output"] : input_filename) + ".out";
    }
    // digits of the snapshot suffix: digits(integration_time / interval_time) + 1
    // (reference: OptionsParser.cpp:395-401)
    length_output_number = 0;
    unsigned int tmp = (unsigned int)(integration_time / interval_time);
    do { ++length_output_number; tmp /= 10; } while (tmp);
    length_output_number += 1;
    return true;
}

}  // namespace gravidy
