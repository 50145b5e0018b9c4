#include "Logger.hpp"

#include <chrono>
#include <cmath>
#include <cstdio>
#include <ctime>
#include <iomanip>
#include <iostream>
#include <sstream>
#include <sys/stat.h>

namespace gravidy {

namespace {
const char *BLUE = "\033[0;34m", *RED = "\033[0;31m", *GREEN = "\033[0;32m", *YELLOW = "\033[0;33m",
           *CYAN = "\033[0;36m", *PLAIN = "\033[m";
bool exists(const std::string &f) { struct stat st; return stat(f.c_str(), &st) != -1; }
double now_s()
{
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
void info_line(std::ostream &o, const std::string &key, const std::string &value, bool left_value = false)
{
    o << std::setw(2) << std::left << "#" << std::setw(20) << std::left << key;
    if (left_value) o << std::setw(20) << std::left << value;
    else o << std::setw(20) << std::right << value;
    o << std::endl;
}
template <class T> std::string str(const T &v) { std::ostringstream s; s << v; return s.str(); }
}  // namespace

Logger::Logger(NbodySystem *ns_) : ns(ns_), print_screen(ns_->ops.print_screen != 0), ofname(ns_->output_filename)
{
    ofname_info = ns->resume ? ns->resume_filename : ofname + ".info";
    wall0 = now_s();
}

// "[YYYY-MM-DD HH:MM:SS]"; the reference adds one to hour, minute and second (Logger.cpp:69-71) — kept.
std::string Logger::get_timestamp()
{
    time_t now = time(nullptr);
    tm *lt = localtime(&now);
    char buf[96];
    snprintf(buf, sizeof buf, "[%04d-%02d-%02d %02d:%02d:%02d]", 1900 + lt->tm_year, 1 + lt->tm_mon, lt->tm_mday,
             1 + lt->tm_hour, 1 + lt->tm_min, 1 + lt->tm_sec);
    return buf;
}

void Logger::write_snapshot(unsigned int snapshot_number, double ITIME)
{
    std::ostringstream s;
    s << std::setw(ns->length_snap) << std::setfill('0') << snapshot_number;
    std::ofstream out((ofname + ".snapshot_" + s.str()).c_str(), std::ios::out);
    out << std::fixed << std::setprecision(4) << "# Time:" << std::setw(6) << ITIME << std::endl;
    out << std::scientific << std::setprecision(snapshot_digits());
    for (unsigned int i = 0; i < ns->n; ++i)
    {
        out << std::setw(6) << std::right << ns->h_id[i];
        out << std::setw(15) << std::right << ns->h_r[i].w;
        out << std::setw(15) << std::right << ns->h_r[i].x << std::setw(15) << std::right << ns->h_r[i].y
            << std::setw(15) << std::right << ns->h_r[i].z;
        out << std::setw(15) << std::right << ns->h_v[i].x << std::setw(15) << std::right << ns->h_v[i].y
            << std::setw(15) << std::right << ns->h_v[i].z;
        out << std::endl;
    }
}

void Logger::add_info(const std::string &key, const std::string &value)
{
    std::ofstream info(ofname_info.c_str(), std::ios_base::app);
    info_line(info, key, value);
}

// An existing info file is moved aside to "<info>.old", "<info>.old0", "<info>.old1", ... first.
void Logger::write_info()
{
    if (exists(ofname_info))
    {
        std::string newname = ofname_info + ".old";
        int ext = 0;
        while (exists(newname)) newname = ofname_info + ".old" + str(ext++);
        if (std::rename(ofname_info.c_str(), newname.c_str()))
            std::cerr << "gravidy: Failed to rename info file from " << ofname_info << " to " << newname << std::endl;
    }
    std::ofstream info(ofname_info.c_str(), std::ios::out);
    info_line(info, "NumberParticles:", str(ns->n));
    info_line(info, "Softening:", str(std::sqrt(ns->e2)));
    info_line(info, "EtaTimestep:", str(ns->eta));
    info_line(info, "IntegrationTime:", str(ns->integration_time));
    info_line(info, "PrintScreen:", str(print_screen));
    info_line(info, "InputFilename:", ns->input_filename);
    info_line(info, "OutputFilename:", ns->output_filename, true);
}

// Truncates <out>.log when the summary goes to a file (the reference's print_info does only that).
void Logger::print_info()
{
    if (!print_screen) std::ofstream((ofname + ".log").c_str(), std::ios::out);
}

void Logger::print_lagrange_radii(double ITIME, const std::vector<double> &radii)
{
    std::ofstream file;
    if (!print_screen) file.open((ofname + ".radii").c_str(), std::ios::out | std::ios::app);
    std::ostream &o = print_screen ? static_cast<std::ostream &>(std::cout) : file;
    o << "01 " << std::fixed << ITIME << " " << std::setprecision(6);
    for (int k = 0; k < N_LAGRANGE; ++k) o << std::setw(6) << std::right << radii[k] << " ";
    o << std::endl;
}

void Logger::print_all(double ITIME, unsigned int snapshot_number)
{
    std::ofstream file;
    if (!print_screen)
    {
        std::ostringstream s;
        s << std::setw(4) << std::setfill('0') << snapshot_number;
        file.open((ofname + ".all.snap_" + s.str()).c_str(), std::ios::out);
    }
    std::ostream &o = print_screen ? static_cast<std::ostream &>(std::cout) : file;
    for (unsigned int i = 0; i < ns->n; ++i)
    {
        o << std::fixed << std::setprecision(4) << std::setw(6) << ITIME;
        o << std::setw(6) << i;
        o << std::scientific << std::setprecision(6);
        o << std::setw(15) << std::right << ns->h_r[i].w;
        o << std::setw(15) << std::right << ns->h_r[i].x << std::setw(15) << std::right << ns->h_r[i].y
          << std::setw(15) << std::right << ns->h_r[i].z;
        o << std::setw(15) << std::right << ns->h_v[i].x << std::setw(15) << std::right << ns->h_v[i].y
          << std::setw(15) << std::right << ns->h_v[i].z;
        for (int k = 0; k < 3; ++k) o << std::setw(15) << std::right << ns->h_f[i].a[k];
        for (int k = 0; k < 3; ++k) o << std::setw(15) << std::right << ns->h_f[i].a1[k];
        o << std::setw(15) << std::right << ns->h_dt[i];
        o << std::endl;
    }
}

// One row of the energy table: IteTime Iter Nsteps Energy RelE CumE ElapsedTime GFLOPS.
// RelE is against the previous row, CumE against t = 0; GFLOPS = 48e-9 * interactions / update seconds
// (reference: Logger.cpp:324-391, common.hpp:331).
void Logger::print_energy_log(double ITIME, unsigned int iterations, long long interactions,
                              unsigned int nsteps, double new_energy)
{
    ns->en.end = new_energy;
    const double rel_error = std::abs((ns->en.end - ns->en.tmp) / ns->en.tmp);
    const double cum_error = std::abs((ns->en.end - ns->en.ini) / ns->en.ini);
    ns->en.tmp = ns->en.end;
    const float elapsed = (float)(now_s() - ns->gtime.integration_ini);
    if (ITIME != 0.0) ns->gtime.gflops = KERNEL_GFLOP * (interactions / ns->gtime.update_end);

    std::ofstream file;
    if (!print_screen) file.open((ofname + ".log").c_str(), std::ios::out | std::ios::app);
    std::ostream &o = print_screen ? static_cast<std::ostream &>(std::cout) : file;
    if (ITIME == 0.0)
    {
        o << std::setw(2) << std::left << "#" << std::setw(10) << std::right << "IteTime" << std::setw(15)
          << std::right << "Iter" << std::setw(15) << std::right << "Nsteps" << std::setw(20) << std::right
          << "Energy" << std::setw(15) << std::right << "RelE" << std::setw(15) << std::right << "CumE"
          << std::setw(16) << std::right << "ElapsedTime" << std::setw(12) << std::right << "GFLOPS" << std::endl;
    }
    o << std::fixed << std::setw(2) << std::right << "00" << std::setprecision(3) << std::setw(10) << std::right
      << ITIME << std::setw(15) << std::right << iterations << std::setw(15) << std::right << nsteps;
    o << std::scientific << std::setprecision(10) << std::setw(20) << std::right << ns->en.end
      << std::setprecision(6) << std::setw(15) << std::right << rel_error << std::setw(15) << std::right
      << cum_error << std::setw(16) << std::right << elapsed;
    o << std::fixed << std::setprecision(3) << std::setw(12) << std::right << ns->gtime.gflops << std::endl;
}

void Logger::log(int type, const std::string &msg)
{
    static const char *colour[] = {RED, YELLOW, GREEN, CYAN};
    static const char *label[] = {" [ERROR] ", " [WARNING] ", " [SUCCESS] ", " [INFO] "};
    if (type < 0 || type > 3) return;
    std::cerr << BLUE << get_timestamp() << colour[type] << label[type] << PLAIN << msg << std::endl;
}

}  // namespace gravidy
