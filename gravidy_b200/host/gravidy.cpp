// gravidy-b200 — entry point of the host program; same sequence as the reference's main
// (src/gravidy.cpp:53-115): options -> NbodySystem -> input -> Logger, NbodyUtils -> back end ->
// integration().
#include "Hermite4B200.hpp"

using namespace gravidy;

int main(int argc, char *argv[])
{
    OptionsParser op(argc, argv);
    if (!op.check_options()) return 1;

    NbodySystem ns(op);
    ns.read_input_file();
    ns.alloc_base_attributes(0);
    ns.copy_input_data();

    Logger logger(&ns);
    NbodyUtils nu(&ns);

    if (ns.n == 0)
    {
        logger.log_error("no particles read from " + (ns.resume ? ns.snapshot_filename : ns.input_filename));
        return 1;
    }

    Hermite4B200 h4(&ns, &logger, &nu);
    h4.integration();

    logger.log_success("Finishing...");
    return 0;
}
