// TEST INFRASTRUCTURE ONLY (oracle build).  Stand-in for <boost/program_options.hpp>.
//
// boost is not installed on this image and there is no network.  The reference's
// OptionsParser.hpp (reference: src/include/utils/OptionsParser.hpp:39,72-73) includes this
// header and declares two members of boost types; every reference TU includes it transitively.
// The oracle build never parses a command line through boost: oracle/standin_options.cpp fills
// the public fields of OptionsParser (OptionsParser.hpp:55-70) directly.  So two empty types in
// the right namespace are all that is needed for the unmodified reference sources to compile.
#ifndef GDY_ORACLE_BOOST_STANDIN_HPP
#define GDY_ORACLE_BOOST_STANDIN_HPP
namespace boost { namespace program_options {
class variables_map {};
class options_description {};
} }
#endif
