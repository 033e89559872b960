/* TEST INFRASTRUCTURE ONLY (oracle/_ref harness).
 * Link-time stand-ins for reference members that live in translation units which cannot be
 * compiled here (src/stats.cxx needs NetCDF, src/fields.cxx needs Boost).  They are only
 * reachable through virtual members the harness never calls (Pres_*::exec/create); each
 * aborts loudly if it is ever reached. */
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <string>

#include "defines.h"
#include "stats.h"
#include "fields.h"
#include "field3d.h"

namespace
{
    [[noreturn]] void unreachable(const char* what)
    {
        std::fprintf(stderr, "oracle harness: %s is a link stub and must not be called\n", what);
        std::abort();
    }
}

template<> void Stats<double>::add_tendency(const Field3d<double>&, const std::string&, const std::string&,
                                            const std::string&, const std::string&) { unreachable("Stats::add_tendency"); }
template<> void Stats<double>::calc_tend(Field3d<double>&, const std::string&) { unreachable("Stats::calc_tend"); }
template<> void Stats<float>::add_tendency(const Field3d<float>&, const std::string&, const std::string&,
                                           const std::string&, const std::string&) { unreachable("Stats::add_tendency"); }
template<> void Stats<float>::calc_tend(Field3d<float>&, const std::string&) { unreachable("Stats::calc_tend"); }

template<> std::shared_ptr<Field3d<double>> Fields<double>::get_tmp() { unreachable("Fields::get_tmp"); }
template<> void Fields<double>::release_tmp(std::shared_ptr<Field3d<double>>&) { unreachable("Fields::release_tmp"); }
template<> std::shared_ptr<Field3d<float>> Fields<float>::get_tmp() { unreachable("Fields::get_tmp"); }
template<> void Fields<float>::release_tmp(std::shared_ptr<Field3d<float>>&) { unreachable("Fields::release_tmp"); }
