/* TEST INFRASTRUCTURE ONLY.  Compiles the reference's src/fields.cxx IN PLACE (Boost and NetCDF replaced by the shims of
 * oracle/shim) and exposes the three .out diagnostics of Fields::check_momentum/tke/mass:
 * calc_momentum_2nd :144, calc_tke_2nd :175, calc_mass :209 (the 4th-order grid uses the same 2nd-order sums,
 * src/fields.cxx:1175-1217). */
#include "fields.cxx"
#include "world.h"
#define API extern "C" __attribute__((visibility("default")))

namespace
{
    struct Args { const Grid_data<double>& gd; Master& master; double denom; };
    Args args(World* W)
    {
        const Grid_data<double>& gd = W->grid->get_grid_data();
        return Args{gd, W->master, static_cast<double>(gd.itot) * gd.jtot * gd.zsize};
    }
}

API double ref_check_momentum(void* h, const double* u, const double* v, const double* w)
{
    World* W = static_cast<World*>(h);
    Args a = args(W);
    return calc_momentum_2nd<double>(u, v, w, a.gd.dz.data(), a.denom, a.gd.istart, a.gd.iend, a.gd.jstart, a.gd.jend,
                                     a.gd.kstart, a.gd.kend, a.gd.icells, a.gd.ijcells, a.master);
}

API double ref_check_tke(void* h, const double* u, const double* v, const double* w)
{
    World* W = static_cast<World*>(h);
    Args a = args(W);
    return calc_tke_2nd<double>(u, v, w, a.gd.dz.data(), a.denom, a.gd.istart, a.gd.iend, a.gd.jstart, a.gd.jend,
                                a.gd.kstart, a.gd.kend, a.gd.icells, a.gd.ijcells, a.master);
}

API double ref_check_mass(void* h, const double* s)
{
    World* W = static_cast<World*>(h);
    Args a = args(W);
    return calc_mass<double>(s, a.gd.dz.data(), a.denom, a.gd.istart, a.gd.iend, a.gd.jstart, a.gd.jend,
                             a.gd.kstart, a.gd.kend, a.gd.icells, a.gd.ijcells, a.master);
}
