/* TEST INFRASTRUCTURE ONLY (oracle/_ref harness): calc_N2 and calc_buoyancy_tend_2nd/4th of
 * src/thermo_dry.cxx:66-79,165-198 compiled in place. */
#include "thermo_dry.cxx"
#include "world.h"
#define API extern "C" __attribute__((visibility("default")))

API void ref_thermo_N2(void* h, double* N2, const double* th, const double* thref)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    calc_N2<double>(N2, th, gd.dzi.data(), thref, gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend,
            gd.icells, gd.ijcells, gd.kcells);
}

API void ref_thermo_buoyancy_tend(void* h, double* wt, const double* th, const double* threfh)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    if (W->order == 2)
        calc_buoyancy_tend_2nd<double>(wt, th, threfh, gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend, gd.icells, gd.ijcells);
    else
        calc_buoyancy_tend_4th<double>(wt, th, threfh, gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend, gd.icells, gd.ijcells);
}
