/* TEST INFRASTRUCTURE ONLY (oracle/_ref harness): rk3/rk4 of src/timeloop.cxx:250,288 compiled in place. */
#include "timeloop.cxx"
#include "world.h"
#define API extern "C" __attribute__((visibility("default")))

API void ref_rk(void* h, int rkorder, double* a, double* at, int substep, double dt)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    if (rkorder == 3)
        rk3<double>(a, at, substep, dt, gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend,
                gd.icells, gd.ijcells, gd.ncells);
    else
        rk4<double>(a, at, substep, dt, gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend,
                gd.icells, gd.ijcells, gd.ncells);
}

API double ref_rk_subdt(int rkorder, double dt, int substep)
{
    return (rkorder == 3) ? rk3subdt<double>(dt, substep) : rk4subdt<double>(dt, substep);
}
