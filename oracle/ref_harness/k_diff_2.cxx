/* TEST INFRASTRUCTURE ONLY (oracle/_ref harness): diff_c/diff_w of src/diff_2.cxx:39,64 compiled in place. */
#include "diff_2.cxx"
#include "world.h"
#define API extern "C" __attribute__((visibility("default")))

API void ref_diff2_c(void* h, double* at, const double* a, double visc)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    diff_c<double>(at, a, visc, gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend,
            gd.icells, gd.ijcells, gd.dx, gd.dy, gd.dzi.data(), gd.dzhi.data());
}

API void ref_diff2_w(void* h, double* wt, const double* w, double visc)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    diff_w<double>(wt, w, visc, gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend,
            gd.icells, gd.ijcells, gd.dx, gd.dy, gd.dzi.data(), gd.dzhi.data());
}
