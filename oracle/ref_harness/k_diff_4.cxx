/* TEST INFRASTRUCTURE ONLY (oracle/_ref harness): diff_c/diff_w<TF,dim3> of src/diff_4.cxx:42,109 compiled in place.
 * dim3 is chosen as Diff_4<TF>::exec does (src/diff_4.cxx:253-300): false when jtot == 1. */
#include "diff_4.cxx"
#include "world.h"
#define API extern "C" __attribute__((visibility("default")))

API void ref_diff4_c(void* h, double* at, const double* a, double visc)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    if (gd.jtot == 1)
        diff_c<double,0>(at, a, visc, gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend,
                gd.icells, gd.ijcells, gd.dx, gd.dy, gd.dzi4.data(), gd.dzhi4.data());
    else
        diff_c<double,1>(at, a, visc, gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend,
                gd.icells, gd.ijcells, gd.dx, gd.dy, gd.dzi4.data(), gd.dzhi4.data());
}

API void ref_diff4_w(void* h, double* wt, const double* w, double visc)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    if (gd.jtot == 1)
        diff_w<double,0>(wt, w, visc, gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend,
                gd.icells, gd.ijcells, gd.dx, gd.dy, gd.dzi4.data(), gd.dzhi4.data());
    else
        diff_w<double,1>(wt, w, visc, gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend,
                gd.icells, gd.ijcells, gd.dx, gd.dy, gd.dzi4.data(), gd.dzhi4.data());
}
