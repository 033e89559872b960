/* TEST INFRASTRUCTURE ONLY (oracle/_ref harness): the reference's Advec_4 kernels compiled in place.
 * advec_u/v/w/s<TF,dim3> and calc_cfl are the anonymous-namespace functions of src/advec_4.cxx:51-590;
 * dim3 is chosen as Advec_4<TF>::exec does (src/advec_4.cxx:592-650): false when jtot == 1. */
#include "advec_4.cxx"
#include "world.h"
#define API extern "C" __attribute__((visibility("default")))

#define GDARGS gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend, gd.icells, gd.ijcells

API void ref_advec4_uvw(void* h, double* ut, double* vt, double* wt, const double* u, const double* v, double* w)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    if (gd.jtot == 1)
    {
        advec_u<double,false>(ut, u, v, w, gd.dzi4.data(), gd.dx, gd.dy, GDARGS);
        advec_v<double,false>(vt, u, v, w, gd.dzi4.data(), gd.dx, gd.dy, GDARGS);
        advec_w<double,false>(wt, u, v, w, gd.dzhi4.data(), gd.dx, gd.dy, GDARGS);
    }
    else
    {
        advec_u<double,true>(ut, u, v, w, gd.dzi4.data(), gd.dx, gd.dy, GDARGS);
        advec_v<double,true>(vt, u, v, w, gd.dzi4.data(), gd.dx, gd.dy, GDARGS);
        advec_w<double,true>(wt, u, v, w, gd.dzhi4.data(), gd.dx, gd.dy, GDARGS);
    }
}

API void ref_advec4_s(void* h, double* st, const double* s, const double* u, const double* v, const double* w)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    if (gd.jtot == 1)
        advec_s<double,false>(st, s, u, v, w, gd.dzi4.data(), gd.dx, gd.dy, GDARGS);
    else
        advec_s<double,true>(st, s, u, v, w, gd.dzi4.data(), gd.dx, gd.dy, GDARGS);
}

API double ref_advec4_cfl(void* h, const double* u, const double* v, const double* w, double dt)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    return calc_cfl<double>(u, v, w, gd.dzi.data(), gd.dx, gd.dy, dt, W->master, GDARGS);
}
