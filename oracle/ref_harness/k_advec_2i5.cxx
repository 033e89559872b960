/* TEST INFRASTRUCTURE ONLY (oracle/_ref harness): the reference's Advec_2i5 kernels compiled in place.
 * advec_u/v/w/s and calc_cfl are the anonymous-namespace functions of src/advec_2i5.cxx:59-728;
 * advec_s_lim is include/advec_monotonic.h:77 (used for scalars in [advec] fluxlimit_list, src/advec_2i5.cxx:1014-1040). */
#include "advec_2i5.cxx"
#include "world.h"
#define API extern "C" __attribute__((visibility("default")))

#define GDARGS gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend, gd.icells, gd.ijcells

API void ref_advec2i5_uvw(void* h, double* ut, double* vt, double* wt, const double* u, const double* v, const double* w)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    const double* rho = W->fields->rhoref.data();
    const double* rhoh = W->fields->rhorefh.data();
    advec_u<double>(ut, u, v, w, gd.dzi.data(), gd.dx, gd.dy, rho, rhoh, GDARGS);
    advec_v<double>(vt, u, v, w, gd.dzi.data(), gd.dx, gd.dy, rho, rhoh, GDARGS);
    advec_w<double>(wt, u, v, w, gd.dzhi.data(), gd.dx, gd.dy, rho, rhoh, GDARGS);
}

API void ref_advec2i5_s(void* h, double* st, const double* s, const double* u, const double* v, const double* w, int limit)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    const double* rho = W->fields->rhoref.data();
    const double* rhoh = W->fields->rhorefh.data();
    if (limit)
        advec_s_lim<double>(st, s, u, v, w, gd.dzi.data(), gd.dx, gd.dy, rho, rhoh, GDARGS);
    else
        advec_s<double>(st, s, u, v, w, gd.dzi.data(), gd.dx, gd.dy, rho, rhoh, GDARGS);
}

API double ref_advec2i5_cfl(void* h, const double* u, const double* v, const double* w, double dt)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    return calc_cfl<double>(u, v, w, gd.dzi.data(), gd.dx, gd.dy, dt, W->master, GDARGS);
}
