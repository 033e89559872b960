/* TEST INFRASTRUCTURE ONLY (oracle/_ref harness): the reference's Advec_2 kernels compiled in place.
 * advec_u/v/w/s and calc_cfl are the anonymous-namespace functions of src/advec_2.cxx:51-203. */
#include "advec_2.cxx"
#include "world.h"
#define API extern "C" __attribute__((visibility("default")))

API void ref_advec2_uvw(void* h, double* ut, double* vt, double* wt, const double* u, const double* v, double* w)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    const double* rho = W->fields->rhoref.data();
    const double* rhoh = W->fields->rhorefh.data();
    advec_u<double>(ut, u, v, w, gd.dzi.data(), gd.dx, gd.dy, rho, rhoh,
            gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend, gd.icells, gd.ijcells);
    advec_v<double>(vt, u, v, w, gd.dzi.data(), gd.dx, gd.dy, rho, rhoh,
            gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend, gd.icells, gd.ijcells);
    advec_w<double>(wt, u, v, w, gd.dzhi.data(), gd.dx, gd.dy, rho, rhoh,
            gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend, gd.icells, gd.ijcells);
}

API void ref_advec2_s(void* h, double* st, const double* s, const double* u, const double* v, const double* w)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    advec_s<double>(st, s, u, v, w, gd.dzi.data(), gd.dx, gd.dy, W->fields->rhoref.data(), W->fields->rhorefh.data(),
            gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend, gd.icells, gd.ijcells);
}

API double ref_advec2_cfl(void* h, const double* u, const double* v, const double* w, double dt)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    return calc_cfl<double>(u, v, w, gd.dzi.data(), gd.dx, gd.dy, dt, W->master,
            gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend, gd.icells, gd.ijcells);
}
