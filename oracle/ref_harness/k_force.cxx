/* TEST INFRASTRUCTURE ONLY (oracle/_ref harness): the anonymous-namespace kernels of src/force.cxx:46-292
 * (add_pressure_force, enforce_fixed_flux, calc_coriolis_2nd/_4th, calc_large_scale_source,
 * calc_nudging_tendency, advec_wls_2nd_mean/local/local_w) compiled in place. */
#include "force.cxx"
#include "world.h"
#define API extern "C" __attribute__((visibility("default")))
#define GD const Grid_data<double>& gd = static_cast<World*>(h)->grid->get_grid_data()
#define BOX gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend, gd.icells, gd.ijcells

API void ref_force_pressure(void* h, double* ut, double fbody)
{ GD; add_pressure_force<double>(ut, fbody, BOX); }

API void ref_force_fixed_flux(void* h, double* ut, const double* u, double uflux, double utrans, double dt)
{
    World* W = static_cast<World*>(h);
    GD;
    const double u_mean = W->ops->calc_mean(u);
    const double ut_mean = W->ops->calc_mean(ut);
    enforce_fixed_flux<double>(ut, uflux, u_mean, ut_mean, utrans, dt, BOX);
}

API void ref_force_coriolis(void* h, double* ut, double* vt, const double* u, const double* v, const double* ug,
                            const double* vg, double fc, double ugrid, double vgrid)
{
    GD;
    if (static_cast<World*>(h)->order == 2) calc_coriolis_2nd<double>(ut, vt, u, v, ug, vg, fc, ugrid, vgrid, BOX);
    else                                    calc_coriolis_4th<double>(ut, vt, u, v, ug, vg, fc, ugrid, vgrid, BOX);
}

API void ref_force_ls_source(void* h, double* st, const double* sls)
{ GD; calc_large_scale_source<double>(st, sls, BOX); }

API void ref_force_nudge(void* h, double* at, const double* fldmean, const double* ref, const double* factor)
{ GD; calc_nudging_tendency<double>(at, fldmean, ref, factor, BOX); }

API void ref_force_wls(void* h, int mode, double* st, const double* s, const double* wls)
{
    GD;
    if (mode == 0)      advec_wls_2nd_mean<double>(st, s, wls, gd.dzhi.data(), BOX);
    else if (mode == 1) advec_wls_2nd_local<double>(st, s, wls, gd.dzhi.data(), BOX);
    else                advec_wls_2nd_local_w<double>(st, s, wls, gd.dzi.data(), BOX);
}
