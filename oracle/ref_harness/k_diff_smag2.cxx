/* TEST INFRASTRUCTURE ONLY (oracle/_ref harness): Diff_smag2 kernels of src/diff_smag2.cxx compiled in place:
 * calc_strain2 :48, calc_evisc_neutral :158, calc_evisc :255, diff_u :370, diff_v :472, diff_w :575,
 * diff_c :620, calc_dnmul :712. */
#include "diff_smag2.cxx"
#include "world.h"
#define API extern "C" __attribute__((visibility("default")))

// surface_model: 1 = MO surface model (swboundary != default), 0 = resolved walls.
API void ref_smag2_strain2(void* h, int surface_model, double* strain2, const double* u, const double* v, const double* w,
                           const double* dudz, const double* dvdz)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    if (surface_model)
        calc_strain2<double, Surface_model::Enabled>(strain2, u, v, w, dudz, dvdz, gd.z.data(), gd.dzi.data(), gd.dzhi.data(),
                1./gd.dx, 1./gd.dy, gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend, gd.icells, gd.ijcells);
    else
        calc_strain2<double, Surface_model::Disabled>(strain2, u, v, w, nullptr, nullptr, gd.z.data(), gd.dzi.data(), gd.dzhi.data(),
                1./gd.dx, 1./gd.dy, gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend, gd.icells, gd.ijcells);
}

// evisc holds strain2 on entry. Includes the trailing boundary_cyclic.exec(evisc).
API void ref_smag2_evisc(void* h, int surface_model, double* evisc, const double* u, const double* v, const double* w,
                         const double* N2, const double* dbdz, const double* z0m, double cs, double tPr)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    if (surface_model)
        calc_evisc<double, Surface_model::Enabled>(evisc, u, v, w, N2, dbdz, gd.z.data(), gd.dz.data(), gd.dzi.data(), z0m,
                gd.dx, gd.dy, cs, tPr, gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend,
                gd.icells, gd.jcells, gd.ijcells, *W->cyclic);
    else
        calc_evisc<double, Surface_model::Disabled>(evisc, u, v, w, N2, nullptr, gd.z.data(), gd.dz.data(), gd.dzi.data(), nullptr,
                gd.dx, gd.dy, cs, tPr, gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend,
                gd.icells, gd.jcells, gd.ijcells, *W->cyclic);
}

API void ref_smag2_evisc_neutral(void* h, int surface_model, double* evisc, const double* u, const double* v, const double* w,
                                 const double* z0m, double cs, double visc)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    if (surface_model)
        calc_evisc_neutral<double, Surface_model::Enabled>(evisc, u, v, w, nullptr, nullptr, gd.z.data(), gd.dz.data(), gd.dzhi.data(), z0m,
                gd.dx, gd.dy, gd.zsize, cs, visc, gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend,
                gd.icells, gd.jcells, gd.ijcells, *W->cyclic);
    else
        calc_evisc_neutral<double, Surface_model::Disabled>(evisc, u, v, w, nullptr, nullptr, gd.z.data(), gd.dz.data(), gd.dzhi.data(), nullptr,
                gd.dx, gd.dy, gd.zsize, cs, visc, gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend,
                gd.icells, gd.jcells, gd.ijcells, *W->cyclic);
}

API void ref_smag2_diff_uvw(void* h, int surface_model, double* ut, double* vt, double* wt,
                            const double* u, const double* v, const double* w, const double* evisc,
                            const double* ufluxbot, const double* ufluxtop, const double* vfluxbot, const double* vfluxtop,
                            double visc)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    const double* rho = W->fields->rhoref.data();
    const double* rhoh = W->fields->rhorefh.data();
    #define ARGS_TAIL rho, rhoh, visc, gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend, gd.icells, gd.ijcells
    if (surface_model)
    {
        diff_u<double, Surface_model::Enabled>(ut, u, v, w, gd.dzi.data(), gd.dzhi.data(), 1./gd.dx, 1./gd.dy, evisc, ufluxbot, ufluxtop, ARGS_TAIL);
        diff_v<double, Surface_model::Enabled>(vt, u, v, w, gd.dzi.data(), gd.dzhi.data(), 1./gd.dx, 1./gd.dy, evisc, vfluxbot, vfluxtop, ARGS_TAIL);
    }
    else
    {
        diff_u<double, Surface_model::Disabled>(ut, u, v, w, gd.dzi.data(), gd.dzhi.data(), 1./gd.dx, 1./gd.dy, evisc, ufluxbot, ufluxtop, ARGS_TAIL);
        diff_v<double, Surface_model::Disabled>(vt, u, v, w, gd.dzi.data(), gd.dzhi.data(), 1./gd.dx, 1./gd.dy, evisc, vfluxbot, vfluxtop, ARGS_TAIL);
    }
    diff_w<double>(wt, u, v, w, gd.dzi.data(), gd.dzhi.data(), 1./gd.dx, 1./gd.dy, evisc, ARGS_TAIL);
    #undef ARGS_TAIL
}

API void ref_smag2_diff_c(void* h, int surface_model, double* at, const double* a, const double* evisc,
                          const double* fluxbot, const double* fluxtop, double tPr, double visc)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    const double* rho = W->fields->rhoref.data();
    const double* rhoh = W->fields->rhorefh.data();
    if (surface_model)
        diff_c<double, Surface_model::Enabled>(at, a, gd.dzi.data(), gd.dzhi.data(), 1./(gd.dx*gd.dx), 1./(gd.dy*gd.dy), evisc,
                fluxbot, fluxtop, rho, rhoh, tPr, visc, gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend, gd.icells, gd.ijcells);
    else
        diff_c<double, Surface_model::Disabled>(at, a, gd.dzi.data(), gd.dzhi.data(), 1./(gd.dx*gd.dx), 1./(gd.dy*gd.dy), evisc,
                fluxbot, fluxtop, rho, rhoh, tPr, visc, gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend, gd.icells, gd.ijcells);
}

API double ref_smag2_dnmul(void* h, const double* evisc, double tPr)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    return calc_dnmul<double>(evisc, gd.dzi.data(), 1./(gd.dx*gd.dx), 1./(gd.dy*gd.dy), tPr,
            gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend, gd.icells, gd.ijcells);
}
