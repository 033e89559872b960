/* TEST INFRASTRUCTURE ONLY (oracle/_ref harness): the Monin-Obukhov surface layer of src/boundary_surface.cxx
 * (stability :54-133, stability_neutral :136-176, surfm :178-292, surfs :294-343) and of
 * include/boundary_surface_kernels.h (prepare_lut :78-138, calc_dutot :140-188, calc_duvdz_mo :190-228,
 * calc_dbdz_mo :230-247, the lookup solver :249-287) compiled in place, plus Thermo_dry's surface buoyancy
 * (calc_buoyancy_bot / calc_buoyancy_fluxbot, src/thermo_dry.cxx:129-160).  One call = Boundary_surface::exec
 * (src/boundary_surface.cxx:827-981) for swconstantz0=true, no Charnock. */
#include "boundary_surface.cxx"
#include "world.h"
#define API extern "C" __attribute__((visibility("default")))

namespace
{
    const double kGrav = 9.81;
    std::vector<float> lut_zL, lut_f;
}

API void ref_surface_lut(void* h, int mbcbot, int thermobc, double z0m, double z0h, float* zL_out, float* f_out)
{
    const Grid_data<double>& gd = static_cast<World*>(h)->grid->get_grid_data();
    lut_zL.assign(nzL_lut, 0.f); lut_f.assign(nzL_lut, 0.f);
    bsk::prepare_lut<double>(lut_zL.data(), lut_f.data(), z0m, z0h, gd.z[gd.kstart], nzL_lut,
                             static_cast<Boundary_type>(mbcbot), static_cast<Boundary_type>(thermobc));
    if (zL_out) for (int n = 0; n < nzL_lut; ++n) { zL_out[n] = lut_zL[n]; f_out[n] = lut_f[n]; }
}

/* thermo: 0 = neutral, 1 = dry (th, thbot, thfluxbot, thgradbot with bc type thermobc; thref/threfh profiles) */
/* iterative != 0: swconstantz0=false (stability<TF,false>: the iterative Obukhov solver on the z0m / z0h planes);
 * charnock != 0: calc_z0_charnock from the previous ustar first, with its two cyclic fills (src/boundary_surface.cxx:836-851) */
API void ref_surface_exec_z0(void* h, int mbcbot, int thermo, int thermobc,
                             const double* u, const double* v, double* ubot, double* vbot,
                             double* ufluxbot, double* vfluxbot, double* ugradbot, double* vgradbot,
                             const double* th, double* thbot, double* thgradbot, double* thfluxbot,
                             const double* thref, const double* threfh,
                             double* ustar, double* obuk, int* nobuk, double* z0m, double* z0h,
                             double* dutot, double* dudz, double* dvdz, double* dbdz,
                             int iterative, int charnock, double alpha_m, double alpha_ch, double alpha_h)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    const Boundary_type mb = static_cast<Boundary_type>(mbcbot), tb = static_cast<Boundary_type>(thermobc);
    if (charnock)
    {
        calc_z0_charnock<double>(z0m, z0h, ustar, alpha_m, alpha_ch, alpha_h, gd.istart, gd.iend, gd.jstart, gd.jend, gd.icells);
        W->cyclic->exec_2d(z0m);
        W->cyclic->exec_2d(z0h);
    }
    bsk::calc_dutot<double>(dutot, u, v, ubot, vbot, gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart,
                            gd.icells, gd.jcells, gd.ijcells, *W->cyclic);
    std::vector<double> b, bbot(gd.ijcells), bflux(gd.ijcells);
    if (!thermo)
        stability_neutral<double>(ustar, obuk, dutot, gd.z.data(), z0m, gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart,
                                  gd.icells, gd.jcells, gd.ijcells, mb, *W->cyclic);
    else
    {
        b.assign(gd.ncells, 0.);
        for (int ij = 0; ij < gd.ijcells; ++ij)       /* calc_buoyancy_bot + calc_buoyancy_fluxbot, src/thermo_dry.cxx:129-160 */
        {
            const int ijk = ij + gd.kstart * gd.ijcells;
            bbot[ij] = kGrav / threfh[gd.kstart] * (thbot[ij] - threfh[gd.kstart]);
            b[ijk] = kGrav / thref[gd.kstart] * (th[ijk] - thref[gd.kstart]);
            bflux[ij] = kGrav / threfh[gd.kstart] * thfluxbot[ij];
        }
        const double db_ref = kGrav / thref[gd.kstart] * (thref[gd.kstart] - threfh[gd.kstart]);   /* get_db_ref :629 */
        if (iterative)
            stability<double, false>(ustar, obuk, bflux.data(), b.data(), bbot.data(), dutot, gd.z.data(), z0m, z0h,
                                     lut_zL.data(), lut_f.data(), nobuk, db_ref, gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart,
                                     gd.icells, gd.jcells, gd.ijcells, mb, tb, *W->cyclic);
        else
            stability<double, true>(ustar, obuk, bflux.data(), b.data(), bbot.data(), dutot, gd.z.data(), z0m, z0h,
                                    lut_zL.data(), lut_f.data(), nobuk, db_ref, gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart,
                                    gd.icells, gd.jcells, gd.ijcells, mb, tb, *W->cyclic);
    }
    surfm<double>(ufluxbot, vfluxbot, ugradbot, vgradbot, ustar, obuk, u, ubot, v, vbot, z0m, gd.z[gd.kstart], mb,
                  gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.icells, gd.jcells, gd.ijcells, *W->cyclic);
    if (thermo)
    {
        surfs<double>(thbot, thgradbot, thfluxbot, ustar, obuk, th, z0h, gd.z[gd.kstart], tb,
                      gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.icells, gd.jcells, gd.ijcells, *W->cyclic);
    }
    bsk::calc_duvdz_mo<double>(dudz, dvdz, u, v, ubot, vbot, ufluxbot, vfluxbot, ustar, obuk, z0m, gd.z[gd.kstart],
                               gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.icells, gd.ijcells);
    if (thermo)
    {
        for (int ij = 0; ij < gd.ijcells; ++ij)
            bflux[ij] = kGrav / threfh[gd.kstart] * thfluxbot[ij];      /* second get_buoyancy_fluxbot, :966 */
        bsk::calc_dbdz_mo<double>(dbdz, bflux.data(), ustar, obuk, gd.z[gd.kstart], gd.istart, gd.iend, gd.jstart, gd.jend, gd.icells);
    }
}

API void ref_surface_exec(void* h, int mbcbot, int thermo, int thermobc,
                          const double* u, const double* v, double* ubot, double* vbot,
                          double* ufluxbot, double* vfluxbot, double* ugradbot, double* vgradbot,
                          const double* th, double* thbot, double* thgradbot, double* thfluxbot,
                          const double* thref, const double* threfh,
                          double* ustar, double* obuk, int* nobuk, const double* z0m, const double* z0h,
                          double* dutot, double* dudz, double* dvdz, double* dbdz)
{
    ref_surface_exec_z0(h, mbcbot, thermo, thermobc, u, v, ubot, vbot, ufluxbot, vfluxbot, ugradbot, vgradbot, th, thbot, thgradbot,
                        thfluxbot, thref, threfh, ustar, obuk, nobuk, const_cast<double*>(z0m), const_cast<double*>(z0h),
                        dutot, dudz, dvdz, dbdz, 0, 0, 0., 0., 0.);
}

/* surfs for one more scalar (src/boundary_surface.cxx:935-946) */
API void ref_surface_scalar(void* h, int bcbot, const double* var, double* varbot, double* vargradbot, double* varfluxbot,
                            const double* ustar, const double* obuk, const double* z0h)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    surfs<double>(varbot, vargradbot, varfluxbot, ustar, obuk, var, z0h, gd.z[gd.kstart], static_cast<Boundary_type>(bcbot),
                  gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.icells, gd.jcells, gd.ijcells, *W->cyclic);
}
