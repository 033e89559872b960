/* TEST / MEASUREMENT INFRASTRUCTURE ONLY: the reference's Smagorinsky kernels (src/diff_smag2.cu: calc_strain2_g :49,
 * evisc_g :141, diff_uvw_g :284, diff_c_g :428) compiled in place for sm_100a, launched as Diff_smag2::exec_viscosity / exec do. */
#define USECUDA
#include "diff_smag2.cu"
#include "refcu.h"

API void refcu_smag2_strain2(const RefCuGrid* g, double* strain2, double* u, double* v, double* w, double* dudz, double* dvdz,
                             double* dzi, double* dzhi, cudaStream_t s)
{
    dim3 gr, bl; refcu_shape(*g, g->kcells, gr, bl);        // exec_viscosity launches kcells levels (src/diff_smag2.cu:562)
    calc_strain2_g<double, Surface_model::Enabled><<<gr, bl, 0, s>>>(strain2, u, v, w, dudz, dvdz, dzi, dzhi, 1. / g->dx, 1. / g->dy,
        g->istart, g->iend, g->jstart, g->jend, g->kstart, g->kend, g->icells, g->ijcells);
}
API void refcu_smag2_evisc(const RefCuGrid* g, double* evisc, double* N2, double* dbdz, double* mlen, double* z0m, double* z, double tPri,
                           cudaStream_t s)
{
    dim3 gr, bl; refcu_shape(*g, g->kcells, gr, bl);
    evisc_g<double, Surface_model::Enabled><<<gr, bl, 0, s>>>(evisc, N2, dbdz, mlen, z0m, z, tPri,
        g->istart, g->iend, g->jstart, g->jend, g->kstart, g->kend, g->icells, g->ijcells);
}
API void refcu_smag2_diff_uvw(const RefCuGrid* g, double* ut, double* vt, double* wt, double* evisc, double* u, double* v, double* w,
                              double* fbu, double* ftu, double* fbv, double* ftv, double* dzi, double* dzhi, double* rhoref, double* rhorefh,
                              double visc, cudaStream_t s)
{
    dim3 gr, bl; refcu_shape(*g, g->ktot, gr, bl);
    diff_uvw_g<double, Surface_model::Enabled><<<gr, bl, 0, s>>>(ut, vt, wt, evisc, u, v, w, fbu, ftu, fbv, ftv, dzi, dzhi,
        1. / g->dx, 1. / g->dy, rhoref, rhorefh, visc, g->istart, g->jstart, g->kstart, g->iend, g->jend, g->kend, g->icells, g->ijcells);
}
API void refcu_smag2_diff_c(const RefCuGrid* g, double* at, double* a, double* evisc, double* fb, double* ft, double* dzi, double* dzhi,
                            double* rhoref, double* rhorefh, double tPri, double visc, cudaStream_t s)
{
    dim3 gr, bl; refcu_shape(*g, g->ktot, gr, bl);
    diff_c_g<double, Surface_model::Enabled><<<gr, bl, 0, s>>>(at, a, evisc, fb, ft, dzi, dzhi, 1. / (g->dx * g->dx), 1. / (g->dy * g->dy),
        rhoref, rhorefh, tPri, visc, g->istart, g->jstart, g->kstart, g->iend, g->jend, g->kend, g->icells, g->ijcells);
}
