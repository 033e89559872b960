/* TEST / MEASUREMENT INFRASTRUCTURE ONLY: the reference's pres_2 kernels (src/pres_2.cu: pres_in_g :43, solve_in_g :116,
 * tdma_g :161, solve_out_g :93, pres_out_g :71) compiled in place for sm_100a, launched as Pres_2::exec does (:255-340).
 * The transforms in between (cuFFT + transposes, src/pres.cu) are not part of this harness. */
#define USECUDA
#include "pres_2.cu"
#include "refcu.h"

API void refcu_pres2_in(const RefCuGrid* g, double* p, double* u, double* v, double* w, double* ut, double* vt, double* wt,
                        double* dzi, double* rhoref, double* rhorefh, double dti, cudaStream_t s)
{
    dim3 gr, bl; refcu_shape(*g, g->ktot, gr, bl);
    pres_in_g<double><<<gr, bl, 0, s>>>(p, u, v, w, ut, vt, wt, dzi, rhoref, rhorefh, 1. / g->dx, 1. / g->dy, dti,
        g->icells, g->ijcells, g->itot, g->itot * g->jtot, g->itot, g->jtot, g->ktot, g->igc, g->jgc, g->kgc);
}
API void refcu_pres2_solve_in(const RefCuGrid* g, double* p, double* work3d, double* b, double* a, double* c, double* dz, double* rhoref,
                              double* bmati, double* bmatj, cudaStream_t s)
{
    dim3 gr, bl; refcu_shape(*g, g->ktot, gr, bl);
    solve_in_g<double><<<gr, bl, 0, s>>>(p, work3d, b, a, c, dz, rhoref, bmati, bmatj, g->itot, g->itot * g->jtot, g->itot, g->jtot, g->ktot, g->kstart);
}
API void refcu_pres2_tdma(const RefCuGrid* g, double* a, double* b, double* c, double* p, double* work3d, cudaStream_t s)
{
    dim3 gr, bl; refcu_shape(*g, 1, gr, bl);
    tdma_g<double><<<gr, bl, 0, s>>>(a, b, c, p, work3d, g->itot, g->itot * g->jtot, g->itot, g->jtot, g->ktot);
}
API void refcu_pres2_solve_out(const RefCuGrid* g, double* p, double* work3d, cudaStream_t s)
{
    dim3 gr, bl; refcu_shape(*g, g->ktot, gr, bl);
    solve_out_g<double><<<gr, bl, 0, s>>>(p, work3d, g->itot, g->itot * g->jtot, g->icells, g->ijcells, g->istart, g->jstart, g->kstart,
                                           g->itot, g->jtot, g->ktot);
}
API void refcu_pres2_out(const RefCuGrid* g, double* ut, double* vt, double* wt, double* p, double* dzhi, cudaStream_t s)
{
    dim3 gr, bl; refcu_shape(*g, g->ktot, gr, bl);
    pres_out_g<double><<<gr, bl, 0, s>>>(ut, vt, wt, p, dzhi, 1. / g->dx, 1. / g->dy, g->icells, g->ijcells,
                                          g->istart, g->jstart, g->kstart, g->iend, g->jend, g->kend);
}
