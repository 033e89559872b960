/* TEST / MEASUREMENT INFRASTRUCTURE ONLY.  The reference's own CUDA kernels (written for sm_35..sm_70) compiled IN PLACE for
 * sm_100a and launched with the reference's launch shape (Grid::prepare_device, src/grid.cu:31-32: 256 x 1 threads for
 * itot >= 256), so that bench tools can time them on the same B200 beside this repository's kernels.  Never linked into the
 * product. */
#define USECUDA
#include "advec_2.cu"
#include "refcu.h"
#define shape refcu_shape

API void refcu_advec2_uvw(const RefCuGrid* g, double* ut, double* vt, double* wt, double* u, double* v, double* w,
                          double* rhoref, double* rhorefh, double* dzi, double* dzhi, cudaStream_t s)
{
    dim3 gr, bl; shape(*g, g->ktot, gr, bl);
    advec_uvw_g<double><<<gr, bl, 0, s>>>(ut, vt, wt, u, v, w, rhoref, rhorefh, dzi, dzhi, 1. / g->dx, 1. / g->dy,
                                           g->icells, g->ijcells, g->istart, g->jstart, g->kstart, g->iend, g->jend, g->kend);
}
API void refcu_advec2_s(const RefCuGrid* g, double* st, double* s_, double* u, double* v, double* w,
                        double* rhoref, double* rhorefh, double* dzi, cudaStream_t s)
{
    dim3 gr, bl; shape(*g, g->ktot, gr, bl);
    advec_s_g<double><<<gr, bl, 0, s>>>(st, s_, u, v, w, rhoref, rhorefh, dzi, 1. / g->dx, 1. / g->dy,
                                         g->icells, g->ijcells, g->istart, g->jstart, g->kstart, g->iend, g->jend, g->kend);
}
