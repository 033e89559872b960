// Constant-viscosity 2nd-order diffusion (src/diff_2.cxx:39-87), u/v/w fused in one launch.
#include "sys.h"
#include "ctx.h"
#ifndef MHH_DIFF2_UNROLL_N
#define MHH_DIFF2_UNROLL_N 1
#endif

namespace mhh
{
namespace
{
    constexpr int KCHUNK = MHH_KCHUNK;

    // diff_c, src/diff_2.cxx:39-62
    __device__ __forceinline__ double lap_c(const double* __restrict__ a, int ijk, int jj, int kk,
                                            double visc, double dxidxi, double dyidyi,
                                            double dzhi_t, double dzhi_c, double dzi_c)
    {
        return visc * (
                + ( (a[ijk + 1] - a[ijk]) - (a[ijk] - a[ijk - 1]) ) * dxidxi
                + ( (a[ijk + jj] - a[ijk]) - (a[ijk] - a[ijk - jj]) ) * dyidyi
                + ( (a[ijk + kk] - a[ijk]) * dzhi_t - (a[ijk] - a[ijk - kk]) * dzhi_c ) * dzi_c );
    }

    __global__ void __launch_bounds__(256, MHH_MINB)
    diff2_uvw_kernel(double* __restrict__ ut, double* __restrict__ vt, double* __restrict__ wt,
                     const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ w,
                     const double visc, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x + gd.istart;
        const int j = blockIdx.y * blockDim.y + threadIdx.y + gd.jstart;
        if (i >= gd.iend || j >= gd.jend) return;
        const int jj = gd.icells, kk = gd.ijcells;
        const int k0 = gd.kstart + blockIdx.z * KCHUNK;
        const int k1 = min(k0 + KCHUNK, gd.kend);
        const double dxidxi = 1 / (gd.dx * gd.dx), dyidyi = 1 / (gd.dy * gd.dy);

        MHH_PRAGMA_UNROLL(MHH_DIFF2_UNROLL_N)
        for (int k = k0; k < k1; ++k)
        {
            const int ijk = i + j * jj + k * kk;
            const double dzhi_t = gd.dzhi[k + 1], dzhi_c = gd.dzhi[k], dzi_c = gd.dzi[k];
            ut[ijk] += lap_c(u, ijk, jj, kk, visc, dxidxi, dyidyi, dzhi_t, dzhi_c, dzi_c);
            vt[ijk] += lap_c(v, ijk, jj, kk, visc, dxidxi, dyidyi, dzhi_t, dzhi_c, dzi_c);
            if (k > gd.kstart)   // diff_w, src/diff_2.cxx:64-87
                wt[ijk] += visc * (
                        + ( (w[ijk + 1] - w[ijk]) - (w[ijk] - w[ijk - 1]) ) * dxidxi
                        + ( (w[ijk + jj] - w[ijk]) - (w[ijk] - w[ijk - jj]) ) * dyidyi
                        + ( (w[ijk + kk] - w[ijk]) * dzi_c - (w[ijk] - w[ijk - kk]) * gd.dzi[k - 1] ) * dzhi_c );
        }
    }

    __global__ void __launch_bounds__(256, MHH_MINB)
    diff2_c_kernel(double* __restrict__ at, const double* __restrict__ a, const double visc, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x + gd.istart;
        const int j = blockIdx.y * blockDim.y + threadIdx.y + gd.jstart;
        if (i >= gd.iend || j >= gd.jend) return;
        const int jj = gd.icells, kk = gd.ijcells;
        const int k0 = gd.kstart + blockIdx.z * KCHUNK;
        const int k1 = min(k0 + KCHUNK, gd.kend);
        const double dxidxi = 1 / (gd.dx * gd.dx), dyidyi = 1 / (gd.dy * gd.dy);
        MHH_PRAGMA_UNROLL(MHH_DIFF2_UNROLL_N)
        for (int k = k0; k < k1; ++k)
        {
            const int ijk = i + j * jj + k * kk;
            at[ijk] += lap_c(a, ijk, jj, kk, visc, dxidxi, dyidyi, gd.dzhi[k + 1], gd.dzhi[k], gd.dzi[k]);
        }
    }
}  // namespace

int diff2_exec(Ctx* c, const double* u, const double* v, const double* w, double* ut, double* vt, double* wt,
               double visc, int nscalars, const double* const* s, double* const* st, const double* svisc)
{
    const GridD& gd = c->gd;
    Launch3 l = launch_interior(gd.imax, gd.jmax, (gd.kmax + KCHUNK - 1) / KCHUNK);
    MHH_LAUNCH(c, "diff2_uvw_kernel", diff2_uvw_kernel<<<l.grid, l.block, 0, c->stream>>>(ut, vt, wt, u, v, w, visc, gd));
    MHH_LAUNCH_CHECK();
    for (int n = 0; n < nscalars; ++n)
    {
        MHH_LAUNCH(c, "diff2_c_kernel", diff2_c_kernel<<<l.grid, l.block, 0, c->stream>>>(st[n], s[n], svisc[n], gd));
        MHH_LAUNCH_CHECK();
    }
    return MHH_OK;
}

}  // namespace mhh
