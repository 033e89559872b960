// Constant-viscosity 4th-order diffusion for sm_100a (reference: src/diff_4.cxx:42-226).
//
// One thread per (i,j) column marches a chunk of k with the seven values a[k-3..k+3] of its own
// column in a register queue, so the vertical 7-point stencil costs ONE new global load per level;
// the x and y neighbours (3 each side) come from L1/L2 (rows are contiguous in i).
// Arithmetic keeps the reference's expression trees: three separate `at +=` terms (x, y, z), sums
// evaluated left to right, biased bg*/tg* rows at the walls.
#include "sys.h"
#include "ctx.h"

namespace mhh
{
using namespace fd;

namespace
{
    constexpr int KCHUNK = MHH_KCHUNK;

    __device__ __forceinline__ double g4(double a, double b, double c, double d)
    { return cg0 * a + cg1 * b + cg2 * c + cg3 * d; }

    // IS_W = false: diff_c (src/diff_4.cxx:42-107), rows kstart..kend-1, inner metric dzhi4[k-1..k+2], outer dzi4[k]
    // IS_W = true : diff_w (src/diff_4.cxx:109-174), rows kstart+1..kend-1, inner metric dzi4[k-2..k+1], outer dzhi4[k]
    template<bool DIM3, bool IS_W>
    __global__ void __launch_bounds__(256, MHH_MINB)
    diff4_kernel(double* __restrict__ at, const double* __restrict__ a, const double visc, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x + gd.istart;
        const int j = blockIdx.y * blockDim.y + threadIdx.y + gd.jstart;
        if (i >= gd.iend || j >= gd.jend) return;
        const int jj = gd.icells, kk = gd.ijcells;
        const int kfirst = IS_W ? gd.kstart + 1 : gd.kstart;
        const int k0 = kfirst + blockIdx.z * KCHUNK;
        const int k1 = min(k0 + KCHUNK, gd.kend);
        if (k0 >= k1) return;
        const double dxidxi = 1. / (gd.dx * gd.dx), dyidyi = 1. / (gd.dy * gd.dy);
        const double* __restrict__ min_ = IS_W ? gd.dzi4 : gd.dzhi4;    // inner metric
        const double* __restrict__ mout = IS_W ? gd.dzhi4 : gd.dzi4;    // outer metric
        const int ms = IS_W ? -1 : 0;                                   // inner metric index shift

        int ijk = i + j * jj + k0 * kk;
        double am3 = a[ijk - 3 * kk], am2 = a[ijk - 2 * kk], am1 = a[ijk - kk], a0 = a[ijk];
        double ap1 = a[ijk + kk], ap2 = a[ijk + 2 * kk];
        for (int k = k0; k < k1; ++k, ijk += kk)
        {
            // the row above kend+2 does not exist for the last level; the top row never uses ap3
            const double ap3 = (k + 3 < gd.kcells) ? a[ijk + 3 * kk] : 0.;
            if (k + 4 < gd.kcells) prefetch_l2(a + ijk + 4 * kk);
            double t = at[ijk];
            t += visc * (cdg3 * a[ijk - 3] + cdg2 * a[ijk - 2] + cdg1 * a[ijk - 1] + cdg0 * a0
                         + cdg1 * a[ijk + 1] + cdg2 * a[ijk + 2] + cdg3 * a[ijk + 3]) * dxidxi;
            if (DIM3)
                t += visc * (cdg3 * a[ijk - 3 * jj] + cdg2 * a[ijk - 2 * jj] + cdg1 * a[ijk - jj] + cdg0 * a0
                             + cdg1 * a[ijk + jj] + cdg2 * a[ijk + 2 * jj] + cdg3 * a[ijk + 3 * jj]) * dyidyi;
            const double f0 = (k == kfirst)
                                  ? (bg0 * am2 + bg1 * am1 + bg2 * a0 + bg3 * ap1)
                                  : g4(am3, am2, am1, a0);
            const double f3 = (k == gd.kend - 1)
                                  ? (tg0 * am1 + tg1 * a0 + tg2 * ap1 + tg3 * ap2)
                                  : g4(a0, ap1, ap2, ap3);
            t += visc * ( cg0 * f0 * min_[k - 1 + ms]
                        + cg1 * g4(am2, am1, a0, ap1) * min_[k + ms]
                        + cg2 * g4(am1, a0, ap1, ap2) * min_[k + 1 + ms]
                        + cg3 * f3 * min_[k + 2 + ms] )
                 * mout[k];
            at[ijk] = t;
            am3 = am2; am2 = am1; am1 = a0; a0 = ap1; ap1 = ap2; ap2 = ap3;
        }
    }

    template<bool IS_W>
    int launch(Ctx* c, double* at, const double* a, double visc)
    {
        const GridD& gd = c->gd;
        const int nk = IS_W ? gd.kmax - 1 : gd.kmax;
        Launch3 l = launch_interior(gd.imax, gd.jmax, (nk + KCHUNK - 1) / KCHUNK);
        if (gd.jtot == 1)
            MHH_LAUNCH(c, IS_W ? "diff4_w_kernel" : "diff4_c_kernel",
                       (diff4_kernel<false, IS_W><<<l.grid, l.block, 0, c->stream>>>(at, a, visc, gd)));
        else
            MHH_LAUNCH(c, IS_W ? "diff4_w_kernel" : "diff4_c_kernel",
                       (diff4_kernel<true, IS_W><<<l.grid, l.block, 0, c->stream>>>(at, a, visc, gd)));
        MHH_LAUNCH_CHECK();
        return MHH_OK;
    }
}  // namespace

// Diff_4<TF>::exec, src/diff_4.cxx:253-300
int diff4_exec(Ctx* c, const double* u, const double* v, const double* w, double* ut, double* vt, double* wt,
               double visc, int nscalars, const double* const* s, double* const* st, const double* svisc)
{
    if (c->gd.kgc < 3 || c->gd.igc < 3 || (c->gd.jtot > 1 && c->gd.jgc < 3) || c->gd.kmax < 4)
    {
        c->last_error = "diff_4 needs 3 ghost cells in every direction (swspatialorder=4) and ktot >= 4";
        return MHH_ERR_ARG;
    }
    int e;
    if ((e = launch<false>(c, ut, u, visc))) return e;
    if ((e = launch<false>(c, vt, v, visc))) return e;
    if ((e = launch<true>(c, wt, w, visc))) return e;
    for (int n = 0; n < nscalars; ++n)
        if ((e = launch<false>(c, st[n], s[n], svisc[n]))) return e;
    return MHH_OK;
}

// The part of Diff_4::exec that mhh_advec4_diff4_exec does not fuse into the advection pass: the scalars.
int diff4_exec_scalars(Ctx* c, int nscalars, const double* const* s, double* const* st, const double* svisc)
{
    int e;
    for (int n = 0; n < nscalars; ++n)
        if ((e = launch<false>(c, st[n], s[n], svisc[n]))) return e;
    return MHH_OK;
}

}  // namespace mhh
