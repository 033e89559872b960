// 2nd-order grid / 5th-6th-order Wicker-Skamarock interpolated advection for sm_100a
// (reference: src/advec_2i5.cxx:59-728, flux limiter include/advec_monotonic.h:10-199).
//
// flux = vel * interp6_ws(6 pts) - |vel| * interp5_ws(6 pts); the vertical order falls 6 -> 4 -> 2 towards
// the walls.  The reference spells the reduced-order rows out one by one; they are all instances of ONE
// face rule (order of the flux through a given face depends only on the face's distance to the wall), so
// the kernels compute each vertical face flux once while marching k and reuse the top flux of level k as
// the bottom flux of level k+1 (bit-identical operands, bit-identical result).  Rows in which the
// reference drops a term get an exact 0 for it, which leaves the sum unchanged to the last bit.
// The horizontal and the vertical contributions are added to the tendency in two steps like the
// reference's two passes.
#include "sys.h"
#include <cfloat>
#ifndef MHH_ADVEC2I5_UNROLL_N
#define MHH_ADVEC2I5_UNROLL_N 2
#endif

#include "ctx.h"

namespace mhh
{
using namespace fd;

namespace
{
    constexpr int KCHUNK = MHH_KCHUNK;

    __device__ __forceinline__ double i6(const double* __restrict__ a, int o, int s)
    { return interp6_ws(a[o], a[o + s], a[o + 2 * s], a[o + 3 * s], a[o + 4 * s], a[o + 5 * s]); }
    __device__ __forceinline__ double i5(const double* __restrict__ a, int o, int s)
    { return interp5_ws(a[o], a[o + s], a[o + 2 * s], a[o + 3 * s], a[o + 4 * s], a[o + 5 * s]); }

    // Vertical flux pair (centred part F, upwind part Fa) through one face.
    //  COMP 0/1/3 (u, v, scalar): face = half level kf, ijkf = index of level kf; velocity from w.
    //  COMP 2 (w): face = cell centre c = kf, ijkf = index of level c; velocity interp2(w[c], w[c+1]).
    template<int COMP>
    __device__ __forceinline__ void vflux(const double* __restrict__ a, const double* __restrict__ w, const int ijkf,
                                          const int kf, const int sa, const int kk, const GridD& gd, double& F, double& Fa)
    {
        if (COMP == 2)
        {
            const double wv = interp2(w[ijkf], w[ijkf + kk]);
            const double rho = gd.rhoref[kf];
            if (kf == gd.kstart || kf == gd.kend - 1)
            {
                F = rho * wv * interp2(w[ijkf], w[ijkf + kk]);
                Fa = 0.;
            }
            else if (kf == gd.kstart + 1 || kf == gd.kend - 2)
            {
                F = rho * wv * interp4_ws(w[ijkf - kk], w[ijkf], w[ijkf + kk], w[ijkf + 2 * kk]);
                Fa = rho * fabs(wv) * interp3_ws(w[ijkf - kk], w[ijkf], w[ijkf + kk], w[ijkf + 2 * kk]);
            }
            else
            {
                F = rho * wv * i6(w, ijkf - 2 * kk, kk);
                Fa = rho * fabs(wv) * i5(w, ijkf - 2 * kk, kk);
            }
            return;
        }
        if (kf <= gd.kstart || kf >= gd.kend)
        {
            F = 0.; Fa = 0.;
            return;
        }
        const double wv = (COMP == 3) ? w[ijkf] : interp2(w[ijkf - sa], w[ijkf]);
        const double rh = gd.rhorefh[kf];
        if (kf == gd.kstart + 1 || kf == gd.kend - 1)
        {
            F = rh * wv * interp2(a[ijkf - kk], a[ijkf]);
            Fa = 0.;
        }
        else if (kf == gd.kstart + 2 || kf == gd.kend - 2)
        {
            F = rh * wv * interp4_ws(a[ijkf - 2 * kk], a[ijkf - kk], a[ijkf], a[ijkf + kk]);
            Fa = rh * fabs(wv) * interp3_ws(a[ijkf - 2 * kk], a[ijkf - kk], a[ijkf], a[ijkf + kk]);
        }
        else
        {
            F = rh * wv * i6(a, ijkf - 3 * kk, kk);
            Fa = rh * fabs(wv) * i5(a, ijkf - 3 * kk, kk);
        }
    }

    // COMP 0 = advec_u (:151), 1 = advec_v (:303), 2 = advec_w (:453), 3 = advec_s (:582) of src/advec_2i5.cxx.
    template<int COMP>
    __global__ void __launch_bounds__(256, 2)
    advec2i5_kernel(double* __restrict__ at, const double* __restrict__ a, const double* __restrict__ u,
                    const double* __restrict__ v, const double* __restrict__ w, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x + gd.istart;
        const int j = blockIdx.y * blockDim.y + threadIdx.y + gd.jstart;
        if (i >= gd.iend || j >= gd.jend) return;
        const int ii = 1, jj = gd.icells, kk = gd.ijcells;
        const int kfirst = (COMP == 2) ? gd.kstart + 1 : gd.kstart;
        const int k0 = kfirst + blockIdx.z * KCHUNK;
        const int k1 = min(k0 + KCHUNK, gd.kend);
        if (k0 >= k1) return;
        const double dxi = 1. / gd.dx, dyi = 1. / gd.dy;
        const int sa = (COMP == 0) ? ii : (COMP == 1) ? jj : (COMP == 2) ? kk : 0;

        int ijk = i + j * jj + k0 * kk;
        double Fb, Fab;     // flux through the lower face of level k
        if (COMP == 2) vflux<2>(a, w, ijk - kk, k0 - 1, sa, kk, gd, Fb, Fab);
        else           vflux<COMP>(a, w, ijk, k0, sa, kk, gd, Fb, Fab);

        MHH_PRAGMA_UNROLL(MHH_ADVEC2I5_UNROLL_N)
        for (int k = k0; k < k1; ++k, ijk += kk)
        {
            double t = at[ijk];
            // ---- horizontal terms ----
            double uR, uL, vR, vL;
            if (COMP == 3)
            {
                uR = u[ijk + ii]; uL = u[ijk];
                vR = v[ijk + jj]; vL = v[ijk];
            }
            else
            {
                uR = interp2(u[ijk - sa + ii], u[ijk + ii]); uL = interp2(u[ijk - sa], u[ijk]);
                vR = interp2(v[ijk - sa + jj], v[ijk + jj]); vL = interp2(v[ijk - sa], v[ijk]);
            }
            t += - ( uR * i6(a, ijk - 2 * ii, ii) - uL * i6(a, ijk - 3 * ii, ii) ) * dxi
                 + ( fabs(uR) * i5(a, ijk - 2 * ii, ii) - fabs(uL) * i5(a, ijk - 3 * ii, ii) ) * dxi
                 - ( vR * i6(a, ijk - 2 * jj, jj) - vL * i6(a, ijk - 3 * jj, jj) ) * dyi
                 + ( fabs(vR) * i5(a, ijk - 2 * jj, jj) - fabs(vL) * i5(a, ijk - 3 * jj, jj) ) * dyi;

            // ---- vertical terms ----
            double Ft, Fat;
            if (COMP == 2) vflux<2>(a, w, ijk, k, sa, kk, gd, Ft, Fat);
            else           vflux<COMP>(a, w, ijk + kk, k + 1, sa, kk, gd, Ft, Fat);
            if (COMP == 2)
                t += - ( Ft - Fb ) / gd.rhorefh[k] * gd.dzhi[k] + ( Fat - Fab ) / gd.rhorefh[k] * gd.dzhi[k];
            else
                t += - ( Ft - Fb ) / gd.rhoref[k] * gd.dzi[k] + ( Fat - Fab ) / gd.rhoref[k] * gd.dzi[k];
            at[ijk] = t;
            Fb = Ft; Fab = Fat;
        }
    }

    // ---- Koren (1993) limiter, include/advec_monotonic.h:10-74 ----
    __device__ __forceinline__ double koren_phi(const double num, const double den)
    {
        const double denom = copysign(1., den) * fmax(fabs(den), DBL_EPSILON);
        const double two_r = 2. * num / denom;
        return fmax(0., fmin(two_r, fmin((1. / 3.) * (1. + two_r), 2.)));
    }
    // MODE 0: flux_lim, 1: flux_lim_bot, 2: flux_lim_top
    template<int MODE>
    __device__ __forceinline__ double flux_lim(const double vel, const double sm2, const double sm1, const double sp1, const double sp2)
    {
        if (vel >= 0.)
        {
            if (MODE == 1) return vel * sm1;
            const double phi = koren_phi(sp1 - sm1, sm1 - sm2);
            return vel * (sm1 + 0.5 * phi * (sm1 - sm2));
        }
        else
        {
            if (MODE == 2) return vel * sp1;
            const double phi = koren_phi(sm1 - sp1, sp1 - sp2);
            return vel * (sp1 + 0.5 * phi * (sp1 - sp2));
        }
    }

    __device__ __forceinline__ double vflux_lim(const double* __restrict__ s, const double* __restrict__ w, const int ijkf,
                                                const int kf, const int kk, const GridD& gd)
    {
        if (kf <= gd.kstart || kf >= gd.kend) return 0.;
        const double sm2 = s[ijkf - 2 * kk], sm1 = s[ijkf - kk], sp1 = s[ijkf], sp2 = s[ijkf + kk];
        const double f = (kf == gd.kstart + 1) ? flux_lim<1>(w[ijkf], sm2, sm1, sp1, sp2)
                       : (kf == gd.kend - 1)   ? flux_lim<2>(w[ijkf], sm2, sm1, sp1, sp2)
                                               : flux_lim<0>(w[ijkf], sm2, sm1, sp1, sp2);
        return gd.rhorefh[kf] * f;
    }

    // advec_s_lim, include/advec_monotonic.h:77-199
    __global__ void __launch_bounds__(256, 2)
    advec2i5_s_lim_kernel(double* __restrict__ st, const double* __restrict__ s, const double* __restrict__ u,
                          const double* __restrict__ v, const double* __restrict__ w, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x + gd.istart;
        const int j = blockIdx.y * blockDim.y + threadIdx.y + gd.jstart;
        if (i >= gd.iend || j >= gd.jend) return;
        const int ii = 1, jj = gd.icells, kk = gd.ijcells;
        const int k0 = gd.kstart + blockIdx.z * KCHUNK;
        const int k1 = min(k0 + KCHUNK, gd.kend);
        const double dxi = 1. / gd.dx, dyi = 1. / gd.dy;
        int ijk = i + j * jj + k0 * kk;
        double Fb = vflux_lim(s, w, ijk, k0, kk, gd);
        MHH_PRAGMA_UNROLL(MHH_ADVEC2I5_UNROLL_N)
        for (int k = k0; k < k1; ++k, ijk += kk)
        {
            const double Ft = vflux_lim(s, w, ijk + kk, k + 1, kk, gd);
            st[ijk] +=
                - ( flux_lim<0>(u[ijk + ii], s[ijk - ii], s[ijk], s[ijk + ii], s[ijk + 2 * ii])
                  - flux_lim<0>(u[ijk], s[ijk - 2 * ii], s[ijk - ii], s[ijk], s[ijk + ii]) ) * dxi
                - ( flux_lim<0>(v[ijk + jj], s[ijk - jj], s[ijk], s[ijk + jj], s[ijk + 2 * jj])
                  - flux_lim<0>(v[ijk], s[ijk - 2 * jj], s[ijk - jj], s[ijk], s[ijk + jj]) ) * dyi
                - ( Ft - Fb ) / gd.rhoref[k] * gd.dzi[k];
            Fb = Ft;
        }
    }

    // calc_cfl, src/advec_2i5.cxx:59-148 (vertical order 2/4/6 by row), fused with the max reduction.
    __global__ void __launch_bounds__(256, MHH_MINB)
    advec2i5_cfl_kernel(const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ w,
                        const GridD gd, unsigned long long* __restrict__ out)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x + gd.istart;
        const int j = blockIdx.y * blockDim.y + threadIdx.y + gd.jstart;
        const int ii = 1, jj = gd.icells, kk = gd.ijcells;
        const int k0 = gd.kstart + blockIdx.z * KCHUNK;
        const int k1 = min(k0 + KCHUNK, gd.kend);
        const double dxi = 1. / gd.dx, dyi = 1. / gd.dy;
        unsigned long long m = 0ULL;
        if (i < gd.iend && j < gd.jend)
            for (int k = k0; k < k1; ++k)
            {
                const int ijk = i + j * jj + k * kk;
                double wc;
                if (k == gd.kstart || k == gd.kend - 1)
                    wc = interp2(w[ijk], w[ijk + kk]);
                else if (k == gd.kstart + 1 || k == gd.kend - 2)
                    wc = interp4_ws(w[ijk - kk], w[ijk], w[ijk + kk], w[ijk + 2 * kk]);
                else
                    wc = i6(w, ijk - 2 * kk, kk);
                const double c = fabs(i6(u, ijk - 2 * ii, ii)) * dxi + fabs(i6(v, ijk - 2 * jj, jj)) * dyi + fabs(wc) * gd.dzi[k];
                const unsigned long long b = abs_bits(c);
                m = b > m ? b : m;
            }
        block_max_bits(m, out);
    }

    int check_grid(Ctx* c)
    {
        const GridD& gd = c->gd;
        if (gd.igc < 3 || gd.jgc < 3 || gd.kgc < 1 || gd.kmax < 6)
        {
            c->last_error = "advec_2i5 needs igc = jgc = 3, kgc >= 1 (src/advec_2i5.cxx:42-45) and ktot >= 6";
            return MHH_ERR_ARG;
        }
        return MHH_OK;
    }
}  // namespace

// Advec_2i5<TF>::exec, src/advec_2i5.cxx:977-1045
int advec2i5_exec(Ctx* c, const double* u, const double* v, const double* w, double* ut, double* vt, double* wt,
                  int nscalars, const double* const* s, double* const* st, const int* limit_flags)
{
    if (int e = check_grid(c)) return e;
    const GridD& gd = c->gd;
    Launch3 l = launch_interior(gd.imax, gd.jmax, (gd.kmax + KCHUNK - 1) / KCHUNK);
    Launch3 lw = launch_interior(gd.imax, gd.jmax, (gd.kmax - 1 + KCHUNK - 1) / KCHUNK);
    MHH_LAUNCH(c, "advec2i5_u_kernel", advec2i5_kernel<0><<<l.grid, l.block, 0, c->stream>>>(ut, u, u, v, w, gd));
    MHH_LAUNCH_CHECK();
    MHH_LAUNCH(c, "advec2i5_v_kernel", advec2i5_kernel<1><<<l.grid, l.block, 0, c->stream>>>(vt, v, u, v, w, gd));
    MHH_LAUNCH_CHECK();
    MHH_LAUNCH(c, "advec2i5_w_kernel", advec2i5_kernel<2><<<lw.grid, lw.block, 0, c->stream>>>(wt, w, u, v, w, gd));
    MHH_LAUNCH_CHECK();
    for (int n = 0; n < nscalars; ++n)
    {
        if (limit_flags && limit_flags[n])
        {
            if (gd.kgc < 2)
            {
                c->last_error = "flux-limited scalars need kgc >= 2 (src/advec_2i5.cxx:44)";
                return MHH_ERR_ARG;
            }
            MHH_LAUNCH(c, "advec2i5_s_lim_kernel", advec2i5_s_lim_kernel<<<l.grid, l.block, 0, c->stream>>>(st[n], s[n], u, v, w, gd));
        }
        else
            MHH_LAUNCH(c, "advec2i5_s_kernel", advec2i5_kernel<3><<<l.grid, l.block, 0, c->stream>>>(st[n], s[n], u, v, w, gd));
        MHH_LAUNCH_CHECK();
    }
    return MHH_OK;
}

int advec2i5_cfl_bits(Ctx* c, const double* u, const double* v, const double* w, unsigned long long* slot)
{
    if (int e = check_grid(c)) return e;
    const GridD& gd = c->gd;
    Launch3 l = launch_interior(gd.imax, gd.jmax, (gd.kmax + KCHUNK - 1) / KCHUNK);
    MHH_CUDA_CHECK(cudaMemsetAsync(slot, 0, sizeof(unsigned long long), c->stream));
    MHH_LAUNCH(c, "advec2i5_cfl_kernel", advec2i5_cfl_kernel<<<l.grid, l.block, 0, c->stream>>>(u, v, w, gd, slot));
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

}  // namespace mhh
