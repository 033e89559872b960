// Smagorinsky-Lilly diffusion for sm_100a (reference: src/diff_smag2.cxx).
//  - strain2 + evisc are fused: |S|^2 never touches HBM (calc_strain2 :48 + calc_evisc :255 /
//    calc_evisc_neutral :158 write/read the evisc field twice in the reference).
//  - std::pow(x, n) with n = 2 (thermo) or 1 (neutral) and the outer pow(., 1/n) are evaluated as
//    x*x / sqrt (resp. identity): same values to < 1 ulp, no FP64 pow on the hot path.  The per-level
//    Smagorinsky length cs*pow(dx*dy*dz[k], 1/3) is tabulated on the host with std::pow.
//  - the u,v,w stress divergences are fused in one pass (diff_u :370, diff_v :472, diff_w :575).
// Expression trees are kept term by term for FP64 parity (built with -fmad=false).
#include <cmath>

#include "ctx.h"

namespace mhh
{
namespace
{
    constexpr int KCHUNK = MHH_KCHUNK;

    __device__ __forceinline__ double pow2(double x) { return x * x; }

    // |S|^2 at one cell; src/diff_smag2.cxx:120-154 (interior) and :80-114 (MO bottom row).
    template<bool bottom_mo>
    __device__ __forceinline__ double strain2_at(const double* __restrict__ u, const double* __restrict__ v,
                                                 const double* __restrict__ w, int ijk, int jj, int kk,
                                                 double dxi, double dyi, double dzi_k, double dzhi_k, double dzhi_kp,
                                                 double ugradbot, double vgradbot)
    {
        const int ii = 1;
        double s;
        if (bottom_mo)
        {
            s = 2. * (
                + pow2((u[ijk + ii] - u[ijk]) * dxi)
                + pow2((v[ijk + jj] - v[ijk]) * dyi)
                + pow2((w[ijk + kk] - w[ijk]) * dzi_k)
                + 0.125 * pow2((u[ijk] - u[ijk - jj]) * dyi + (v[ijk] - v[ijk - ii]) * dxi)
                + 0.125 * pow2((u[ijk + ii] - u[ijk + ii - jj]) * dyi + (v[ijk + ii] - v[ijk]) * dxi)
                + 0.125 * pow2((u[ijk + jj] - u[ijk]) * dyi + (v[ijk + jj] - v[ijk - ii + jj]) * dxi)
                + 0.125 * pow2((u[ijk + ii + jj] - u[ijk + ii]) * dyi + (v[ijk + ii + jj] - v[ijk + jj]) * dxi)
                + 0.5 * pow2(ugradbot)
                + 0.125 * pow2((w[ijk] - w[ijk - ii]) * dxi)
                + 0.125 * pow2((w[ijk + ii] - w[ijk]) * dxi)
                + 0.125 * pow2((w[ijk + kk] - w[ijk - ii + kk]) * dxi)
                + 0.125 * pow2((w[ijk + ii + kk] - w[ijk + kk]) * dxi)
                + 0.5 * pow2(vgradbot)
                + 0.125 * pow2((w[ijk] - w[ijk - jj]) * dyi)
                + 0.125 * pow2((w[ijk + jj] - w[ijk]) * dyi)
                + 0.125 * pow2((w[ijk + kk] - w[ijk - jj + kk]) * dyi)
                + 0.125 * pow2((w[ijk + jj + kk] - w[ijk + kk]) * dyi));
        }
        else
        {
            s = 2. * (
                + pow2((u[ijk + ii] - u[ijk]) * dxi)
                + pow2((v[ijk + jj] - v[ijk]) * dyi)
                + pow2((w[ijk + kk] - w[ijk]) * dzi_k)
                + 0.125 * pow2((u[ijk] - u[ijk - jj]) * dyi + (v[ijk] - v[ijk - ii]) * dxi)
                + 0.125 * pow2((u[ijk + ii] - u[ijk + ii - jj]) * dyi + (v[ijk + ii] - v[ijk]) * dxi)
                + 0.125 * pow2((u[ijk + jj] - u[ijk]) * dyi + (v[ijk + jj] - v[ijk - ii + jj]) * dxi)
                + 0.125 * pow2((u[ijk + ii + jj] - u[ijk + ii]) * dyi + (v[ijk + ii + jj] - v[ijk + jj]) * dxi)
                + 0.125 * pow2((u[ijk] - u[ijk - kk]) * dzhi_k + (w[ijk] - w[ijk - ii]) * dxi)
                + 0.125 * pow2((u[ijk + ii] - u[ijk + ii - kk]) * dzhi_k + (w[ijk + ii] - w[ijk]) * dxi)
                + 0.125 * pow2((u[ijk + kk] - u[ijk]) * dzhi_kp + (w[ijk + kk] - w[ijk - ii + kk]) * dxi)
                + 0.125 * pow2((u[ijk + ii + kk] - u[ijk + ii]) * dzhi_kp + (w[ijk + ii + kk] - w[ijk + kk]) * dxi)
                + 0.125 * pow2((v[ijk] - v[ijk - kk]) * dzhi_k + (w[ijk] - w[ijk - jj]) * dyi)
                + 0.125 * pow2((v[ijk + jj] - v[ijk + jj - kk]) * dzhi_k + (w[ijk + jj] - w[ijk]) * dyi)
                + 0.125 * pow2((v[ijk + kk] - v[ijk]) * dzhi_kp + (w[ijk + kk] - w[ijk - jj + kk]) * dyi)
                + 0.125 * pow2((v[ijk + jj + kk] - v[ijk + jj]) * dzhi_kp + (w[ijk + jj + kk] - w[ijk + kk]) * dyi));
        }
        return s + kDsmall;
    }

    // MODE 0: thermo on (calc_evisc), 1: neutral (calc_evisc_neutral), 2: strain2 only (stage test)
    template<bool SURF, int MODE>
    __global__ void __launch_bounds__(256, MHH_MINB)
    evisc_kernel(double* __restrict__ evisc,
                 const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ w,
                 const double* __restrict__ N2, const double* __restrict__ dudz, const double* __restrict__ dvdz,
                 const double* __restrict__ dbdz, const double* __restrict__ z0m,
                 const double* __restrict__ mlen0, const double tPr, const double visc, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x + gd.istart;
        const int j = blockIdx.y * blockDim.y + threadIdx.y + gd.jstart;
        if (i >= gd.iend || j >= gd.jend) return;
        const int jj = gd.icells, kk = gd.ijcells;
        const int ij = i + j * jj;
        const int k0 = gd.kstart + blockIdx.z * KCHUNK;
        const int k1 = min(k0 + KCHUNK, gd.kend);
        const double dxi = 1. / gd.dx, dyi = 1. / gd.dy;

        double z0 = 0.;
        if (SURF && MODE != 2) z0 = z0m[ij];

        // van Driest friction velocities (resolved walls, neutral): src/diff_smag2.cxx:199-204
        double u_tau_bot = 0., u_tau_top = 0.;
        if (!SURF && MODE == 1)
        {
            const int ijk_bot = ij + gd.kstart * kk;
            const int ijk_top = ij + gd.kend * kk;
            u_tau_bot = sqrt(sqrt(pow2(visc * (u[ijk_bot] - u[ijk_bot - kk]) * gd.dzhi[gd.kstart])
                                + pow2(visc * (v[ijk_bot] - v[ijk_bot - kk]) * gd.dzhi[gd.kstart])));
            u_tau_top = sqrt(sqrt(pow2(visc * (u[ijk_top] - u[ijk_top - kk]) * gd.dzhi[gd.kend])
                                + pow2(visc * (v[ijk_top] - v[ijk_top - kk]) * gd.dzhi[gd.kend])));
        }

        for (int k = k0; k < k1; ++k)
        {
            const int ijk = ij + k * kk;
            if (k + MHH_PF < gd.kcells - 1)
            {
                const int pf = ijk + (MHH_PF + 1) * kk;
                prefetch_l2(u + pf); prefetch_l2(v + pf); prefetch_l2(w + pf);
                if (MODE == 0) prefetch_l2(N2 + pf - kk);
            }
            const bool bot = SURF && (k == gd.kstart);
            const double s2 = bot
                ? strain2_at<true>(u, v, w, ijk, jj, kk, dxi, dyi, gd.dzi[k], 0., 0., dudz[ij], dvdz[ij])
                : strain2_at<false>(u, v, w, ijk, jj, kk, dxi, dyi, gd.dzi[k], gd.dzhi[k], gd.dzhi[k + 1], 0., 0.);

            double ev;
            if (MODE == 2)
                ev = s2;
            else if (MODE == 0)
            {
                // src/diff_smag2.cxx:291-295 / :332-338 / :355-361
                double rit = (bot ? dbdz[ij] : N2[ijk]) / s2 / tPr;
                rit = fmin(rit, 1. - kDsmall);
                if (SURF)
                {
                    const double kz = kKappa * (gd.z[k] + z0);
                    const double m0 = mlen0[k];
                    const double mlen = sqrt(1. / (1. / (m0 * m0) + 1. / (kz * kz)));
                    ev = pow2(mlen) * sqrt(s2) * sqrt(1. - rit);
                }
                else
                    ev = pow2(mlen0[k]) * sqrt(s2) * sqrt(1. - rit);
            }
            else
            {
                if (SURF)
                {
                    // Mason length with n = 1, src/diff_smag2.cxx:243-247
                    const double mlen = 1. / (1. / mlen0[k] + 1. / (kKappa * (gd.z[k] + z0)));
                    ev = pow2(mlen) * sqrt(s2);
                }
                else
                {
                    // van Driest damping, src/diff_smag2.cxx:206-211
                    const double fac_bot = 1. - exp(-(gd.z[k] * u_tau_bot) / (26. * visc));
                    const double fac_top = 1. - exp(-((gd.zsize - gd.z[k]) * u_tau_top) / (26. * visc));
                    const double fac = fmin(fac_bot, fac_top);
                    ev = pow2(fac * mlen0[k]) * sqrt(s2);
                }
            }
            evisc[ijk] = ev;
        }
    }

    // mirrored wall rows for resolved walls, src/diff_smag2.cxx:214-226, :300-312
    __global__ void evisc_mirror_kernel(double* __restrict__ evisc, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        const int j = blockIdx.y;
        if (i >= gd.icells) return;
        const int ij = i + j * gd.icells;
        evisc[ij + (gd.kstart - 1) * gd.ijcells] = evisc[ij + gd.kstart * gd.ijcells];
        evisc[ij + gd.kend * gd.ijcells] = evisc[ij + (gd.kend - 1) * gd.ijcells];
    }

    template<bool SURF>
    __global__ void __launch_bounds__(256, MHH_MINB)
    smag2_diff_uvw_kernel(double* __restrict__ ut, double* __restrict__ vt, double* __restrict__ wt,
                          const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ w,
                          const double* __restrict__ evisc,
                          const double* __restrict__ ufluxbot, const double* __restrict__ ufluxtop,
                          const double* __restrict__ vfluxbot, const double* __restrict__ vfluxtop,
                          const double visc, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x + gd.istart;
        const int j = blockIdx.y * blockDim.y + threadIdx.y + gd.jstart;
        if (i >= gd.iend || j >= gd.jend) return;
        const int ii = 1, jj = gd.icells, kk = gd.ijcells;
        const int ij = i + j * jj;
        const int k0 = gd.kstart + blockIdx.z * KCHUNK;
        const int k1 = min(k0 + KCHUNK, gd.kend);
        const double dxi = 1. / gd.dx, dyi = 1. / gd.dy;

        for (int k = k0; k < k1; ++k)
        {
            const int ijk = ij + k * kk;
            if (k + MHH_PF < gd.kcells - 1)
            {
                const int pf = ijk + (MHH_PF + 1) * kk;
                prefetch_l2(u + pf); prefetch_l2(v + pf); prefetch_l2(w + pf); prefetch_l2(evisc + pf);
                prefetch_l2(ut + pf - kk); prefetch_l2(vt + pf - kk); prefetch_l2(wt + pf - kk);
            }
            const bool bot = SURF && (k == gd.kstart);
            const bool top = SURF && (k == gd.kend - 1);
            const double rho = gd.rhoref[k], rhoh_t = gd.rhorefh[k + 1], rhoh_b = gd.rhorefh[k];
            const double dzi = gd.dzi[k], dzhi_t = gd.dzhi[k + 1], dzhi_b = gd.dzhi[k];

            const double e_c = evisc[ijk], e_w = evisc[ijk - ii], e_s = evisc[ijk - jj];
            const double e_e = evisc[ijk + ii], e_n = evisc[ijk + jj];
            const double e_t = evisc[ijk + kk], e_b = evisc[ijk - kk];

            // ---- u: src/diff_smag2.cxx:395-467
            {
                const double evisce = e_c + visc;
                const double eviscw = e_w + visc;
                const double eviscn = 0.25 * (e_w + e_c + evisc[ijk - ii + jj] + e_n) + visc;
                const double eviscs = 0.25 * (evisc[ijk - ii - jj] + e_s + e_w + e_c) + visc;
                const double evisct = 0.25 * (e_w + e_c + evisc[ijk - ii + kk] + e_t) + visc;
                const double eviscb = 0.25 * (evisc[ijk - ii - kk] + e_b + e_w + e_c) + visc;

                const double hor =
                    + ( evisce * (u[ijk + ii] - u[ijk]) * dxi
                      - eviscw * (u[ijk] - u[ijk - ii]) * dxi ) * 2. * dxi
                    + ( eviscn * ((u[ijk + jj] - u[ijk]) * dyi + (v[ijk + jj] - v[ijk - ii + jj]) * dxi)
                      - eviscs * ((u[ijk] - u[ijk - jj]) * dyi + (v[ijk] - v[ijk - ii]) * dxi) ) * dyi;
                const double grad_t = (u[ijk + kk] - u[ijk]) * dzhi_t + (w[ijk + kk] - w[ijk - ii + kk]) * dxi;
                const double grad_b = (u[ijk] - u[ijk - kk]) * dzhi_b + (w[ijk] - w[ijk - ii]) * dxi;
                double ver;
                if (bot)      ver = ( rhoh_t * evisct * grad_t + rhoh_b * ufluxbot[ij] ) / rho * dzi;
                else if (top) ver = ( -rhoh_t * ufluxtop[ij] - rhoh_b * eviscb * grad_b ) / rho * dzi;
                else          ver = ( rhoh_t * evisct * grad_t - rhoh_b * eviscb * grad_b ) / rho * dzi;
                ut[ijk] += hor + ver;
            }
            // ---- v: src/diff_smag2.cxx:497-569
            {
                const double evisce = 0.25 * (e_s + e_c + evisc[ijk + ii - jj] + e_e) + visc;
                const double eviscw = 0.25 * (evisc[ijk - ii - jj] + e_w + e_s + e_c) + visc;
                const double eviscn = e_c + visc;
                const double eviscs = e_s + visc;
                const double evisct = 0.25 * (e_s + e_c + evisc[ijk + kk - jj] + e_t) + visc;
                const double eviscb = 0.25 * (evisc[ijk - kk - jj] + e_b + e_s + e_c) + visc;

                const double hor =
                    + ( evisce * ((v[ijk + ii] - v[ijk]) * dxi + (u[ijk + ii] - u[ijk + ii - jj]) * dyi)
                      - eviscw * ((v[ijk] - v[ijk - ii]) * dxi + (u[ijk] - u[ijk - jj]) * dyi) ) * dxi
                    + ( eviscn * (v[ijk + jj] - v[ijk]) * dyi
                      - eviscs * (v[ijk] - v[ijk - jj]) * dyi ) * 2. * dyi;
                const double grad_t = (v[ijk + kk] - v[ijk]) * dzhi_t + (w[ijk + kk] - w[ijk - jj + kk]) * dyi;
                const double grad_b = (v[ijk] - v[ijk - kk]) * dzhi_b + (w[ijk] - w[ijk - jj]) * dyi;
                double ver;
                if (bot)      ver = ( rhoh_t * evisct * grad_t + rhoh_b * vfluxbot[ij] ) / rho * dzi;
                else if (top) ver = ( -rhoh_t * vfluxtop[ij] - rhoh_b * eviscb * grad_b ) / rho * dzi;
                else          ver = ( rhoh_t * evisct * grad_t - rhoh_b * eviscb * grad_b ) / rho * dzi;
                vt[ijk] += hor + ver;
            }
            // ---- w: src/diff_smag2.cxx:594-614 (k from kstart+1)
            if (k > gd.kstart)
            {
                const double evisce = 0.25 * (e_b + e_c + evisc[ijk + ii - kk] + e_e) + visc;
                const double eviscw = 0.25 * (evisc[ijk - ii - kk] + e_w + e_b + e_c) + visc;
                const double eviscn = 0.25 * (e_b + e_c + evisc[ijk + jj - kk] + e_n) + visc;
                const double eviscs = 0.25 * (evisc[ijk - jj - kk] + e_s + e_b + e_c) + visc;
                const double evisct = e_c + visc;
                const double eviscb = e_b + visc;
                wt[ijk] +=
                    + ( evisce * ((w[ijk + ii] - w[ijk]) * dxi + (u[ijk + ii] - u[ijk + ii - kk]) * dzhi_b)
                      - eviscw * ((w[ijk] - w[ijk - ii]) * dxi + (u[ijk] - u[ijk - kk]) * dzhi_b) ) * dxi
                    + ( eviscn * ((w[ijk + jj] - w[ijk]) * dyi + (v[ijk + jj] - v[ijk + jj - kk]) * dzhi_b)
                      - eviscs * ((w[ijk] - w[ijk - jj]) * dyi + (v[ijk] - v[ijk - kk]) * dzhi_b) ) * dyi
                    + ( gd.rhoref[k] * evisct * (w[ijk + kk] - w[ijk]) * dzi
                      - gd.rhoref[k - 1] * eviscb * (w[ijk] - w[ijk - kk]) * gd.dzi[k - 1] ) / rhoh_b * 2. * dzhi_b;
            }
        }
    }

    // diff_c, src/diff_smag2.cxx:620-708
    template<bool SURF>
    __global__ void __launch_bounds__(256, MHH_MINB)
    smag2_diff_c_kernel(double* __restrict__ at, const double* __restrict__ a, const double* __restrict__ evisc,
                        const double* __restrict__ fluxbot, const double* __restrict__ fluxtop,
                        const double tPr, const double visc, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x + gd.istart;
        const int j = blockIdx.y * blockDim.y + threadIdx.y + gd.jstart;
        if (i >= gd.iend || j >= gd.jend) return;
        const int ii = 1, jj = gd.icells, kk = gd.ijcells;
        const int ij = i + j * jj;
        const int k0 = gd.kstart + blockIdx.z * KCHUNK;
        const int k1 = min(k0 + KCHUNK, gd.kend);
        const double dxidxi = 1. / (gd.dx * gd.dx), dyidyi = 1. / (gd.dy * gd.dy);

        int ijk = ij + k0 * kk;
        double a_b = a[ijk - kk], a_c = a[ijk];
        double e_b = evisc[ijk - kk], e_c = evisc[ijk];
        for (int k = k0; k < k1; ++k, ijk += kk)
        {
            if (k + MHH_PF < gd.kcells - 1)
            {
                const int pf = ijk + (MHH_PF + 1) * kk;
                prefetch_l2(a + pf); prefetch_l2(evisc + pf); prefetch_l2(at + pf - kk);
            }
            const double a_t = a[ijk + kk], e_t = evisc[ijk + kk];
            const bool bot = SURF && (k == gd.kstart);
            const bool top = SURF && (k == gd.kend - 1);
            const double evisce = 0.5 * (e_c + evisc[ijk + ii]) / tPr + visc;
            const double eviscw = 0.5 * (evisc[ijk - ii] + e_c) / tPr + visc;
            const double eviscn = 0.5 * (e_c + evisc[ijk + jj]) / tPr + visc;
            const double eviscs = 0.5 * (evisc[ijk - jj] + e_c) / tPr + visc;
            const double evisct = 0.5 * (e_c + e_t) / tPr + visc;
            const double eviscb = 0.5 * (e_b + e_c) / tPr + visc;

            const double hor =
                + ( evisce * (a[ijk + ii] - a_c) - eviscw * (a_c - a[ijk - ii]) ) * dxidxi
                + ( eviscn * (a[ijk + jj] - a_c) - eviscs * (a_c - a[ijk - jj]) ) * dyidyi;
            const double rhoh_t = gd.rhorefh[k + 1], rhoh_b = gd.rhorefh[k];
            double ver;
            if (bot)      ver = ( rhoh_t * evisct * (a_t - a_c) * gd.dzhi[k + 1] + rhoh_b * fluxbot[ij] ) / gd.rhoref[k] * gd.dzi[k];
            else if (top) ver = ( -rhoh_t * fluxtop[ij] - rhoh_b * eviscb * (a_c - a_b) * gd.dzhi[k] ) / gd.rhoref[k] * gd.dzi[k];
            else          ver = ( rhoh_t * evisct * (a_t - a_c) * gd.dzhi[k + 1]
                                - rhoh_b * eviscb * (a_c - a_b) * gd.dzhi[k] ) / gd.rhoref[k] * gd.dzi[k];
            at[ijk] += hor + ver;
            a_b = a_c; a_c = a_t; e_b = e_c; e_c = e_t;
        }
    }

    // calc_dnmul, src/diff_smag2.cxx:712-735, fused with the max reduction
    __global__ void __launch_bounds__(256, MHH_MINB)
    smag2_dnmul_kernel(const double* __restrict__ evisc, const double tPrfac_i, const GridD gd,
                       unsigned long long* __restrict__ out)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x + gd.istart;
        const int j = blockIdx.y * blockDim.y + threadIdx.y + gd.jstart;
        const int k0 = gd.kstart + blockIdx.z * KCHUNK;
        const int k1 = min(k0 + KCHUNK, gd.kend);
        const double dxidxi = 1. / (gd.dx * gd.dx), dyidyi = 1. / (gd.dy * gd.dy);
        unsigned long long m = 0ULL;
        if (i < gd.iend && j < gd.jend)
            for (int k = k0; k < k1; ++k)
            {
                const int ijk = i + j * gd.icells + k * gd.ijcells;
                const unsigned long long b = abs_bits(evisc[ijk] * tPrfac_i * (dxidxi + dyidyi + gd.dzi[k] * gd.dzi[k]));
                m = b > m ? b : m;
            }
        block_max_bits(m, out);
    }
}  // namespace

int cyclic_exec(Ctx* c, double* fld, int edge);   // boundary.cu

static int smag2_upload_mlen0(Ctx* c, double cs)
{
    // mlen0[k] = cs * pow(dx*dy*dz[k], 1/3): src/diff_smag2.cxx:192,237,282,319,346 (host std::pow, as the CPU build)
    const GridD& gd = c->gd;
    if (!c->mlen0_d) MHH_CUDA_CHECK(cudaMalloc(&c->mlen0_d, gd.kcells * sizeof(double)));
    MHH_CUDA_CHECK(cudaStreamSynchronize(c->stream));   // host_pinned may still be in flight
    double* h = c->host_pinned;
    for (int k = 0; k < gd.kcells; ++k)
        h[k] = cs * std::pow(gd.dx * gd.dy * c->dz[k], 1. / 3.);
    MHH_CUDA_CHECK(cudaMemcpyAsync(c->mlen0_d, h, gd.kcells * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    MHH_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    c->mlen0_cs = cs;
    return MHH_OK;
}

int smag2_exec_viscosity(Ctx* c, int surface_model, const double* u, const double* v, const double* w,
                         const double* N2, const double* dudz, const double* dvdz, const double* dbdz,
                         const double* z0m, double cs, double tPr, double visc, double* evisc, int strain_only)
{
    const GridD& gd = c->gd;
    if (c->mlen0_cs != cs || c->mlen0_d == nullptr)
        if (int e = smag2_upload_mlen0(c, cs)) return e;
    Launch3 l = launch_interior(gd.imax, gd.jmax, (gd.kmax + KCHUNK - 1) / KCHUNK);
    #define EV_ARGS evisc, u, v, w, N2, dudz, dvdz, dbdz, z0m, c->mlen0_d, tPr, visc, gd
    if (strain_only)
    {
        if (surface_model) MHH_LAUNCH(c, "evisc_kernel", evisc_kernel<true, 2><<<l.grid, l.block, 0, c->stream>>>(EV_ARGS));
        else               MHH_LAUNCH(c, "evisc_kernel", evisc_kernel<false, 2><<<l.grid, l.block, 0, c->stream>>>(EV_ARGS));
        MHH_LAUNCH_CHECK();
        return MHH_OK;
    }
    if (N2 != nullptr)
    {
        if (surface_model) MHH_LAUNCH(c, "evisc_kernel", evisc_kernel<true, 0><<<l.grid, l.block, 0, c->stream>>>(EV_ARGS));
        else               MHH_LAUNCH(c, "evisc_kernel", evisc_kernel<false, 0><<<l.grid, l.block, 0, c->stream>>>(EV_ARGS));
    }
    else
    {
        if (surface_model) MHH_LAUNCH(c, "evisc_kernel", evisc_kernel<true, 1><<<l.grid, l.block, 0, c->stream>>>(EV_ARGS));
        else               MHH_LAUNCH(c, "evisc_kernel", evisc_kernel<false, 1><<<l.grid, l.block, 0, c->stream>>>(EV_ARGS));
    }
    #undef EV_ARGS
    MHH_LAUNCH_CHECK();
    if (!surface_model)
    {
        dim3 b(128), g((gd.icells + 127) / 128, gd.jcells);
        MHH_LAUNCH(c, "evisc_mirror_kernel", evisc_mirror_kernel<<<g, b, 0, c->stream>>>(evisc, gd));
        MHH_LAUNCH_CHECK();
    }
    return cyclic_exec(c, evisc, 2);
}

int smag2_exec(Ctx* c, int surface_model, const double* u, const double* v, const double* w, const double* evisc,
               double* ut, double* vt, double* wt,
               const double* ufluxbot, const double* ufluxtop, const double* vfluxbot, const double* vfluxtop,
               double visc, int nscalars, const double* const* s, double* const* st,
               const double* const* sfluxbot, const double* const* sfluxtop, const double* svisc, double tPr)
{
    const GridD& gd = c->gd;
    Launch3 l = launch_interior(gd.imax, gd.jmax, (gd.kmax + KCHUNK - 1) / KCHUNK);
    if (surface_model)
        MHH_LAUNCH(c, "smag2_diff_uvw_kernel", smag2_diff_uvw_kernel<true><<<l.grid, l.block, 0, c->stream>>>(ut, vt, wt, u, v, w, evisc,
                ufluxbot, ufluxtop, vfluxbot, vfluxtop, visc, gd));
    else
        MHH_LAUNCH(c, "smag2_diff_uvw_kernel", smag2_diff_uvw_kernel<false><<<l.grid, l.block, 0, c->stream>>>(ut, vt, wt, u, v, w, evisc,
                nullptr, nullptr, nullptr, nullptr, visc, gd));
    MHH_LAUNCH_CHECK();
    for (int n = 0; n < nscalars; ++n)
    {
        if (surface_model)
            MHH_LAUNCH(c, "smag2_diff_c_kernel", smag2_diff_c_kernel<true><<<l.grid, l.block, 0, c->stream>>>(st[n], s[n], evisc, sfluxbot[n], sfluxtop[n], tPr, svisc[n], gd));
        else
            MHH_LAUNCH(c, "smag2_diff_c_kernel", smag2_diff_c_kernel<false><<<l.grid, l.block, 0, c->stream>>>(st[n], s[n], evisc, nullptr, nullptr, tPr, svisc[n], gd));
        MHH_LAUNCH_CHECK();
    }
    return MHH_OK;
}

int smag2_dnmul_bits(Ctx* c, const double* evisc, double tPr, unsigned long long* slot)
{
    const GridD& gd = c->gd;
    const double tPrfac_i = 1. / std::fmin(1., tPr);
    Launch3 l = launch_interior(gd.imax, gd.jmax, (gd.kmax + KCHUNK - 1) / KCHUNK);
    MHH_CUDA_CHECK(cudaMemsetAsync(slot, 0, sizeof(unsigned long long), c->stream));
    MHH_LAUNCH(c, "smag2_dnmul_kernel", smag2_dnmul_kernel<<<l.grid, l.block, 0, c->stream>>>(evisc, tPrfac_i, gd, slot));
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

}  // namespace mhh
