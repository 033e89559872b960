// Smagorinsky-Lilly diffusion for sm_100a (reference: src/diff_smag2.cxx).
//  - strain2 + evisc are fused: |S|^2 never touches HBM (calc_strain2 :48 + calc_evisc :255 /
//    calc_evisc_neutral :158 write/read the evisc field twice in the reference).
//  - std::pow(x, n) with n = 2 (thermo) or 1 (neutral) and the outer pow(., 1/n) are evaluated as
//    x*x / sqrt (resp. identity): same values to < 1 ulp, no FP64 pow on the hot path.  The per-level
//    Smagorinsky length cs*pow(dx*dy*dz[k], 1/3) is tabulated on the host with std::pow.
//  - the u,v,w stress divergences are fused in one pass (diff_u :370, diff_v :472, diff_w :575).
// Expression trees are kept term by term for FP64 parity (built with -fmad=false).
#include "sys.h"
#include <cmath>

#include "ctx.h"
#include "tma.cuh"

// k levels unrolled in the momentum kernel's march (the register queues rotate by renaming instead of by moves)
#ifndef MHH_UVW_UNROLL_N
#define MHH_UVW_UNROLL_N 2
#endif
#define MHH_UVW_UNROLL MHH_PRAGMA_UNROLL(MHH_UVW_UNROLL_N)
// 1: the momentum kernel re-reads the seven level k-1 neighbour values of the w stencils from the plane ring (held one level
// longer) instead of carrying them in registers
#ifndef MHH_UVW_REREAD_B
#define MHH_UVW_REREAD_B 0
#endif
#ifndef MHH_EVISC_UNROLL_N
#define MHH_EVISC_UNROLL_N 3
#endif
// the LIM variant of the evisc pass (once per step) carries two running maxima more: its own unroll / occupancy knobs
#ifndef MHH_EVISC_LIM_UNROLL_N
#define MHH_EVISC_LIM_UNROLL_N 1
#endif
#ifndef MHH_EVISC_LIM_MINB
#define MHH_EVISC_LIM_MINB 2
#endif
#ifndef MHH_SCALAR_UNROLL_N
#define MHH_SCALAR_UNROLL_N 6
#endif

namespace mhh
{
namespace
{
    constexpr int KCHUNK = MHH_TILE_KCHUNK;
#ifndef MHH_TILE_X
#define MHH_TILE_X 32
#endif
// slots of the momentum kernel's plane ring: 6 x 4 fields x (34 x 6) doubles = 39 KB keeps four CTAs on an SM
constexpr int UVW_SLOTS = 6;
#ifndef MHH_EVISC_MINB
#define MHH_EVISC_MINB 3
#endif
#ifndef MHH_SCALAR_MINB
#define MHH_SCALAR_MINB 4
#endif
// 1: the momentum kernel stages its planes by TMA; 0 (default): per-thread cp.async.  Measured at 512^3 (r02b): TMA makes
// the scalar pass 1.63 -> 1.29 ms and the evisc pass 1.71 -> 1.54 ms, but the momentum pass 3.12 -> 3.76 ms — it runs without
// a per-level __syncthreads() and at its register limit; the one issuing thread's wait / issue sequence sits on the critical
// path of its warp, which every other warp then waits for.
#ifndef MHH_UVW_TMA
#define MHH_UVW_TMA 0
#endif
#ifndef MHH_UVW_TILE_Y
#define MHH_UVW_TILE_Y 4      // 32 x 4 columns per CTA, four CTAs per SM: shorter barrier waits than 32 x 8 (4.91 -> 4.66 ms)
#endif

    __device__ __forceinline__ double pow2(double x) { return x * x; }

    // load on a once-per-column path (wall rows): the index passes through an opaque asm so that the compiler does not
    // hoist the address out of the k loop, where it would cost two registers (or a spill) on every level
    __device__ __forceinline__ double rare_load(const double* __restrict__ p, int idx)
    {
        asm volatile("" : "+r"(idx));
        return p[idx];
    }

    // MODE 0: thermo on (calc_evisc), 1: neutral (calc_evisc_neutral), 2: strain2 only (stage test),
    // MODE 3: as 0 with N2 = g/thref * dth/dz (calc_N2, src/thermo_dry.cxx:66-79) evaluated on the fly from th: the N2 pointer
    //         then carries th and `mlen0 + kcells` the g/thref table (saves writing and re-reading the N2 field)
    //
    // |S|^2 (src/diff_smag2.cxx:120-154 interior, :80-114 MO bottom row) needs u at (i,i+1) x (j-1,j,j+1) x (k-1,k,k+1),
    // v at (i-1,i,i+1) x (j,j+1) x (k-1,k,k+1) and w at the 5-point star on levels k and k+1.  Marching k, the columns
    // u_c, u_e, v_c, v_n (three levels) and the w star (two levels) stay in registers: 18 global loads per point instead
    // of 31 (ncu r01: L1 68 %, DRAM 30 %).
    // Shared-memory tiles: the (TX+2) x (TY+2) planes of u,v,w sit in a ring of four slots filled by cp.async two levels
    // ahead (see TileMap); the columns u_c, u_e, v_c, v_n (levels k-1, k) and the w star (level k) stay in registers.
    // LIM: the time-step limits of the step ride along (the reference evaluates them right after exec_viscosity,
    // src/model.cxx:354-357): lim[0] = max of calc_cfl's sum (src/advec_2.cxx:51-78, before the * dt), lim[1] = max of
    // calc_dnmul (src/diff_smag2.cxx:712-735) of the evisc just computed — both read values this pass already holds.
    // TMA: the planes are staged by one elected thread with cp.async.bulk.tensor (tma.cuh); `full[slot]` then carries the
    // transaction bytes and replaces cp.async.wait_group; the per-level __syncthreads() still frees the slot of plane k-1.
    template<bool SURF, int MODE, int TX, int TY, bool LIM = false, bool TMA = false>
    __global__ void __launch_bounds__(TX * TY, (LIM ? MHH_EVISC_LIM_MINB : MHH_EVISC_MINB))
    evisc_kernel(double* __restrict__ evisc,
                 const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ w,
                 const double* __restrict__ N2, const double* __restrict__ dudz, const double* __restrict__ dvdz,
                 const double* __restrict__ dbdz, const double* __restrict__ z0m,
                 const double* __restrict__ mlen0, const double tPr, const double visc, const GridD gd,
                 unsigned long long* __restrict__ lim, const double tPrfac_i, const __grid_constant__ TmaMaps4 maps)
    {
        using Map = TileMap<TX, TY, 1>;
        constexpr int W = Map::W, PS = TMA ? tma_plane_pitch(Map::W, Map::HT) : Map::N, NS = 4;
        __shared__ __align__(128) double sm[NS][3][PS];  // [slot][u,v,w][tile plane]
        __shared__ unsigned long long full[NS];
        __shared__ double met[5][KCHUNK + 2];            // dzi, dzhi, z, mlen0 (thermo: mlen0^2 or its inverse), g/thref at k0-1 .. k1

        const int tid = threadIdx.y * TX + threadIdx.x;
        const int i0 = blockIdx.x * TX + gd.istart, j0 = blockIdx.y * TY + gd.jstart;
        const int i = i0 + threadIdx.x, j = j0 + threadIdx.y;
        const bool live = (i < gd.iend) && (j < gd.jend);
        const int jj = gd.icells, kk = gd.ijcells;
        const int ij = min(i, gd.icells - 1) + min(j, gd.jcells - 1) * jj;
        const int k0 = gd.kstart + blockIdx.z * KCHUNK;
        const int k1 = min(k0 + KCHUNK, gd.kend);
        const double dxi = 1. / gd.dx, dyi = 1. / gd.dy;
        const double tPr_i = 1. / tPr;
        const double dxidxi = 1. / (gd.dx * gd.dx), dyidyi = 1. / (gd.dy * gd.dy);   // LIM only
        unsigned long long m_cfl = 0ULL, m_dn = 0ULL;

        Map map;
        if (!TMA) map.init(i0, j0, gd.icells, gd.jcells, tid);
        if (TMA && tid == 0)
        {
            for (int sl = 0; sl < NS; ++sl) mbar_init(&full[sl], 1);
            fence_mbar_init();
        }
        auto stage = [&](const int k)                    // plane k -> slot (k - k0 + 1) & 3; always one commit group
        {
            if (TMA)
            {
                if (tid == 0 && k <= k1)
                {
                    const int sl = (k - k0 + 1) & (NS - 1);
                    mbar_expect_tx(&full[sl], 3u * Map::N * (unsigned)sizeof(double));
                    tma_load_plane(sm[sl][0], &maps.m[0], i0 - 1, j0 - 1, k, &full[sl]);
                    tma_load_plane(sm[sl][1], &maps.m[1], i0 - 1, j0 - 1, k, &full[sl]);
                    tma_load_plane(sm[sl][2], &maps.m[2], i0 - 1, j0 - 1, k, &full[sl]);
                }
                return;
            }
            if (k <= k1)
            {
                const int sl = (k - k0 + 1) & (NS - 1);
                const long off = (long)k * kk;
                map.issue(sm[sl][0], u + off, tid); map.issue(sm[sl][1], v + off, tid); map.issue(sm[sl][2], w + off, tid);
            }
            cp_async_commit();
        };
        // plane k is the ((k - k0 + 1) >> 2)-th fill of its slot
        auto landed = [&](const int k) { mbar_wait(&full[(k - k0 + 1) & (NS - 1)], ((k - k0 + 1) >> 2) & 1); };
        stage(k0 - 1); stage(k0); stage(k0 + 1); stage(k0 + 2);
        if (tid < KCHUNK + 2)
        {
            const int k = min(k0 - 1 + tid, gd.kcells - 1);
            met[0][tid] = gd.dzi[k]; met[1][tid] = gd.dzhi[k]; met[2][tid] = gd.z[k];
            if (MODE == 1) met[3][tid] = mlen0[k];
            if (MODE == 0 || MODE == 3) met[3][tid] = SURF ? 1. / (mlen0[k] * mlen0[k]) : mlen0[k] * mlen0[k];
            if (MODE == 3) met[4][tid] = mlen0[gd.kcells + k];
        }

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

        const int c = (threadIdx.y + 1) * W + threadIdx.x + 1;
        const int cw = c - 1, ce = c + 1, cs = c - W, cn = c + W, cwn = c + W - 1, ces = c - W + 1, cen = c + W + 1;

        int ijk = ij + k0 * kk;
        // centre-only columns come straight from global memory, one level ahead of their use
        double th_b = 0., th_c = 0., th_n = 0., n2_n = 0.;
        if (MODE == 3) { th_b = N2[ijk - kk]; th_c = N2[ijk]; th_n = N2[ijk + kk]; }
        if (MODE == 0) n2_n = N2[ijk];

        if (!TMA) cp_async_wait<2>();                    // planes k0-1 and k0 have landed
        __syncthreads();
        if (TMA) { landed(k0 - 1); landed(k0); }
        double u_b, u_c, u_eb, u_e, v_b, v_c, v_nb, v_n, w_c, w_w, w_e, w_s, w_n;
        {
            const double* B = sm[0][0]; const double* C = sm[1][0];
            u_b = B[c]; u_eb = B[ce]; u_c = C[c]; u_e = C[ce];
            v_b = B[PS + c]; v_nb = B[PS + cn]; v_c = C[PS + c]; v_n = C[PS + cn];
            w_c = C[2 * PS + c]; w_w = C[2 * PS + cw]; w_e = C[2 * PS + ce]; w_s = C[2 * PS + cs]; w_n = C[2 * PS + cn];
        }
        double xz_c, xz_e, yz_c, yz_n;                  // lower-face vertical cross terms of the level being computed
        {
            const double dzhi_k = met[1][1];
            xz_c = 0.125 * pow2((u_c - u_b) * dzhi_k + (w_c - w_w) * dxi);
            xz_e = 0.125 * pow2((u_e - u_eb) * dzhi_k + (w_e - w_c) * dxi);
            yz_c = 0.125 * pow2((v_c - v_b) * dzhi_k + (w_c - w_s) * dyi);
            yz_n = 0.125 * pow2((v_n - v_nb) * dzhi_k + (w_n - w_c) * dyi);
        }

        MHH_PRAGMA_UNROLL((LIM ? MHH_EVISC_LIM_UNROLL_N : MHH_EVISC_UNROLL_N))
        for (int k = k0; k < k1; ++k, ijk += kk)
        {
            if (TMA) landed(k + 1);
            else cp_async_wait<1>();                     // plane k+1 has landed (k+2 may still be in flight)
            __syncthreads();                             // ... for every thread; everyone is done with plane k-1
            stage(k + 3);                                // into the slot of plane k-1
            const double th_t = th_n, n2_c = n2_n;
            if (MODE == 3 && k + 1 < k1) th_n = N2[ijk + 2 * kk];
            if (MODE == 0 && k + 1 < k1) n2_n = N2[ijk + kk];

            const double* C = sm[(k - k0 + 1) & (NS - 1)][0];
            const double* T = sm[(k - k0 + 2) & (NS - 1)][0];
            const double u_t = T[c], u_et = T[ce];
            const double u_s = C[cs], u_es = C[ces], u_n = C[cn], u_en = C[cen];
            const double v_t = T[PS + c], v_nt = T[PS + cn];
            const double v_w = C[PS + cw], v_e = C[PS + ce], v_wn = C[PS + cwn], v_en = C[PS + cen];
            const double w_t = T[2 * PS + c], w_wt = T[2 * PS + cw], w_et = T[2 * PS + ce];
            const double w_st = T[2 * PS + cs], w_nt = T[2 * PS + cn];
            const int km = k - k0 + 1;                   // row of level k in met[]
            const double dzi_k = met[0][km], dzhi_kp = met[1][km + 1];

            const bool bot = SURF && (k == gd.kstart);
            // the four vertical cross terms on the upper face of level k are the lower-face terms of level k+1: they are
            // evaluated once and carried (same expression, same value)
            const double xz_t  = 0.125 * pow2((u_t - u_c) * dzhi_kp + (w_t - w_wt) * dxi);
            const double xz_et = 0.125 * pow2((u_et - u_e) * dzhi_kp + (w_et - w_t) * dxi);
            const double yz_t  = 0.125 * pow2((v_t - v_c) * dzhi_kp + (w_t - w_st) * dyi);
            const double yz_nt = 0.125 * pow2((v_nt - v_n) * dzhi_kp + (w_nt - w_t) * dyi);
            // left-to-right sum of src/diff_smag2.cxx:120-154 (interior) / :80-114 (MO bottom row)
            double acc = pow2((u_e - u_c) * dxi)
                       + pow2((v_n - v_c) * dyi)
                       + pow2((w_t - w_c) * dzi_k)
                       + 0.125 * pow2((u_c - u_s) * dyi + (v_c - v_w) * dxi)
                       + 0.125 * pow2((u_e - u_es) * dyi + (v_e - v_c) * dxi)
                       + 0.125 * pow2((u_n - u_c) * dyi + (v_n - v_wn) * dxi)
                       + 0.125 * pow2((u_en - u_e) * dyi + (v_en - v_n) * dxi);
            if (bot)
            {
                acc = acc + 0.5 * pow2(rare_load(dudz, ij));
                acc = acc + 0.125 * pow2((w_c - w_w) * dxi);
                acc = acc + 0.125 * pow2((w_e - w_c) * dxi);
                acc = acc + 0.125 * pow2((w_t - w_wt) * dxi);
                acc = acc + 0.125 * pow2((w_et - w_t) * dxi);
                acc = acc + 0.5 * pow2(rare_load(dvdz, ij));
                acc = acc + 0.125 * pow2((w_c - w_s) * dyi);
                acc = acc + 0.125 * pow2((w_n - w_c) * dyi);
                acc = acc + 0.125 * pow2((w_t - w_st) * dyi);
                acc = acc + 0.125 * pow2((w_nt - w_t) * dyi);
            }
            else
            {
                acc = acc + xz_c; acc = acc + xz_e; acc = acc + xz_t; acc = acc + xz_et;
                acc = acc + yz_c; acc = acc + yz_n; acc = acc + yz_t; acc = acc + yz_nt;
            }
            double s2 = 2. * acc;
            xz_c = xz_t; xz_e = xz_et; yz_c = yz_t; yz_n = yz_nt;
            s2 = s2 + kDsmall;

            double ev;
            if (MODE == 2)
                ev = s2;
            else if (MODE == 0 || MODE == 3)
            {
                // src/diff_smag2.cxx:291-295 / :332-338 / :355-361
                double n2;
                if (MODE == 3)
                {
                    n2 = met[4][km] * 0.5 * (th_t - th_b) * dzi_k;
                    th_b = th_c; th_c = th_t;
                }
                else
                    n2 = bot ? 0. : n2_c;
                double rit = div_by((bot ? rare_load(dbdz, ij) : n2) * fast_rcp(s2), tPr, tPr_i);
                rit = fmin(rit, 1. - kDsmall);
                // mlen^2 * sqrt(s2) * sqrt(1 - rit) with mlen^2 = 1 / (1/mlen0^2 + 1/(kappa (z+z0))^2) taken without the
                // sqrt-then-square round trip, the two roots merged, and the remaining quotients / root through fast_rcp /
                // fast_sqrt (~1 ulp each): a few ulp from the reference expression (tolerance 1e-12) at a fraction of the six
                // IEEE divide / sqrt sequences per point (the kernel is FP64-issue bound, ncu r01d)
                if (SURF)
                {
                    const double kz = kKappa * (met[2][km] + z0);
                    const double kz2 = kz * kz;
                    const double mlen2 = kz2 * fast_rcp(met[3][km] * kz2 + 1.);   // met[3] = 1 / mlen0^2; one reciprocal, not two divisions
                    ev = mlen2 * fast_sqrt(s2 * (1. - rit));
                }
                else
                    ev = met[3][km] * fast_sqrt(s2 * (1. - rit));                 // met[3] = mlen0^2
            }
            else
            {
                if (SURF)
                {
                    // Mason length with n = 1, src/diff_smag2.cxx:243-247
                    const double mlen = 1. / (1. / met[3][km] + 1. / (kKappa * (met[2][km] + z0)));
                    ev = pow2(mlen) * sqrt(s2);
                }
                else
                {
                    // van Driest damping, src/diff_smag2.cxx:206-211
                    const double fac_bot = 1. - exp(-(met[2][km] * u_tau_bot) / (26. * visc));
                    const double fac_top = 1. - exp(-((gd.zsize - met[2][km]) * u_tau_top) / (26. * visc));
                    const double fac = fmin(fac_bot, fac_top);
                    ev = pow2(fac * met[3][km]) * sqrt(s2);
                }
            }
            if (live) evisc[ijk] = ev;
            if (LIM && live)
            {
                const double cf = fabs(fd::interp2(u_c, u_e)) * dxi + fabs(fd::interp2(v_c, v_n)) * dyi
                                + fabs(fd::interp2(w_c, w_t)) * dzi_k;
                const unsigned long long bc = abs_bits(cf);
                m_cfl = bc > m_cfl ? bc : m_cfl;
                const unsigned long long bd = abs_bits(ev * tPrfac_i * (dxidxi + dyidyi + dzi_k * dzi_k));
                m_dn = bd > m_dn ? bd : m_dn;
            }
            u_c = u_t; u_e = u_et;
            v_c = v_t; v_n = v_nt;
            w_c = w_t; w_w = w_wt; w_e = w_et; w_s = w_st; w_n = w_nt;
        }
        if (!TMA) cp_async_wait<0>();
        if (LIM)
        {
            block_max_bits(m_cfl, lim);
            __syncthreads();                             // block_max_bits reuses its shared scratch
            block_max_bits(m_dn, lim + 1);
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

    // u, v, w stress divergences in one pass (diff_u :370, diff_v :472, diff_w :575).  A CTA owns a TX x TY tile of columns
    // and marches k through its chunk; the (TX+2) x (TY+2) tile planes of evisc,u,v,w live in a ring of four shared-memory
    // slots filled by cp.async two levels ahead (TileMap, mhh_common.cuh).  Per point and level: 25 shared-memory loads
    // and the 3 tendency read-modify-writes; no global load sits on the dependent path (ncu r01b: the register-queue
    // version of this kernel was long-scoreboard bound at 2 CTAs/SM).
    //
    // ADVEC: the kernel first adds the 2nd-order advection tendencies (advec_u/v/w, src/advec_2.cxx:98-171), which need
    //        exactly the u,v,w neighbours the stresses already hold; BUOY: before that, the dry buoyancy tendency on w
    //        (src/thermo_dry.cxx:43-62).  The sums are formed in the time loop's order, ((t + buoyancy) + advection) +
    //        diffusion, so the result is bit-identical to Thermo::exec, Advec::exec, Diff::exec run one after the other,
    //        at 11 instead of 22 field passes over HBM.
    // FORCE: the terms that follow Diff::exec for u,v,w in Model::exec ride in the pass, applied to the finished tendency of a
    //        cell in the time loop's order (so the result is bit-identical to Buffer::exec and Force::exec run afterwards):
    //   sig != NULL : Buffer::exec, t -= sigma(z_k) (a - abuf[k]) for k >= bufk (w: sigma(zh_k), bufkh)   (src/buffer.cxx:37-58)
    //   coriolis    : Force::exec swlspres=geo, calc_coriolis_2nd                                         (src/force.cxx:76-107)
    struct ForceTail
    {
        const double* sig;       // sigma(z[k]) (kcells) followed by sigma(zh[k]) (kcells), or NULL
        const double* ubuf; const double* vbuf; const double* wbuf;
        int bufk, bufkh;
        int coriolis;
        double fc, ugrid, vgrid;
        const double* ug; const double* vg;
    };

    // TMA: the tile planes are staged by cp.async.bulk.tensor (one elected thread, tma.cuh) instead of per-thread cp.async
    template<bool SURF, bool ADVEC, bool BUOY, int TX, int TY, bool TMA = false, bool FORCE = false>
#ifndef MHH_UVW_MINB
#define MHH_UVW_MINB (384 / (MHH_TILE_X * MHH_UVW_TILE_Y))
#endif
    __global__ void __launch_bounds__(TX * TY, MHH_UVW_MINB)
    smag2_diff_uvw_tile_kernel(double* __restrict__ ut, double* __restrict__ vt, double* __restrict__ wt,
                               const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ w,
                               const double* __restrict__ evisc,
                               const double* __restrict__ ufluxbot, const double* __restrict__ ufluxtop,
                               const double* __restrict__ vfluxbot, const double* __restrict__ vfluxtop,
                               const double visc, const double* __restrict__ th, const double* __restrict__ threfh,
                               const double* __restrict__ g_threfh, const GridD gd, const int fresh,
                               const ForceTail ft, const __grid_constant__ TmaMaps4 maps)
    {
        using Map = TileMap<TX, TY, 1>;
        // PS: doubles per staged plane; the TMA destinations must stay 128-byte aligned
        constexpr int W = Map::W, PS = TMA ? tma_plane_pitch(Map::W, Map::HT) : Map::N, NS = UVW_SLOTS, NT = TX * TY;
        extern __shared__ __align__(128) double sm[];    // [slot][evisc,u,v,w][tile plane]: NS * 4 * PS doubles (dynamic)
        // rhoref, rhorefh, 1/rhoref, 1/rhorefh, dzi, dzhi, threfh, g/threfh, dzi/4, dzhi/4, 2 dzhi at k0-1 .. k1 (one array per
        // metric: 16-byte records read as pairs cost this kernel registers it does not have, 3.20 -> 3.74 ms);
        // FORCE: + sigma(z), sigma(zh), ubuf, vbuf, wbuf, vg, ug
        __shared__ double met[FORCE ? 18 : 11][KCHUNK + 2];
        __shared__ unsigned long long full[NS], empty[NS];

        const int tid = threadIdx.y * TX + threadIdx.x;
        const int i0 = blockIdx.x * TX + gd.istart, j0 = blockIdx.y * TY + gd.jstart;
        const int i = i0 + threadIdx.x, j = j0 + threadIdx.y;
        const bool live = (i < gd.iend) && (j < gd.jend);
        const bool ldt = live && !fresh;                 // fresh: the tendency arrays hold no data (logically zero), start from 0
        const int jj = gd.icells, kk = gd.ijcells;
        const int ij = min(i, gd.icells - 1) + min(j, gd.jcells - 1) * jj;
        const int k0 = gd.kstart + blockIdx.z * KCHUNK;
        const int k1 = min(k0 + KCHUNK, gd.kend);
        const double dxi = 1. / gd.dx, dyi = 1. / gd.dy;
        // Scalings by powers of two commute with every rounding (no underflow at these magnitudes), so the reference's
        // interp2(a,b) = 0.5 (a+b) is carried as the plain sum a+b and the factors are collected in the metric:
        // (0.5 s1)(0.5 s2) - (0.5 s3)(0.5 s4) = 0.25 (s1 s2 - s3 s4), and (0.25 X) dxi = X (0.25 dxi), bit for bit.
        // Likewise (X * 2.) * dxi = X * (2 dxi).
        const double qdxi = 0.25 * dxi, qdyi = 0.25 * dyi, dxi2 = 2. * dxi, dyi2 = 2. * dyi;

        Map map;
        if (!TMA) map.init(i0, j0, gd.icells, gd.jcells, tid);
        // Plane ring of NS slots.  Plane k has ordinal q = k - (k0-1) and lives in slot q & (NS-1); it is the (q / NS)-th fill
        // of that slot.  full[slot] completes when all copies of a fill have landed (cp.async: every thread's arrival; TMA: the
        // issuing thread's arrival plus the transaction bytes of the four bulk copies), empty[slot] when all NT threads have
        // read the fill for the last time (as level k of iteration q-1).
        if (tid == 0)
        {
            for (int sl = 0; sl < NS; ++sl) { mbar_init(&full[sl], TMA ? 1 : NT); mbar_init(&empty[sl], NT); }
            if (TMA)
            {
                fence_mbar_init();
                tma_prefetch_map(&maps.m[0]); tma_prefetch_map(&maps.m[1]); tma_prefetch_map(&maps.m[2]); tma_prefetch_map(&maps.m[3]);
            }
        }
        __syncthreads();
        auto stage = [&](const int q)
        {
            const int k = k0 - 1 + q;
            if (k > k1) return;
            const int sl = q % NS;
            double* dst = sm + sl * 4 * PS;
            if (TMA)
            {
                // the issuing role rotates over the warps, level by level: no warp carries the whole cost of the waits
                if (tid == ((q % (NT / 32)) << 5))
                {
                    if (q >= NS) mbar_wait(&empty[sl], ((q / NS) - 1) & 1);     // the previous fill has been read by everyone
                    mbar_expect_tx(&full[sl], 4u * Map::N * (unsigned)sizeof(double));
                    tma_load_plane(dst, &maps.m[0], i0 - 1, j0 - 1, k, &full[sl]);
                    tma_load_plane(dst + PS, &maps.m[1], i0 - 1, j0 - 1, k, &full[sl]);
                    tma_load_plane(dst + 2 * PS, &maps.m[2], i0 - 1, j0 - 1, k, &full[sl]);
                    tma_load_plane(dst + 3 * PS, &maps.m[3], i0 - 1, j0 - 1, k, &full[sl]);
                }
                return;
            }
            if (q >= NS) mbar_wait(&empty[sl], ((q / NS) - 1) & 1);     // the previous fill has been read by everyone
            const long off = (long)k * kk;
            map.issue(dst, evisc + off, tid);      map.issue(dst + PS, u + off, tid);
            map.issue(dst + 2 * PS, v + off, tid); map.issue(dst + 3 * PS, w + off, tid);
            mbar_arrive_on_cp_async(&full[sl]);
        };
        auto plane = [&](const int q) -> const double*   // waits until plane q is complete in its slot
        {
            mbar_wait(&full[q % NS], (q / NS) & 1);
            return sm + (q % NS) * 4 * PS;
        };
        stage(0); stage(1); stage(2); stage(3);
        for (int t = tid; t < KCHUNK + 2; t += NT)       // (the chunk may be longer than the CTA has threads)
        {
            const int k = min(k0 - 1 + t, gd.kcells - 1);
            met[0][t] = gd.rhoref[k]; met[1][t] = gd.rhorefh[k]; met[2][t] = gd.rhoref_i[k];
            met[3][t] = gd.rhorefh_i[k]; met[4][t] = gd.dzi[k]; met[5][t] = gd.dzhi[k];
            if (BUOY) { met[6][t] = threfh[k]; met[7][t] = g_threfh[k]; }
            met[8][t] = 0.25 * gd.dzi[k]; met[9][t] = 0.25 * gd.dzhi[k]; met[10][t] = 2. * gd.dzhi[k];
            if (FORCE)
            {
                const bool sp = ft.sig != nullptr;
                met[11][t] = (sp && k >= ft.bufk) ? ft.sig[k] : 0.;
                met[12][t] = (sp && k >= ft.bufkh) ? ft.sig[gd.kcells + k] : 0.;
                met[13][t] = sp ? ft.ubuf[k] : 0.; met[14][t] = sp ? ft.vbuf[k] : 0.; met[15][t] = sp ? ft.wbuf[k] : 0.;
                met[16][t] = ft.coriolis ? ft.vg[k] : 0.;
                met[17][t] = ft.coriolis ? ft.ug[k] : 0.;
            }
        }

        // tile-plane offsets of the centre and its neighbours
        const int c = (threadIdx.y + 1) * W + threadIdx.x + 1;
        const int cw = c - 1, ce = c + 1, cs = c - W, cn = c + W, cwn = c + W - 1, cws = c - W - 1, ces = c - W + 1;

        __syncthreads();                                 // met[] is complete
        // Column state carried up the k march.  Besides the centre values of levels k-1 (…_b) and k (…_c):
        //   f?_b, a?_b : diffusive / advective flux through the lower face (= the upper-face flux of the level below);
        //   gu_b, gv_b : the vertical shear at the lower west / south edge, (u_c-u_b) dzhi + (w_c-w_w) dxi.  Diff u/v evaluate
        //                it as the upper-edge shear of the level below and diff_w reads the same two products in the other
        //                order (a + b = b + a), so it is evaluated once;
        //   su_b, sv_b, sw_w, sw_s : the sums u_b+u_c, v_b+v_c, w_w+w_c, w_s+w_c of the advecting/advected interpolations,
        //                which were u_c+u_t, v_c+v_t, w_wt+w_t, w_st+w_t one level below.
        double e_b, e_c, e_wb, e_w, e_sb, e_s, e_eb, e_nb, u_c, u_eb, v_c, v_nb, w_c;
        double fu_b, fv_b, fw_b, gu_b, gv_b, au_b = 0., av_b = 0., aw_b = 0., su_b = 0., sv_b = 0., sw_w = 0., sw_s = 0.;
        {
            const double* B = plane(0); const double* C = plane(1);
            e_b = B[c]; e_wb = B[cw]; e_sb = B[cs]; e_eb = B[ce]; e_nb = B[cn];
            e_c = C[c]; e_w = C[cw]; e_s = C[cs];
            const double u_b = B[PS + c]; u_eb = B[PS + ce]; u_c = C[PS + c];
            const double v_b = B[2 * PS + c]; v_nb = B[2 * PS + cn]; v_c = C[2 * PS + c];
            const double w_b = B[3 * PS + c]; w_c = C[3 * PS + c];
            const double w_w = C[3 * PS + cw], w_s = C[3 * PS + cs];
            const double rhoh_b = met[1][1], dzhi_b = met[5][1];
            gu_b = (u_c - u_b) * dzhi_b + (w_c - w_w) * dxi;
            gv_b = (v_c - v_b) * dzhi_b + (w_c - w_s) * dyi;
            fu_b = rhoh_b * (0.25 * (e_wb + e_b + e_w + e_c) + visc) * gu_b;
            fv_b = rhoh_b * (0.25 * (e_sb + e_b + e_s + e_c) + visc) * gv_b;
            fw_b = met[0][0] * (e_b + visc) * (w_c - w_b) * met[4][0];
            if (ADVEC)
            {
                su_b = u_b + u_c; sv_b = v_b + v_c; sw_w = w_w + w_c; sw_s = w_s + w_c;
                au_b = rhoh_b * sw_w * su_b;             // 4 x the reference's flux: see qdxi above
                av_b = rhoh_b * sw_s * sv_b;
                aw_b = met[0][0] * (w_b + w_c) * (w_b + w_c);
            }
        }
        if (!MHH_UVW_REREAD_B) mbar_arrive(&empty[0]);   // plane k0-1 is not read again

        int ijk = ij + k0 * kk;
        double ut_c = 0., vt_c = 0., wt_c = 0.;
        if (ldt) { ut_c = ut[ijk]; vt_c = vt[ijk]; wt_c = wt[ijk]; }
        double th_b = 0., th_c = 0., th_n = 0.;          // theta column (centre only), one level ahead from global memory
        if (BUOY) { th_b = th[ijk - kk]; th_c = th[ijk]; th_n = th[ijk + kk]; }
        MHH_UVW_UNROLL
        for (int k = k0; k < k1; ++k, ijk += kk)
        {
            const int it = k - k0;                       // level k: planes it+1 (this level) and it+2 (the level above)
            stage(it + 4);                               // plane k+3, into the slot last read NS - 4 levels ago
            // tendencies of the next level: in flight during this level's arithmetic
            double ut_n = 0., vt_n = 0., wt_n = 0.;
            if (ldt && k + 1 < k1) { ut_n = ut[ijk + kk]; vt_n = vt[ijk + kk]; wt_n = wt[ijk + kk]; }
            const double th_t = th_n;
            if (BUOY && k + 1 < k1) th_n = th[ijk + 2 * kk];

            const double* C = sm + ((it + 1) % NS) * 4 * PS;      // complete: it was waited for as T one level ago
            const double* T = plane(it + 2);
            const bool bot = SURF && (k == gd.kstart);
            const bool top = SURF && (k == gd.kend - 1);
            const int km = k - k0 + 1;                   // row of level k in met[]
            const double rho = met[0][km], rhoh_t = met[1][km + 1], rhoh_b = met[1][km];
            const double rho_i = met[2][km];
            const double dzi = met[4][km], dzhi_t = met[5][km + 1], dzhi_b = met[5][km];

            const double e_t = T[c], e_wt = T[cw], e_st = T[cs];
            const double e_e = C[ce], e_n = C[cn], e_wn = C[cwn], e_ws = C[cws], e_es = C[ces];
            const double u_t = T[PS + c], u_e = C[PS + ce], u_w = C[PS + cw], u_n = C[PS + cn], u_s = C[PS + cs];
            const double u_es = C[PS + ces];
            const double v_t = T[2 * PS + c], v_n = C[2 * PS + cn], v_e = C[2 * PS + ce], v_w = C[2 * PS + cw], v_s = C[2 * PS + cs];
            const double v_wn = C[2 * PS + cwn];
            const double w_t = T[3 * PS + c], w_wt = T[3 * PS + cw], w_st = T[3 * PS + cs];
            const double w_e = C[3 * PS + ce], w_n = C[3 * PS + cn];
            if (MHH_UVW_REREAD_B)
            {
                const double* B = sm + (it % NS) * 4 * PS;       // plane k-1
                e_b = B[c]; e_wb = B[cw]; e_sb = B[cs]; e_eb = B[ce]; e_nb = B[cn];
                u_eb = B[PS + ce]; v_nb = B[2 * PS + cn];
                mbar_arrive(&empty[it % NS]);         // this thread's last read of plane k-1
            }
            else
                mbar_arrive(&empty[(it + 1) % NS]);   // this thread's last read of plane k

            if (BUOY && k > gd.kstart)      // calc_buoyancy_tend_2nd, src/thermo_dry.cxx:43-62
                wt_c = wt_c + met[7][km] * (fd::interp2(th_b, th_c) - met[6][km]);
            if (ADVEC)
            {
                const double qdzi = met[8][km], qdzhi_b = met[9][km];
                // sums of the interpolation pairs (interp2 without its 0.5)
                const double su_t = u_c + u_t, sv_t = v_c + v_t, sw_wt = w_wt + w_t, sw_st = w_st + w_t, sw_t = w_c + w_t;
                const double su_e = u_c + u_e, su_w = u_w + u_c, su_n = u_c + u_n, su_s = u_s + u_c;
                const double sv_e = v_c + v_e, sv_w = v_w + v_c, sv_n = v_c + v_n, sv_s = v_s + v_c;
                // fluxes through the upper faces (the lower-face ones were the upper-face ones of the level below)
                const double au_t = rhoh_t * sw_wt * su_t;
                const double av_t = rhoh_t * sw_st * sv_t;
                const double aw_t = rho * sw_t * sw_t;
                // advec_u, src/advec_2.cxx:98-109
                ut_c = ut_c +
                    ( - ( su_e * su_e
                        - su_w * su_w ) * qdxi
                      - ( (v_wn + v_n) * su_n
                        - sv_w * su_s ) * qdyi
                      - div_by( au_t - au_b, rho, rho_i ) * qdzi );
                // advec_v, src/advec_2.cxx:129-140
                vt_c = vt_c +
                    ( - ( (u_es + u_e) * sv_e
                        - su_s * sv_w ) * qdxi
                      - ( sv_n * sv_n
                        - sv_s * sv_s ) * qdyi
                      - div_by( av_t - av_b, rho, rho_i ) * qdzi );
                // advec_w, src/advec_2.cxx:160-171 (k from kstart+1)
                if (k > gd.kstart)
                    wt_c = wt_c +
                        ( - ( (u_eb + u_e) * (w_c + w_e)
                            - su_b * sw_w ) * qdxi
                          - ( (v_nb + v_n) * (w_c + w_n)
                            - sv_b * sw_s ) * qdyi
                          - div_by( aw_t - aw_b, rhoh_b, met[3][km] ) * qdzhi_b );
                au_b = au_t; av_b = av_t; aw_b = aw_t;
                su_b = su_t; sv_b = sv_t; sw_w = sw_wt; sw_s = sw_st;
            }
            // ---- u: src/diff_smag2.cxx:395-467
            double gu_t, gv_t;
            {
                const double evisce = e_c + visc;
                const double eviscw = e_w + visc;
                const double eviscn = 0.25 * (e_w + e_c + e_wn + e_n) + visc;
                const double eviscs = 0.25 * (e_ws + e_s + e_w + e_c) + visc;
                const double evisct = 0.25 * (e_w + e_c + e_wt + e_t) + visc;

                const double hor =
                    + ( evisce * (u_e - u_c) * dxi
                      - eviscw * (u_c - u_w) * dxi ) * dxi2
                    + ( eviscn * ((u_n - u_c) * dyi + (v_n - v_wn) * dxi)
                      - eviscs * ((u_c - u_s) * dyi + (v_c - v_w) * dxi) ) * dyi;
                gu_t = (u_t - u_c) * dzhi_t + (w_t - w_wt) * dxi;
                const double fu_t = rhoh_t * evisct * gu_t;
                double ver;
                if (bot)      ver = div_by( fu_t + rhoh_b * rare_load(ufluxbot, ij), rho, rho_i ) * dzi;
                else if (top) ver = div_by( -rhoh_t * rare_load(ufluxtop, ij) - fu_b, rho, rho_i ) * dzi;
                else          ver = div_by( fu_t - fu_b, rho, rho_i ) * dzi;
                double tn = ut_c + (hor + ver);
                if (FORCE)
                {
                    if (ft.sig != nullptr) tn = tn - met[11][km] * (u_c - met[13][km]);
                    if (ft.coriolis) tn = tn + ft.fc * (0.25 * (v_w + v_c + v_wn + v_n) + ft.vgrid - met[16][km]);
                }
                if (live) ut[ijk] = tn;
                fu_b = fu_t;
            }
            // ---- v: src/diff_smag2.cxx:497-569
            {
                const double evisce = 0.25 * (e_s + e_c + e_es + e_e) + visc;
                const double eviscw = 0.25 * (e_ws + e_w + e_s + e_c) + visc;
                const double eviscn = e_c + visc;
                const double eviscs = e_s + visc;
                const double evisct = 0.25 * (e_s + e_c + e_st + e_t) + visc;

                const double hor =
                    + ( evisce * ((v_e - v_c) * dxi + (u_e - u_es) * dyi)
                      - eviscw * ((v_c - v_w) * dxi + (u_c - u_s) * dyi) ) * dxi
                    + ( eviscn * (v_n - v_c) * dyi
                      - eviscs * (v_c - v_s) * dyi ) * dyi2;
                gv_t = (v_t - v_c) * dzhi_t + (w_t - w_st) * dyi;
                const double fv_t = rhoh_t * evisct * gv_t;
                double ver;
                if (bot)      ver = div_by( fv_t + rhoh_b * rare_load(vfluxbot, ij), rho, rho_i ) * dzi;
                else if (top) ver = div_by( -rhoh_t * rare_load(vfluxtop, ij) - fv_b, rho, rho_i ) * dzi;
                else          ver = div_by( fv_t - fv_b, rho, rho_i ) * dzi;
                double tn = vt_c + (hor + ver);
                if (FORCE)
                {
                    if (ft.sig != nullptr) tn = tn - met[11][km] * (v_c - met[14][km]);
                    if (ft.coriolis) tn = tn - ft.fc * (0.25 * (u_s + u_c + u_es + u_e) + ft.ugrid - met[17][km]);
                }
                if (live) vt[ijk] = tn;
                fv_b = fv_t;
            }
            // ---- w: src/diff_smag2.cxx:594-614 (k from kstart+1)
            const double fw_t = rho * (e_c + visc) * (w_t - w_c) * dzi;
            if (k > gd.kstart)
            {
                const double evisce = 0.25 * (e_b + e_c + e_eb + e_e) + visc;
                const double eviscw = 0.25 * (e_wb + e_w + e_b + e_c) + visc;
                const double eviscn = 0.25 * (e_b + e_c + e_nb + e_n) + visc;
                const double eviscs = 0.25 * (e_sb + e_s + e_b + e_c) + visc;
                // west / south edge shear: (w_c - w_w) dxi + (u_c - u_b) dzhi_b = gu_b, (w_c - w_s) dyi + (v_c - v_b) dzhi_b = gv_b
                const double wd =
                    + ( evisce * ((w_e - w_c) * dxi + (u_e - u_eb) * dzhi_b)
                      - eviscw * gu_b ) * dxi
                    + ( eviscn * ((w_n - w_c) * dyi + (v_n - v_nb) * dzhi_b)
                      - eviscs * gv_b ) * dyi
                    + div_by( fw_t - fw_b, rhoh_b, met[3][km] ) * met[10][km];
                double tn = wt_c + wd;
                if (FORCE && ft.sig != nullptr) tn = tn - met[12][km] * (w_c - met[15][km]);
                if (live) wt[ijk] = tn;
            }
            else if (fresh && live) wt[ijk] = 0.;      // the row the pass never writes
            fw_b = fw_t; gu_b = gu_t; gv_b = gv_t;
            if (BUOY) { th_b = th_c; th_c = th_t; }
            if (MHH_UVW_REREAD_B) { e_c = e_t; e_w = e_wt; e_s = e_st; u_c = u_t; v_c = v_t; }
            else
            {
                e_b = e_c; e_c = e_t; e_wb = e_w; e_w = e_wt; e_sb = e_s; e_s = e_st; e_eb = e_e; e_nb = e_n;
                u_c = u_t; u_eb = u_e;
                v_c = v_t; v_nb = v_n;
            }
            w_c = w_t;
            ut_c = ut_n; vt_c = vt_n; wt_c = wt_n;
        }
    }

    // diff_c, src/diff_smag2.cxx:620-708, on shared-memory tiles (the scalar and evisc; with ADVEC also u and v).
    // ADVEC: the kernel first adds the 2nd-order scalar advection (advec_s, src/advec_2.cxx:191-202), so that the scalar,
    //        its tendency and evisc cross HBM once for both operators: (t + advection) + diffusion, bit-identical to
    //        Advec::exec followed by Diff::exec.
    // Optional tail of the scalar pass (mhh_advec2_smag2_exec_rk), applied to the finished tendency t of a cell before it is
    // stored — the terms that follow Diff::exec for a scalar in Model::exec when nothing else touches its tendency:
    //   sig != NULL : Buffer::exec, t -= sigma(z_k) (a - abuf[k]) for k >= bufk            (src/buffer.cxx:37-58)
    //   rk_mode 1/2 : Timeloop::exec, a_new = a + cB dt t ; t = cA t (mode 2: last substep, the caller zeroes the
    //                 tendency array afterwards)                                            (src/timeloop.cxx:250-285)
    // a_new is a SECOND array (neighbouring CTAs still read a); the caller swaps the two pointers.
    struct ScalarTail
    {
        double* a_new;
        double cB_dt, cA_next;
        int rk_mode;
        const double* sig;
        const double* abuf;
        int bufk;
        int fresh;       // the tendency array holds no data (lazy-tendency mode): start from 0 instead of reading it
    };

    template<bool SURF, bool ADVEC, int TX, int TY, bool TMA = false>
    __global__ void __launch_bounds__(TX * TY, MHH_SCALAR_MINB)
    smag2_scalar_tile_kernel(double* __restrict__ at, const double* __restrict__ a, const double* __restrict__ evisc,
                             const double* __restrict__ fluxbot, const double* __restrict__ fluxtop,
                             const double tPr, const double visc,
                             const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ w,
                             const ScalarTail tail, const GridD gd, const __grid_constant__ TmaMaps4 maps, const int kc0)
    {
        using Map = TileMap<TX, TY, 1>;
        constexpr int W = Map::W, PS = TMA ? tma_plane_pitch(Map::W, Map::HT) : Map::N, NS = 4, NF = ADVEC ? 4 : 2;
        __shared__ __align__(128) double sm[NS][NF][PS]; // [slot][a, evisc(, u, v)][tile plane]
        __shared__ unsigned long long full[NS];
        // per-level metric records at k0-1 .. k1, read as 16-byte pairs: {rhoref, 1/rhoref} {dzi, dzi/2} {rhorefh, dzhi}
        constexpr int MR = 6;
        __shared__ __align__(16) double met[(KCHUNK + 2) * MR];
        auto met2 = [&](const int row, const int pair) -> double2
        { return *reinterpret_cast<const double2*>(&met[row * MR + 2 * pair]); };

        const int tid = threadIdx.y * TX + threadIdx.x;
        const int i0 = blockIdx.x * TX + gd.istart, j0 = blockIdx.y * TY + gd.jstart;
        const int i = i0 + threadIdx.x, j = j0 + threadIdx.y;
        const bool live = (i < gd.iend) && (j < gd.jend);
        const int jj = gd.icells, kk = gd.ijcells;
        const int ij = min(i, gd.icells - 1) + min(j, gd.jcells - 1) * jj;
        const int k0 = gd.kstart + (blockIdx.z + kc0) * KCHUNK;     // kc0: first chunk of this call (k-range parts)
        const int k1 = min(k0 + KCHUNK, gd.kend);
        const double dxi = 1. / gd.dx, dyi = 1. / gd.dy;
        const double dxidxi = 1. / (gd.dx * gd.dx), dyidyi = 1. / (gd.dy * gd.dy);
        const double tPr_i = 1. / tPr;
        // Powers of two commute with every rounding: (0.5 s) / tPr = s / (2 tPr) through the same reciprocal sequence, and
        // u (0.5 s) dxi = (u s) (0.5 dxi), bit for bit — the interp2 halvings are collected in the constants.
        const double tPr2 = 2. * tPr, tPr_hi = 0.5 * tPr_i, hdxi = 0.5 * dxi, hdyi = 0.5 * dyi;

        Map map;
        if (!TMA) map.init(i0, j0, gd.icells, gd.jcells, tid);
        if (TMA && tid == 0)
        {
            for (int sl = 0; sl < NS; ++sl) mbar_init(&full[sl], 1);
            fence_mbar_init();
        }
        auto stage = [&](const int k)                    // plane k -> slot (k - k0 + 1) & 3; always one commit group
        {
            if (TMA)
            {
                if (tid == 0 && k <= k1)
                {
                    const int sl = (k - k0 + 1) & (NS - 1);
                    mbar_expect_tx(&full[sl], (unsigned)NF * Map::N * (unsigned)sizeof(double));
                    #pragma unroll
                    for (int q = 0; q < NF; ++q) tma_load_plane(sm[sl][q], &maps.m[q], i0 - 1, j0 - 1, k, &full[sl]);
                }
                return;
            }
            if (k <= k1)
            {
                const int sl = (k - k0 + 1) & (NS - 1);
                const long off = (long)k * kk;
                map.issue(sm[sl][0], a + off, tid); map.issue(sm[sl][1], evisc + off, tid);
                if (ADVEC) { map.issue(sm[sl][NF - 2], u + off, tid); map.issue(sm[sl][NF - 1], v + off, tid); }
            }
            cp_async_commit();
        };
        auto landed = [&](const int k) { mbar_wait(&full[(k - k0 + 1) & (NS - 1)], ((k - k0 + 1) >> 2) & 1); };
        stage(k0 - 1); stage(k0); stage(k0 + 1); stage(k0 + 2);
        if (tid < KCHUNK + 2)
        {
            const int k = min(k0 - 1 + tid, gd.kcells - 1);
            double* m = met + tid * MR;
            m[0] = gd.rhoref[k]; m[1] = gd.rhoref_i[k]; m[2] = gd.dzi[k]; m[3] = 0.5 * gd.dzi[k];
            m[4] = gd.rhorefh[k]; m[5] = gd.dzhi[k];
        }

        const int c = (threadIdx.y + 1) * W + threadIdx.x + 1;
        const int cw = c - 1, ce = c + 1, cs = c - W, cn = c + W;

        int ijk = ij + k0 * kk;
        const bool ldt = live && !tail.fresh;
        double at_c = 0.;
        if (ldt) at_c = at[ijk];
        double w_c = 0., w_n = 0.;                       // w column (centre only), one level ahead from global memory
        if (ADVEC) { w_c = w[ijk]; w_n = w[ijk + kk]; }

        if (!TMA) cp_async_wait<2>();                    // planes k0-1 and k0 have landed
        __syncthreads();
        if (TMA) { landed(k0 - 1); landed(k0); }
        double a_c = sm[1][0][c], e_c = sm[1][1][c];
        // vertical fluxes through the lower face of the first level; in the loop the upper-face flux of level k is carried
        // as the lower-face flux of level k+1 (same operands, same operations)
        double fd_b, fa_b = 0.;
        {
            const double a_b = sm[0][0][c], e_b = sm[0][1][c];
            fd_b = met[MR + 4] * (div_by(e_b + e_c, tPr2, tPr_hi) + visc) * (a_c - a_b) * met[MR + 5];
            if (ADVEC) fa_b = met[MR + 4] * w_c * (a_b + a_c);     // 2 x the reference's flux
        }

        MHH_PRAGMA_UNROLL(MHH_SCALAR_UNROLL_N)
        for (int k = k0; k < k1; ++k, ijk += kk)
        {
            if (TMA) landed(k + 1);
            else cp_async_wait<1>();                     // plane k+1 has landed (k+2 may still be in flight)
            __syncthreads();                             // ... for every thread; everyone is done with plane k-1
            stage(k + 3);                                // into the slot of plane k-1
            double at_n = 0.;
            if (ldt && k + 1 < k1) at_n = at[ijk + kk];
            const double w_t = w_n;
            if (ADVEC && k + 1 < k1) w_n = w[ijk + 2 * kk];

            const double* C = sm[(k - k0 + 1) & (NS - 1)][0];
            const double* T = sm[(k - k0 + 2) & (NS - 1)][0];
            const int km = k - k0 + 1;                   // row of level k in met[]
            const double2 m_c = met2(km, 0), m_z = met2(km, 1), m_ht = met2(km + 1, 2);
            const double rho = m_c.x, rho_i = m_c.y, dzi = m_z.x, rhoh_t = m_ht.x, dzhi_t = m_ht.y, rhoh_b = met[km * MR + 4];
            const bool bot = SURF && (k == gd.kstart);
            const bool top = SURF && (k == gd.kend - 1);

            const double a_t = T[c], a_e = C[ce], a_w = C[cw], a_n = C[cn], a_s = C[cs];
            const double e_t = T[PS + c], e_e = C[PS + ce], e_w = C[PS + cw], e_n = C[PS + cn], e_s = C[PS + cs];

            if (ADVEC)
            {
                const double u_c = C[2 * PS + c], u_e = C[2 * PS + ce], v_c = C[3 * PS + c], v_n = C[3 * PS + cn];
                const double fa_t = rhoh_t * w_t * (a_c + a_t);
                // advec_s, src/advec_2.cxx:191-202
                at_c = at_c +
                    ( - ( u_e * (a_c + a_e)
                        - u_c * (a_w + a_c) ) * hdxi
                      - ( v_n * (a_c + a_n)
                        - v_c * (a_s + a_c) ) * hdyi
                      - div_by( fa_t - fa_b, rho, rho_i ) * m_z.y );
                fa_b = fa_t;
            }

            const double evisce = div_by(e_c + e_e, tPr2, tPr_hi) + visc;
            const double eviscw = div_by(e_w + e_c, tPr2, tPr_hi) + visc;
            const double eviscn = div_by(e_c + e_n, tPr2, tPr_hi) + visc;
            const double eviscs = div_by(e_s + e_c, tPr2, tPr_hi) + visc;
            const double evisct = div_by(e_c + e_t, tPr2, tPr_hi) + visc;

            const double hor =
                + ( evisce * (a_e - a_c) - eviscw * (a_c - a_w) ) * dxidxi
                + ( eviscn * (a_n - a_c) - eviscs * (a_c - a_s) ) * dyidyi;
            const double fd_t = rhoh_t * evisct * (a_t - a_c) * dzhi_t;
            double ver;
            if (bot)      ver = div_by( fd_t + rhoh_b * rare_load(fluxbot, ij), rho, rho_i ) * dzi;
            else if (top) ver = div_by( -rhoh_t * rare_load(fluxtop, ij) - fd_b, rho, rho_i ) * dzi;
            else          ver = div_by( fd_t - fd_b, rho, rho_i ) * dzi;
            double tn = at_c + (hor + ver);
            if (tail.sig != nullptr && k >= tail.bufk) tn -= tail.sig[k] * (a_c - tail.abuf[k]);
            if (live)
            {
                if (tail.rk_mode == 0) at[ijk] = tn;
                else
                {
                    tail.a_new[ijk] = a_c + tail.cB_dt * tn;
                    if (tail.rk_mode == 1) at[ijk] = tn * tail.cA_next;
                }
            }
            fd_b = fd_t;
            a_c = a_t; e_c = e_t;
            at_c = at_n;
        }
        if (!TMA) cp_async_wait<0>();
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
    if (!c->mlen0_d) MHH_CUDA_CHECK(cudaMalloc(&c->mlen0_d, 2 * gd.kcells * sizeof(double)));
    std::vector<double> h(gd.kcells);
    for (int k = 0; k < gd.kcells; ++k)
        h[k] = cs * std::pow(gd.dx * gd.dy * c->dz[k], 1. / 3.);
    MHH_CUDA_CHECK(cudaMemcpyAsync(c->mlen0_d, h.data(), gd.kcells * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    MHH_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    c->mlen0_cs = cs;
    return MHH_OK;
}

int smag2_exec_viscosity(Ctx* c, int surface_model, const double* u, const double* v, const double* w,
                         const double* N2, const double* dudz, const double* dvdz, const double* dbdz,
                         const double* z0m, double cs, double tPr, double visc, double* evisc, int strain_only,
                         unsigned long long* lim)
{
    const GridD& gd = c->gd;
    if (lim != nullptr && strain_only != 2 && !(strain_only == 0 && N2 != nullptr)) return MHH_ERR_UNSUPPORTED;
    if (lim != nullptr) MHH_CUDA_CHECK(cudaMemsetAsync(lim, 0, 2 * sizeof(unsigned long long), c->stream));
    const double tPrfac_i = 1. / std::fmin(1., tPr);
    if (c->mlen0_cs != cs || c->mlen0_d == nullptr)
        if (int e = smag2_upload_mlen0(c, cs)) return e;
    if (strain_only == 2)
    {
        // fused dry-thermo variant: N2 carries th; the g/thref table sits behind mlen0
        if (!c->thref_d) { c->last_error = "thermo reference profiles not set"; return MHH_ERR_STATE; }
        MHH_CUDA_CHECK(cudaMemcpyAsync(c->mlen0_d + gd.kcells, c->thref_d + 2 * gd.kcells, gd.kcells * sizeof(double),
                                       cudaMemcpyDeviceToDevice, c->stream));
    }
    constexpr int TX = MHH_TILE_X, TY = 256 / MHH_TILE_X;
    const dim3 tb(TX, TY, 1), tg((gd.imax + TX - 1) / TX, (gd.jmax + TY - 1) / TY, (gd.kmax + KCHUNK - 1) / KCHUNK);
    TmaMaps4 maps;
    bool tma = tma_planes_supported(c, TX + 2);
    if (tma)
    {
        const double* flds[3] = {u, v, w};
        for (int q = 0; q < 3 && tma; ++q)
        {
            const int e = tma_plane_map(c, flds[q], TX + 2, TY + 2, &maps.m[q]);
            if (e == MHH_ERR_CUDA) return e;
            tma = (e == MHH_OK);
        }
    }
    #define EV_ARGS evisc, u, v, w, N2, dudz, dvdz, dbdz, z0m, c->mlen0_d, tPr, visc, gd, lim, tPrfac_i, maps
    #define EV_LAUNCH(S, M, L) do { \
        if (tma) MHH_LAUNCH(c, "evisc_kernel", (evisc_kernel<S, M, TX, TY, L, true><<<tg, tb, 0, c->stream>>>(EV_ARGS))); \
        else     MHH_LAUNCH(c, "evisc_kernel", (evisc_kernel<S, M, TX, TY, L, false><<<tg, tb, 0, c->stream>>>(EV_ARGS))); } while (0)
    if (strain_only == 1)
    {
        if (surface_model) EV_LAUNCH(true, 2, false);
        else               EV_LAUNCH(false, 2, false);
        MHH_LAUNCH_CHECK();
        return MHH_OK;
    }
    if (strain_only == 2)
    {
        if (lim != nullptr)
        {
            if (surface_model) EV_LAUNCH(true, 3, true);
            else               EV_LAUNCH(false, 3, true);
        }
        else if (surface_model) EV_LAUNCH(true, 3, false);
        else                    EV_LAUNCH(false, 3, false);
    }
    else if (N2 != nullptr)
    {
        if (lim != nullptr)
        {
            if (surface_model) EV_LAUNCH(true, 0, true);
            else               EV_LAUNCH(false, 0, true);
        }
        else if (surface_model) EV_LAUNCH(true, 0, false);
        else                    EV_LAUNCH(false, 0, false);
    }
    else
    {
        if (surface_model) EV_LAUNCH(true, 1, false);
        else               EV_LAUNCH(false, 1, false);
    }
    #undef EV_LAUNCH
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

// Momentum tendencies.  advec == 0: Diff_smag2::exec for u,v,w.  advec == 1: preceded in the same pass by Advec_2::exec
// for u,v,w and, when th != nullptr, by the dry buoyancy tendency on w (bit-identical to the separate calls, see kernel).
int smag2_tend_uvw(Ctx* c, int surface_model, int advec, const double* u, const double* v, const double* w, const double* evisc,
                   double* ut, double* vt, double* wt,
                   const double* ufluxbot, const double* ufluxtop, const double* vfluxbot, const double* vfluxtop,
                   double visc, const double* th)
{
    const GridD& gd = c->gd;
    if (th != nullptr && (!advec || !c->thref_d)) { c->last_error = "fused buoyancy needs advec = 1 and the thermo reference profiles"; return MHH_ERR_STATE; }
    constexpr int TX = MHH_TILE_X, TY = MHH_UVW_TILE_Y;
    const dim3 tb(TX, TY, 1), tg((gd.imax + TX - 1) / TX, (gd.jmax + TY - 1) / TY, (gd.kmax + KCHUNK - 1) / KCHUNK);
    const double* threfh = c->thref_d ? c->thref_d + gd.kcells : nullptr;
    const double* g_threfh = c->thref_d ? c->thref_d + 3 * gd.kcells : nullptr;
    // lazy-tendency mode (mhh_set_lazy_tendencies): the arrays the last fused pressure tail left undefined start from 0
    const int fresh = (c->fresh_uvw[0] == ut && c->fresh_uvw[1] == vt && c->fresh_uvw[2] == wt) ? 1 : 0;
    c->fresh_uvw[0] = c->fresh_uvw[1] = c->fresh_uvw[2] = nullptr;
    // the Buffer / Coriolis tail armed by mhh_advec2_smag2_set_force_tail rides in the fused advection + diffusion pass only
    ForceTail ft{nullptr, nullptr, nullptr, nullptr, 0, 0, 0, 0., 0., 0., nullptr, nullptr};
    const bool force = advec && (c->uvw_tail.buffer || c->uvw_tail.coriolis);
    if (force)
    {
        const auto& t = c->uvw_tail;
        if (t.buffer)
        {
            if (!c->buffer_sigma_d) { c->last_error = "buffer not planned (mhh_buffer_plan)"; return MHH_ERR_STATE; }
            ft.sig = c->buffer_sigma_d; ft.ubuf = t.ubuf; ft.vbuf = t.vbuf; ft.wbuf = t.wbuf;
            ft.bufk = c->buffer_kstart; ft.bufkh = c->buffer_kstarth;
        }
        if (t.coriolis) { ft.coriolis = 1; ft.fc = t.fc; ft.ugrid = t.ugrid; ft.vgrid = t.vgrid; ft.ug = t.ug; ft.vg = t.vg; }
    }
    // plane staging: TMA bulk tensor copies where the layout allows (icells even), else per-thread cp.async
    TmaMaps4 maps;
    bool tma = MHH_UVW_TMA && tma_planes_supported(c, TX + 2);
    if (tma)
    {
        const double* flds[4] = {evisc, u, v, w};
        for (int q = 0; q < 4 && tma; ++q)
        {
            const int e = tma_plane_map(c, flds[q], TX + 2, TY + 2, &maps.m[q]);
            if (e == MHH_ERR_CUDA) return e;
            tma = (e == MHH_OK);
        }
    }
    #define UVW_ARGS ut, vt, wt, u, v, w, evisc, ufluxbot, ufluxtop, vfluxbot, vfluxtop, visc, th, threfh, g_threfh, gd, fresh, ft, maps
    constexpr size_t smem_ldgsts = (size_t)UVW_SLOTS * 4 * (TX + 2) * (TY + 2) * sizeof(double);
    constexpr size_t smem_tma = (size_t)UVW_SLOTS * 4 * tma_plane_pitch(TX + 2, TY + 2) * sizeof(double);
    #define UVW_LAUNCH1(NAME, S, A, B, T, F) do { \
        static bool attr = false; \
        const size_t smem = T ? smem_tma : smem_ldgsts; \
        if (!attr) { cudaFuncSetAttribute(smag2_diff_uvw_tile_kernel<S, A, B, TX, TY, T, F>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); attr = true; } \
        MHH_LAUNCH(c, NAME, (smag2_diff_uvw_tile_kernel<S, A, B, TX, TY, T, F><<<tg, tb, smem, c->stream>>>(UVW_ARGS))); } while (0)
    #define UVW_LAUNCH(NAME, S, A, B) do { \
        if (force) { if (tma) UVW_LAUNCH1(NAME, S, A, B, true, A); else UVW_LAUNCH1(NAME, S, A, B, false, A); } \
        else       { if (tma) UVW_LAUNCH1(NAME, S, A, B, true, false); else UVW_LAUNCH1(NAME, S, A, B, false, false); } } while (0)
    if (!advec)
    {
        if (surface_model) UVW_LAUNCH("smag2_diff_uvw_kernel", true, false, false);
        else               UVW_LAUNCH("smag2_diff_uvw_kernel", false, false, false);
    }
    else if (th == nullptr)
    {
        if (surface_model) UVW_LAUNCH("advec2_smag2_uvw_kernel", true, true, false);
        else               UVW_LAUNCH("advec2_smag2_uvw_kernel", false, true, false);
    }
    else
    {
        if (surface_model) UVW_LAUNCH("advec2_smag2_uvw_kernel", true, true, true);
        else               UVW_LAUNCH("advec2_smag2_uvw_kernel", false, true, true);
    }
    #undef UVW_LAUNCH
    #undef UVW_LAUNCH1
    #undef UVW_ARGS
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

// Scalar tendency.  advec == 0: Diff_smag2's diff_c.  advec == 1: preceded in the same pass by Advec_2's advec_s.
void rk_coefs(int rkorder, int substep, double dt, double* cB_dt, double* cA_next, int* zero_all);

// sbuf: Buffer profile of this scalar or NULL; a_new: RK target or NULL (then rkorder/substep/dt are unused)
int smag2_tend_scalar_tail(Ctx* c, int surface_model, int advec, double* at, const double* a, const double* evisc,
                           const double* fluxbot, const double* fluxtop, double tPr, double visc,
                           const double* u, const double* v, const double* w,
                           const double* sbuf, double* a_new, int rkorder, int substep, double dt)
{
    const GridD& gd = c->gd;
    ScalarTail tail{nullptr, 0., 0., 0, nullptr, nullptr, 0, 0};
    // k-range parts (mhh_advec2_smag2_exec_rk_parts): chunks [kc0, kc0 + nkc) of the column; bookkeeping that concerns the
    // whole tendency array (lazy-tendency flags, the zeroing after the last substep) happens with the part that ends the column
    const int nchunks = (gd.kmax + KCHUNK - 1) / KCHUNK;
    const int kc0 = c->sc_kc0 < 0 ? 0 : (c->sc_kc0 > nchunks ? nchunks : c->sc_kc0);
    const int nkc = (c->sc_nkc <= 0 || kc0 + c->sc_nkc > nchunks) ? nchunks - kc0 : c->sc_nkc;
    const bool last_part = kc0 + nkc >= nchunks;
    if (nkc <= 0) return MHH_OK;
    for (auto& fp : c->fresh_s)
        if (fp == at && at != nullptr) { tail.fresh = 1; if (last_part) fp = nullptr; }
    if (sbuf)
    {
        if (!c->buffer_sigma_d) { c->last_error = "buffer not planned (mhh_buffer_plan)"; return MHH_ERR_STATE; }
        tail.sig = c->buffer_sigma_d; tail.abuf = sbuf; tail.bufk = c->buffer_kstart;
    }
    if (a_new)
    {
        int zero_all = 0;
        rk_coefs(rkorder, substep, dt, &tail.cB_dt, &tail.cA_next, &zero_all);
        tail.a_new = a_new; tail.rk_mode = zero_all ? 2 : 1;
    }
    constexpr int TX = MHH_TILE_X, TY = 256 / MHH_TILE_X;
    const dim3 tb(TX, TY, 1), tg((gd.imax + TX - 1) / TX, (gd.jmax + TY - 1) / TY, nkc);
    TmaMaps4 maps;
    bool tma = tma_planes_supported(c, TX + 2);
    if (tma)
    {
        const double* flds[4] = {a, evisc, u, v};
        for (int q = 0; q < (advec ? 4 : 2) && tma; ++q)
        {
            const int e = tma_plane_map(c, flds[q], TX + 2, TY + 2, &maps.m[q]);
            if (e == MHH_ERR_CUDA) return e;
            tma = (e == MHH_OK);
        }
    }
    #define SC_LAUNCH(NAME, S, A) do { \
        if (tma) MHH_LAUNCH(c, NAME, (smag2_scalar_tile_kernel<S, A, TX, TY, true><<<tg, tb, 0, c->stream>>>(at, a, evisc, fluxbot, fluxtop, tPr, visc, u, v, w, tail, gd, maps, kc0))); \
        else     MHH_LAUNCH(c, NAME, (smag2_scalar_tile_kernel<S, A, TX, TY, false><<<tg, tb, 0, c->stream>>>(at, a, evisc, fluxbot, fluxtop, tPr, visc, u, v, w, tail, gd, maps, kc0))); } while (0)
    if (!advec)
    {
        if (surface_model) SC_LAUNCH("smag2_diff_c_kernel", true, false);
        else               SC_LAUNCH("smag2_diff_c_kernel", false, false);
    }
    else
    {
        if (surface_model) SC_LAUNCH("advec2_smag2_c_kernel", true, true);
        else               SC_LAUNCH("advec2_smag2_c_kernel", false, true);
    }
    #undef SC_LAUNCH
    MHH_LAUNCH_CHECK();
    if (tail.rk_mode == 2 && last_part)      // last substep: the whole tendency array (ghost cells too) is zero afterwards, src/timeloop.cu:34-70
    {
        bool lazy = false;
        if (c->lazy_tend)       // ... or declared undefined until the next fused tendency pass, which then starts from 0
            for (auto& fp : c->fresh_s)
                if (fp == nullptr) { fp = at; lazy = true; break; }
        if (!lazy) MHH_CUDA_CHECK(cudaMemsetAsync(at, 0, sizeof(double) * gd.ncells, c->stream));
    }
    return MHH_OK;
}

int smag2_tend_scalar(Ctx* c, int surface_model, int advec, double* at, const double* a, const double* evisc,
                      const double* fluxbot, const double* fluxtop, double tPr, double visc,
                      const double* u, const double* v, const double* w)
{
    return smag2_tend_scalar_tail(c, surface_model, advec, at, a, evisc, fluxbot, fluxtop, tPr, visc, u, v, w, nullptr, nullptr, 0, 0, 0.);
}

// mhh_advec2_smag2_exec_rk: as advec2_smag2_exec; per scalar optionally the Buffer term and the RK update (see ScalarTail)
int advec2_smag2_exec_rk(Ctx* c, int surface_model, const double* u, const double* v, const double* w, const double* evisc,
                         double* ut, double* vt, double* wt,
                         const double* ufluxbot, const double* ufluxtop, const double* vfluxbot, const double* vfluxtop,
                         double visc, const double* th_buoy, int nscalars, const double* const* s, double* const* st,
                         const double* const* sfluxbot, const double* const* sfluxtop, const double* svisc, double tPr,
                         const double* const* sbuf, double* const* s_new, int rkorder, int substep, double dt)
{
    if (int e = smag2_tend_uvw(c, surface_model, 1, u, v, w, evisc, ut, vt, wt, ufluxbot, ufluxtop, vfluxbot, vfluxtop, visc, th_buoy)) return e;
    for (int n = 0; n < nscalars; ++n)
        if (int e = smag2_tend_scalar_tail(c, surface_model, 1, st[n], s[n], evisc, surface_model ? sfluxbot[n] : nullptr,
                                           surface_model ? sfluxtop[n] : nullptr, tPr, svisc[n], u, v, w,
                                           sbuf ? sbuf[n] : nullptr, s_new ? s_new[n] : nullptr, rkorder, substep, dt)) return e;
    return MHH_OK;
}

// mhh_advec2_smag2_exec_rk_parts: the momentum pass and / or the scalar passes restricted to the 64-level chunks [kc0, kc0 + nkc)
int advec2_smag2_exec_rk_parts(Ctx* c, int surface_model, const double* u, const double* v, const double* w, const double* evisc,
                               double* ut, double* vt, double* wt,
                               const double* ufluxbot, const double* ufluxtop, const double* vfluxbot, const double* vfluxtop,
                               double visc, const double* th_buoy, int nscalars, const double* const* s, double* const* st,
                               const double* const* sfluxbot, const double* const* sfluxtop, const double* svisc, double tPr,
                               const double* const* sbuf, double* const* s_new, int rkorder, int substep, double dt,
                               int do_uvw, int do_scalars, int kc0, int nkc)
{
    if (do_uvw)
        if (int e = smag2_tend_uvw(c, surface_model, 1, u, v, w, evisc, ut, vt, wt, ufluxbot, ufluxtop, vfluxbot, vfluxtop, visc, th_buoy)) return e;
    if (!do_scalars) return MHH_OK;
    c->sc_kc0 = kc0; c->sc_nkc = nkc;
    int err = MHH_OK;
    for (int n = 0; n < nscalars && !err; ++n)
        err = smag2_tend_scalar_tail(c, surface_model, 1, st[n], s[n], evisc, surface_model ? sfluxbot[n] : nullptr,
                                     surface_model ? sfluxtop[n] : nullptr, tPr, svisc[n], u, v, w,
                                     sbuf ? sbuf[n] : nullptr, s_new ? s_new[n] : nullptr, rkorder, substep, dt);
    c->sc_kc0 = 0; c->sc_nkc = 0;
    return err;
}

// Thermo_dry::exec (th_buoy != nullptr) + Advec_2::exec + Diff_smag2::exec: one pass per prognostic field
int advec2_smag2_exec(Ctx* c, int surface_model, const double* u, const double* v, const double* w, const double* evisc,
                      double* ut, double* vt, double* wt,
                      const double* ufluxbot, const double* ufluxtop, const double* vfluxbot, const double* vfluxtop,
                      double visc, const double* th_buoy, int nscalars, const double* const* s, double* const* st,
                      const double* const* sfluxbot, const double* const* sfluxtop, const double* svisc, double tPr)
{
    if (int e = smag2_tend_uvw(c, surface_model, 1, u, v, w, evisc, ut, vt, wt, ufluxbot, ufluxtop, vfluxbot, vfluxtop, visc, th_buoy)) return e;
    for (int n = 0; n < nscalars; ++n)
        if (int e = smag2_tend_scalar(c, surface_model, 1, st[n], s[n], evisc, surface_model ? sfluxbot[n] : nullptr,
                                      surface_model ? sfluxtop[n] : nullptr, tPr, svisc[n], u, v, w)) return e;
    return MHH_OK;
}

int smag2_exec(Ctx* c, int surface_model, const double* u, const double* v, const double* w, const double* evisc,
               double* ut, double* vt, double* wt,
               const double* ufluxbot, const double* ufluxtop, const double* vfluxbot, const double* vfluxtop,
               double visc, int nscalars, const double* const* s, double* const* st,
               const double* const* sfluxbot, const double* const* sfluxtop, const double* svisc, double tPr)
{
    if (int e = smag2_tend_uvw(c, surface_model, 0, u, v, w, evisc, ut, vt, wt, ufluxbot, ufluxtop, vfluxbot, vfluxtop, visc, nullptr)) return e;
    for (int n = 0; n < nscalars; ++n)
        if (int e = smag2_tend_scalar(c, surface_model, 0, st[n], s[n], evisc, surface_model ? sfluxbot[n] : nullptr,
                                      surface_model ? sfluxtop[n] : nullptr, tPr, svisc[n], nullptr, nullptr, nullptr)) return e;
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
