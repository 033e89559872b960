// 2nd-order advection for sm_100a.  Arithmetic follows the reference expression trees
// (src/advec_2.cxx:81-203) term by term so FP64 results match the CPU build without FMA
// contraction; the three momentum components are fused in one pass over u,v,w.
//
// Layout: thread (i,j) of a 2-D tile marches a chunk of k with the k-1/k/k+1 values of
// u,v,w held in registers (each array element is loaded from global once per thread for
// the centre column; the i+-1 / j+-1 neighbours come from L1/L2).
#include "sys.h"
#include "ctx.h"

namespace mhh
{
using fd::interp2;

namespace
{
    constexpr int KCHUNK = MHH_KCHUNK;
    constexpr int KT = MHH_TILE_KCHUNK;          // advec2_uvw_kernel (shared-memory tiles)

    // A CTA owns a TX x TY tile of columns and marches k through its chunk.  The (TX+2) x (TY+2) tile planes of u,v,w sit in
    // a ring of four shared-memory slots filled by cp.async two levels ahead (TileMap, mhh_common.cuh); the centre columns
    // at k-1,k and the cross-term columns (u_e, v_n at k-1; w_w, w_s at k) stay in registers.  Per point and level: 17
    // shared-memory loads plus the three tendency read-modify-writes, whose loads are issued one level early.
    template<int TX, int TY>
    __global__ void __launch_bounds__(TX * TY, 3)
    advec2_uvw_kernel(double* __restrict__ ut, double* __restrict__ vt, double* __restrict__ wt,
                      const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ w,
                      const GridD gd)
    {
        using Map = TileMap<TX, TY, 1>;
        constexpr int W = Map::W, PS = Map::N, NS = 4;
        __shared__ double sm[NS][3][PS];                 // [slot][u,v,w][tile plane]
        __shared__ double met[6][KT + 2];            // rhoref, rhorefh, 1/rhoref, 1/rhorefh, dzi, dzhi at k0-1 .. k1

        const int tid = threadIdx.y * TX + threadIdx.x;
        const int i0 = blockIdx.x * TX + gd.istart, j0 = blockIdx.y * TY + gd.jstart;
        const int i = i0 + threadIdx.x, j = j0 + threadIdx.y;
        const bool live = (i < gd.iend) && (j < gd.jend);
        const int jj = gd.icells, kk = gd.ijcells;
        const int ij = min(i, gd.icells - 1) + min(j, gd.jcells - 1) * jj;
        const int k0 = gd.kstart + blockIdx.z * KT;
        const int k1 = min(k0 + KT, gd.kend);
        const double dxi = 1. / gd.dx, dyi = 1. / gd.dy;

        Map map;
        map.init(i0, j0, gd.icells, gd.jcells, tid);
        auto stage = [&](const int k)                    // plane k -> slot (k - k0 + 1) & 3; always one commit group
        {
            if (k <= k1)
            {
                const int sl = (k - k0 + 1) & (NS - 1);
                const long off = (long)k * kk;
                map.issue(sm[sl][0], u + off, tid); map.issue(sm[sl][1], v + off, tid); map.issue(sm[sl][2], w + off, tid);
            }
            cp_async_commit();
        };
        stage(k0 - 1); stage(k0); stage(k0 + 1); stage(k0 + 2);
        if (tid < KT + 2)
        {
            const int k = min(k0 - 1 + tid, gd.kcells - 1);
            met[0][tid] = gd.rhoref[k]; met[1][tid] = gd.rhorefh[k]; met[2][tid] = gd.rhoref_i[k];
            met[3][tid] = gd.rhorefh_i[k]; met[4][tid] = gd.dzi[k]; met[5][tid] = gd.dzhi[k];
        }

        const int c = (threadIdx.y + 1) * W + threadIdx.x + 1;
        const int cw = c - 1, ce = c + 1, cs = c - W, cn = c + W, cwn = c + W - 1, ces = c - W + 1;

        int ijk = ij + k0 * kk;
        double ut_c = 0., vt_c = 0., wt_c = 0.;
        if (live) { ut_c = ut[ijk]; vt_c = vt[ijk]; wt_c = wt[ijk]; }

        cp_async_wait<2>();                              // planes k0-1 and k0 have landed
        __syncthreads();
        double u_b, u_c, u_eb, v_b, v_c, v_nb, w_b, w_c, w_w, w_s;
        {
            const double* B = sm[0][0]; const double* C = sm[1][0];
            u_b = B[c]; u_eb = B[ce]; u_c = C[c];
            v_b = B[PS + c]; v_nb = B[PS + cn]; v_c = C[PS + c];
            w_b = B[2 * PS + c]; w_c = C[2 * PS + c]; w_w = C[2 * PS + cw]; w_s = C[2 * PS + cs];
        }

        for (int k = k0; k < k1; ++k, ijk += kk)
        {
            cp_async_wait<1>();                          // plane k+1 has landed (k+2 may still be in flight)
            __syncthreads();                             // ... for every thread; everyone is done with plane k-1
            stage(k + 3);                                // into the slot of plane k-1
            double ut_n = 0., vt_n = 0., wt_n = 0.;
            if (live && k + 1 < k1) { ut_n = ut[ijk + kk]; vt_n = vt[ijk + kk]; wt_n = wt[ijk + kk]; }

            const double* C = sm[(k - k0 + 1) & (NS - 1)][0];
            const double* T = sm[(k - k0 + 2) & (NS - 1)][0];
            const int km = k - k0 + 1;                   // row of level k in met[]
            const double rhoh_t = met[1][km + 1], rhoh_b = met[1][km], rho = met[0][km], dzi = met[4][km];
            const double rho_i = met[2][km];

            const double u_t = T[c], u_e = C[ce], u_w = C[cw], u_n = C[cn], u_s = C[cs], u_es = C[ces];
            const double v_t = T[PS + c], v_n = C[PS + cn], v_e = C[PS + ce], v_w = C[PS + cw], v_s = C[PS + cs], v_wn = C[PS + cwn];
            const double w_t = T[2 * PS + c], w_wt = T[2 * PS + cw], w_st = T[2 * PS + cs];
            const double w_e = C[2 * PS + ce], w_n = C[2 * PS + cn];

            // advec_u, src/advec_2.cxx:98-109
            const double ud =
                - ( interp2(u_c, u_e) * interp2(u_c, u_e)
                  - interp2(u_w, u_c) * interp2(u_w, u_c) ) * dxi
                - ( interp2(v_wn, v_n) * interp2(u_c, u_n)
                  - interp2(v_w, v_c) * interp2(u_s, u_c) ) * dyi
                - div_by( rhoh_t * interp2(w_wt, w_t) * interp2(u_c, u_t)
                        - rhoh_b * interp2(w_w, w_c) * interp2(u_b, u_c), rho, rho_i ) * dzi;
            if (live) ut[ijk] = ut_c + ud;

            // advec_v, src/advec_2.cxx:129-140
            const double vd =
                - ( interp2(u_es, u_e) * interp2(v_c, v_e)
                  - interp2(u_s, u_c) * interp2(v_w, v_c) ) * dxi
                - ( interp2(v_c, v_n) * interp2(v_c, v_n)
                  - interp2(v_s, v_c) * interp2(v_s, v_c) ) * dyi
                - div_by( rhoh_t * interp2(w_st, w_t) * interp2(v_c, v_t)
                        - rhoh_b * interp2(w_s, w_c) * interp2(v_b, v_c), rho, rho_i ) * dzi;
            if (live) vt[ijk] = vt_c + vd;

            // advec_w, src/advec_2.cxx:160-171 (k from kstart+1)
            if (k > gd.kstart)
            {
                const double wd =
                    - ( interp2(u_eb, u_e) * interp2(w_c, w_e)
                      - interp2(u_b, u_c) * interp2(w_w, w_c) ) * dxi
                    - ( interp2(v_nb, v_n) * interp2(w_c, w_n)
                      - interp2(v_b, v_c) * interp2(w_s, w_c) ) * dyi
                    - div_by( rho * interp2(w_c, w_t) * interp2(w_c, w_t)
                            - met[0][km - 1] * interp2(w_b, w_c) * interp2(w_b, w_c), rhoh_b, met[3][km] ) * met[5][km];
                if (live) wt[ijk] = wt_c + wd;
            }
            u_b = u_c; u_c = u_t; u_eb = u_e;
            v_b = v_c; v_c = v_t; v_nb = v_n;
            w_b = w_c; w_c = w_t; w_w = w_wt; w_s = w_st;
            ut_c = ut_n; vt_c = vt_n; wt_c = wt_n;
        }
        cp_async_wait<0>();
    }

    __global__ void __launch_bounds__(256, MHH_MINB)
    advec2_s_kernel(double* __restrict__ st, const double* __restrict__ s,
                    const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ w,
                    const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x + gd.istart;
        const int j = blockIdx.y * blockDim.y + threadIdx.y + gd.jstart;
        if (i >= gd.iend || j >= gd.jend) return;

        const int ii = 1, jj = gd.icells, kk = gd.ijcells;
        const int k0 = gd.kstart + blockIdx.z * KCHUNK;
        const int k1 = min(k0 + KCHUNK, gd.kend);
        const double dxi = 1. / gd.dx, dyi = 1. / gd.dy;

        int ijk = i + j * jj + k0 * kk;
        double s_b = s[ijk - kk], s_c = s[ijk];
        double w_c = w[ijk];
        for (int k = k0; k < k1; ++k, ijk += kk)
        {
            if (k + MHH_PF < gd.kcells - 1)
            {
                const int pf = ijk + (MHH_PF + 1) * kk;
                prefetch_l2(s + pf); prefetch_l2(w + pf); prefetch_l2(u + pf - kk); prefetch_l2(v + pf - kk); prefetch_l2(st + pf - kk);
            }
            const double s_t = s[ijk + kk];
            const double w_t = w[ijk + kk];
            // advec_s, src/advec_2.cxx:191-202
            st[ijk] +=
                - ( u[ijk + ii] * interp2(s_c, s[ijk + ii])
                  - u[ijk] * interp2(s[ijk - ii], s_c) ) * dxi
                - ( v[ijk + jj] * interp2(s_c, s[ijk + jj])
                  - v[ijk] * interp2(s[ijk - jj], s_c) ) * dyi
                - div_by( gd.rhorefh[k + 1] * w_t * interp2(s_c, s_t)
                        - gd.rhorefh[k] * w_c * interp2(s_b, s_c), gd.rhoref[k], gd.rhoref_i[k] ) * gd.dzi[k];
            s_b = s_c; s_c = s_t; w_c = w_t;
        }
    }

    // calc_cfl, src/advec_2.cxx:51-78, fused with the max reduction.
    __global__ void __launch_bounds__(256, MHH_MINB)
    advec2_cfl_kernel(const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ w,
                      const GridD gd, unsigned long long* __restrict__ out)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x + gd.istart;
        const int j = blockIdx.y * blockDim.y + threadIdx.y + gd.jstart;
        const int jj = gd.icells, kk = gd.ijcells;
        const int k0 = gd.kstart + blockIdx.z * KCHUNK;
        const int k1 = min(k0 + KCHUNK, gd.kend);
        const double dxi = 1. / gd.dx, dyi = 1. / gd.dy;

        unsigned long long m = 0ULL;
        if (i < gd.iend && j < gd.jend)
        {
            int ijk = i + j * jj + k0 * kk;
            double w_c = w[ijk];
            for (int k = k0; k < k1; ++k, ijk += kk)
            {
                const double w_t = w[ijk + kk];
                const double c = fabs(interp2(u[ijk], u[ijk + 1])) * dxi + fabs(interp2(v[ijk], v[ijk + jj])) * dyi
                               + fabs(interp2(w_c, w_t)) * gd.dzi[k];
                const unsigned long long b = abs_bits(c);
                m = b > m ? b : m;
                w_c = w_t;
            }
        }
        block_max_bits(m, out);
    }
}  // namespace

int advec2_exec(Ctx* c, const double* u, const double* v, const double* w, double* ut, double* vt, double* wt,
                int nscalars, const double* const* s, double* const* st)
{
    const GridD& gd = c->gd;
    Launch3 l = launch_interior(gd.imax, gd.jmax, (gd.kmax + KCHUNK - 1) / KCHUNK);
    {
        constexpr int TX = 32, TY = 8;
        const dim3 tb(TX, TY, 1), tg((gd.imax + TX - 1) / TX, (gd.jmax + TY - 1) / TY, (gd.kmax + KT - 1) / KT);
        MHH_LAUNCH(c, "advec2_uvw_kernel", (advec2_uvw_kernel<TX, TY><<<tg, tb, 0, c->stream>>>(ut, vt, wt, u, v, w, gd)));
    }
    MHH_LAUNCH_CHECK();
    for (int n = 0; n < nscalars; ++n)
    {
        MHH_LAUNCH(c, "advec2_s_kernel", advec2_s_kernel<<<l.grid, l.block, 0, c->stream>>>(st[n], s[n], u, v, w, gd));
        MHH_LAUNCH_CHECK();
    }
    return MHH_OK;
}

int advec2_cfl_bits(Ctx* c, const double* u, const double* v, const double* w, unsigned long long* slot)
{
    const GridD& gd = c->gd;
    Launch3 l = launch_interior(gd.imax, gd.jmax, (gd.kmax + KCHUNK - 1) / KCHUNK);
    MHH_CUDA_CHECK(cudaMemsetAsync(slot, 0, sizeof(unsigned long long), c->stream));
    MHH_LAUNCH(c, "advec2_cfl_kernel", advec2_cfl_kernel<<<l.grid, l.block, 0, c->stream>>>(u, v, w, gd, slot));
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

}  // namespace mhh
