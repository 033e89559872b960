// 2nd-order advection for sm_100a.  Arithmetic follows the reference expression trees
// (src/advec_2.cxx:81-203) term by term so FP64 results match the CPU build without FMA
// contraction; the three momentum components are fused in one pass over u,v,w.
//
// Layout: thread (i,j) of a 2-D tile marches a chunk of k with the k-1/k/k+1 values of
// u,v,w held in registers (each array element is loaded from global once per thread for
// the centre column; the i+-1 / j+-1 neighbours come from L1/L2).
#include "ctx.h"

namespace mhh
{
using fd::interp2;

namespace
{
    constexpr int KCHUNK = MHH_KCHUNK;

    // Register queues while marching k: the centre columns of u,v,w at k-1,k,k+1 and the columns the cross terms
    // need at the neighbouring level (u_e and v_n at k-1; w_w and w_s at k+1).  20 global loads per point instead
    // of 30 (ncu r01: the kernel was held by L1 wavefront throughput at 80 %, not by HBM).
    __global__ void __launch_bounds__(256, 3)
    advec2_uvw_kernel(double* __restrict__ ut, double* __restrict__ vt, double* __restrict__ wt,
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
        double u_b = u[ijk - kk], u_c = u[ijk], u_eb = u[ijk + ii - kk];
        double v_b = v[ijk - kk], v_c = v[ijk], v_nb = v[ijk + jj - kk];
        double w_b = w[ijk - kk], w_c = w[ijk], w_w = w[ijk - ii], w_s = w[ijk - jj];

        for (int k = k0; k < k1; ++k, ijk += kk)
        {
            if (k + MHH_PF < gd.kcells - 1)
            {
                const int pf = ijk + (MHH_PF + 1) * kk;
                prefetch_l2(u + pf); prefetch_l2(v + pf); prefetch_l2(w + pf);
                prefetch_l2(ut + pf - kk); prefetch_l2(vt + pf - kk); prefetch_l2(wt + pf - kk);
            }
            const double rhoh_t = gd.rhorefh[k + 1], rhoh_b = gd.rhorefh[k], rho = gd.rhoref[k], dzi = gd.dzi[k];
            const double rho_i = gd.rhoref_i[k];

            const double u_t = u[ijk + kk], u_e = u[ijk + ii], u_w = u[ijk - ii], u_n = u[ijk + jj], u_s = u[ijk - jj];
            const double v_t = v[ijk + kk], v_n = v[ijk + jj], v_e = v[ijk + ii], v_w = v[ijk - ii], v_s = v[ijk - jj];
            const double w_t = w[ijk + kk], w_wt = w[ijk - ii + kk], w_st = w[ijk - jj + kk];
            const double w_e = w[ijk + ii], w_n = w[ijk + jj];

            // advec_u, src/advec_2.cxx:98-109
            ut[ijk] +=
                - ( interp2(u_c, u_e) * interp2(u_c, u_e)
                  - interp2(u_w, u_c) * interp2(u_w, u_c) ) * dxi
                - ( interp2(v[ijk - ii + jj], v_n) * interp2(u_c, u_n)
                  - interp2(v_w, v_c) * interp2(u_s, u_c) ) * dyi
                - div_by( rhoh_t * interp2(w_wt, w_t) * interp2(u_c, u_t)
                        - rhoh_b * interp2(w_w, w_c) * interp2(u_b, u_c), rho, rho_i ) * dzi;

            // advec_v, src/advec_2.cxx:129-140
            vt[ijk] +=
                - ( interp2(u[ijk + ii - jj], u_e) * interp2(v_c, v_e)
                  - interp2(u_s, u_c) * interp2(v_w, v_c) ) * dxi
                - ( interp2(v_c, v_n) * interp2(v_c, v_n)
                  - interp2(v_s, v_c) * interp2(v_s, v_c) ) * dyi
                - div_by( rhoh_t * interp2(w_st, w_t) * interp2(v_c, v_t)
                        - rhoh_b * interp2(w_s, w_c) * interp2(v_b, v_c), rho, rho_i ) * dzi;

            // advec_w, src/advec_2.cxx:160-171 (k from kstart+1)
            if (k > gd.kstart)
            {
                wt[ijk] +=
                    - ( interp2(u_eb, u_e) * interp2(w_c, w_e)
                      - interp2(u_b, u_c) * interp2(w_w, w_c) ) * dxi
                    - ( interp2(v_nb, v_n) * interp2(w_c, w_n)
                      - interp2(v_b, v_c) * interp2(w_s, w_c) ) * dyi
                    - div_by( rho * interp2(w_c, w_t) * interp2(w_c, w_t)
                            - gd.rhoref[k - 1] * interp2(w_b, w_c) * interp2(w_b, w_c), rhoh_b, gd.rhorefh_i[k] ) * gd.dzhi[k];
            }
            u_b = u_c; u_c = u_t; u_eb = u_e;
            v_b = v_c; v_c = v_t; v_nb = v_n;
            w_b = w_c; w_c = w_t; w_w = w_wt; w_s = w_st;
        }
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
    MHH_LAUNCH(c, "advec2_uvw_kernel", advec2_uvw_kernel<<<l.grid, l.block, 0, c->stream>>>(ut, vt, wt, u, v, w, gd));
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
