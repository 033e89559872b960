// 2nd-order Poisson pressure solver for sm_100a (reference: src/pres_2.cxx, src/fft.cxx).
//
//   input   : rho-weighted divergence of (u/dt + ut) into a ghost-free work array      (:156-196)
//   forward : x then y real->half-complex FFTs (fft.cu)                                 (fft.cxx:335-389)
//   solve   : per-(kx,ky) tridiagonal system over k; the diagonal b and the dz^2 scaling of the
//             RHS (solve loop :292-304, BCs :306-324) are computed on the fly inside the Thomas
//             sweep (tdma :202-263), so only p and the c' work array touch HBM
//   backward: y then x half-complex->real, normalised; the x pass scatters straight into the
//             ghost-celled p (replaces the copy loop :340-348)
//   output  : bottom Neumann ghost row, cyclic halos, ut/vt/wt -= grad p                 (:352-387)
#include "sys.h"
#include <cmath>
#ifndef MHH_PRES_OUT_UNROLL_N
#define MHH_PRES_OUT_UNROLL_N 1
#endif
#ifndef MHH_PRES_IN_UNROLL_N
#define MHH_PRES_IN_UNROLL_N 1
#endif

#include "ctx.h"
#include "fft.h"

namespace mhh
{
int cyclic_exec(Ctx* c, double* fld, int edge);

namespace
{
    constexpr int KCHUNK = MHH_KCHUNK;

    // Pres_2::input, src/pres_2.cxx:183-195
    __global__ void __launch_bounds__(256, MHH_MINB)
    pres2_in_kernel(double* __restrict__ p, const double* __restrict__ u, const double* __restrict__ v,
                    const double* __restrict__ w, const double* __restrict__ ut, const double* __restrict__ vt,
                    const double* __restrict__ wt, const double dti, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        const int j = blockIdx.y * blockDim.y + threadIdx.y;
        if (i >= gd.imax || j >= gd.jmax) return;
        const int ii = 1, jj = gd.icells, kk = gd.ijcells;
        const int jjp = gd.imax, kkp = gd.imax * gd.jmax;
        const int k0 = blockIdx.z * KCHUNK;
        const int k1 = min(k0 + KCHUNK, gd.kmax);
        const double dxi = 1. / gd.dx, dyi = 1. / gd.dy;

        int ijk = i + gd.igc + (j + gd.jgc) * jj + (k0 + gd.kgc) * kk;
        double wflux_b = wt[ijk] + w[ijk] * dti;
        MHH_PRAGMA_UNROLL(MHH_PRES_IN_UNROLL_N)
        for (int k = k0; k < k1; ++k, ijk += kk)
        {
            const int ijkp = i + j * jjp + k * kkp;
            const double rho = gd.rhoref[k + gd.kgc];
            const double wflux_t = wt[ijk + kk] + w[ijk + kk] * dti;
            p[ijkp] = rho * ((ut[ijk + ii] + u[ijk + ii] * dti) - (ut[ijk] + u[ijk] * dti)) * dxi
                    + rho * ((vt[ijk + jj] + v[ijk + jj] * dti) - (vt[ijk] + v[ijk] * dti)) * dyi
                    + (gd.rhorefh[k + gd.kgc + 1] * wflux_t - gd.rhorefh[k + gd.kgc] * wflux_b) * gd.dzi[k + gd.kgc];
            wflux_b = wflux_t;
        }
    }

    // solve_in + tdma fused (src/pres_2.cxx:267-349, tdma :202-232): one thread per (kx,ky) column, k sequential, i-fastest
    // coalescing.  The reference keeps c' = c[k-1]/work2d for the back substitution in a full 3-D work array; here only the
    // reciprocal pivot 1/work2d is checkpointed every TDMA_M levels, and the back substitution recomputes the c' of one
    // block of levels from its checkpoint into registers: 4 + 2/TDMA_M field passes over HBM instead of 6.  Each level
    // costs one reciprocal (fast_rcp, ~1 ulp) per sweep instead of the reference's two divisions: the kernel is otherwise
    // FP64-issue bound on B200 (the IEEE division sequence is ~20 instructions on a k-sequential chain).
    constexpr int TDMA_M = 16;

    __global__ void __launch_bounds__(128)
    pres2_tdma_kernel(double* __restrict__ p, double* __restrict__ ckpt,
                      const double* __restrict__ a, const double* __restrict__ c,
                      const double* __restrict__ bmati, const double* __restrict__ bmatj,
                      const int iblock, const int jblock, const int ioff, const int joff, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        const int j = blockIdx.y;
        if (i >= iblock) return;
        const int kmax = gd.kmax, kgc = gd.kgc;
        const int kk = iblock * jblock;
        const int ij = i + j * iblock;
        const int iindex = ioff + i, jindex = joff + j;
        const double bm = bmati[iindex] + bmatj[jindex];
        const bool mode0 = (iindex == 0 && jindex == 0);

        // b at level k with the boundary substitutions of src/pres_2.cxx:306-324
        auto bdiag = [&](int k) -> double
        {
            const double dz = gd.dz[k + kgc];
            double b = dz * dz * gd.rhoref[k + kgc] * bm - (a[k] + c[k]);
            if (k == 0) b += a[0];
            if (k == kmax - 1) { if (mode0) b -= c[kmax - 1]; else b += c[kmax - 1]; }
            return b;
        };
        auto rhs = [&](int k) -> double
        {
            const double dz = gd.dz[k + kgc];
            return dz * dz * p[ij + k * kk];
        };

        // forward elimination; r = 1/work2d of levels 0, M, 2M, ... goes to the checkpoint array
        double r = fast_rcp(bdiag(0));
        double pprev = rhs(0) * r;
        p[ij] = pprev;
        ckpt[ij] = r;
        for (int k = 1; k < kmax; ++k)
        {
            if (k + 4 < kmax) prefetch_l2(p + ij + (k + 4) * kk);
            const double w3 = c[k - 1] * r;
            r = fast_rcp(bdiag(k) - a[k] * w3);
            double pk = rhs(k);
            pk -= a[k] * pprev;
            pk *= r;
            p[ij + k * kk] = pk;
            pprev = pk;
            if ((k & (TDMA_M - 1)) == 0) ckpt[ij + (k / TDMA_M) * kk] = r;
        }
        // back substitution, block by block from the top: p[k] -= w3[k+1] * p[k+1]
        for (int kb = ((kmax - 1) / TDMA_M) * TDMA_M; kb >= 0; kb -= TDMA_M)
        {
            // w3 of levels kb+1 .. kb+M recomputed from 1/work2d(kb) exactly as the forward sweep formed them
            double w3[TDMA_M];
            double rr = ckpt[ij + (kb / TDMA_M) * kk];
            #pragma unroll
            for (int m = 0; m < TDMA_M; ++m)
            {
                const int k = kb + m + 1;
                if (k < kmax)
                {
                    w3[m] = c[k - 1] * rr;
                    if (m + 1 < TDMA_M) rr = fast_rcp(bdiag(k) - a[k] * w3[m]);
                }
                else
                    w3[m] = 0.;
            }
            if (kb >= TDMA_M) prefetch_l2(ckpt + ij + (kb / TDMA_M - 1) * kk);
            #pragma unroll
            for (int m = TDMA_M - 1; m >= 0; --m)
            {
                const int k = kb + m;
                if (k < kmax - 1)
                {
                    if (k >= 4) prefetch_l2(p + ij + (k - 4) * kk);
                    double pk = p[ij + k * kk];
                    pk -= w3[m] * pprev;
                    p[ij + k * kk] = pk;
                    pprev = pk;
                }
            }
        }
    }

    // The same sweep, software-pipelined for launches that cannot fill the GPU (kx sub-blocks of the y-slab solver, long
    // columns on few ranks): there the sweep runs at the latency of its dependent chain, and in pres2_tdma_kernel every level's
    // loads (p from L2, the level's metrics) are issued only after the previous level's store — the compiler may not move them
    // across it.  Here the loads of level k+2 (p) and k+1 (metrics) are issued before the chain of level k, and the back
    // substitution fetches the 16 p values of a block before it recomputes the block's c'.  Same operations in the same order
    // on every value: bit-identical to pres2_tdma_kernel.  More registers, so the GPU-filling launches keep the other kernel.
    __global__ void __launch_bounds__(128)
    pres2_tdma_pipe_kernel(double* __restrict__ p, double* __restrict__ ckpt,
                           const double* __restrict__ a, const double* __restrict__ c,
                           const double* __restrict__ bmati, const double* __restrict__ bmatj,
                           const int iblock, const int jblock, const int ioff, const int joff, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        const int j = blockIdx.y;
        if (i >= iblock) return;
        const int kmax = gd.kmax, kgc = gd.kgc;
        const long kk = (long)iblock * jblock;
        const long ij = i + (long)j * iblock;
        const int iindex = ioff + i, jindex = joff + j;
        const double bm = bmati[iindex] + bmatj[jindex];
        const bool mode0 = (iindex == 0 && jindex == 0);
        const double* __restrict__ dzp = gd.dz + kgc;
        const double* __restrict__ rhop = gd.rhoref + kgc;

        // level record: dz, rho, a, c of level k (read-only path: free to move across the stores)
        struct Lev { double dz, rho, a, c; };
        auto lev = [&](int k) -> Lev { return Lev{__ldg(dzp + k), __ldg(rhop + k), __ldg(a + k), __ldg(c + k)}; };
        auto bdiag = [&](int k, const Lev& l) -> double
        {
            double b = l.dz * l.dz * l.rho * bm - (l.a + l.c);
            if (k == 0) b += l.a;
            if (k == kmax - 1) { if (mode0) b -= l.c; else b += l.c; }
            return b;
        };

        Lev l0 = lev(0);
        double r = fast_rcp(bdiag(0, l0));
        double pprev = (l0.dz * l0.dz * p[ij]) * r;
        p[ij] = pprev;
        ckpt[ij] = r;
        double cprev = l0.c;
        Lev ln = lev(kmax > 1 ? 1 : 0);
        double p1 = kmax > 1 ? p[ij + kk] : 0., p2 = kmax > 2 ? p[ij + 2 * kk] : 0.;
        for (int k = 1; k < kmax; ++k)
        {
            const Lev l = ln;
            const double praw = p1;
            if (k + 1 < kmax) ln = lev(k + 1);
            p1 = p2;
            if (k + 2 < kmax) p2 = p[ij + (k + 2) * kk];
            if (k + 6 < kmax) prefetch_l2(p + ij + (k + 6) * kk);
            const double w3 = cprev * r;
            r = fast_rcp(bdiag(k, l) - l.a * w3);
            double pk = l.dz * l.dz * praw;
            pk -= l.a * pprev;
            pk *= r;
            p[ij + k * kk] = pk;
            pprev = pk;
            cprev = l.c;
            if ((k & (TDMA_M - 1)) == 0) ckpt[ij + (k / TDMA_M) * kk] = r;
        }
        for (int kb = ((kmax - 1) / TDMA_M) * TDMA_M; kb >= 0; kb -= TDMA_M)
        {
            // the block's p values first: their latency hides behind the recomputation of c'
            double pv[TDMA_M];
            #pragma unroll
            for (int m = 0; m < TDMA_M; ++m)
                pv[m] = (kb + m < kmax - 1) ? p[ij + (kb + m) * kk] : 0.;
            double w3[TDMA_M];
            double rr = ckpt[ij + (kb / TDMA_M) * kk];
            if (kb >= TDMA_M) prefetch_l2(ckpt + ij + (kb / TDMA_M - 1) * kk);
            #pragma unroll
            for (int m = 0; m < TDMA_M; ++m)
            {
                const int k = kb + m + 1;
                if (k < kmax)
                {
                    const Lev l = lev(k);
                    w3[m] = __ldg(c + k - 1) * rr;
                    if (m + 1 < TDMA_M) rr = fast_rcp(bdiag(k, l) - l.a * w3[m]);
                }
                else
                    w3[m] = 0.;
            }
            #pragma unroll
            for (int m = TDMA_M - 1; m >= 0; --m)
            {
                const int k = kb + m;
                if (k < kmax - 1)
                {
                    double pk = pv[m];
                    pk -= w3[m] * pprev;
                    p[ij + k * kk] = pk;
                    pprev = pk;
                }
            }
        }
    }

    // p[kstart-1] = p[kstart] on interior columns, src/pres_2.cxx:352-358
    __global__ void pres2_bottom_ghost_kernel(double* __restrict__ p, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x + gd.istart;
        const int j = blockIdx.y + gd.jstart;
        if (i >= gd.iend) return;
        const int ijk = i + j * gd.icells + gd.kstart * gd.ijcells;
        p[ijk - gd.ijcells] = p[ijk];
    }

    // ghost-free -> ghost-celled copy (stage API only; the fused path scatters from the FFT kernel)
    __global__ void pres_scatter_kernel(double* __restrict__ p, const double* __restrict__ src, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        const int j = blockIdx.y;
        const int k = blockIdx.z;
        if (i >= gd.imax) return;
        p[i + gd.igc + (j + gd.jgc) * gd.icells + (k + gd.kgc) * gd.ijcells] = src[i + j * gd.imax + k * gd.imax * gd.jmax];
    }

    // Pres_2::output, src/pres_2.cxx:365-387
    __global__ void __launch_bounds__(256, MHH_MINB)
    pres2_out_kernel(double* __restrict__ ut, double* __restrict__ vt, double* __restrict__ wt,
                     const double* __restrict__ p, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x + gd.istart;
        const int j = blockIdx.y * blockDim.y + threadIdx.y + gd.jstart;
        if (i >= gd.iend || j >= gd.jend) return;
        const int jj = gd.icells, kk = gd.ijcells;
        const int k0 = gd.kstart + blockIdx.z * KCHUNK;
        const int k1 = min(k0 + KCHUNK, gd.kend);
        const double dxi = 1. / gd.dx, dyi = 1. / gd.dy;
        int ijk = i + j * jj + k0 * kk;
        double p_b = p[ijk - kk];
        MHH_PRAGMA_UNROLL(MHH_PRES_OUT_UNROLL_N)
        for (int k = k0; k < k1; ++k, ijk += kk)
        {
            const double p_c = p[ijk];
            ut[ijk] -= (p_c - p[ijk - 1]) * dxi;
            vt[ijk] -= (p_c - p[ijk - jj]) * dyi;
            wt[ijk] -= (p_c - p_b) * gd.dzhi[k];
            p_b = p_c;
        }
    }

    // Pres_2::output (src/pres_2.cxx:365-387) fused with the Runge-Kutta update of u,v,w that follows it in the substep
    // (Timeloop::exec, src/timeloop.cxx:358-411): the corrected tendencies are consumed while they are in registers, which
    // saves one read and one write of ut,vt,wt per substep (13 instead of 19 field passes).  Same operations in the same
    // order as pres2_out_kernel followed by rk_kernel, so the results are bit-identical.  One thread per cell of the
    // ghost-celled array, as rk_kernel (which also clears the tendency ghost cells at the end of a full step).
    __global__ void __launch_bounds__(256)
    pres2_out_rk_kernel(double* __restrict__ ut, double* __restrict__ vt, double* __restrict__ wt,
                        double* __restrict__ u, double* __restrict__ v, double* __restrict__ w,
                        const double* __restrict__ p, const double cB_dt, const double cA_next, const int zero_all,
                        const GridD gd, const int keep_t)
    {
        const int ijk = blockIdx.x * blockDim.x + threadIdx.x;
        if (ijk >= gd.ncells) return;
        const int i = ijk % gd.icells;
        const int row = ijk / gd.icells;       // j + k*jcells
        const int j = row % gd.jcells, k = row / gd.jcells;
        const bool interior = i >= gd.istart && i < gd.iend && j >= gd.jstart && j < gd.jend
                           && k >= gd.kstart && k < gd.kend;
        if (interior)
        {
            const double dxi = 1. / gd.dx, dyi = 1. / gd.dy;
            const double p_c = p[ijk], p_w = p[ijk - 1], p_s = p[ijk - gd.icells], p_b = p[ijk - gd.ijcells];
            const double tu0 = ut[ijk], tv0 = vt[ijk], tw0 = wt[ijk];
            const double u0 = u[ijk], v0 = v[ijk], w0 = w[ijk];
            const double tu = tu0 - (p_c - p_w) * dxi;
            const double tv = tv0 - (p_c - p_s) * dyi;
            const double tw = tw0 - (p_c - p_b) * gd.dzhi[k];
            u[ijk] = u0 + cB_dt * tu;
            v[ijk] = v0 + cB_dt * tv;
            w[ijk] = w0 + cB_dt * tw;
            if (!keep_t)         // keep_t (lazy-tendency mode, last substep): the tendency arrays are left undefined
            {
                ut[ijk] = zero_all ? 0. : tu * cA_next;
                vt[ijk] = zero_all ? 0. : tv * cA_next;
                wt[ijk] = zero_all ? 0. : tw * cA_next;
            }
        }
        else if (zero_all && !keep_t)
        {
            ut[ijk] = 0.; vt[ijk] = 0.; wt[ijk] = 0.;
        }
    }

    // Pres_2::calc_divergence, src/pres_2.cxx:391-422
    __global__ void __launch_bounds__(256, MHH_MINB)
    pres2_div_kernel(const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ w,
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
            for (int k = k0; k < k1; ++k)
            {
                const int ijk = i + j * jj + k * kk;
                const double div = gd.rhoref[k] * ((u[ijk + 1] - u[ijk]) * dxi + (v[ijk + jj] - v[ijk]) * dyi)
                                 + (gd.rhorefh[k + 1] * w[ijk + kk] - gd.rhorefh[k] * w[ijk]) * gd.dzi[k];
                const unsigned long long b = abs_bits(div);
                m = b > m ? b : m;
            }
        block_max_bits(m, out);
    }
}  // namespace

// Pres_2::set_values, src/pres_2.cxx:125-153
int pres2_set_values(Ctx* c)
{
    const GridD& gd = c->gd;
    const double dxidxi = 1. / (gd.dx * gd.dx);
    const double dyidyi = 1. / (gd.dy * gd.dy);
    const double pi = std::acos(-1.);
    c->bmati.assign(gd.itot, 0.);
    c->bmatj.assign(gd.jtot, 0.);
    for (int j = 0; j < gd.jtot / 2 + 1; ++j)
        c->bmatj[j] = 2. * (std::cos(2. * pi * (double)j / (double)gd.jtot) - 1.) * dyidyi;
    for (int j = gd.jtot / 2 + 1; j < gd.jtot; ++j)
        c->bmatj[j] = c->bmatj[gd.jtot - j];
    for (int i = 0; i < gd.itot / 2 + 1; ++i)
        c->bmati[i] = 2. * (std::cos(2. * pi * (double)i / (double)gd.itot) - 1.) * dxidxi;
    for (int i = gd.itot / 2 + 1; i < gd.itot; ++i)
        c->bmati[i] = c->bmati[gd.itot - i];
    c->pres_a.assign(gd.kmax, 0.);
    c->pres_c.assign(gd.kmax, 0.);
    for (int k = 0; k < gd.kmax; ++k)
    {
        c->pres_a[k] = c->dz[k + gd.kgc] * c->rhorefh[k + gd.kgc] * c->dzhi[k + gd.kgc];
        c->pres_c[k] = c->dz[k + gd.kgc] * c->rhorefh[k + gd.kgc + 1] * c->dzhi[k + gd.kgc + 1];
    }
    return MHH_OK;
}

int pres2_input(Ctx* c, const double* u, const double* v, const double* w,
                const double* ut, const double* vt, const double* wt, double dt, double* work)
{
    const GridD& gd = c->gd;
    Launch3 l = launch_interior(gd.imax, gd.jmax, (gd.kmax + KCHUNK - 1) / KCHUNK);
    MHH_LAUNCH(c, "pres2_in_kernel", pres2_in_kernel<<<l.grid, l.block, 0, c->stream>>>(work, u, v, w, ut, vt, wt, 1. / dt, gd));
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

int pres2_solve_spectral_ws(Ctx* c, double* data, double* work3d)
{
    const GridD& gd = c->gd;
    dim3 b(128), g((c->iblock + 127) / 128, c->jblock);
    MHH_LAUNCH(c, "pres2_tdma_kernel", pres2_tdma_kernel<<<g, b, 0, c->stream>>>(data, work3d, c->pres_a_d, c->pres_c_d, c->bmati_d, c->bmatj_d,
                                              c->iblock, c->jblock, c->rank_y * c->iblock, 0, gd));
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

// kx sub-block b of nsub (capi.cu: mhh_slab_*_sub): Y_b = [k][j < jtot][ii < ib] at b * kmax * jtot * ib, columns
// kx = rank * iblock + b * ib + ii (src/pres_2.cxx:297-299); work: kmax / 16 + 1 planes of jtot * ib
int pres2_solve_spectral_sub(Ctx* c, double* data, double* work, int b, int nsub)
{
    const GridD& gd = c->gd;
    const int ib = c->iblock / nsub;
    const int bx = ib >= 128 ? 128 : (ib >= 64 ? 64 : 32);
    dim3 blk(bx), g((ib + bx - 1) / bx, c->jblock);
    // launches of few columns run at the latency of the sweep's chain: the software-pipelined kernel (0.54 instead of 0.84 ms
    // for 32 x 1024 columns of 1024 levels, tools/bench_tdma_sub.py); GPU-filling ones keep the leaner kernel (0.95 against 1.17)
    static const int pipe_env = []() { const char* e = getenv("MHH_TDMA_PIPE"); return e ? atoi(e) : -1; }();
    const bool pipe = pipe_env >= 0 ? pipe_env != 0 : (long)ib * c->jblock <= 40000;
    double* d = data + (long)b * gd.kmax * c->jblock * ib;
    const int ioff = c->rank_y * c->iblock + b * ib;
    if (pipe)
        MHH_LAUNCH(c, "pres2_tdma_kernel", pres2_tdma_pipe_kernel<<<g, blk, 0, c->stream>>>(d, work, c->pres_a_d, c->pres_c_d, c->bmati_d, c->bmatj_d,
                                                                                              ib, c->jblock, ioff, 0, gd));
    else
        MHH_LAUNCH(c, "pres2_tdma_kernel", pres2_tdma_kernel<<<g, blk, 0, c->stream>>>(d, work, c->pres_a_d, c->pres_c_d, c->bmati_d, c->bmatj_d,
                                                                                         ib, c->jblock, ioff, 0, gd));
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

int pres2_solve_spectral(Ctx* c, double* data) { return pres2_solve_spectral_ws(c, data, c->work_c); }

int pres_scatter(Ctx* c, double* p, const double* src)
{
    const GridD& gd = c->gd;
    dim3 b(128), g((gd.imax + 127) / 128, gd.jmax, gd.kmax);
    MHH_LAUNCH(c, "pres_scatter_kernel", pres_scatter_kernel<<<g, b, 0, c->stream>>>(p, src, gd));
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

int pres2_output_stage(Ctx* c, int stage, double* ut, double* vt, double* wt, double* p);

// p already holds the interior (scattered); finish ghost rows/halos and apply the gradient.
int pres2_output(Ctx* c, double* ut, double* vt, double* wt, double* p)
{
    if (int e = pres2_output_stage(c, 0, ut, vt, wt, p)) return e;
    return pres2_output_stage(c, 1, ut, vt, wt, p);
}

// stage 0: bottom ghost row + cyclic halos (east/west only on a y-slab rank); stage 1: gradient.
int pres2_output_stage(Ctx* c, int stage, double* ut, double* vt, double* wt, double* p)
{
    const GridD& gd = c->gd;
    if (stage == 0)
    {
        dim3 b(128), g((gd.imax + 127) / 128, gd.jmax);
        MHH_LAUNCH(c, "pres2_bottom_ghost_kernel", pres2_bottom_ghost_kernel<<<g, b, 0, c->stream>>>(p, gd));
        MHH_LAUNCH_CHECK();
    }
    if (stage == 0) return cyclic_exec(c, p, 2);
    Launch3 l = launch_interior(gd.imax, gd.jmax, (gd.kmax + KCHUNK - 1) / KCHUNK);
    MHH_LAUNCH(c, "pres2_out_kernel", pres2_out_kernel<<<l.grid, l.block, 0, c->stream>>>(ut, vt, wt, p, gd));
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

void rk_coefs(int rkorder, int substep, double dt, double* cB_dt, double* cA_next, int* zero_all);

// gradient stage fused with the RK substep of u,v,w; p must have its ghost row and halos (pres2_output_stage 0)
int pres2_output_rk(Ctx* c, double* ut, double* vt, double* wt, double* p, double* u, double* v, double* w,
                    int rkorder, int substep, double dt)
{
    if ((rkorder != 3 && rkorder != 4) || substep < 0 || substep >= (rkorder == 3 ? 3 : 5)) return MHH_ERR_ARG;
    const GridD& gd = c->gd;
    double cB_dt, cA_next; int zero_all;
    rk_coefs(rkorder, substep, dt, &cB_dt, &cA_next, &zero_all);
    dim3 b(256), g((gd.ncells + 255) / 256);
    const int keep_t = (zero_all && c->lazy_tend) ? 1 : 0;
    MHH_LAUNCH(c, "pres2_out_rk_kernel", pres2_out_rk_kernel<<<g, b, 0, c->stream>>>(ut, vt, wt, u, v, w, p, cB_dt, cA_next, zero_all, gd, keep_t));
    if (keep_t) { c->fresh_uvw[0] = ut; c->fresh_uvw[1] = vt; c->fresh_uvw[2] = wt; }
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

int pres2_divergence_bits(Ctx* c, const double* u, const double* v, const double* w, unsigned long long* slot)
{
    const GridD& gd = c->gd;
    Launch3 l = launch_interior(gd.imax, gd.jmax, (gd.kmax + KCHUNK - 1) / KCHUNK);
    MHH_CUDA_CHECK(cudaMemsetAsync(slot, 0, sizeof(unsigned long long), c->stream));
    MHH_LAUNCH(c, "pres2_div_kernel", pres2_div_kernel<<<l.grid, l.block, 0, c->stream>>>(u, v, w, gd, slot));
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

}  // namespace mhh
