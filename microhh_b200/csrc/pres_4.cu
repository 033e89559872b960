// 4th-order Poisson pressure solver for sm_100a (reference: src/pres_4.cxx).
//
//   set_values : 4th-order modified wavenumbers and the seven k-only bands m1..m7            (:179-252)
//   input      : mirror wt over the walls, 4th-order divergence of (u/dt + ut) into a ghost-free array (:256-317)
//   solve      : FFTs (fft.cu) + per-(kx,ky) hepta-diagonal system on kmax+4 rows                      (:320-505, hdma :574-729)
//   output     : two p ghost rows per wall, cyclic halos, ut/vt/wt -= 4th-order grad p                (:508-571)
//
// hdma on B200: the reference materialises eight (kmax+4)-row work slabs per column and therefore solves in
// j-slices.  Here one thread owns a (kx,ky) column: the bands are k-only constants (only m4 gets
// + bmati + bmatj), so the LU factorisation runs on the fly with the three previous rows of U (and of y)
// in registers, fused with the forward substitution; only U (m4..m7 after elimination) of the kmax interior
// rows and y go to HBM, and the backward substitution streams them back.  The four boundary rows never leave
// registers.  All rows use the reference's general elimination formulas; on the rows where the reference
// drops terms the dropped operands are exact zeros, so the results are the same to the last bit.
#include "sys.h"
#include <cmath>
#ifndef MHH_PRES_OUT_UNROLL_N
#define MHH_PRES_OUT_UNROLL_N 1
#endif
#ifndef MHH_PRES_IN_UNROLL_N
#define MHH_PRES_IN_UNROLL_N 2
#endif

#include "ctx.h"
#include "fft.h"

namespace mhh
{
int cyclic_exec(Ctx* c, double* fld, int edge);
using namespace fd;

namespace
{
    constexpr int KCHUNK = MHH_KCHUNK;

    __device__ __forceinline__ double G4(double a, double b, double c, double d)
    { return cg0 * a + cg1 * b + cg2 * c + cg3 * d; }

    // wt[kstart-1] = -wt[kstart+1], wt[kend+1] = -wt[kend-1] on interior columns, src/pres_4.cxx:288-303
    __global__ void pres4_wt_mirror_kernel(double* __restrict__ wt, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x + gd.istart;
        const int j = blockIdx.y + gd.jstart;
        if (i >= gd.iend) return;
        const int kk = gd.ijcells;
        const int ijb = i + j * gd.icells + gd.kstart * kk;
        wt[ijb - kk] = -wt[ijb + kk];
        const int ijt = i + j * gd.icells + gd.kend * kk;
        wt[ijt + kk] = -wt[ijt - kk];
    }

    // Pres_4::input, src/pres_4.cxx:305-316
    template<bool DIM3>
    __global__ void __launch_bounds__(256, MHH_MINB)
    pres4_in_kernel(double* __restrict__ p, const double* __restrict__ u, const double* __restrict__ v,
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
        // k-marching queue of (wt + w*dti) at levels k-1, k, k+1 (k+2 is loaded per level)
        double f0 = wt[ijk - kk] + w[ijk - kk] * dti;
        double f1 = wt[ijk] + w[ijk] * dti;
        double f2 = wt[ijk + kk] + w[ijk + kk] * dti;
        MHH_PRAGMA_UNROLL(MHH_PRES_IN_UNROLL_N)
        for (int k = k0; k < k1; ++k, ijk += kk)
        {
            const double f3 = wt[ijk + 2 * kk] + w[ijk + 2 * kk] * dti;
            double r = G4(ut[ijk - ii] + u[ijk - ii] * dti, ut[ijk] + u[ijk] * dti,
                          ut[ijk + ii] + u[ijk + ii] * dti, ut[ijk + 2 * ii] + u[ijk + 2 * ii] * dti) * dxi;
            if (DIM3)
                r += G4(vt[ijk - jj] + v[ijk - jj] * dti, vt[ijk] + v[ijk] * dti,
                        vt[ijk + jj] + v[ijk + jj] * dti, vt[ijk + 2 * jj] + v[ijk + 2 * jj] * dti) * dyi;
            r += G4(f0, f1, f2, f3) * gd.dzi4[k + gd.kgc];
            p[i + j * jjp + k * kkp] = r;
            f0 = f1; f1 = f2; f2 = f3;
        }
    }

    // A row of U after elimination; the pivot is kept as its reciprocal (one fast_rcp per row instead of the reference's
    // four divisions by it: three in the rows below, one in the back substitution)
    struct Row { double u4i, u5, u6, u7, y; };

    // One elimination step of hdma (src/pres_4.cxx:636-647 and the reduced rows :597-634,:658-686) fused with
    // Ly = p (:691-708): a,b,c are rows r-1, r-2, r-3 after elimination.
    __device__ __forceinline__ Row eliminate(double m1, double m2, double m3, double m4, double m5, double m6, double m7,
                                             double rhs, const Row& a, const Row& b, const Row& c)
    {
        Row n;
        const double l1 = m1 * c.u4i;
        const double l2 = (m2 - l1 * c.u5) * b.u4i;
        const double l3 = (m3 - l2 * b.u5 - l1 * c.u6) * a.u4i;
        n.u4i = fast_rcp(m4 - l3 * a.u5 - l2 * b.u6 - l1 * c.u7);
        n.u5 = m5 - l3 * a.u6 - l2 * b.u7;
        n.u6 = m6 - l3 * a.u7;
        n.u7 = m7;
        n.y = rhs - a.y * l3 - b.y * l2 - c.y * l1;
        return n;
    }

    // bands: m1..m7, kmax each; fac: 3 arrays (1/u4, u5, u6) of iblock*jblock*kmax (u7 = m7 is a per-level constant and is
    // re-read from the bands).  A checkpoint-and-recompute variant as in pres2_tdma_kernel (5.5 instead of 10 field
    // passes) was measured SLOWER here (1.59 vs 1.39 ms at 512x512x256): one elimination step is ~30 dependent FP64
    // operations, and rebuilding the rows doubles the k-sequential chain this kernel is bound by.
    __global__ void __launch_bounds__(128)
    pres4_hdma_kernel(double* __restrict__ p, double* __restrict__ fac, const double* __restrict__ bands,
                      const double* __restrict__ bmati, const double* __restrict__ bmatj,
                      const int iblock, const int jblock, const int ioff, const int joff, const int kmax)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        const int j = blockIdx.y;
        if (i >= iblock) return;
        const size_t kk = (size_t)iblock * jblock;
        const size_t n = kk * kmax;
        const size_t ij = i + (size_t)j * iblock;
        const int iindex = ioff + i, jindex = joff + j;
        const double bm = bmati[iindex] + bmatj[jindex];
        const bool mode0 = (iindex == 0 && jindex == 0);

        const Row phantom = {1., 0., 0., 0., 0.};
        // rows 0 and 1: zero-gradient bottom BC (:364-391)
        Row c = eliminate(0., 0., 0., 1., 0., 0., -1., 0., phantom, phantom, phantom);
        Row b = c;
        Row a = eliminate(0., 0., 0., 1., -1., 0., 0., 0., c, phantom, phantom);
        c = phantom;
        // now a = row 1, b = row 0, c = phantom (row -1)
        for (int k = 0; k < kmax; ++k)
        {
            const size_t o = ij + k * kk;
            if (k + 4 < kmax) prefetch_l2(p + o + 4 * kk);
            const Row r = eliminate(bands[k], bands[kmax + k], bands[2 * kmax + k], bands[3 * kmax + k] + bm,
                                    bands[4 * kmax + k], bands[5 * kmax + k], bands[6 * kmax + k], p[o], a, b, c);
            fac[o] = r.u4i; fac[n + o] = r.u5; fac[2 * n + o] = r.u6;
            p[o] = r.y;
            c = b; b = a; a = r;
        }
        // top rows kmax+2, kmax+3 (:422-468): mode (0,0) pins the mean, the others set dp/dz = 0
        Row t2, t3;
        if (mode0)
        {
            t2 = eliminate(0., -1. / 3., 2., 1., 0., 0., 0., 0., a, b, c);
            t3 = eliminate(-2., 9., 0., 1., 0., 0., 0., 0., t2, a, b);
        }
        else
        {
            t2 = eliminate(0., 0., -1., 1., 0., 0., 0., 0., a, b, c);
            t3 = eliminate(-1., 0., 0., 1., 0., 0., 0., 0., t2, a, b);
        }
        // Ux = y backward (:711-728). x3 = x[r+3], x2 = x[r+2], x1 = x[r+1]
        double x3 = t3.y * t3.u4i;
        double x2 = (t2.y - x3 * t2.u5) * t2.u4i;
        double x1 = (a.y - x2 * a.u5 - x3 * a.u6) * a.u4i;     // row kmax+1 = level kmax-1
        p[ij + (size_t)(kmax - 1) * kk] = x1;
        for (int k = kmax - 2; k >= 0; --k)
        {
            const size_t o = ij + k * kk;
            if (k >= 4) { prefetch_l2(p + o - 4 * kk); prefetch_l2(fac + o - 4 * kk); prefetch_l2(fac + n + o - 4 * kk); prefetch_l2(fac + 2 * n + o - 4 * kk); }
            const double x = (p[o] - x1 * fac[n + o] - x2 * fac[2 * n + o] - x3 * bands[6 * kmax + k]) * fac[o];
            p[o] = x;
            x3 = x2; x2 = x1; x1 = x;
        }
    }

    // two zero-gradient ghost rows per wall on interior columns, src/pres_4.cxx:508-525
    __global__ void pres4_ghost_kernel(double* __restrict__ p, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x + gd.istart;
        const int j = blockIdx.y + gd.jstart;
        if (i >= gd.iend) return;
        const int kk = gd.ijcells;
        const int ijb = i + j * gd.icells + gd.kstart * kk;
        p[ijb - kk] = p[ijb];
        p[ijb - 2 * kk] = p[ijb + kk];
        const int ijt = i + j * gd.icells + (gd.kend - 1) * kk;
        p[ijt + kk] = p[ijt];
        p[ijt + 2 * kk] = p[ijt - kk];
    }

    // Pres_4::output, src/pres_4.cxx:533-571
    template<bool DIM3>
    __global__ void __launch_bounds__(256, MHH_MINB)
    pres4_out_kernel(double* __restrict__ ut, double* __restrict__ vt, double* __restrict__ wt,
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
        double pm2 = p[ijk - 2 * kk], pm1 = p[ijk - kk], p0 = p[ijk];
        MHH_PRAGMA_UNROLL(MHH_PRES_OUT_UNROLL_N)
        for (int k = k0; k < k1; ++k, ijk += kk)
        {
            const double pp1 = p[ijk + kk];
            ut[ijk] -= G4(p[ijk - 2], p[ijk - 1], p0, p[ijk + 1]) * dxi;
            if (DIM3)
                vt[ijk] -= G4(p[ijk - 2 * jj], p[ijk - jj], p0, p[ijk + jj]) * dyi;
            if (k > gd.kstart)
                wt[ijk] -= G4(pm2, pm1, p0, pp1) * gd.dzhi4[k];
            pm2 = pm1; pm1 = p0; p0 = pp1;
        }
    }

    // Pres_4::output (src/pres_4.cxx:533-571) fused with the Runge-Kutta update of u,v,w (Timeloop::exec,
    // src/timeloop.cxx:358-411), as pres2_out_rk_kernel: the corrected tendencies are consumed in registers, one thread per
    // cell of the ghost-celled array, same operations in the same order as pres4_out_kernel followed by rk_kernel.
    template<bool DIM3>
    __global__ void __launch_bounds__(256)
    pres4_out_rk_kernel(double* __restrict__ ut, double* __restrict__ vt, double* __restrict__ wt,
                        double* __restrict__ u, double* __restrict__ v, double* __restrict__ w,
                        const double* __restrict__ p, const double cB_dt, const double cA_next, const int zero_all,
                        const GridD gd)
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
            const int jj = gd.icells, kk = gd.ijcells;
            const double dxi = 1. / gd.dx, dyi = 1. / gd.dy;
            const double p0 = p[ijk];
            const double tu0 = ut[ijk], tv0 = vt[ijk], tw0 = wt[ijk];
            const double u0 = u[ijk], v0 = v[ijk], w0 = w[ijk];
            const double tu = tu0 - G4(p[ijk - 2], p[ijk - 1], p0, p[ijk + 1]) * dxi;
            double tv = tv0, tw = tw0;
            if (DIM3) tv = tv0 - G4(p[ijk - 2 * jj], p[ijk - jj], p0, p[ijk + jj]) * dyi;
            if (k > gd.kstart) tw = tw0 - G4(p[ijk - 2 * kk], p[ijk - kk], p0, p[ijk + kk]) * gd.dzhi4[k];
            u[ijk] = u0 + cB_dt * tu;
            v[ijk] = v0 + cB_dt * tv;
            w[ijk] = w0 + cB_dt * tw;
            ut[ijk] = zero_all ? 0. : tu * cA_next;
            vt[ijk] = zero_all ? 0. : tv * cA_next;
            wt[ijk] = zero_all ? 0. : tw * cA_next;
        }
        else if (zero_all)
        {
            ut[ijk] = 0.; vt[ijk] = 0.; wt[ijk] = 0.;
        }
    }

    // Pres_4::calc_divergence, src/pres_4.cxx:733-767 (the y term is always included)
    __global__ void __launch_bounds__(256, MHH_MINB)
    pres4_div_kernel(const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ w,
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
                const double div = G4(u[ijk - 1], u[ijk], u[ijk + 1], u[ijk + 2]) * dxi
                                 + G4(v[ijk - jj], v[ijk], v[ijk + jj], v[ijk + 2 * jj]) * dyi
                                 + G4(w[ijk - kk], w[ijk], w[ijk + kk], w[ijk + 2 * kk]) * gd.dzi4[k];
                const unsigned long long b = abs_bits(div);
                m = b > m ? b : m;
            }
        block_max_bits(m, out);
    }
}  // namespace

// Pres_4::set_values, src/pres_4.cxx:179-252
int pres4_set_values(Ctx* c)
{
    const GridD& gd = c->gd;
    if (gd.igc < 3 || gd.jgc < 3 || gd.kgc < 3 || gd.kmax < 4)
    {
        c->last_error = "pres_4 needs swspatialorder=4 (3 ghost cells) and ktot >= 4";
        return MHH_ERR_ARG;
    }
    const int itot = gd.itot, jtot = gd.jtot, kmax = gd.kmax, kstart = gd.kstart;
    const double dxidxi = 1. / (gd.dx * gd.dx);
    const double dyidyi = 1. / (gd.dy * gd.dy);
    const double pi = std::acos(-1.);
    c->bmati.assign(itot, 0.);
    c->bmatj.assign(jtot, 0.);
    for (int j = 0; j < jtot / 2 + 1; ++j)
        c->bmatj[j] = ( 2. * (1. / 576.)   * std::cos(6. * pi * (double)j / (double)jtot)
                      - 2. * (54. / 576.)  * std::cos(4. * pi * (double)j / (double)jtot)
                      + 2. * (783. / 576.) * std::cos(2. * pi * (double)j / (double)jtot)
                      -      (1460. / 576.) ) * dyidyi;
    for (int j = jtot / 2 + 1; j < jtot; ++j)
        c->bmatj[j] = c->bmatj[jtot - j];
    for (int i = 0; i < itot / 2 + 1; ++i)
        c->bmati[i] = ( 2. * (1. / 576.)   * std::cos(6. * pi * (double)i / (double)itot)
                      - 2. * (54. / 576.)  * std::cos(4. * pi * (double)i / (double)itot)
                      + 2. * (783. / 576.) * std::cos(2. * pi * (double)i / (double)itot)
                      -      (1460. / 576.) ) * dxidxi;
    for (int i = itot / 2 + 1; i < itot; ++i)
        c->bmati[i] = c->bmati[itot - i];

    const std::vector<double>& dzi4 = c->dzi4;
    const std::vector<double>& dzhi4 = c->dzhi4;
    c->bands.assign(7 * (size_t)kmax, 0.);
    double* m1 = c->bands.data();
    double* m2 = m1 + kmax; double* m3 = m2 + kmax; double* m4 = m3 + kmax;
    double* m5 = m4 + kmax; double* m6 = m5 + kmax; double* m7 = m6 + kmax;
    // bottom row: w mirrored over the wall to conserve global momentum
    int k = 0, kc = kstart;
    m1[k] = 0.;
    m2[k] = (1. / 576.) * (                      -  27. * dzhi4[kc]                                              ) * dzi4[kc];
    m3[k] = (1. / 576.) * (  -1. * dzhi4[kc + 1] + 729. * dzhi4[kc] +  27. * dzhi4[kc + 1]                       ) * dzi4[kc];
    m4[k] = (1. / 576.) * (  27. * dzhi4[kc + 1] - 729. * dzhi4[kc] - 729. * dzhi4[kc + 1] -  1. * dzhi4[kc + 2] ) * dzi4[kc];
    m5[k] = (1. / 576.) * ( -27. * dzhi4[kc + 1] +  27. * dzhi4[kc] + 729. * dzhi4[kc + 1] + 27. * dzhi4[kc + 2] ) * dzi4[kc];
    m6[k] = (1. / 576.) * (   1. * dzhi4[kc + 1]                    -  27. * dzhi4[kc + 1] - 27. * dzhi4[kc + 2] ) * dzi4[kc];
    m7[k] = (1. / 576.) * (                                                                +  1. * dzhi4[kc + 2] ) * dzi4[kc];
    for (k = 1; k < kmax - 1; ++k)
    {
        kc = kstart + k;
        m1[k] = (1. / 576.) * (   1. * dzhi4[kc - 1]                                                                  ) * dzi4[kc];
        m2[k] = (1. / 576.) * ( -27. * dzhi4[kc - 1] -  27. * dzhi4[kc]                                               ) * dzi4[kc];
        m3[k] = (1. / 576.) * (  27. * dzhi4[kc - 1] + 729. * dzhi4[kc] +  27. * dzhi4[kc + 1]                        ) * dzi4[kc];
        m4[k] = (1. / 576.) * (  -1. * dzhi4[kc - 1] - 729. * dzhi4[kc] - 729. * dzhi4[kc + 1] -  1. * dzhi4[kc + 2]  ) * dzi4[kc];
        m5[k] = (1. / 576.) * (                      +  27. * dzhi4[kc] + 729. * dzhi4[kc + 1] + 27. * dzhi4[kc + 2]  ) * dzi4[kc];
        m6[k] = (1. / 576.) * (                                         -  27. * dzhi4[kc + 1] - 27. * dzhi4[kc + 2]  ) * dzi4[kc];
        m7[k] = (1. / 576.) * (                                                                +  1. * dzhi4[kc + 2]  ) * dzi4[kc];
    }
    // top row
    k = kmax - 1; kc = kstart + k;
    m1[k] = (1. / 576.) * (   1. * dzhi4[kc - 1]                                                             ) * dzi4[kc];
    m2[k] = (1. / 576.) * ( -27. * dzhi4[kc - 1] -  27. * dzhi4[kc]                       +  1. * dzhi4[kc]  ) * dzi4[kc];
    m3[k] = (1. / 576.) * (  27. * dzhi4[kc - 1] + 729. * dzhi4[kc] +  27. * dzhi4[kc + 1] - 27. * dzhi4[kc] ) * dzi4[kc];
    m4[k] = (1. / 576.) * (  -1. * dzhi4[kc - 1] - 729. * dzhi4[kc] - 729. * dzhi4[kc + 1] + 27. * dzhi4[kc] ) * dzi4[kc];
    m5[k] = (1. / 576.) * (                      +  27. * dzhi4[kc] + 729. * dzhi4[kc + 1] -  1. * dzhi4[kc] ) * dzi4[kc];
    m6[k] = (1. / 576.) * (                                         -  27. * dzhi4[kc + 1]                   ) * dzi4[kc];
    m7[k] = 0.;
    c->pres_a.clear();
    c->pres_c.clear();
    return MHH_OK;
}

int pres4_input(Ctx* c, const double* u, const double* v, const double* w,
                const double* ut, const double* vt, double* wt, double dt, double* work)
{
    const GridD& gd = c->gd;
    {
        dim3 b(128), g((gd.imax + 127) / 128, gd.jmax);
        MHH_LAUNCH(c, "pres4_wt_mirror_kernel", pres4_wt_mirror_kernel<<<g, b, 0, c->stream>>>(wt, gd));
        MHH_LAUNCH_CHECK();
    }
    Launch3 l = launch_interior(gd.imax, gd.jmax, (gd.kmax + KCHUNK - 1) / KCHUNK);
    if (gd.jtot == 1)
        MHH_LAUNCH(c, "pres4_in_kernel", (pres4_in_kernel<false><<<l.grid, l.block, 0, c->stream>>>(work, u, v, w, ut, vt, wt, 1. / dt, gd)));
    else
        MHH_LAUNCH(c, "pres4_in_kernel", (pres4_in_kernel<true><<<l.grid, l.block, 0, c->stream>>>(work, u, v, w, ut, vt, wt, 1. / dt, gd)));
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

int pres4_solve_spectral(Ctx* c, double* data)
{
    const GridD& gd = c->gd;
    const size_t n = (size_t)c->iblock * c->jblock * gd.kmax;
    if (!c->work_h)
        MHH_CUDA_CHECK(cudaMalloc(&c->work_h, sizeof(double) * 3 * n));
    dim3 b(128), g((c->iblock + 127) / 128, c->jblock);
    MHH_LAUNCH(c, "pres4_hdma_kernel", pres4_hdma_kernel<<<g, b, 0, c->stream>>>(data, c->work_h, c->bands_d, c->bmati_d, c->bmatj_d,
                                              c->iblock, c->jblock, c->rank_y * c->iblock, 0, gd.kmax));
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

// stage 0 of Pres_4's tail: the two ghost rows of p at each wall (src/pres_4.cxx:508-525) and the halo fill — on a y-slab
// rank (npy > 1) only its east/west part; the caller then exchanges the three north/south rows before the gradient stage
int pres4_output_ghosts(Ctx* c, double* p)
{
    const GridD& gd = c->gd;
    dim3 b(128), g((gd.imax + 127) / 128, gd.jmax);
    MHH_LAUNCH(c, "pres4_ghost_kernel", pres4_ghost_kernel<<<g, b, 0, c->stream>>>(p, gd));
    MHH_LAUNCH_CHECK();
    return cyclic_exec(c, p, 2);
}

// p already holds the interior (scattered by the backward x FFT); ghost rows, halos (with_ghosts), gradient.
int pres4_output(Ctx* c, double* ut, double* vt, double* wt, double* p, const bool with_ghosts)
{
    const GridD& gd = c->gd;
    if (with_ghosts)
        if (int e = pres4_output_ghosts(c, p)) return e;
    Launch3 l = launch_interior(gd.imax, gd.jmax, (gd.kmax + KCHUNK - 1) / KCHUNK);
    if (gd.jtot == 1)
        MHH_LAUNCH(c, "pres4_out_kernel", (pres4_out_kernel<false><<<l.grid, l.block, 0, c->stream>>>(ut, vt, wt, p, gd)));
    else
        MHH_LAUNCH(c, "pres4_out_kernel", (pres4_out_kernel<true><<<l.grid, l.block, 0, c->stream>>>(ut, vt, wt, p, gd)));
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

void rk_coefs(int rkorder, int substep, double dt, double* cB_dt, double* cA_next, int* zero_all);
int ghost_cells_w_exec(Ctx* c, double* w, int type);

// ghost rows + halos of p, then w's ghost rows in their Normal state (the reference sets them between Pres::exec and
// Timeloop::exec, src/model.cxx:412, from the not yet updated w), then the gradient fused with the RK substep of u,v,w
int pres4_output_rk(Ctx* c, double* ut, double* vt, double* wt, double* p, double* u, double* v, double* w,
                    int rkorder, int substep, double dt, const bool with_ghosts)
{
    if ((rkorder != 3 && rkorder != 4) || substep < 0 || substep >= (rkorder == 3 ? 3 : 5)) return MHH_ERR_ARG;
    const GridD& gd = c->gd;
    if (with_ghosts)
        if (int e = pres4_output_ghosts(c, p)) return e;
    if (int e = ghost_cells_w_exec(c, w, 0 /* MHH_BCW_NORMAL */)) return e;
    double cB_dt, cA_next; int zero_all;
    rk_coefs(rkorder, substep, dt, &cB_dt, &cA_next, &zero_all);
    dim3 b(256), g((gd.ncells + 255) / 256);
    if (gd.jtot == 1)
        MHH_LAUNCH(c, "pres4_out_rk_kernel", (pres4_out_rk_kernel<false><<<g, b, 0, c->stream>>>(ut, vt, wt, u, v, w, p, cB_dt, cA_next, zero_all, gd)));
    else
        MHH_LAUNCH(c, "pres4_out_rk_kernel", (pres4_out_rk_kernel<true><<<g, b, 0, c->stream>>>(ut, vt, wt, u, v, w, p, cB_dt, cA_next, zero_all, gd)));
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

int pres4_divergence_bits(Ctx* c, const double* u, const double* v, const double* w, unsigned long long* slot)
{
    const GridD& gd = c->gd;
    Launch3 l = launch_interior(gd.imax, gd.jmax, (gd.kmax + KCHUNK - 1) / KCHUNK);
    MHH_CUDA_CHECK(cudaMemsetAsync(slot, 0, sizeof(unsigned long long), c->stream));
    MHH_LAUNCH(c, "pres4_div_kernel", pres4_div_kernel<<<l.grid, l.block, 0, c->stream>>>(u, v, w, gd, slot));
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

}  // namespace mhh
