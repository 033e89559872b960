// 4th-order advection for sm_100a (reference: src/advec_4.cxx:51-590; Advec_4<TF>::exec :592).
//
// flux-form divergence  cg0..3 * [ interp(velocity) * interp(advected) ]  with ci0..3 centred
// interpolations and the biased bi*/ti* interpolations in the first/last row (no density weighting).
// One thread per (i,j) column marches a chunk of k; every 4-point interpolation is written out with
// the reference's left-to-right summation order and the x, y, z contributions are subtracted from the
// tendency one after the other, as the reference does, so results agree to rounding with the CPU build.
#include "sys.h"
#include "ctx.h"
#ifndef MHH_ADVEC4_MINB
#define MHH_ADVEC4_MINB 2
#endif
#ifndef MHH_ADVEC4_UNROLL_N
#define MHH_ADVEC4_UNROLL_N 1
#endif

namespace mhh
{
using namespace fd;

namespace
{
    constexpr int KCHUNK = MHH_KCHUNK;

    // ci-interpolation of p[o], p[o+s], p[o+2s], p[o+3s]
    __device__ __forceinline__ double I4(const double* __restrict__ p, int o, int s)
    { return ci0 * p[o] + ci1 * p[o + s] + ci2 * p[o + 2 * s] + ci3 * p[o + 3 * s]; }
    __device__ __forceinline__ double B4(const double* __restrict__ p, int o, int s)
    { return bi0 * p[o] + bi1 * p[o + s] + bi2 * p[o + 2 * s] + bi3 * p[o + 3 * s]; }
    __device__ __forceinline__ double T4(const double* __restrict__ p, int o, int s)
    { return ti0 * p[o] + ti1 * p[o + s] + ti2 * p[o + 2 * s] + ti3 * p[o + 3 * s]; }
    // 16 x the ci / bi / ti interpolations, same roundings (see advec4_mom_kernel)
    __device__ __forceinline__ double CS(double a, double b, double c, double d) { return ((9. * b - a) + 9. * c) - d; }
    __device__ __forceinline__ double BS(double a, double b, double c, double d) { return ((5. * a + 15. * b) - 5. * c) + d; }
    __device__ __forceinline__ double TS(double a, double b, double c, double d) { return ((a - 5. * b) + 15. * c) + 5. * d; }
    __device__ __forceinline__ double I4S(const double* __restrict__ p, int o, int s)
    { return CS(p[o], p[o + s], p[o + 2 * s], p[o + 3 * s]); }
    __device__ __forceinline__ double G4(double a, double b, double c, double d)
    { return cg0 * a + cg1 * b + cg2 * c + cg3 * d; }

    // Momentum component COMP (0 = u, 1 = v, 2 = w): src/advec_4.cxx:89 (advec_u), :189 (advec_v), :289 (advec_w).
    // `a` is the advected component itself; the advecting velocity in direction d is interpolated along the
    // staggering direction of `a` (ci over 4 points starting 2 below), exactly as in the reference.
    // DIFF: Diff_4's diff_c (src/diff_4.cxx:42-107) of the same component appended as three more `+=` terms — every value
    // it reads (a at i-3..i+3, j-3..j+3 and the k-3..k+3 column) is already in flight for the advection stencil, so the
    // fused pass costs no extra global traffic.  For w, diff_w (src/diff_4.cxx:109-174) reads Normal-type ghost rows while
    // the advection reads the Conservation-type ones (src/model.cxx:387-392): the two Normal-type values a column needs
    // (rows kstart-1 and kend+1, src/boundary.cxx:838-870) are evaluated in the kernel and substituted where the diffusion
    // stencil touches those rows.
    template<int COMP, bool DIM3, bool DIFF>
    __global__ void __launch_bounds__(256, MHH_ADVEC4_MINB)
    advec4_mom_kernel(double* __restrict__ at, const double* __restrict__ u, const double* __restrict__ v,
                      const double* __restrict__ w, const double visc, const GridD gd)
    {

        const int i = blockIdx.x * blockDim.x + threadIdx.x + gd.istart;
        const int j = blockIdx.y * blockDim.y + threadIdx.y + gd.jstart;
        if (i >= gd.iend || j >= gd.jend) return;
        const int ii = 1, jj = gd.icells, kk = gd.ijcells;
        const int kfirst = (COMP == 2) ? gd.kstart + 1 : gd.kstart;
        const int k0 = kfirst + blockIdx.z * KCHUNK;
        const int k1 = min(k0 + KCHUNK, gd.kend);
        // The interpolation weights are k/16 with k = -1, 9, 9, -1 (ci), 5, 15, -5, 1 (bi), 1, -5, 15, 5 (ti): a power-of-two
        // scaling commutes with every rounding, so ci0*a + ci1*b + ci2*c + ci3*d = (((9b - a) + 9c) - d) / 16 bit for bit.  The
        // march works on the unscaled sums (5 instead of 7 operations each), every flux is a product of two of them, and the
        // 1/256 is collected in the metric: G4(P/256 ...) * dxi = G4(P ...) * (dxi / 256).
        const double dxi = 1. / gd.dx, dyi = 1. / gd.dy;
        const double sdxi = dxi * (1. / 256.), sdyi = dyi * (1. / 256.);
        const double* __restrict__ a = (COMP == 0) ? u : (COMP == 1) ? v : w;
        const int sa = (COMP == 0) ? ii : (COMP == 1) ? jj : kk;     // staggering direction of a
        const double* __restrict__ mz = (COMP == 2) ? gd.dzhi4 : gd.dzi4;

        // ---- z direction: every face product  F(l) = interp_sa(w)(l) * interp_z(a)(l)  (w: interp_z(w)(l)^2) is shared by
        // the four levels l-2..l+1, so the march carries three of them and the a-column in registers: one new face per
        // level (4+1 loads, 15 flops instead of 32 loads, 60 flops).  The first / last row swap the outermost
        // interpolation for the biased one; those faces are used once (they leave the queue first / enter it last).
        const int ij = i + j * jj;
        auto face = [&](const int l, const double A) -> double
        { return (COMP == 2) ? A * A : I4S(w, ij + l * kk - 2 * sa, sa) * A; };
        double qa = a[ij + (k0 - 1) * kk], qb = a[ij + k0 * kk], qc = a[ij + (k0 + 1) * kk], qd = a[ij + (k0 + 2) * kk];
        double f0, f1, f2;
        double qm2 = 0., qm3 = 0.;      // a[k-2], a[k-3]: only the diffusion stencil keeps them
        double gbot = 0., gtop = 0.;    // Normal-type w ghost values of this column (COMP == 2 && DIFF)
        if (COMP == 2 && DIFF)
        {
            const int b = ij + gd.kstart * kk, tp = ij + gd.kend * kk;
            if (k0 <= gd.kstart + 2) gbot = -6. * w[b + kk] + 4. * w[b + 2 * kk] - w[b + 3 * kk];
            if (k1 >= gd.kend - 2)   gtop = -6. * w[tp - kk] + 4. * w[tp - 2 * kk] - w[tp - 3 * kk];
        }
        {
            const double am2 = a[ij + (k0 - 2) * kk], am3 = a[ij + (k0 - 3) * kk];
            if (DIFF) { qm2 = am2; qm3 = am3; }
            const double A0 = (k0 == kfirst) ? BS(am2, qa, qb, qc) : CS(am3, am2, qa, qb);
            f0 = face(k0 - 1, A0);
            f1 = face(k0, CS(am2, qa, qb, qc));
            f2 = face(k0 + 1, CS(qa, qb, qc, qd));
        }

        MHH_PRAGMA_UNROLL(MHH_ADVEC4_UNROLL_N)
        for (int k = k0; k < k1; ++k)
        {
            const int ijk = i + j * jj + k * kk;
            double t = at[ijk];

            // ---- x direction ----
            if (COMP == 0)
            {
                const double f0 = I4S(u, ijk - 3 * ii, ii), f1 = I4S(u, ijk - 2 * ii, ii);
                const double f2 = I4S(u, ijk - ii, ii), f3 = I4S(u, ijk, ii);
                t -= G4(f0 * f0, f1 * f1, f2 * f2, f3 * f3) * sdxi;
            }
            else
            {
                t -= G4(I4S(u, ijk - ii - 2 * sa, sa)     * I4S(a, ijk - 3 * ii, ii),
                        I4S(u, ijk - 2 * sa, sa)          * I4S(a, ijk - 2 * ii, ii),
                        I4S(u, ijk + ii - 2 * sa, sa)     * I4S(a, ijk - ii, ii),
                        I4S(u, ijk + 2 * ii - 2 * sa, sa) * I4S(a, ijk, ii)) * sdxi;
            }

            // ---- y direction ----
            if (DIM3)
            {
                if (COMP == 1)
                {
                    const double f0 = I4S(v, ijk - 3 * jj, jj), f1 = I4S(v, ijk - 2 * jj, jj);
                    const double f2 = I4S(v, ijk - jj, jj), f3 = I4S(v, ijk, jj);
                    t -= G4(f0 * f0, f1 * f1, f2 * f2, f3 * f3) * sdyi;
                }
                else
                {
                    t -= G4(I4S(v, ijk - jj - 2 * sa, sa)     * I4S(a, ijk - 3 * jj, jj),
                            I4S(v, ijk - 2 * sa, sa)          * I4S(a, ijk - 2 * jj, jj),
                            I4S(v, ijk + jj - 2 * sa, sa)     * I4S(a, ijk - jj, jj),
                            I4S(v, ijk + 2 * jj - 2 * sa, sa) * I4S(a, ijk, jj)) * sdyi;
                }
            }

            // ---- z direction ----
            {
                const double qe = a[ijk + 3 * kk];
                const double A3 = (k == gd.kend - 1) ? TS(qa, qb, qc, qd) : CS(qb, qc, qd, qe);
                const double f3 = face(k + 2, A3);
                t -= G4(f0, f1, f2, f3) * (mz[k] * (1. / 256.));
                f0 = f1; f1 = f2; f2 = f3;
                if (DIFF)
                {
                    const double dxidxi = 1. / (gd.dx * gd.dx), dyidyi = 1. / (gd.dy * gd.dy);
                    t += visc * (cdg3 * a[ijk - 3] + cdg2 * a[ijk - 2] + cdg1 * a[ijk - 1] + cdg0 * qb
                                 + cdg1 * a[ijk + 1] + cdg2 * a[ijk + 2] + cdg3 * a[ijk + 3]) * dxidxi;
                    if (DIM3)
                        t += visc * (cdg3 * a[ijk - 3 * jj] + cdg2 * a[ijk - 2 * jj] + cdg1 * a[ijk - jj] + cdg0 * qb
                                     + cdg1 * a[ijk + jj] + cdg2 * a[ijk + 2 * jj] + cdg3 * a[ijk + 3 * jj]) * dyidyi;
                    // rows of the column as the diffusion sees them (w: Normal-type ghost rows)
                    const bool isw = (COMP == 2);
                    const double e3 = (isw && k == gd.kstart + 2) ? gbot : qm3;
                    const double e2 = (isw && k == gd.kstart + 1) ? gbot : qm2;
                    const double ed = (isw && k == gd.kend - 1) ? gtop : qd;
                    const double ee = (isw && k == gd.kend - 2) ? gtop : qe;
                    const double* __restrict__ min_ = isw ? gd.dzi4 : gd.dzhi4;     // inner / outer metric, diff_w / diff_c
                    const double* __restrict__ mout = isw ? gd.dzhi4 : gd.dzi4;
                    const int ms = isw ? -1 : 0;
                    const double d0 = (k == kfirst) ? (bg0 * e2 + bg1 * qa + bg2 * qb + bg3 * qc) : G4(e3, e2, qa, qb);
                    const double d3 = (k == gd.kend - 1) ? (tg0 * qa + tg1 * qb + tg2 * qc + tg3 * ed) : G4(qb, qc, ed, ee);
                    t += visc * ( cg0 * d0 * min_[k - 1 + ms]
                                + cg1 * G4(e2, qa, qb, qc) * min_[k + ms]
                                + cg2 * G4(qa, qb, qc, ed) * min_[k + 1 + ms]
                                + cg3 * d3 * min_[k + 2 + ms] )
                         * mout[k];
                    qm3 = qm2; qm2 = qa;
                }
                qa = qb; qb = qc; qc = qd; qd = qe;
            }
            at[ijk] = t;
        }
    }

    // advec_s, src/advec_4.cxx:389-497: face velocities are used directly.
    template<bool DIM3>
    __global__ void __launch_bounds__(256, 2)
    advec4_s_kernel(double* __restrict__ st, const double* __restrict__ s, const double* __restrict__ u,
                    const double* __restrict__ v, const double* __restrict__ w, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x + gd.istart;
        const int j = blockIdx.y * blockDim.y + threadIdx.y + gd.jstart;
        if (i >= gd.iend || j >= gd.jend) return;
        const int ii = 1, jj = gd.icells, kk = gd.ijcells;
        const int k0 = gd.kstart + blockIdx.z * KCHUNK;
        const int k1 = min(k0 + KCHUNK, gd.kend);
        const double dxi = 1. / gd.dx, dyi = 1. / gd.dy;

        for (int k = k0; k < k1; ++k)
        {
            const int ijk = i + j * jj + k * kk;
            double t = st[ijk];
            t -= G4(u[ijk - ii] * I4(s, ijk - 3 * ii, ii), u[ijk] * I4(s, ijk - 2 * ii, ii),
                    u[ijk + ii] * I4(s, ijk - ii, ii), u[ijk + 2 * ii] * I4(s, ijk, ii)) * dxi;
            if (DIM3)
                t -= G4(v[ijk - jj] * I4(s, ijk - 3 * jj, jj), v[ijk] * I4(s, ijk - 2 * jj, jj),
                        v[ijk + jj] * I4(s, ijk - jj, jj), v[ijk + 2 * jj] * I4(s, ijk, jj)) * dyi;
            const double s0 = (k == gd.kstart) ? B4(s, ijk - 2 * kk, kk) : I4(s, ijk - 3 * kk, kk);
            const double s3 = (k == gd.kend - 1) ? T4(s, ijk - kk, kk) : I4(s, ijk, kk);
            t -= G4(w[ijk - kk] * s0, w[ijk] * I4(s, ijk - 2 * kk, kk),
                    w[ijk + kk] * I4(s, ijk - kk, kk), w[ijk + 2 * kk] * s3) * gd.dzi4[k];
            st[ijk] = t;
        }
    }

    // calc_cfl, src/advec_4.cxx:51-86 (note: dzi, not dzi4), fused with the max reduction.
    __global__ void __launch_bounds__(256, MHH_MINB)
    advec4_cfl_kernel(const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ w,
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
                const double c = fabs(interp4c(u[ijk - 1], u[ijk], u[ijk + 1], u[ijk + 2])) * dxi
                               + fabs(interp4c(v[ijk - jj], v[ijk], v[ijk + jj], v[ijk + 2 * jj])) * dyi
                               + fabs(interp4c(w[ijk - kk], w[ijk], w[ijk + kk], w[ijk + 2 * kk])) * gd.dzi[k];
                const unsigned long long b = abs_bits(c);
                m = b > m ? b : m;
            }
        block_max_bits(m, out);
    }

    // fuse_diff: u, v, w get Diff_4::diff_c / diff_w in the same pass; the caller then runs the scalars' diff_c.
    template<bool DIM3>
    int exec_t(Ctx* c, const double* u, const double* v, const double* w, double* ut, double* vt, double* wt,
               int nscalars, const double* const* s, double* const* st, bool fuse_diff, double visc)
    {
        const GridD& gd = c->gd;
        Launch3 l = launch_interior(gd.imax, gd.jmax, (gd.kmax + KCHUNK - 1) / KCHUNK);
        Launch3 lw = launch_interior(gd.imax, gd.jmax, (gd.kmax - 1 + KCHUNK - 1) / KCHUNK);
        if (fuse_diff)
        {
            MHH_LAUNCH(c, "advec4_diff4_u_kernel", (advec4_mom_kernel<0, DIM3, true><<<l.grid, l.block, 0, c->stream>>>(ut, u, v, w, visc, gd)));
            MHH_LAUNCH_CHECK();
            MHH_LAUNCH(c, "advec4_diff4_v_kernel", (advec4_mom_kernel<1, DIM3, true><<<l.grid, l.block, 0, c->stream>>>(vt, u, v, w, visc, gd)));
            MHH_LAUNCH_CHECK();
        }
        else
        {
            MHH_LAUNCH(c, "advec4_u_kernel", (advec4_mom_kernel<0, DIM3, false><<<l.grid, l.block, 0, c->stream>>>(ut, u, v, w, 0., gd)));
            MHH_LAUNCH_CHECK();
            MHH_LAUNCH(c, "advec4_v_kernel", (advec4_mom_kernel<1, DIM3, false><<<l.grid, l.block, 0, c->stream>>>(vt, u, v, w, 0., gd)));
            MHH_LAUNCH_CHECK();
        }
        if (fuse_diff) MHH_LAUNCH(c, "advec4_diff4_w_kernel", (advec4_mom_kernel<2, DIM3, true><<<lw.grid, lw.block, 0, c->stream>>>(wt, u, v, w, visc, gd)));
        else           MHH_LAUNCH(c, "advec4_w_kernel", (advec4_mom_kernel<2, DIM3, false><<<lw.grid, lw.block, 0, c->stream>>>(wt, u, v, w, 0., gd)));
        MHH_LAUNCH_CHECK();
        for (int n = 0; n < nscalars; ++n)
        {
            MHH_LAUNCH(c, "advec4_s_kernel", (advec4_s_kernel<DIM3><<<l.grid, l.block, 0, c->stream>>>(st[n], s[n], u, v, w, gd)));
            MHH_LAUNCH_CHECK();
        }
        return MHH_OK;
    }
}  // namespace

static int advec4_check(Ctx* c)
{
    const GridD& gd = c->gd;
    if (gd.igc < 3 || gd.kgc < 3 || (gd.jtot > 1 && gd.jgc < 3) || gd.kmax < 4)
    {
        c->last_error = "advec_4 needs 3 ghost cells in every direction (swspatialorder=4) and ktot >= 4";
        return MHH_ERR_ARG;
    }
    return MHH_OK;
}

int advec4_exec(Ctx* c, const double* u, const double* v, const double* w, double* ut, double* vt, double* wt,
                int nscalars, const double* const* s, double* const* st)
{
    if (int e = advec4_check(c)) return e;
    return c->gd.jtot == 1 ? exec_t<false>(c, u, v, w, ut, vt, wt, nscalars, s, st, false, 0.)
                           : exec_t<true>(c, u, v, w, ut, vt, wt, nscalars, s, st, false, 0.);
}

// Advec_4::exec with Diff_4's diff_c / diff_w of u, v, w in the same pass (the scalars' diff_c follows in capi.cu).
int advec4_exec_diff_uv(Ctx* c, const double* u, const double* v, const double* w, double* ut, double* vt, double* wt,
                        int nscalars, const double* const* s, double* const* st, double visc)
{
    if (int e = advec4_check(c)) return e;
    return c->gd.jtot == 1 ? exec_t<false>(c, u, v, w, ut, vt, wt, nscalars, s, st, true, visc)
                           : exec_t<true>(c, u, v, w, ut, vt, wt, nscalars, s, st, true, visc);
}

int advec4_cfl_bits(Ctx* c, const double* u, const double* v, const double* w, unsigned long long* slot)
{
    const GridD& gd = c->gd;
    Launch3 l = launch_interior(gd.imax, gd.jmax, (gd.kmax + KCHUNK - 1) / KCHUNK);
    MHH_CUDA_CHECK(cudaMemsetAsync(slot, 0, sizeof(unsigned long long), c->stream));
    MHH_LAUNCH(c, "advec4_cfl_kernel", advec4_cfl_kernel<<<l.grid, l.block, 0, c->stream>>>(u, v, w, gd, slot));
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

}  // namespace mhh
