// Buffer<TF>::exec (sponge layer, src/buffer.cxx:37-58,101-125,163-203) and Force<TF>::exec (large-scale forcings,
// src/force.cxx:46-292,586-720) — SURVEY 8(f) rank 3: the remaining per-substep tendency terms of the named LES/DNS
// configurations (drycblles/gabls1: swbuffer=1; gabls1: swlspres=geo; moser180: swlspres=uflux).
//
// All of them are per-level axpy passes: one read-modify-write of the tendency (+ one read of the field where the term
// depends on it), so they are HBM-bound at 24 B/pt (buffer, only above zstart), 16 B/pt (constant / profile sources) and
// 48 B/pt (Coriolis, u and v in one pass).  Per-level scalars (sigma(z), the source profiles) live in small device tables;
// the sponge strength sigma*((z-zstart)/(zsize-zstart))^beta is evaluated ONCE on the host with std::pow, as the CPU
// reference does per level (src/buffer.cxx:49), so the table holds the reference's own values.
#include "sys.h"
#include <cmath>

#include "ctx.h"

namespace mhh
{
using namespace fd;

namespace
{
    constexpr int KCHUNK = MHH_KCHUNK;

    // at -= sigmaz[k]*(a - abuf[k]) for k in [kfirst, kend)   (calc_buffer, src/buffer.cxx:37-58)
    __global__ void __launch_bounds__(256, MHH_MINB)
    buffer_kernel(double* __restrict__ at, const double* __restrict__ a, const double* __restrict__ abuf,
                  const double* __restrict__ sigmaz, const int kfirst, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x + gd.istart;
        const int j = blockIdx.y * blockDim.y + threadIdx.y + gd.jstart;
        if (i >= gd.iend || j >= gd.jend) return;
        const int k0 = kfirst + blockIdx.z * KCHUNK;
        const int k1 = min(k0 + KCHUNK, gd.kend);
        for (int k = k0; k < k1; ++k)
        {
            const int ijk = i + j * gd.icells + k * gd.ijcells;
            at[ijk] -= sigmaz[k] * (a[ijk] - abuf[k]);
        }
    }

    // MODE 0: at += c                       (add_pressure_force, src/force.cxx:46-60)
    // MODE 1: at += prof[k]                 (calc_large_scale_source, :146-161)
    // MODE 2: at += -fac[k]*(mean[k]-ref[k]) (calc_nudging_tendency, :163-183)
    template<int MODE>
    __global__ void __launch_bounds__(256, MHH_MINB)
    level_source_kernel(double* __restrict__ at, const double cst, const double* __restrict__ prof,
                        const double* __restrict__ ref, const double* __restrict__ fac, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x + gd.istart;
        const int j = blockIdx.y * blockDim.y + threadIdx.y + gd.jstart;
        if (i >= gd.iend || j >= gd.jend) return;
        const int k0 = gd.kstart + blockIdx.z * KCHUNK;
        const int k1 = min(k0 + KCHUNK, gd.kend);
        for (int k = k0; k < k1; ++k)
        {
            const int ijk = i + j * gd.icells + k * gd.ijcells;
            double s;
            if (MODE == 0) s = cst;
            else if (MODE == 1) s = prof[k];
            else s = -fac[k] * (prof[k] - ref[k]);
            at[ijk] += s;
        }
    }

    // Coriolis + geostrophic wind (calc_coriolis_2nd / _4th, src/force.cxx:78-144): u and v tendencies in one pass.
    template<int ORDER>
    __global__ void __launch_bounds__(256, MHH_MINB)
    coriolis_kernel(double* __restrict__ ut, double* __restrict__ vt, const double* __restrict__ u,
                    const double* __restrict__ v, const double* __restrict__ ug, const double* __restrict__ vg,
                    const double fc, const double ugrid, const double vgrid, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x + gd.istart;
        const int j = blockIdx.y * blockDim.y + threadIdx.y + gd.jstart;
        if (i >= gd.iend || j >= gd.jend) return;
        const int ii = 1, jj = gd.icells;
        const int k0 = gd.kstart + blockIdx.z * KCHUNK;
        const int k1 = min(k0 + KCHUNK, gd.kend);
        for (int k = k0; k < k1; ++k)
        {
            const int ijk = i + j * jj + k * gd.ijcells;
            if (ORDER == 2)
            {
                ut[ijk] += fc * (0.25 * (v[ijk - ii] + v[ijk] + v[ijk - ii + jj] + v[ijk + jj]) + vgrid - vg[k]);
                vt[ijk] -= fc * (0.25 * (u[ijk - jj] + u[ijk] + u[ijk + ii - jj] + u[ijk + ii]) + ugrid - ug[k]);
            }
            else
            {
                const int ii2 = 2, jj2 = 2 * jj;
                ut[ijk] += fc * ( ( ci0*(ci0*v[ijk-ii2-jj ] + ci1*v[ijk-ii-jj ] + ci2*v[ijk-jj ] + ci3*v[ijk+ii-jj ])
                                  + ci1*(ci0*v[ijk-ii2    ] + ci1*v[ijk-ii    ] + ci2*v[ijk    ] + ci3*v[ijk+ii    ])
                                  + ci2*(ci0*v[ijk-ii2+jj ] + ci1*v[ijk-ii+jj ] + ci2*v[ijk+jj ] + ci3*v[ijk+ii+jj ])
                                  + ci3*(ci0*v[ijk-ii2+jj2] + ci1*v[ijk-ii+jj2] + ci2*v[ijk+jj2] + ci3*v[ijk+ii+jj2]) )
                                + vgrid - vg[k] );
                vt[ijk] -= fc * ( ( ci0*(ci0*u[ijk-ii-jj2] + ci1*u[ijk-jj2] + ci2*u[ijk+ii-jj2] + ci3*u[ijk+ii2-jj2])
                                  + ci1*(ci0*u[ijk-ii-jj ] + ci1*u[ijk-jj ] + ci2*u[ijk+ii-jj ] + ci3*u[ijk+ii2-jj ])
                                  + ci2*(ci0*u[ijk-ii    ] + ci1*u[ijk    ] + ci2*u[ijk+ii    ] + ci3*u[ijk+ii2    ])
                                  + ci3*(ci0*u[ijk-ii+jj ] + ci1*u[ijk+jj ] + ci2*u[ijk+ii+jj ] + ci3*u[ijk+ii2+jj ]) )
                                + ugrid - ug[k] );
            }
        }
    }

    // Large-scale subsidence, upwind in z (src/force.cxx:195-292).
    // MODE 0: advec_wls_2nd_mean (s = mean profile[kcells]); 1: advec_wls_2nd_local; 2: advec_wls_2nd_local_w.
    template<int MODE>
    __global__ void __launch_bounds__(256, MHH_MINB)
    wls_kernel(double* __restrict__ st, const double* __restrict__ s, const double* __restrict__ wls, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x + gd.istart;
        const int j = blockIdx.y * blockDim.y + threadIdx.y + gd.jstart;
        if (i >= gd.iend || j >= gd.jend) return;
        const int kk = gd.ijcells;
        const int kfirst = (MODE == 2) ? gd.kstart + 1 : gd.kstart;
        const int k0 = kfirst + blockIdx.z * KCHUNK;
        const int k1 = min(k0 + KCHUNK, gd.kend);
        for (int k = k0; k < k1; ++k)
        {
            const int ijk = i + j * gd.icells + k * kk;
            if (MODE == 0)
            {
                if (wls[k] > 0.) st[ijk] -= wls[k] * (s[k] - s[k - 1]) * gd.dzhi[k];
                else             st[ijk] -= wls[k] * (s[k + 1] - s[k]) * gd.dzhi[k + 1];
            }
            else if (MODE == 1)
            {
                if (wls[k] > 0.) st[ijk] -= wls[k] * (s[ijk] - s[ijk - kk]) * gd.dzhi[k];
                else             st[ijk] -= wls[k] * (s[ijk + kk] - s[ijk]) * gd.dzhi[k + 1];
            }
            else
            {
                const double wh = interp2(wls[k - 1], wls[k]);
                if (wh > 0.) st[ijk] -= wh * (s[ijk] - s[ijk - kk]) * gd.dzi[k - 1];
                else         st[ijk] -= wh * (s[ijk + kk] - s[ijk]) * gd.dzi[k];
            }
        }
    }
}  // namespace

// Buffer::create (src/buffer.cxx:101-125): first full/half level inside the sponge and the per-level strength tables.
int buffer_plan(Ctx* c, double zstart, double sigma, double beta)
{
    const GridD& gd = c->gd;
    int ks = gd.kstart, ksh = gd.kstart;
    for (int k = gd.kstart; k < gd.kend; ++k)
    {
        if (c->z[k] < zstart) ++ks;
        if (c->zh[k] < zstart) ++ksh;
    }
    if (ksh == gd.kend) { c->last_error = "Buffer is too close to the model top"; return MHH_ERR_ARG; }
    const double zsizebuf = gd.zsize - zstart;
    std::vector<double> h(2 * (size_t)gd.kcells, 0.);
    for (int k = ks; k < gd.kend; ++k)  h[k] = sigma * std::pow((c->z[k] - zstart) / zsizebuf, beta);
    for (int k = ksh; k < gd.kend; ++k) h[gd.kcells + k] = sigma * std::pow((c->zh[k] - zstart) / zsizebuf, beta);
    if (!c->buffer_sigma_d) MHH_CUDA_CHECK(cudaMalloc(&c->buffer_sigma_d, sizeof(double) * 2 * gd.kcells));
    MHH_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    MHH_CUDA_CHECK(cudaMemcpy(c->buffer_sigma_d, h.data(), sizeof(double) * 2 * gd.kcells, cudaMemcpyHostToDevice));
    c->buffer_kstart = ks; c->buffer_kstarth = ksh;
    return MHH_OK;
}

int buffer_exec(Ctx* c, double* at, const double* a, const double* abuf, int is_w)
{
    if (!c->buffer_sigma_d) { c->last_error = "buffer not planned (mhh_buffer_plan)"; return MHH_ERR_STATE; }
    const GridD& gd = c->gd;
    const int kfirst = is_w ? c->buffer_kstarth : c->buffer_kstart;
    const int nk = gd.kend - kfirst;
    if (nk <= 0) return MHH_OK;
    Launch3 l = launch_interior(gd.imax, gd.jmax, (nk + KCHUNK - 1) / KCHUNK);
    MHH_LAUNCH(c, "buffer_kernel", buffer_kernel<<<l.grid, l.block, 0, c->stream>>>(
        at, a, abuf, c->buffer_sigma_d + (is_w ? gd.kcells : 0), kfirst, gd));
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

int force_level_source(Ctx* c, int mode, double* at, double cst, const double* prof, const double* ref, const double* fac)
{
    const GridD& gd = c->gd;
    Launch3 l = launch_interior(gd.imax, gd.jmax, (gd.kmax + KCHUNK - 1) / KCHUNK);
    switch (mode)
    {
        case 0: MHH_LAUNCH(c, "force_const_kernel", level_source_kernel<0><<<l.grid, l.block, 0, c->stream>>>(at, cst, prof, ref, fac, gd)); break;
        case 1: MHH_LAUNCH(c, "force_profile_kernel", level_source_kernel<1><<<l.grid, l.block, 0, c->stream>>>(at, cst, prof, ref, fac, gd)); break;
        case 2: MHH_LAUNCH(c, "force_nudge_kernel", level_source_kernel<2><<<l.grid, l.block, 0, c->stream>>>(at, cst, prof, ref, fac, gd)); break;
        default: c->last_error = "force_level_source: bad mode"; return MHH_ERR_ARG;
    }
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

int force_coriolis(Ctx* c, double* ut, double* vt, const double* u, const double* v, const double* ug, const double* vg,
                   double fc, double ugrid, double vgrid)
{
    const GridD& gd = c->gd;
    Launch3 l = launch_interior(gd.imax, gd.jmax, (gd.kmax + KCHUNK - 1) / KCHUNK);
    if (c->order == 2) MHH_LAUNCH(c, "coriolis_kernel", coriolis_kernel<2><<<l.grid, l.block, 0, c->stream>>>(ut, vt, u, v, ug, vg, fc, ugrid, vgrid, gd));
    else               MHH_LAUNCH(c, "coriolis_kernel", coriolis_kernel<4><<<l.grid, l.block, 0, c->stream>>>(ut, vt, u, v, ug, vg, fc, ugrid, vgrid, gd));
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

int force_wls(Ctx* c, int mode, double* st, const double* s, const double* wls)
{
    const GridD& gd = c->gd;
    const int nk = (mode == 2) ? gd.kmax - 1 : gd.kmax;
    if (nk <= 0) return MHH_OK;
    Launch3 l = launch_interior(gd.imax, gd.jmax, (nk + KCHUNK - 1) / KCHUNK);
    switch (mode)
    {
        case 0: MHH_LAUNCH(c, "wls_mean_kernel", wls_kernel<0><<<l.grid, l.block, 0, c->stream>>>(st, s, wls, gd)); break;
        case 1: MHH_LAUNCH(c, "wls_local_kernel", wls_kernel<1><<<l.grid, l.block, 0, c->stream>>>(st, s, wls, gd)); break;
        case 2: MHH_LAUNCH(c, "wls_local_w_kernel", wls_kernel<2><<<l.grid, l.block, 0, c->stream>>>(st, s, wls, gd)); break;
        default: c->last_error = "force_wls: bad mode"; return MHH_ERR_ARG;
    }
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}
}  // namespace mhh
