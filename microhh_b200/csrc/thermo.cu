// Thermo_dry pieces that sit inside the dynamical-core substep:
// buoyancy tendency on w (src/thermo_dry.cxx:165-198) and N2 = g/thref * dth/dz (:66-79), the
// stability input of Diff_smag2::exec_viscosity (src/diff_smag2.cxx:1143).
#include "sys.h"
#include "ctx.h"

namespace mhh
{
namespace
{
    constexpr int KCHUNK = MHH_KCHUNK;
    constexpr double kGrav = 9.81;   // Constants::grav

    template<int ORDER>
    __global__ void __launch_bounds__(256, MHH_MINB)
    buoyancy_tend_kernel(double* __restrict__ wt, const double* __restrict__ th, const double* __restrict__ threfh,
                         const double* __restrict__ g_threfh, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x + gd.istart;
        const int j = blockIdx.y * blockDim.y + threadIdx.y + gd.jstart;
        if (i >= gd.iend || j >= gd.jend) return;
        const int kk = gd.ijcells;
        const int k0 = max(gd.kstart + (int)blockIdx.z * KCHUNK, gd.kstart + 1);
        const int k1 = min(gd.kstart + ((int)blockIdx.z + 1) * KCHUNK, gd.kend);
        for (int k = k0; k < k1; ++k)
        {
            const int ijk = i + j * gd.icells + k * kk;
            if (ORDER == 2)
                wt[ijk] += g_threfh[k] * (fd::interp2(th[ijk - kk], th[ijk]) - threfh[k]);
            else
                wt[ijk] += g_threfh[k] * (fd::interp4c(th[ijk - 2 * kk], th[ijk - kk], th[ijk], th[ijk + kk]) - threfh[k]);
        }
    }

    __global__ void __launch_bounds__(256, MHH_MINB)
    n2_kernel(double* __restrict__ N2, const double* __restrict__ th, const double* __restrict__ g_thref, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x + gd.istart;
        const int j = blockIdx.y * blockDim.y + threadIdx.y + gd.jstart;
        if (i >= gd.iend || j >= gd.jend) return;
        const int kk = gd.ijcells;
        const int k0 = gd.kstart + blockIdx.z * KCHUNK;
        const int k1 = min(k0 + KCHUNK, gd.kend);
        for (int k = k0; k < k1; ++k)
        {
            const int ijk = i + j * gd.icells + k * kk;
            N2[ijk] = g_thref[k] * 0.5 * (th[ijk + kk] - th[ijk - kk]) * gd.dzi[k];
        }
    }
}  // namespace

int thermo_set_ref(Ctx* c, const double* thref, const double* threfh)
{
    const int kcells = c->gd.kcells;
    // layout: thref | threfh | g/thref | g/threfh.  The two quotient tables are the per-level constants the reference
    // evaluates per point (grav/thref[k], src/thermo_dry.cxx:76,176): same IEEE value, computed once on the host.
    if (!c->thref_d) MHH_CUDA_CHECK(cudaMalloc(&c->thref_d, sizeof(double) * 4 * kcells));
    std::vector<double> h(4 * (size_t)kcells);
    for (int k = 0; k < kcells; ++k)
    {
        h[k] = thref[k]; h[kcells + k] = threfh[k];
        h[2 * kcells + k] = kGrav / thref[k]; h[3 * kcells + k] = kGrav / threfh[k];
    }
    c->thref_ks = thref[c->gd.kstart]; c->threfh_ks = threfh[c->gd.kstart];
    MHH_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    MHH_CUDA_CHECK(cudaMemcpy(c->thref_d, h.data(), sizeof(double) * 4 * kcells, cudaMemcpyHostToDevice));
    return MHH_OK;
}

int thermo_buoyancy_tend(Ctx* c, double* wt, const double* th)
{
    if (!c->thref_d) { c->last_error = "thermo reference profiles not set"; return MHH_ERR_STATE; }
    const GridD& gd = c->gd;
    Launch3 l = launch_interior(gd.imax, gd.jmax, (gd.kmax + KCHUNK - 1) / KCHUNK);
    if (c->order == 2) MHH_LAUNCH(c, "buoyancy_tend_kernel", buoyancy_tend_kernel<2><<<l.grid, l.block, 0, c->stream>>>(wt, th, c->thref_d + gd.kcells, c->thref_d + 3 * gd.kcells, gd));
    else               MHH_LAUNCH(c, "buoyancy_tend_kernel", buoyancy_tend_kernel<4><<<l.grid, l.block, 0, c->stream>>>(wt, th, c->thref_d + gd.kcells, c->thref_d + 3 * gd.kcells, gd));
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

int thermo_n2(Ctx* c, double* N2, const double* th)
{
    if (!c->thref_d) { c->last_error = "thermo reference profiles not set"; return MHH_ERR_STATE; }
    const GridD& gd = c->gd;
    Launch3 l = launch_interior(gd.imax, gd.jmax, (gd.kmax + KCHUNK - 1) / KCHUNK);
    MHH_LAUNCH(c, "n2_kernel", n2_kernel<<<l.grid, l.block, 0, c->stream>>>(N2, th, c->thref_d + 2 * gd.kcells, gd));
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}
}  // namespace mhh
