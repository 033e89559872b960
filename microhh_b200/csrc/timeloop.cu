// Low-storage Runge-Kutta update (src/timeloop.cxx:250-334): a += cB[s]*dt*at ; at *= cA[s+1].
// All prognostic fields go through ONE launch (pointer table in kernel parameters, blockIdx.y = field)
// instead of the reference's launch per field (src/timeloop.cu:138-183).
// CPU semantics are kept: after the last substep EVERY cell of the tendency (ghosts too) is zeroed
// (src/timeloop.cxx:268-273), otherwise only interior tendencies are rescaled.
#include "sys.h"
#include "ctx.h"

namespace mhh
{
namespace
{
    constexpr int MAXF = 16;
    struct FieldTable
    {
        double* a[MAXF];
        double* at[MAXF];
    };

    __global__ void __launch_bounds__(256)
    rk_kernel(const FieldTable tab, const mhh_f64 cB_dt, const double cA_next, const int zero_all, const GridD gd)
    {
        double* __restrict__ a = tab.a[blockIdx.y];
        double* __restrict__ at = tab.at[blockIdx.y];
        const int ijk = blockIdx.x * blockDim.x + threadIdx.x;
        if (ijk >= gd.ncells) return;
        const int i = ijk % gd.icells;
        const int row = ijk / gd.icells;       // j + k*jcells
        const int j = row % gd.jcells, k = row / gd.jcells;
        const bool interior = i >= gd.istart && i < gd.iend && j >= gd.jstart && j < gd.jend
                           && k >= gd.kstart && k < gd.kend;
        if (interior)
        {
            const double t = at[ijk];
            a[ijk] += cB_dt * t;
            at[ijk] = zero_all ? 0. : t * cA_next;
        }
        else if (zero_all)
            at[ijk] = 0.;
    }
}  // namespace

static const mhh_f64 rk3_cA[] = {0., -5. / 9., -153. / 128.};
static const mhh_f64 rk3_cB[] = {1. / 3., 15. / 16., 8. / 15.};
static const mhh_f64 rk4_cA[] = {
    0.,
    -567301805773. / 1357537059087.,
    -2404267990393. / 2016746695238.,
    -3550918686646. / 2091501179385.,
    -1275806237668. / 842570457699.};
static const mhh_f64 rk4_cB[] = {
    1432997174477. / 9575080441755.,
    5161836677717. / 13612068292357.,
    1720146321549. / 2090206949498.,
    3134564353537. / 4481467310338.,
    2277821191437. / 14882151754819.};

mhh_f64 rk_sub_dt(int rkorder, int substep, mhh_f64 dt)
{
    return (rkorder == 3 ? rk3_cB[substep] : rk4_cB[substep]) * dt;
}

// cB[substep]*dt is formed first, then multiplied by at: (cB*dt)*at as in the reference expression
void rk_coefs(int rkorder, int substep, double dt, double* cB_dt, double* cA_next, int* zero_all)
{
    const int nsub = rkorder == 3 ? 3 : 5;
    const int substepn = (substep + 1) % nsub;
    *cB_dt = (rkorder == 3 ? rk3_cB[substep] : rk4_cB[substep]) * dt;
    *cA_next = rkorder == 3 ? rk3_cA[substepn] : rk4_cA[substepn];
    *zero_all = substepn == 0;
}

int rk_substep(Ctx* c, int rkorder, int substep, mhh_f64 dt, int nfields, double* const* a, double* const* at)
{
    if ((rkorder != 3 && rkorder != 4) || substep < 0 || substep >= (rkorder == 3 ? 3 : 5))
        return MHH_ERR_ARG;
    const GridD& gd = c->gd;
    const int nsub = rkorder == 3 ? 3 : 5;
    const int substepn = (substep + 1) % nsub;
    // cB[substep]*dt is formed first, then multiplied by at: (cB*dt)*at as in the reference expression; in double for
    // TF = float too (rk3_g / rk4_g take `double dt`, src/timeloop.cu:35,74)
    const mhh_f64 cB_dt = (rkorder == 3 ? rk3_cB[substep] : rk4_cB[substep]) * dt;
    const double cA_next = rkorder == 3 ? rk3_cA[substepn] : rk4_cA[substepn];
    for (int f0 = 0; f0 < nfields; f0 += MAXF)
    {
        FieldTable tab;
        const int nf = nfields - f0 < MAXF ? nfields - f0 : MAXF;
        for (int f = 0; f < nf; ++f) { tab.a[f] = a[f0 + f]; tab.at[f] = at[f0 + f]; }
        dim3 b(256), g((gd.ncells + 255) / 256, nf);
        MHH_LAUNCH(c, "rk_kernel", rk_kernel<<<g, b, 0, c->stream>>>(tab, cB_dt, cA_next, substepn == 0, gd));
        MHH_LAUNCH_CHECK();
    }
    return MHH_OK;
}

}  // namespace mhh
