// Field3d_operators / Fields diagnostics as one-pass reductions: one CTA per k level, warp-shuffle +
// block reduce with a fixed tree (deterministic), FP64 accumulation as in the reference
// (src/field3d_operators.cxx:46-154, src/fields.cxx:144-236).  The reference's GPU path is a 2-3 kernel
// cascade through a tmp field (src/tools.cu:207-302); here each field is read exactly once.
#include "sys.h"
#include "ctx.h"

namespace mhh
{
namespace
{
    enum { OP_SUM = 0, OP_SUM_DZ = 1, OP_MOMENTUM = 2, OP_TKE = 3, OP_MAX = 4 };

    template<int OP>
    __global__ void __launch_bounds__(256)
    level_reduce_kernel(mhh_f64* __restrict__ out, const double* __restrict__ a, const double* __restrict__ b,
                        const double* __restrict__ cfld, const GridD gd, const int kfirst)
    {
        const int k = blockIdx.x + kfirst;
        const int jj = gd.icells, kk = gd.ijcells;
        const int n = gd.imax * gd.jmax;
        mhh_f64 acc = (OP == OP_MAX) ? -1.7976931348623157e308 : 0.;
        for (int t = threadIdx.x; t < n; t += blockDim.x)
        {
            const int i = t % gd.imax + gd.istart;
            const int j = t / gd.imax + gd.jstart;
            const int ijk = i + j * jj + k * kk;
            if (OP == OP_SUM) acc += a[ijk];
            else if (OP == OP_SUM_DZ) acc += a[ijk] * gd.dz[k];
            else if (OP == OP_MOMENTUM)
                acc += (fd::interp2(a[ijk], a[ijk + 1]) + fd::interp2(b[ijk], b[ijk + jj])
                      + fd::interp2(cfld[ijk], cfld[ijk + kk])) * gd.dz[k];
            else if (OP == OP_TKE)
                acc += (fd::interp2(a[ijk] * a[ijk], a[ijk + 1] * a[ijk + 1])
                      + fd::interp2(b[ijk] * b[ijk], b[ijk + jj] * b[ijk + jj])
                      + fd::interp2(cfld[ijk] * cfld[ijk], cfld[ijk + kk] * cfld[ijk + kk])) * gd.dz[k];
            else
            {
                const mhh_f64 x = a[ijk];
                acc = (x > acc || x != x) ? x : acc;
            }
        }
        if (OP == OP_MAX)
        {
            #pragma unroll
            for (int o = 16; o > 0; o >>= 1)
            {
                const mhh_f64 x = __shfl_xor_sync(0xffffffffu, acc, o);
                acc = (x > acc || x != x) ? x : acc;
            }
            __shared__ mhh_f64 wm[8];
            if ((threadIdx.x & 31) == 0) wm[threadIdx.x >> 5] = acc;
            __syncthreads();
            if (threadIdx.x == 0)
            {
                for (int wi = 1; wi < (blockDim.x >> 5); ++wi)
                {
                    const mhh_f64 x = wm[wi];
                    acc = (x > acc || x != x) ? x : acc;
                }
                out[blockIdx.x] = acc;
            }
        }
        else
        {
            const mhh_f64 s = block_sum(acc);
            if (threadIdx.x == 0) out[blockIdx.x] = s;
        }
    }

    __global__ void __launch_bounds__(256)
    plane_sum_kernel(mhh_f64* __restrict__ out, const double* __restrict__ a, const GridD gd)
    {
        const int n = gd.imax * gd.jmax;
        mhh_f64 acc = 0.;
        for (int t = threadIdx.x; t < n; t += blockDim.x)
            acc += a[(t % gd.imax + gd.istart) + (t / gd.imax + gd.jstart) * gd.icells];
        const mhh_f64 s = block_sum(acc);
        if (threadIdx.x == 0) out[0] = s;
    }

    // Stats profiles over the default mask (every point), src/stats.cxx: calc_mean :264 (m1 == NULL, p1 = 1), calc_moment :342
    // (f2 == NULL), calc_cov :368 — per level the sum of pow(f1 - m1[k] + o1, p1) * pow(f2 - m2[k] + o2, p2)
    __global__ void __launch_bounds__(256)
    stats_level_kernel(mhh_f64* __restrict__ out, const double* __restrict__ f1, const double* __restrict__ m1, const double o1,
                       const int p1, const double* __restrict__ f2, const double* __restrict__ m2, const double o2, const int p2,
                       const GridD gd)
    {
        const int k = blockIdx.x + gd.kstart;
        const int n = gd.imax * gd.jmax;
        const double mean1 = m1 ? m1[k] : 0., mean2 = m2 ? m2[k] : 0.;
        mhh_f64 acc = 0.;
        for (int t = threadIdx.x; t < n; t += blockDim.x)
        {
            const int ijk = (t % gd.imax + gd.istart) + (t / gd.imax + gd.jstart) * gd.icells + k * gd.ijcells;
            double term = m1 ? pow(f1[ijk] - mean1 + o1, (double)p1) : f1[ijk];
            if (f2) term *= pow(f2[ijk] - mean2 + o2, (double)p2);
            acc += term;
        }
        const mhh_f64 s = block_sum(acc);
        if (threadIdx.x == 0) out[blockIdx.x] = s;
    }

    template<int OP>
    int run_levels(Ctx* c, const double* a, const double* b, const double* w, int kfirst, int nk)
    {
        MHH_LAUNCH(c, "level_reduce_kernel", level_reduce_kernel<OP><<<nk, 256, 0, c->stream>>>(c->sum_d, a, b, w, c->gd, kfirst));
        MHH_LAUNCH_CHECK();
        MHH_CUDA_CHECK(cudaMemcpyAsync(c->host_pinned, c->sum_d, nk * sizeof(mhh_f64), cudaMemcpyDeviceToHost, c->stream));
        MHH_CUDA_CHECK(cudaStreamSynchronize(c->stream));
        return MHH_OK;
    }
}  // namespace

// local (this rank's) per-level sums; the caller divides / all-reduces.
int level_sums(Ctx* c, const double* fld, double* sums_host)
{
    if (int e = run_levels<OP_SUM>(c, fld, nullptr, nullptr, 0, c->gd.kcells)) return e;
    for (int k = 0; k < c->gd.kcells; ++k) sums_host[k] = c->host_pinned[k];
    return MHH_OK;
}

int weighted_sum(Ctx* c, int op, const double* a, const double* b, const double* w, double* out)
{
    const GridD& gd = c->gd;
    int e;
    if (op == OP_SUM_DZ) e = run_levels<OP_SUM_DZ>(c, a, b, w, gd.kstart, gd.kmax);
    else if (op == OP_MOMENTUM) e = run_levels<OP_MOMENTUM>(c, a, b, w, gd.kstart, gd.kmax);
    else e = run_levels<OP_TKE>(c, a, b, w, gd.kstart, gd.kmax);
    if (e) return e;
    mhh_f64 s = 0.;
    for (int k = 0; k < gd.kmax; ++k) s += c->host_pinned[k];
    *out = s;
    return MHH_OK;
}

int interior_max(Ctx* c, const double* a, double* out)
{
    const GridD& gd = c->gd;
    if (int e = run_levels<OP_MAX>(c, a, nullptr, nullptr, gd.kstart, gd.kmax)) return e;
    mhh_f64 m = c->host_pinned[0];
    for (int k = 1; k < gd.kmax; ++k)
    {
        const mhh_f64 x = c->host_pinned[k];
        m = (x > m || x != x) ? x : m;
    }
    *out = m;
    return MHH_OK;
}

// levels kstart .. kend (inclusive: half-level fields carry one level more, src/stats.cxx:271); this rank's sums divided by
// the GLOBAL itot*jtot, so that the ranks' profiles add up to the statistic (master.sum in Stats::calc_stats)
int stats_profile(Ctx* c, const double* f1, const double* mean1_host, double o1, int p1,
                  const double* f2, const double* mean2_host, double o2, int p2, double* prof_host)
{
    const GridD& gd = c->gd;
    const int nk = gd.kmax + 1;
    if (!c->stats_d) MHH_CUDA_CHECK(cudaMalloc(&c->stats_d, 2 * gd.kcells * sizeof(double)));
    MHH_CUDA_CHECK(cudaStreamSynchronize(c->stream));      // host_pinned is about to be reused
    double* m1 = nullptr; double* m2 = nullptr;
    if (mean1_host)
    {
        MHH_CUDA_CHECK(cudaMemcpyAsync(c->stats_d, mean1_host, gd.kcells * sizeof(double), cudaMemcpyHostToDevice, c->stream));
        MHH_CUDA_CHECK(cudaStreamSynchronize(c->stream));
        m1 = c->stats_d;
    }
    if (f2 && mean2_host)
    {
        MHH_CUDA_CHECK(cudaMemcpyAsync(c->stats_d + gd.kcells, mean2_host, gd.kcells * sizeof(double), cudaMemcpyHostToDevice, c->stream));
        MHH_CUDA_CHECK(cudaStreamSynchronize(c->stream));
        m2 = c->stats_d + gd.kcells;
    }
    MHH_LAUNCH(c, "stats_level_kernel", stats_level_kernel<<<nk, 256, 0, c->stream>>>(c->sum_d, f1, m1, o1, p1, f2, m2, o2, p2, gd));
    MHH_LAUNCH_CHECK();
    MHH_CUDA_CHECK(cudaMemcpyAsync(c->host_pinned, c->sum_d, nk * sizeof(mhh_f64), cudaMemcpyDeviceToHost, c->stream));
    MHH_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    const mhh_f64 n = (mhh_f64)gd.itot * gd.jtot;
    for (int k = 0; k < gd.kcells; ++k) prof_host[k] = 0.;
    for (int k = 0; k < nk; ++k) prof_host[gd.kstart + k] = c->host_pinned[k] / n;
    return MHH_OK;
}

int plane_sum(Ctx* c, const double* a2d, double* out)
{
    MHH_LAUNCH(c, "plane_sum_kernel", plane_sum_kernel<<<1, 256, 0, c->stream>>>(c->sum_d, a2d, c->gd));
    MHH_LAUNCH_CHECK();
    MHH_CUDA_CHECK(cudaMemcpyAsync(c->host_pinned, c->sum_d, sizeof(mhh_f64), cudaMemcpyDeviceToHost, c->stream));
    MHH_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    *out = c->host_pinned[0];
    return MHH_OK;
}

}  // namespace mhh
