// mhh_mean_static: the volume mean of a field (the quantity of mhh_mean — Field3d_operators<TF>::calc_mean,
// src/field3d_operators.cxx:134-154) for callers that must not allocate: code under CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc.  No counterpart in the reference.
//
// The per-level partial sums — the scratch memory that mhh_mean keeps in the context's cudaMalloc'ed sum_d — live in a
// file-scope device array of MHH_STATIC_SCRATCH_LEVELS (4096) FP64 slots = 32 KiB, which every GPU that loads the
// library holds until the process exits.  The entry point calls no cudaMalloc / cudaFree / cudaMallocHost and no
// allocating library routine; input and output are the caller's DEVICE buffers, both kernels go onto the context's
// stream and nothing is waited for.
//
// Sharing: there is ONE array per device (per precision build).  Every call overwrites slots 0 .. kmax-1 before its
// second kernel reads them, so nothing has to start at zero and the previous call's contents are never read; but two
// calls that overlap on the GPU (different contexts / streams on the same device) race on the array — serialise them.
#include "sys.h"
#include "../../include/mhh_b200.h"
#include "ctx.h"

#define MHH_STATIC_SCRATCH_LEVELS 4096
#define MHH_STATIC_MAX_DEVICES 64

namespace mhh
{
namespace
{
    static __device__ __align__(256) mhh_f64 static_scratch[MHH_STATIC_SCRATCH_LEVELS];

    // Host side: `static_scratch` is only a stand-in for the symbol here, never a device address.  The address is looked
    // up once per device (with that device current) by cudaGetSymbolAddress and cached.
    mhh_f64* scratch_addr[MHH_STATIC_MAX_DEVICES] = {};
    std::mutex scratch_mutex;

    int scratch_for_current_device(Ctx* c, mhh_f64** out)
    {
        int dev = -1;
        MHH_CUDA_CHECK(cudaGetDevice(&dev));
        if (dev < 0 || dev >= MHH_STATIC_MAX_DEVICES)
        { c->last_error = "mhh_mean_static: device ordinal beyond the 64 the scratch table holds"; return MHH_ERR_UNSUPPORTED; }
        std::lock_guard<std::mutex> lock(scratch_mutex);
        if (!scratch_addr[dev])
        {
            mhh_f64* p = nullptr;
            const cudaError_t err = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), static_scratch);
            if (err != cudaSuccess || !p)
            {
                c->last_error = std::string("mhh_mean_static: cudaGetSymbolAddress failed: ") + cudaGetErrorString(err);
                return MHH_ERR_CUDA;
            }
            scratch_addr[dev] = p;
        }
        *out = scratch_addr[dev];
        return MHH_OK;
    }

    // one CTA per interior level: sum of fld * dz[k] over the level, the expression and the fixed reduction tree of
    // level_reduce_kernel<OP_SUM_DZ> (reduce.cu)
    __global__ void __launch_bounds__(256)
    level_sum_dz_kernel(mhh_f64* __restrict__ partial, const double* __restrict__ a, const GridD gd)
    {
        const int k = blockIdx.x + gd.kstart;
        const int n = gd.imax * gd.jmax;
        mhh_f64 acc = 0.;
        for (int t = threadIdx.x; t < n; t += blockDim.x)
        {
            const int ijk = (t % gd.imax + gd.istart) + (t / gd.imax + gd.jstart) * gd.icells + k * gd.ijcells;
            acc += a[ijk] * gd.dz[k];
        }
        const mhh_f64 s = block_sum(acc);
        if (threadIdx.x == 0) partial[blockIdx.x] = s;
    }

    // the levels in ascending order, as weighted_sum (reduce.cu) adds them on the host, then the division of mhh_mean
    __global__ void finish_mean_kernel(double* __restrict__ out, const mhh_f64* __restrict__ partial, const int nk, const double denom)
    {
        if (blockIdx.x == 0 && threadIdx.x == 0)
        {
            mhh_f64 s = 0.;
            for (int k = 0; k < nk; ++k) s += partial[k];
            const double sw = s;
            out[0] = sw / denom;
        }
    }
}  // namespace
}  // namespace mhh

extern "C" int mhh_mean_static(mhh_ctx* ctx, const double* fld_dev, double* mean_out_dev)
{
    using namespace mhh;
    if (ctx == nullptr) return MHH_ERR_ARG;
    Ctx* c = ctx;
    if (!fld_dev || !mean_out_dev) { c->last_error = "mhh_mean_static: null field or output pointer"; return MHH_ERR_ARG; }
    const GridD& gd = c->gd;
    if (gd.kmax > MHH_STATIC_SCRATCH_LEVELS)
    {
        c->last_error = "mhh_mean_static: ktot = " + std::to_string(gd.kmax) + " exceeds the static scratch of "
                      + std::to_string(MHH_STATIC_SCRATCH_LEVELS) + " levels (use mhh_mean)";
        return MHH_ERR_ARG;
    }
    mhh_f64* scratch = nullptr;
    if (int e = scratch_for_current_device(c, &scratch)) return e;
    const double denom = gd.itot * gd.jtot * gd.zsize;      // mhh_mean (capi.cu), src/field3d_operators.cxx:152
    // launched directly, not through MHH_LAUNCH: the live-timing scope creates CUDA events when profiling is on
    c->launches += 2;
    level_sum_dz_kernel<<<gd.kmax, 256, 0, c->stream>>>(scratch, fld_dev, gd);
    MHH_LAUNCH_CHECK();
    finish_mean_kernel<<<1, 32, 0, c->stream>>>(mean_out_dev, scratch, gd.kmax, denom);
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}
