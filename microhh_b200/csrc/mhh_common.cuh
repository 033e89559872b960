// Shared device/host definitions for the B200-native dynamical core.
//
// Data layout (same as the reference so fields are drop-in): flat, i-fastest,
//   ijk = i + j*icells + k*ijcells, ghost-celled icells x jcells x kcells
// (reference: src/grid.cxx:153-166, src/advec_2.cxx:69).
#pragma once

#include "sys.h"
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

// minimum resident CTAs per SM requested from ptxas for the 256-thread stencil kernels
#ifndef MHH_MINB
#define MHH_MINB 4
#endif
// k levels marched by one thread of the stencil kernels
#ifndef MHH_KCHUNK
#define MHH_KCHUNK 32
#endif
// k levels per CTA of the shared-memory tile kernels: their three-plane pipeline fill is amortised over the chunk
#ifndef MHH_TILE_KCHUNK
#define MHH_TILE_KCHUNK 64
#endif

// `#pragma unroll n` with n a macro: the k marches rotate their register queues by renaming instead of by moves
#define MHH_PRAGMA_STR(x) #x
#define MHH_PRAGMA_UNROLL(n) _Pragma(MHH_PRAGMA_STR(unroll n))

namespace mhh
{

// Grid description passed BY VALUE to kernels (mirrors Grid_data<TF>, include/grid.h:50-135).
struct GridD
{
    int itot, jtot, ktot;
    int imax, jmax, kmax;
    int igc, jgc, kgc;
    int icells, jcells, kcells, ijcells, ncells;
    int istart, iend, jstart, jend, kstart, kend;
    double dx, dy, dxi, dyi;
    double xsize, ysize, zsize;
    double dzhi4bot, dzhi4top;
    // device metric arrays, length kcells
    const double* z;
    const double* zh;
    const double* dz;
    const double* dzh;
    const double* dzi;
    const double* dzhi;
    const double* dzi4;
    const double* dzhi4;
    const double* rhoref;
    const double* rhorefh;
    const double* rhoref_i;    // RN(1/rhoref[k]), RN(1/rhorefh[k]): see div_by()
    const double* rhorefh_i;
};

constexpr double kDsmall = 1.e-9;   // Constants::dsmall, include/constants.h
constexpr double kDhuge  = 1.e30;   // Constants::dhuge, include/constants.h:99
constexpr double kKappa  = 0.4;     // Constants::kappa<TF>

// Error codes returned through the C ABI (0 = ok).
enum : int
{
    MHH_OK = 0,
    MHH_ERR_ARG = 1,
    MHH_ERR_CUDA = 2,
    MHH_ERR_UNSUPPORTED = 3,
    MHH_ERR_STATE = 4
};

#define MHH_CUDA_CHECK(expr)                                                            \
    do {                                                                                \
        cudaError_t err__ = (expr);                                                     \
        if (err__ != cudaSuccess) {                                                     \
            std::fprintf(stderr, "mhh_b200: CUDA error %s at %s:%d\n",                  \
                         cudaGetErrorString(err__), __FILE__, __LINE__);                \
            return MHH_ERR_CUDA;                                                        \
        }                                                                               \
    } while (0)

#define MHH_LAUNCH_CHECK() MHH_CUDA_CHECK(cudaGetLastError())

// Launch shape for "one thread per interior point" kernels: x fastest, 128 threads along i.
struct Launch3
{
    dim3 grid, block;
};

inline Launch3 launch_interior(int ni, int nj, int nk)
{
    Launch3 l;
    const int bx = ni >= 128 ? 128 : (ni >= 64 ? 64 : 32);
    const int by = 256 / bx;
    l.block = dim3(bx, by, 1);
    l.grid = dim3((ni + bx - 1) / bx, (nj + by - 1) / by, nk);
    return l;
}

// ---- max reduction on |x| that PROPAGATES NaN/inf (the reference's CFL guard, src/model.cxx:832-836,
// must stay meaningful): non-negative doubles order like their bit patterns, and NaN patterns sort
// above +inf, so the reduction runs on unsigned 64-bit integers.
__device__ __forceinline__ unsigned long long abs_bits(double x)
{
    return static_cast<unsigned long long>(__double_as_longlong(fabs(x))) & 0x7fffffffffffffffULL;
}

__device__ __forceinline__ void block_max_bits(unsigned long long v, unsigned long long* out)
{
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1)
    {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, v, o);
        v = other > v ? other : v;
    }
    __shared__ unsigned long long warp_max[32];
    const int tid = threadIdx.x + threadIdx.y * blockDim.x + threadIdx.z * blockDim.x * blockDim.y;
    const int nwarps = (blockDim.x * blockDim.y * blockDim.z + 31) >> 5;
    if ((tid & 31) == 0) warp_max[tid >> 5] = v;
    __syncthreads();
    if (tid < 32)
    {
        v = tid < nwarps ? warp_max[tid] : 0ULL;
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1)
        {
            const unsigned long long other = __shfl_xor_sync(0xffffffffu, v, o);
            v = other > v ? other : v;
        }
        if (tid == 0 && v != 0ULL) atomicMax(out, v);
    }
}

// Deterministic block sum (fixed tree order) of doubles — mhh_f64 in the single-precision build too: the reference
// accumulates its means in `double` for TF = float (src/field3d_operators.cxx:53,139); result valid in thread 0.
__device__ __forceinline__ mhh_f64 block_sum(mhh_f64 v)
{
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        v += __shfl_down_sync(0xffffffffu, v, o);
    __shared__ mhh_f64 warp_sum[32];
    const int tid = threadIdx.x + threadIdx.y * blockDim.x;
    const int nwarps = (blockDim.x * blockDim.y + 31) >> 5;
    if ((tid & 31) == 0) warp_sum[tid >> 5] = v;
    __syncthreads();
    if (tid < 32)
    {
        v = tid < nwarps ? warp_sum[tid] : 0.;
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1)
            v += __shfl_down_sync(0xffffffffu, v, o);
    }
    return v;
}

// x / c for a divisor whose correctly rounded reciprocal rc = RN(1/c) is already known (per-level density, tPr):
// q = RN(x*rc); r = x - q*c (exact, FMA); RN(q + r*rc) is the correctly rounded quotient (Markstein 1990) unless the
// mantissa of c is all ones, where it can be 1 ulp off.  3 FP64 ops instead of the ~12 of a full IEEE division:
// ncu showed the density / Prandtl divisions holding the diffusion kernels on the FP64 pipe.
__device__ __forceinline__ double div_by(const double x, const double c, const double rc)
{
    const double q = x * rc;
    const double r = fma(-q, c, x);
    return fma(r, rc, q);
}

// L2 prefetch of the line a later k iteration will need (hides DRAM latency of the k-marching loops:
// ncu showed every stencil kernel stalled on long_scoreboard, profiles/r01a_*).
#ifndef MHH_PF
#define MHH_PF 1
#endif
__device__ __forceinline__ void prefetch_l2(const void* p)
{
#if MHH_PF > 0
    asm volatile("prefetch.global.L2 [%0];" :: "l"(p));
#endif
}

// 1/x to within ~1 ulp for normal, non-zero x: hardware seed (rcp.approx, ~20 bits) + two Newton steps (4 FMAs) instead of
// the ~20-instruction IEEE division sequence.  Used where a quotient feeds an iterative solve, not a compared tendency.
__device__ __forceinline__ double fast_rcp(const double x)
{
#ifdef MHH_F32
    return 1.f / x;
#else
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    double e = fma(-x, r, 1.);
    r = fma(r, e, r);
    e = fma(-x, r, 1.);
    return fma(r, e, r);
#endif
}

// sqrt(x) to within ~1 ulp for normal, positive x: hardware seed (rsqrt.approx) + two Newton steps on 1/sqrt(x) and one
// correction of the root, instead of the IEEE sqrt sequence.
__device__ __forceinline__ double fast_sqrt(const double x)
{
#ifdef MHH_F32
    return sqrtf(x);
#else
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    const double hx = 0.5 * x;
    y = fma(y, fma(-hx * y, y, 0.5), y);          // y += y (1/2 - x y^2 / 2)
    y = fma(y, fma(-hx * y, y, 0.5), y);
    const double s = x * y;
    return fma(0.5 * y, fma(-s, s, x), s);        // s += (x - s^2) y / 2
#endif
}

// ---- asynchronous staging of ghost-celled (i,j) tiles into shared memory ---------------------------------------------
// The k-marching stencil kernels keep a ring of tile planes (interior TX x TY plus a halo of H) in shared memory.  Planes
// are fetched with cp.async (LDGSTS, 8-byte elements: ghost-celled rows are only 8-byte aligned) two levels ahead of
// the level being computed, so the global-memory latency overlaps the arithmetic of two full levels without holding
// the loads in registers; horizontal neighbours then come from shared memory instead of the L1 wavefront path.
__device__ __forceinline__ void cp_async8(double* smem, const double* g)
{
    const unsigned s = static_cast<unsigned>(__cvta_generic_to_shared(smem));
#if MHH_REAL_BYTES == 8
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" :: "r"(s), "l"(g) : "memory");
#else
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" :: "r"(s), "l"(g) : "memory");
#endif
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template<int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" :: "n"(N) : "memory"); }

// ---- mbarrier (sm_90+) helpers for the plane ring: `full[slot]` is completed by the hardware when every thread's cp.async
// copies of a plane have landed (cp.async.mbarrier.arrive.noinc), `empty[slot]` when every thread has finished reading it.
// Both are waited on with try_wait, normally long after they completed: warps of a CTA no longer meet at a
// __syncthreads() every level and may drift apart by up to the ring depth.
__device__ __forceinline__ unsigned smem_u32(const void* p) { return static_cast<unsigned>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(unsigned long long* b, const unsigned count)
{ asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(b)), "r"(count) : "memory"); }
__device__ __forceinline__ void mbar_arrive(unsigned long long* b)
{ asm volatile("{\n .reg .b64 st;\n mbarrier.arrive.shared::cta.b64 st, [%0];\n}" :: "r"(smem_u32(b)) : "memory"); }
// arrival triggered when all cp.async issued so far by this thread have completed (counts as this thread's arrival)
__device__ __forceinline__ void mbar_arrive_on_cp_async(unsigned long long* b)
{ asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" :: "r"(smem_u32(b)) : "memory"); }
__device__ __forceinline__ void mbar_wait(unsigned long long* b, const unsigned parity)
{
    asm volatile("{\n .reg .pred p;\n MBAR_WAIT:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n @p bra MBAR_DONE;\n bra MBAR_WAIT;\n MBAR_DONE:\n}"
                 :: "r"(smem_u32(b)), "r"(parity) : "memory");
}

// NTH: threads that share the copy of one plane (default: one per tile column)
template<int TX, int TY, int H, int NTH = TX * TY>
struct TileMap
{
    static constexpr int W = TX + 2 * H, HT = TY + 2 * H, N = W * HT, NT = NTH, R = (N + NT - 1) / NT;
    int g[R];      // offset of this thread's r-th element within a global (i,j) plane (clamped into the plane); -1: none

    // (i0, j0): first interior point of the tile
    __device__ __forceinline__ void init(const int i0, const int j0, const int icells, const int jcells, const int tid)
    {
        #pragma unroll
        for (int r = 0; r < R; ++r)
        {
            const int e = tid + r * NT;
            const int row = e / W, col = e - row * W;
            const int gi = min(i0 - H + col, icells - 1), gj = min(j0 - H + row, jcells - 1);
            g[r] = (e < N) ? gi + gj * icells : -1;
        }
    }
    // one plane of one field: sm[e] = gplane[g(e)]
    __device__ __forceinline__ void issue(double* sm, const double* __restrict__ gplane, const int tid) const
    {
        #pragma unroll
        for (int r = 0; r < R; ++r)
            if (r + 1 < R || g[r] >= 0) cp_async8(sm + tid + r * NT, gplane + g[r]);
    }
};

// ---- finite-difference helpers (restated from include/finite_difference.h:34-156) ----------
namespace fd
{
    __device__ __forceinline__ double interp2(double a, double b) { return 0.5 * (a + b); }

    // O4 coefficients (include/finite_difference.h:57-90)
    constexpr double ci0 = -1. / 16., ci1 = 9. / 16., ci2 = 9. / 16., ci3 = -1. / 16.;
    constexpr double bi0 = 5. / 16., bi1 = 15. / 16., bi2 = -5. / 16., bi3 = 1. / 16.;
    constexpr double ti0 = 1. / 16., ti1 = -5. / 16., ti2 = 15. / 16., ti3 = 5. / 16.;
    constexpr double cg0 = 1. / 24., cg1 = -27. / 24., cg2 = 27. / 24., cg3 = -1. / 24.;
    constexpr double bg0 = -23. / 24., bg1 = 21. / 24., bg2 = 3. / 24., bg3 = -1. / 24.;
    constexpr double tg0 = 1. / 24., tg1 = -3. / 24., tg2 = -21. / 24., tg3 = 23. / 24.;
    constexpr double cdg0 = -1460. / 576., cdg1 = 783. / 576., cdg2 = -54. / 576., cdg3 = 1. / 576.;

    __device__ __forceinline__ double interp22(double a, double b, double c, double d) { return 0.25 * (a + b + c + d); }
    __device__ __forceinline__ double interp4c(double a, double b, double c, double d)
    { return ci0 * (a + d) + ci1 * (b + c); }
    __device__ __forceinline__ double interp4b(double a, double b, double c, double d)
    { return bi0 * a + bi1 * b - bi2 * c + bi3 * d; }   // sign quirk kept; unused by the path
    __device__ __forceinline__ double interp4t(double a, double b, double c, double d)
    { return ti0 * a + ti1 * b + ti2 * c + ti3 * d; }
    __device__ __forceinline__ double interp4_ws(double a, double b, double c, double d)
    { return (7. / 12.) * (b + c) - (1. / 12.) * (a + d); }
    __device__ __forceinline__ double interp3_ws(double a, double b, double c, double d)
    { return (3. / 12.) * (c - b) - (1. / 12.) * (d - a); }
    __device__ __forceinline__ double grad4(double a, double b, double c, double d)
    { return (-cg0 * (d - a) - cg1 * (c - b)); }
    __device__ __forceinline__ double interp6_ws(double a, double b, double c, double d, double e, double f)
    { return (37. / 60.) * (c + d) - (8. / 60.) * (b + e) + (1. / 60.) * (a + f); }
    __device__ __forceinline__ double interp5_ws(double a, double b, double c, double d, double e, double f)
    { return (10. / 60.) * (d - c) - (5. / 60.) * (e - b) + (1. / 60.) * (f - a); }
}

}  // namespace mhh
