// TMA (cp.async.bulk.tensor) staging of ghost-celled (i,j) tile planes for the k-marching stencil kernels.
//
// A field is described to the TMA unit as a rank-3 FP64 tensor (icells, jcells, kcells) in the reference layout
// (ijk = i + j*icells + k*ijcells, src/grid.cxx:153-166).  One elected thread per CTA issues ONE bulk tensor copy per field
// and level — box (TX+2H) x (TY+2H) x 1 starting at (i0-H, j0-H, k) — which lands densely ([row][col]) in shared memory and
// completes the slot's `full` mbarrier through complete_tx; parts of the box outside the array are zero-filled by the
// hardware (they only feed threads that store nothing).  This replaces the per-thread 8-byte cp.async (LDGSTS) copies of
// TileMap and their address arithmetic.  Requirements of the tiled mode: 16-byte aligned base and row pitch (icells even),
// box rows a multiple of 16 bytes (TX + 2H even), box starts on 16-byte boundaries (odd igc), 128-byte aligned shared-memory
// destinations.
#pragma once

#include "sys.h"
#include <cuda.h>

#include "mhh_common.cuh"

namespace mhh
{
struct Ctx;

// up to four fields of one kernel
struct TmaMaps4
{
    CUtensorMap m[4];
};

// Host: encodes (or fetches from the context's cache) the tensor map of `fld` with a boxx x boxy x 1 box.
// Returns MHH_ERR_UNSUPPORTED when the layout does not meet the tiled-mode requirements (the caller then takes the
// cp.async path), MHH_ERR_CUDA when the driver refuses.
int tma_plane_map(Ctx* c, const double* fld, int boxx, int boxy, CUtensorMap* out);
bool tma_planes_supported(const Ctx* c, int boxx);

// doubles per staged plane, padded so that consecutive planes stay 128-byte aligned
__host__ __device__ constexpr int tma_plane_pitch(const int w, const int h) { return ((w * h * 8 + 127) / 128) * 16; }

__device__ __forceinline__ void mbar_expect_tx(unsigned long long* b, const unsigned bytes)
{ asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(b)), "r"(bytes) : "memory"); }

__device__ __forceinline__ void tma_load_plane(double* dst, const CUtensorMap* map, const int x, const int y, const int z,
                                               unsigned long long* bar)
{
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 :: "r"(smem_u32(dst)), "l"(map), "r"(x), "r"(y), "r"(z), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tma_prefetch_map(const CUtensorMap* map)
{ asm volatile("prefetch.tensormap [%0];" :: "l"(map) : "memory"); }
// makes the mbarrier initialisation visible to the async proxy before the first bulk copy signals it
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
}  // namespace mhh
