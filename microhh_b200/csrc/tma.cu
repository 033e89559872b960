// Host side of tma.cuh: tensor-map encoding through the driver entry point (no link-time dependency on libcuda).
#include "sys.h"
#include "ctx.h"
#include "tma.cuh"

#include <cudaTypedefs.h>

namespace mhh
{
namespace
{
    PFN_cuTensorMapEncodeTiled_v12000 encode_fn()
    {
        static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
        static bool tried = false;
        if (!tried)
        {
            tried = true;
            void* p = nullptr;
            cudaDriverEntryPointQueryResult q;
            if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
                fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
        }
        return fn;
    }
}

bool tma_planes_supported(const Ctx* c, const int boxx)
{
#ifdef MHH_F32
    return false;     // 4-byte elements: a 34-wide box row is 136 bytes, not a multiple of 16 -> the cp.async path
#endif
    const GridD& gd = c->gd;
    // the box of tile t starts at element t*TX + igc - 1 of a row: kept 16-byte aligned (odd igc, even TX) — with igc = 2 the
    // copies raised "illegal instruction" on B200 (r02f), so such grids take the cp.async path
    return !c->tma_off && (gd.icells % 2 == 0) && (gd.igc % 2 == 1) && (boxx % 2 == 0) && boxx <= 256 && encode_fn() != nullptr;
}

int tma_plane_map(Ctx* c, const double* fld, const int boxx, const int boxy, CUtensorMap* out)
{
    const GridD& gd = c->gd;
    if (!tma_planes_supported(c, boxx) || (reinterpret_cast<uintptr_t>(fld) & 15u) != 0) return MHH_ERR_UNSUPPORTED;
    for (const auto& e : c->tma_cache)
        if (e.fld == fld && e.boxx == boxx && e.boxy == boxy) { *out = e.map; return MHH_OK; }
    const cuuint64_t dims[3] = {(cuuint64_t)gd.icells, (cuuint64_t)gd.jcells, (cuuint64_t)gd.kcells};
    const cuuint64_t strides[2] = {(cuuint64_t)gd.icells * sizeof(double), (cuuint64_t)gd.ijcells * sizeof(double)};
    const cuuint32_t box[3] = {(cuuint32_t)boxx, (cuuint32_t)boxy, 1u};
    const cuuint32_t estr[3] = {1u, 1u, 1u};
    CUtensorMap m;
    const CUresult r = encode_fn()(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, const_cast<double*>(fld), dims, strides, box, estr,
                                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
    {
        c->last_error = "cuTensorMapEncodeTiled failed with code " + std::to_string((int)r);
        return MHH_ERR_CUDA;
    }
    if (c->tma_cache.size() >= 64) c->tma_cache.erase(c->tma_cache.begin());
    c->tma_cache.push_back({fld, boxx, boxy, m});
    *out = m;
    return MHH_OK;
}
}  // namespace mhh
