// y-slab multi-GPU support kernels (reference semantics: Transpose::exec_xy/exec_yx,
// src/transpose.cxx:163-217, and the north/south part of Boundary_cyclic::exec, src/boundary_cyclic.cxx:115-176).
// The collective itself (NCCL all-to-all / send-recv over NVLink) is issued by the host between these
// pack/unpack kernels; buffers are laid out so each peer's chunk is contiguous.
//
// Layouts (i fastest):  X = [k][j < jmax][i < itot]   (this rank's slab, ghost-free)
//                       Y = [k][j < jtot][ii < iblock] (this rank's block of x wavenumbers, all y)
//                       chunk for peer r: [k][j < jmax][ii < iblock], chunks ordered by r.
#include "sys.h"
#include "ctx.h"

namespace mhh
{
namespace
{
    // mode 0: X -> chunks (send side of exec_xy)     mode 1: chunks -> Y (receive side of exec_xy)
    // mode 2: Y -> chunks (send side of exec_yx)     mode 3: chunks -> X (receive side of exec_yx)
    __global__ void __launch_bounds__(256)
    transpose_copy_kernel(double* __restrict__ dst, const double* __restrict__ src, const int mode,
                          const int itot, const int jmax, const int kmax, const int iblock, const int npy)
    {
        const long n = (long)itot * jmax * kmax;
        const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
        if (t >= n) return;
        // t enumerates the chunk layout: t = ((r*kmax + k)*jmax + j)*iblock + ii
        const int ii = (int)(t % iblock);
        long q = t / iblock;
        const int j = (int)(q % jmax); q /= jmax;
        const int k = (int)(q % kmax);
        const int r = (int)(q / kmax);
        const long x_idx = ((long)k * jmax + j) * itot + (long)r * iblock + ii;
        const long y_idx = ((long)k * (jmax * npy) + (long)r * jmax + j) * iblock + ii;
        if (mode == 0) dst[t] = src[x_idx];
        else if (mode == 1) dst[y_idx] = src[t];
        else if (mode == 2) dst[t] = src[y_idx];
        else dst[x_idx] = src[t];
    }

    // north/south halo rows: buffers are [k < nk][g < jgc][i < icells]
    // pack: to_north <- rows [jend-jgc, jend), to_south <- rows [jstart, jstart+jgc)
    // unpack: rows [0, jgc) <- from_south, rows [jend, jend+jgc) <- from_north
    __global__ void halo_ns_kernel(double* __restrict__ fld, double* __restrict__ north, double* __restrict__ south,
                                   const int unpack, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        const int g = blockIdx.y, k = blockIdx.z;
        if (i >= gd.icells) return;
        const int b = i + (g + k * gd.jgc) * gd.icells;
        const int base = i + k * gd.ijcells;
        if (!unpack)
        {
            north[b] = fld[base + (gd.jend - gd.jgc + g) * gd.icells];
            south[b] = fld[base + (gd.jstart + g) * gd.icells];
        }
        else
        {
            fld[base + g * gd.icells] = south[b];
            fld[base + (gd.jend + g) * gd.icells] = north[b];
        }
    }
}  // namespace

int transpose_copy(Ctx* c, int mode, double* dst, const double* src)
{
    const GridD& gd = c->gd;
    const long n = (long)gd.itot * gd.jmax * gd.kmax;
    const int nb = (int)((n + 255) / 256);
    MHH_LAUNCH(c, mode < 2 ? (mode == 0 ? "transpose_pack_xy" : "transpose_unpack_xy") : (mode == 2 ? "transpose_pack_yx" : "transpose_unpack_yx"),
               transpose_copy_kernel<<<nb, 256, 0, c->stream>>>(dst, src, mode, gd.itot, gd.jmax, gd.kmax, c->iblock, c->npy));
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

int halo_ns(Ctx* c, double* fld, double* north, double* south, int unpack, int nk)
{
    const GridD& gd = c->gd;
    dim3 b(128), g((gd.icells + 127) / 128, gd.jgc, nk);
    MHH_LAUNCH(c, unpack ? "halo_ns_unpack" : "halo_ns_pack", halo_ns_kernel<<<g, b, 0, c->stream>>>(fld, north, south, unpack, gd));
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

}  // namespace mhh
