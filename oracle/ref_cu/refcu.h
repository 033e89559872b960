/* shared by the oracle/ref_cu harness TUs: the grid numbers the reference kernels take and its launch shape
 * (Grid::prepare_device, src/grid.cu:31-32; exec members use (imax/bi, jmax/bj, nk) blocks of bi x bj threads) */
#pragma once
#include <algorithm>
#include <cuda_runtime.h>
struct RefCuGrid { int itot, jtot, ktot, icells, jcells, kcells, ijcells, istart, jstart, kstart, iend, jend, kend, igc, jgc, kgc; double dx, dy, zsize; };
static inline void refcu_shape(const RefCuGrid& g, int nk, dim3& grid, dim3& block)
{
    const int bi = std::min(256, 16 * ((g.itot / 16) + (g.itot % 16 > 0))), bj = 256 / bi;
    block = dim3(bi, bj, 1);
    grid = dim3(g.itot / bi + (g.itot % bi > 0), g.jtot / bj + (g.jtot % bj > 0), nk);
}
#define API extern "C" __attribute__((visibility("default")))
