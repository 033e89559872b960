// Periodic halos (src/boundary_cyclic.cxx:369-443) and vertical ghost rows (src/boundary.cxx:687-907).
// Single-GPU version: in-place copies. East/west first, then north/south including the freshly
// filled east/west ghosts, so corners come out as in the reference.
#include "sys.h"
#include "ctx.h"

namespace mhh
{
namespace
{
    // one thread per (ghost column g, j, k): both sides at once
    __global__ void cyclic_x_kernel(double* __restrict__ a, const GridD gd, const int nk)
    {
        const int n = blockIdx.x * blockDim.x + threadIdx.x;   // over igc * jcells
        const int k = blockIdx.y;
        if (n >= gd.igc * gd.jcells) return;
        const int g = n % gd.igc, j = n / gd.igc;
        const int base = j * gd.icells + k * gd.ijcells;
        const double east = a[base + gd.iend - gd.igc + g], west = a[base + gd.istart + g];   // both loads in flight
        a[base + g] = east;
        a[base + gd.iend + g] = west;
    }

    __global__ void cyclic_y_kernel(double* __restrict__ a, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        const int g = blockIdx.y;   // ghost row
        const int k = blockIdx.z;
        if (i >= gd.icells) return;
        const int base = i + k * gd.ijcells;
        const double north = a[base + (gd.jend - gd.jgc + g) * gd.icells], south = a[base + (gd.jstart + g) * gd.icells];
        a[base + g * gd.icells] = north;
        a[base + (gd.jend + g) * gd.icells] = south;
    }

    // 2-D run (jtot == 1): replicate the single row, interior k only (src/boundary_cyclic.cxx:428-441)
    __global__ void cyclic_y_2drun_kernel(double* __restrict__ a, const GridD gd, const int kfirst)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        const int g = blockIdx.y;
        const int k = blockIdx.z + kfirst;
        if (i >= gd.icells) return;
        const int base = i + k * gd.ijcells;
        const double ref = a[base + gd.jstart * gd.icells];
        a[base + g * gd.icells] = ref;
        a[base + (gd.jend + g) * gd.icells] = ref;
    }

    // ---- vertical ghost rows ------------------------------------------------------------------
    __global__ void ghost_2nd_kernel(double* __restrict__ a, const int bcbot, const int bctop,
                                     double* __restrict__ abot, double* __restrict__ atop,
                                     const double* __restrict__ agradbot, const double* __restrict__ agradtop,
                                     const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        const int j = blockIdx.y;
        if (i >= gd.icells) return;
        const int ij = i + j * gd.icells;
        const int kk = gd.ijcells;
        // bottom, src/boundary.cxx:687-716
        {
            const int ijk = ij + gd.kstart * kk;
            if (bcbot == 0)
                a[ijk - kk] = 2. * abot[ij] - a[ijk];
            else if (bcbot == 1 || bcbot == 2)
                a[ijk - kk] = -agradbot[ij] * gd.dzh[gd.kstart] + a[ijk];
        }
        // top, src/boundary.cxx:719-760
        {
            const int ijk = ij + (gd.kend - 1) * kk;
            if (bctop == 0 || bctop == 4)
            {
                if (bctop == 4)
                    atop[ij] = (3. / 2.) * a[ijk] - (1. / 2.) * a[ijk - kk];
                a[ijk + kk] = 2. * atop[ij] - a[ijk];
            }
            else if (bctop == 1 || bctop == 2)
                a[ijk + kk] = agradtop[ij] * gd.dzh[gd.kend] + a[ijk];
        }
    }

    __global__ void ghost_4th_kernel(double* __restrict__ a, const int bcbot, const int bctop,
                                     const double* __restrict__ abot, const double* __restrict__ atop,
                                     const double* __restrict__ agradbot, const double* __restrict__ agradtop,
                                     const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        const int j = blockIdx.y;
        if (i >= gd.icells) return;
        const int ij = i + j * gd.icells;
        const int kk1 = gd.ijcells, kk2 = 2 * gd.ijcells;
        const double* z = gd.z;
        // bottom, src/boundary.cxx:763-797
        {
            const int ks = gd.kstart;
            const int ijk = ij + ks * kk1;
            if (bcbot == 0)
            {
                a[ijk - kk1] = (8. / 3.) * abot[ij] - 2. * a[ijk] + (1. / 3.) * a[ijk + kk1];
                a[ijk - kk2] = 8. * abot[ij] - 9. * a[ijk] + 2. * a[ijk + kk1];
            }
            else if (bcbot == 1 || bcbot == 2)
            {
                const double g4 = fd::grad4(z[ks - 2], z[ks - 1], z[ks], z[ks + 1]);
                a[ijk - kk1] = -1. * g4 * agradbot[ij] + a[ijk];
                a[ijk - kk2] = -3. * g4 * agradbot[ij] + a[ijk + kk1];
            }
        }
        // top, src/boundary.cxx:800-834
        {
            const int ke = gd.kend;
            const int ijk = ij + (ke - 1) * kk1;
            if (bctop == 0)
            {
                a[ijk + kk1] = (8. / 3.) * atop[ij] - 2. * a[ijk] + (1. / 3.) * a[ijk - kk1];
                a[ijk + kk2] = 8. * atop[ij] - 9. * a[ijk] + 2. * a[ijk - kk1];
            }
            else if (bctop == 1 || bctop == 2)
            {
                const double g4 = fd::grad4(z[ke - 2], z[ke - 1], z[ke], z[ke + 1]);
                a[ijk + kk1] = 1. * g4 * agradtop[ij] + a[ijk];
                a[ijk + kk2] = 3. * g4 * agradtop[ij] + a[ijk - kk1];
            }
        }
    }

    // w ghost rows, 4th order: src/boundary.cxx:838-907
    __global__ void ghost_w_4th_kernel(double* __restrict__ w, const int type, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        const int j = blockIdx.y;
        if (i >= gd.icells) return;
        const int ij = i + j * gd.icells;
        const int kk1 = gd.ijcells, kk2 = 2 * gd.ijcells, kk3 = 3 * gd.ijcells;
        const int b = ij + gd.kstart * kk1;
        const int t = ij + gd.kend * kk1;
        if (type == 0)
        {
            w[b - kk1] = -6. * w[b + kk1] + 4. * w[b + kk2] - w[b + kk3];
            w[t + kk1] = -6. * w[t - kk1] + 4. * w[t - kk2] - w[t - kk3];
        }
        else
        {
            w[b - kk1] = -w[b + kk1];
            w[b - kk2] = -w[b + kk2];
            w[t + kk1] = -w[t - kk1];
            w[t + kk2] = -w[t - kk2];
        }
    }

    // set_bc<TF,true>, src/boundary.cxx:51-91
    __global__ void set_bc_kernel(double* __restrict__ a, double* __restrict__ agrad, double* __restrict__ aflux,
                                  const int sw, const double aval, const double visc, const double offset, const int n)
    {
        const int ij = blockIdx.x * blockDim.x + threadIdx.x;
        if (ij >= n) return;
        if (sw == 0)
            a[ij] = aval - offset;
        else if (sw == 1)
        {
            agrad[ij] = aval;
            aflux[ij] = -aval * visc;
        }
        else if (sw == 2)
        {
            aflux[ij] = aval;
            agrad[ij] = -aval / visc;
        }
    }
}  // namespace

int cyclic_exec_k(Ctx* c, double* fld, int edge, int nk_total, bool is2d)
{
    const GridD& gd = c->gd;
    // On a y-slab rank (npy > 1) only the east/west part is local; the north/south rows are the
    // caller's halo exchange (mhh_halo_ns + NCCL send/recv), done AFTER this call so corners are right.
    if (c->npy != 1)
    {
        if (edge == 1) return MHH_OK;
        edge = 0;
    }
    if (edge == 0 || edge == 2)
    {
        const int n = gd.igc * gd.jcells;
        dim3 b(128), g((n + 127) / 128, nk_total);
        MHH_LAUNCH(c, "cyclic_x_kernel", cyclic_x_kernel<<<g, b, 0, c->stream>>>(fld, gd, nk_total));
        MHH_LAUNCH_CHECK();
    }
    if (edge == 1 || edge == 2)
    {
        dim3 b(128);
        if (gd.jtot > 1)
        {
            dim3 g((gd.icells + 127) / 128, gd.jgc, nk_total);
            MHH_LAUNCH(c, "cyclic_y_kernel", cyclic_y_kernel<<<g, b, 0, c->stream>>>(fld, gd));
        }
        else if (is2d)
        {
            // exec_2d with jtot == 1 (src/boundary_cyclic.cxx:~480): single slice
            dim3 g((gd.icells + 127) / 128, gd.jgc, 1);
            MHH_LAUNCH(c, "cyclic_y_2drun_kernel", cyclic_y_2drun_kernel<<<g, b, 0, c->stream>>>(fld, gd, 0));
        }
        else
        {
            dim3 g((gd.icells + 127) / 128, gd.jgc, gd.kmax);
            MHH_LAUNCH(c, "cyclic_y_2drun_kernel", cyclic_y_2drun_kernel<<<g, b, 0, c->stream>>>(fld, gd, gd.kstart));
        }
        MHH_LAUNCH_CHECK();
    }
    return MHH_OK;
}

int cyclic_exec(Ctx* c, double* fld, int edge) { return cyclic_exec_k(c, fld, edge, c->gd.kcells, false); }
int cyclic_exec_2d(Ctx* c, double* fld) { return cyclic_exec_k(c, fld, 2, 1, true); }

int ghost_cells_exec(Ctx* c, double* a, int bcbot, int bctop, double* abot, double* atop,
                     const double* agradbot, const double* agradtop)
{
    const GridD& gd = c->gd;
    dim3 b(128), g((gd.icells + 127) / 128, gd.jcells);
    if (c->order == 2)
        MHH_LAUNCH(c, "ghost_2nd_kernel", ghost_2nd_kernel<<<g, b, 0, c->stream>>>(a, bcbot, bctop, abot, atop, agradbot, agradtop, gd));
    else
        MHH_LAUNCH(c, "ghost_4th_kernel", ghost_4th_kernel<<<g, b, 0, c->stream>>>(a, bcbot, bctop, abot, atop, agradbot, agradtop, gd));
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

int ghost_cells_w_exec(Ctx* c, double* w, int type)
{
    if (c->order != 4) return MHH_OK;   // src/boundary.cxx:990
    const GridD& gd = c->gd;
    dim3 b(128), g((gd.icells + 127) / 128, gd.jcells);
    MHH_LAUNCH(c, "ghost_w_4th_kernel", ghost_w_4th_kernel<<<g, b, 0, c->stream>>>(w, type, gd));
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

int set_bc_exec(Ctx* c, double* a, double* agrad, double* aflux, int sw, double aval, double visc, double offset)
{
    const int n = c->gd.ijcells;
    MHH_LAUNCH(c, "set_bc_kernel", set_bc_kernel<<<(n + 255) / 256, 256, 0, c->stream>>>(a, agrad, aflux, sw, aval, visc, offset, n));
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

}  // namespace mhh
