// extern "C" surface declared in include/mhh_b200.h.
#include "sys.h"
#include <cmath>
#include <cstring>
#include <new>
#include <string>

#include "../../include/mhh_b200.h"
#include "ctx.h"
#include "fft.h"

namespace mhh
{
// ctx.cu
int prof_collect(Ctx*, std::vector<double>&, std::vector<long>&);
int ctx_init_grid(Ctx*, int, int, int, int, int, int, double, double, double, int, int, int);
int ctx_upload_metrics(Ctx*);
int ctx_alloc_scratch(Ctx*);
void ctx_free(Ctx*);
// advec_2.cu
int advec2_exec(Ctx*, const double*, const double*, const double*, double*, double*, double*, int, const double* const*, double* const*);
int advec2_cfl_bits(Ctx*, const double*, const double*, const double*, unsigned long long*);
// advec_2i5.cu / advec_4.cu
int advec2i5_exec(Ctx*, const double*, const double*, const double*, double*, double*, double*, int, const double* const*, double* const*, const int*);
int advec2i5_cfl_bits(Ctx*, const double*, const double*, const double*, unsigned long long*);
int advec4_exec(Ctx*, const double*, const double*, const double*, double*, double*, double*, int, const double* const*, double* const*);
int advec4_cfl_bits(Ctx*, const double*, const double*, const double*, unsigned long long*);
int advec4_exec_diff_uv(Ctx*, const double*, const double*, const double*, double*, double*, double*, int, const double* const*, double* const*, double);
int diff4_exec_scalars(Ctx*, int, const double* const*, double* const*, const double*);
// diff_2.cu / diff_4.cu / diff_smag2.cu
int diff2_exec(Ctx*, const double*, const double*, const double*, double*, double*, double*, double, int, const double* const*, double* const*, const double*);
int diff4_exec(Ctx*, const double*, const double*, const double*, double*, double*, double*, double, int, const double* const*, double* const*, const double*);
int smag2_exec_viscosity(Ctx*, int, const double*, const double*, const double*, const double*, const double*, const double*,
                         const double*, const double*, double, double, double, double*, int, unsigned long long* = nullptr);
int smag2_exec(Ctx*, int, const double*, const double*, const double*, const double*, double*, double*, double*,
               const double*, const double*, const double*, const double*, double, int, const double* const*, double* const*,
               const double* const*, const double* const*, const double*, double);
int advec2_smag2_exec(Ctx*, int, const double*, const double*, const double*, const double*, double*, double*, double*,
                      const double*, const double*, const double*, const double*, double, const double*, int,
                      const double* const*, double* const*, const double* const*, const double* const*, const double*, double);
int advec2_smag2_exec_rk(Ctx*, int, const double*, const double*, const double*, const double*, double*, double*, double*,
                         const double*, const double*, const double*, const double*, double, const double*, int,
                         const double* const*, double* const*, const double* const*, const double* const*, const double*, double,
                         const double* const*, double* const*, int, int, double);
int advec2_smag2_exec_rk_parts(Ctx*, int, const double*, const double*, const double*, const double*, double*, double*, double*,
                               const double*, const double*, const double*, const double*, double, const double*, int,
                               const double* const*, double* const*, const double* const*, const double* const*, const double*, double,
                               const double* const*, double* const*, int, int, double, int, int, int, int);
int smag2_dnmul_bits(Ctx*, const double*, double, unsigned long long*);
// boundary.cu
int cyclic_exec(Ctx*, double*, int);
int cyclic_exec_2d(Ctx*, double*);
int ghost_cells_exec(Ctx*, double*, int, int, double*, double*, const double*, const double*);
int ghost_cells_w_exec(Ctx*, double*, int);
int set_bc_exec(Ctx*, double*, double*, double*, int, double, double, double);
// timeloop.cu
int rk_substep(Ctx*, int, int, mhh_f64, int, double* const*, double* const*);
mhh_f64 rk_sub_dt(int, int, mhh_f64);
// reduce.cu
int level_sums(Ctx*, const double*, double*);
int weighted_sum(Ctx*, int, const double*, const double*, const double*, double*);
int interior_max(Ctx*, const double*, double*);
int plane_sum(Ctx*, const double*, double*);
int stats_profile(Ctx*, const double*, const double*, double, int, const double*, const double*, double, int, double*);
// thermo.cu
int thermo_set_ref(Ctx*, const double*, const double*);
int thermo_buoyancy_tend(Ctx*, double*, const double*);
int thermo_n2(Ctx*, double*, const double*);
// force.cu
int buffer_plan(Ctx*, double, double, double);
int buffer_exec(Ctx*, double*, const double*, const double*, int);
int force_level_source(Ctx*, int, double*, double, const double*, const double*, const double*);
int force_coriolis(Ctx*, double*, double*, const double*, const double*, const double*, const double*, double, double, double);
int force_wls(Ctx*, int, double*, const double*, const double*);
// tma.cu
bool tma_planes_supported(const Ctx*, int);
// peer.cu
int peer_halo_push(Ctx*, int, double* const*, int);
int peer_barrier(Ctx*);
int peer_error(Ctx*, int*);
#define MHH_TILE_X_PUB 32
// surface.cu
struct SurfaceArgs
{
    int mbcbot, thermo, thermobc;
    const double *u, *v;
    double *ubot, *vbot, *ufluxbot, *vfluxbot, *ugradbot, *vgradbot;
    const double* th;
    double *thbot, *thgradbot, *thfluxbot;
    double *ustar, *obuk;
    int* nobuk;
    const double *z0m, *z0h;
    double *dutot, *dudz, *dvdz, *dbdz;
    int constant_z0 = 1, charnock = 0;
    double alpha_m = 0., alpha_ch = 0., alpha_h = 0.;
};
int surface_lut(Ctx*, int, int, double, double);
int surface_exec(Ctx*, const SurfaceArgs&, int stages);
int surface_scalar(Ctx*, int, const double*, double*, double*, double*, const double*, const double*, const double*);
// pres_2.cu / pres_4.cu
int pres2_set_values(Ctx*);
int pres2_input(Ctx*, const double*, const double*, const double*, const double*, const double*, const double*, double, double*);
int pres2_solve_spectral(Ctx*, double*);
int pres2_output(Ctx*, double*, double*, double*, double*);
int pres2_divergence_bits(Ctx*, const double*, const double*, const double*, unsigned long long*);
int pres_scatter(Ctx*, double*, const double*);
int pres4_set_values(Ctx*);
int pres4_input(Ctx*, const double*, const double*, const double*, const double*, const double*, double*, double, double*);
int pres4_solve_spectral(Ctx*, double*);
int pres4_output(Ctx*, double*, double*, double*, double*, bool with_ghosts = true);
int pres4_output_rk(Ctx*, double*, double*, double*, double*, double*, double*, double*, int, int, double, bool with_ghosts = true);
int pres4_output_ghosts(Ctx*, double*);
int pres4_divergence_bits(Ctx*, const double*, const double*, const double*, unsigned long long*);

static int read_max(Ctx* c, unsigned long long* slot, double* out)
{
    unsigned long long* h = reinterpret_cast<unsigned long long*>(c->host_pinned);
    MHH_CUDA_CHECK(cudaMemcpyAsync(h, slot, sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
    MHH_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    mhh_f64 d;                                   // the slot holds the bit pattern of a double in either build
    std::memcpy(&d, h, sizeof(mhh_f64));
    *out = d;
    return MHH_OK;
}

// Lazy-tendency guard: a tendency array the fused last-substep tails left UNDEFINED (mhh_set_lazy_tendencies) is logically zero.
// Only mhh_advec2_smag2_exec[_rk] / mhh_smag2_exec start from 0 without reading it; every other entry that reads or accumulates
// into a tendency calls this first, which writes the zeros on demand (u,v,w tendencies as a group: the fused pass tests all three).
static int materialize_tend(Ctx* c, const double* p)
{
    if (!c->lazy_tend || p == nullptr) return MHH_OK;
    const size_t bytes = sizeof(double) * c->gd.ncells;
    if (c->fresh_uvw[0] != nullptr && (p == c->fresh_uvw[0] || p == c->fresh_uvw[1] || p == c->fresh_uvw[2]))
    {
        for (auto& fp : c->fresh_uvw)
        {
            MHH_CUDA_CHECK(cudaMemsetAsync(const_cast<double*>(fp), 0, bytes, c->stream));
            fp = nullptr;
        }
    }
    for (auto& fp : c->fresh_s)
        if (fp == p)
        {
            MHH_CUDA_CHECK(cudaMemsetAsync(const_cast<double*>(fp), 0, bytes, c->stream));
            fp = nullptr;
        }
    return MHH_OK;
}
#define MHH_TEND(p) do { if (int e__ = mhh::materialize_tend(c, p)) return e__; } while (0)
#define MHH_TENDS(arr, n) do { if (arr) for (int q__ = 0; q__ < (n); ++q__) MHH_TEND((arr)[q__]); } while (0)

// FFT stages (src/fft.cxx:335-448 serial, :452-583 MPI). x lines live in the slab layout
// X = [k][j<jmax][i<itot]; y lines in Y = [k][j<jtot][ii<iblock] (identical when npy == 1).
static int fft_x(Ctx* c, int dir, double* data, double* p)
{
    const GridD& gd = c->gd;
    FftLayout L;
    L.nlines = gd.jmax * gd.kmax; L.n1 = L.nlines; L.s1 = gd.itot; L.s2 = 0; L.sn = 1; L.base_in = 0; L.base_out = 0;
    L.n1o = L.n1; L.s1o = L.s1; L.s2o = L.s2; L.sno = L.sn;
    L.scale = dir < 0 ? 1. : 1. / gd.itot;
    if (dir > 0 && p != nullptr)
    {
        // scatter straight into the ghost-celled pressure field (replaces src/pres_2.cxx:340-348)
        L.n1o = gd.jmax; L.s1o = gd.icells; L.s2o = gd.ijcells; L.sno = 1;
        L.base_out = gd.igc + (long)gd.jgc * gd.icells + (long)gd.kgc * gd.ijcells;
        return fft_lines(c, c->plan_i, +1, data, p, L);
    }
    return fft_lines(c, c->plan_i, dir, data, data, L);
}

static int fft_y(Ctx* c, int dir, double* data)
{
    const GridD& gd = c->gd;
    FftLayout L;
    L.nlines = c->iblock * gd.kmax; L.n1 = c->iblock; L.s1 = 1; L.s2 = (long)c->iblock * gd.jtot; L.sn = c->iblock;
    L.base_in = 0; L.base_out = 0;
    L.n1o = L.n1; L.s1o = L.s1; L.s2o = L.s2; L.sno = L.sn;
    L.scale = dir < 0 ? 1. : 1. / gd.jtot;
    return fft_lines(c, c->plan_j, dir, data, data, L);
}

// y-slab variants that read / write the all-to-all chunk layout directly (slab.cu: chunk for peer r =
// [k][j < jmax][ii < iblock], chunks ordered by r), so that no pack / unpack pass is needed around the collectives.
//   x, dir < 0: X -> chunks (send side of exec_xy)       x, dir > 0: chunks (received, exec_yx) -> X or p
//   y, dir < 0: chunks (received, exec_xy) -> Y          y, dir > 0: Y -> chunks (send side of exec_yx)
// [k0, k0+nk): the k range handled by this call.  The chunk buffer of a k range is laid out on its own,
// [k0 .. ][r][k' < nk][j < jmax][ii < iblock], so that the all-to-all of one k range can run while the transforms of the
// next are computed (nk = kmax: one collective for the whole slab).
// peers != nullptr (send sides only): peers[r] = rank r's receive buffer as mapped in this process (peer.cu); the transform
// then writes segment r of every line into slot `me` of rank r's buffer instead of into a local chunk buffer.
static void peer_segtab(Ctx* c, FftLayout& L, double* const* peers, long R)
{
    L.use_segtab_out = true;
    for (int r = 0; r < c->npy && r < 8; ++r)
        L.segtab_out[r] = (long)(peers[r] - peers[c->rank_y]) + (long)c->rank_y * R;
}

static int fft_x_chunks(Ctx* c, int dir, const double* in, double* out, double* p, int k0, int nk, double* const* peers = nullptr)
{
    const GridD& gd = c->gd;
    if (nk <= 0) { k0 = 0; nk = gd.kmax; }
    if (k0 < 0 || k0 + nk > gd.kmax) return MHH_ERR_ARG;
    const long R = (long)nk * gd.jmax * c->iblock;            // one peer's chunk of this k range
    const long cbase = (long)k0 * c->npy * gd.jmax * c->iblock, xbase = (long)k0 * gd.jmax * gd.itot;
    FftLayout L;
    L.nlines = gd.jmax * nk;
    L.scale = dir < 0 ? 1. : 1. / gd.itot;
    if (dir < 0)
    {
        L.n1 = L.nlines; L.s1 = gd.itot; L.s2 = 0; L.sn = 1; L.base_in = xbase;
        L.n1o = L.nlines; L.s1o = c->iblock; L.s2o = 0; L.sno = 1; L.esplit_out = c->iblock; L.eseg_out = R; L.base_out = cbase;
        if (peers) { peer_segtab(c, L, peers, R); out = peers[c->rank_y]; }
        return fft_lines(c, c->plan_i, -1, in, out, L);
    }
    L.n1 = L.nlines; L.s1 = c->iblock; L.s2 = 0; L.sn = 1; L.esplit_in = c->iblock; L.eseg_in = R; L.base_in = cbase;
    L.n1o = L.nlines; L.s1o = gd.itot; L.s2o = 0; L.sno = 1; L.base_out = xbase;
    if (p != nullptr)
    {
        L.n1o = gd.jmax; L.s1o = gd.icells; L.s2o = gd.ijcells;
        L.base_out = gd.igc + (long)gd.jgc * gd.icells + (long)(gd.kgc + k0) * gd.ijcells;
        out = p;
    }
    return fft_lines(c, c->plan_i, +1, in, out, L);
}

static int fft_y_chunks(Ctx* c, int dir, const double* in, double* out, int k0, int nk, double* const* peers = nullptr)
{
    const GridD& gd = c->gd;
    if (nk <= 0) { k0 = 0; nk = gd.kmax; }
    if (k0 < 0 || k0 + nk > gd.kmax) return MHH_ERR_ARG;
    const long R = (long)nk * gd.jmax * c->iblock;
    const long cbase = (long)k0 * c->npy * gd.jmax * c->iblock, ybase = (long)k0 * c->iblock * gd.jtot;
    FftLayout L;
    L.nlines = c->iblock * nk;
    L.scale = dir < 0 ? 1. : 1. / gd.jtot;
    // Y side: lines (ii, k), element j at stride iblock;  chunk side: element j = (r, j') at r*R + j'*iblock
    L.n1 = c->iblock; L.s1 = 1; L.sn = c->iblock; L.n1o = c->iblock; L.s1o = 1; L.sno = c->iblock;
    if (dir < 0)
    {
        L.s2 = (long)gd.jmax * c->iblock; L.esplit_in = gd.jmax; L.eseg_in = R; L.base_in = cbase;
        L.s2o = (long)c->iblock * gd.jtot; L.base_out = ybase;
    }
    else
    {
        L.s2 = (long)c->iblock * gd.jtot; L.base_in = ybase;
        L.s2o = (long)gd.jmax * c->iblock; L.esplit_out = gd.jmax; L.eseg_out = R; L.base_out = cbase;
        if (peers) { peer_segtab(c, L, peers, R); out = peers[c->rank_y]; }
    }
    return fft_lines(c, c->plan_j, dir, in, out, L);
}

// ---- the same four transforms on kx SUB-BLOCKS of a rank's block (nsub per rank, ib = iblock / nsub columns each), so that
// the y transforms and the vertical solve of sub-block b run while the copy engines move sub-blocks b+1 (forward) and b-1
// (return): the solve stops being a join between the two all-to-alls.  Segment v = r * nsub + b of a line (rank r's
// sub-block b) is [k][j < jmax][ii < ib], Rb = kmax * jmax * ib elements; in the receive / staging buffers segment v sits at
// v * Rb (v counted over the SOURCE ranks on the receiving side).  Y_b = [k][j < jtot][ii < ib] at b * kmax * jtot * ib.
static int sub_check(Ctx* c, int nsub, int b)
{
    if (nsub < 1 || c->iblock % nsub != 0 || c->npy * nsub > FFT_SEGTAB || b < 0 || b >= nsub)
    { c->last_error = "kx sub-blocks: nsub must divide iblock, ranks * nsub <= 32, 0 <= b < nsub"; return MHH_ERR_ARG; }
    return MHH_OK;
}
static void ptr_segtab(FftLayout& L, double* const* ptrs, int n)
{
    L.use_segtab_out = true;
    for (int v = 0; v < n; ++v) L.segtab_out[v] = (long)(ptrs[v] - ptrs[0]);
}
// X -> segments; segptrs[v], v < ranks * nsub: where segment v of every line goes (staging array, or straight into a buffer)
static int fft_x_fwd_sub(Ctx* c, const double* in, double* const* segptrs, int nsub)
{
    const GridD& gd = c->gd;
    const int ib = c->iblock / nsub;
    FftLayout L;
    L.nlines = gd.jmax * gd.kmax; L.scale = 1.;
    L.n1 = L.nlines; L.s1 = gd.itot; L.s2 = 0; L.sn = 1; L.base_in = 0;
    L.n1o = L.nlines; L.s1o = ib; L.s2o = 0; L.sno = 1; L.esplit_out = ib; L.eseg_out = (long)gd.kmax * gd.jmax * ib; L.base_out = 0;
    ptr_segtab(L, segptrs, c->npy * nsub);
    return fft_lines(c, c->plan_i, -1, in, segptrs[0], L);
}
// received segments of sub-block b (from every rank) -> Y_b
static int fft_y_fwd_sub(Ctx* c, const double* in, double* out, int b, int nsub)
{
    const GridD& gd = c->gd;
    const int ib = c->iblock / nsub;
    const long Rb = (long)gd.kmax * gd.jmax * ib;
    FftLayout L;
    L.nlines = ib * gd.kmax; L.scale = 1.;
    L.n1 = ib; L.s1 = 1; L.sn = ib; L.s2 = (long)gd.jmax * ib; L.esplit_in = gd.jmax; L.eseg_in = (long)nsub * Rb; L.base_in = (long)b * Rb;
    L.n1o = ib; L.s1o = 1; L.sno = ib; L.s2o = (long)ib * gd.jtot; L.base_out = (long)b * gd.kmax * gd.jtot * ib;
    return fft_lines(c, c->plan_j, -1, in, out, L);
}
// Y_b -> segments by destination rank; segptrs[r], r < ranks
static int fft_y_bwd_sub(Ctx* c, const double* in, double* const* segptrs, int b, int nsub)
{
    const GridD& gd = c->gd;
    const int ib = c->iblock / nsub;
    FftLayout L;
    L.nlines = ib * gd.kmax; L.scale = 1. / gd.jtot;
    L.n1 = ib; L.s1 = 1; L.sn = ib; L.s2 = (long)ib * gd.jtot; L.base_in = (long)b * gd.kmax * gd.jtot * ib;
    L.n1o = ib; L.s1o = 1; L.sno = ib; L.s2o = (long)gd.jmax * ib; L.esplit_out = gd.jmax; L.eseg_out = 0; L.base_out = 0;
    ptr_segtab(L, segptrs, c->npy);
    return fft_lines(c, c->plan_j, +1, in, segptrs[0], L);
}
// received segments (all sub-blocks, segment v at v * Rb) -> the ghost-celled p
static int fft_x_bwd_sub(Ctx* c, const double* in, double* p, int nsub)
{
    const GridD& gd = c->gd;
    const int ib = c->iblock / nsub;
    FftLayout L;
    L.nlines = gd.jmax * gd.kmax; L.scale = 1. / gd.itot;
    L.n1 = L.nlines; L.s1 = ib; L.s2 = 0; L.sn = 1; L.esplit_in = ib; L.eseg_in = (long)gd.kmax * gd.jmax * ib; L.base_in = 0;
    L.n1o = gd.jmax; L.s1o = gd.icells; L.s2o = gd.ijcells; L.sno = 1;
    L.base_out = gd.igc + (long)gd.jgc * gd.icells + (long)gd.kgc * gd.ijcells;
    return fft_lines(c, c->plan_i, +1, in, p, L);
}

int pres2_in_fft_x(Ctx*, const FftPlan&, const double*, const double*, const double*, const double*, const double*, const double*,
                   double, double*, const FftLayout&, int);   // fft_pow2.cu

// Pres::input + FFT::exec_forward of the single-GPU pipeline: halos of the tendencies, then either the fused input + x pass
// (2nd order, power-of-two itot) or the two separate stages; then the y pass.  The spectrum lands in c->work_p.
static int fft_y(Ctx* c, int dir, double* data);
static int pres_input_fft_forward(Ctx* c, const double* u, const double* v, const double* w,
                                  double* ut, double* vt, double* wt, mhh_double dt)
{
    const GridD& gd = c->gd;
    if (int e = cyclic_exec(c, ut, MHH_EDGE_EAST_WEST)) return e;       // src/pres_2.cxx:180-181 / src/pres_4.cxx:284-286
    if (int e = cyclic_exec(c, vt, MHH_EDGE_NORTH_SOUTH)) return e;
    int e = MHH_ERR_UNSUPPORTED;
    if (c->order == 2)
    {
        FftLayout L;
        L.nlines = gd.jmax * gd.kmax; L.n1 = L.nlines; L.s1 = gd.itot; L.s2 = 0; L.sn = 1; L.base_in = 0; L.base_out = 0;
        L.n1o = L.n1; L.s1o = L.s1; L.s2o = L.s2; L.sno = L.sn; L.scale = 1.;
        e = pres2_in_fft_x(c, c->plan_i, u, v, w, ut, vt, wt, dt, c->work_p, L, 0);
    }
    if (e == MHH_ERR_UNSUPPORTED)
    {
        e = (c->order == 2) ? pres2_input(c, u, v, w, ut, vt, wt, dt, c->work_p) : pres4_input(c, u, v, w, ut, vt, wt, dt, c->work_p);
        if (e) return e;
        e = fft_x(c, -1, c->work_p, nullptr);
    }
    if (e) return e;
    return fft_y(c, -1, c->work_p);
}

static int fft_forward_inplace(Ctx* c, double* data)
{
    if (c->npy != 1) return MHH_ERR_STATE;
    if (int e = fft_x(c, -1, data, nullptr)) return e;
    return fft_y(c, -1, data);
}

static int fft_backward_impl(Ctx* c, double* data, double* p)
{
    if (c->npy != 1) return MHH_ERR_STATE;
    if (int e = fft_y(c, +1, data)) return e;
    return fft_x(c, +1, data, p);
}
// slab.cu
int transpose_copy(Ctx*, int, double*, const double*);
int halo_ns(Ctx*, double*, double*, double*, int, int);
int pres2_solve_spectral_ws(Ctx*, double*, double*);
int pres2_solve_spectral_sub(Ctx*, double*, double*, int, int);
int pres2_output_stage(Ctx*, int, double*, double*, double*, double*);
int pres2_output_rk(Ctx*, double*, double*, double*, double*, double*, double*, double*, int, int, double);
}  // namespace mhh

using namespace mhh;

#define CTX_CHECK(ctx) do { if ((ctx) == nullptr) return MHH_ERR_ARG; } while (0)

extern "C" {

int mhh_create(const mhh_grid_desc* d, const double* z_host, mhh_ctx** out)
{
    if (!d || !z_host || !out) return MHH_ERR_ARG;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    {
        std::fprintf(stderr, "mhh_b200: no CUDA device: the B200 path has no CPU fallback\n");
        return MHH_ERR_CUDA;
    }
    mhh_ctx* c = new (std::nothrow) mhh_ctx;
    if (!c) return MHH_ERR_ARG;
    int e = ctx_init_grid(c, d->itot, d->jtot, d->ktot, d->igc, d->jgc, d->kgc, d->xsize, d->ysize, d->zsize,
                          d->order, d->npy < 1 ? 1 : d->npy, d->rank_y);
    if (!e) e = ctx_set_metrics(c, z_host);
    if (!e)
    {
        c->rhoref.assign(c->gd.kcells, 1.);
        c->rhorefh.assign(c->gd.kcells, 1.);
        e = ctx_upload_metrics(c);
    }
    if (!e) e = ctx_alloc_scratch(c);
    if (e)
    {
        std::fprintf(stderr, "mhh_create: %s\n", c->last_error.c_str());
        ctx_free(c);
        delete c;
        return e;
    }
    *out = c;
    return MHH_OK;
}

int mhh_destroy(mhh_ctx* c)
{
    CTX_CHECK(c);
    cudaStreamSynchronize(c->stream);
    ctx_free(c);
    delete c;
    return MHH_OK;
}

const char* mhh_last_error(const mhh_ctx* c) { return c ? c->last_error.c_str() : "null context"; }

int mhh_set_stream(mhh_ctx* c, void* s) { CTX_CHECK(c); c->stream = static_cast<cudaStream_t>(s); return MHH_OK; }

int mhh_set_rhoref(mhh_ctx* c, const double* rhoref, const double* rhorefh)
{
    CTX_CHECK(c);
    c->rhoref.assign(rhoref, rhoref + c->gd.kcells);
    c->rhorefh.assign(rhorefh, rhorefh + c->gd.kcells);
    MHH_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    if (int e = ctx_upload_metrics(c)) return e;
    c->pres_ready = false;
    return MHH_OK;
}

int mhh_get_dims(const mhh_ctx* c, int* out)
{
    CTX_CHECK(c);
    const GridD& gd = c->gd;
    const int v[20] = {gd.itot, gd.jtot, gd.ktot, gd.imax, gd.jmax, gd.kmax, gd.igc, gd.jgc, gd.kgc,
                       gd.icells, gd.jcells, gd.kcells, gd.ijcells, gd.ncells,
                       gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend};
    std::memcpy(out, v, sizeof(v));
    return MHH_OK;
}

static const std::vector<double>* metric_host(const mhh_ctx* c, const char* name)
{
    const std::string n(name);
    if (n == "z") return &c->z; if (n == "zh") return &c->zh; if (n == "dz") return &c->dz;
    if (n == "dzh") return &c->dzh; if (n == "dzi") return &c->dzi; if (n == "dzhi") return &c->dzhi;
    if (n == "dzi4") return &c->dzi4; if (n == "dzhi4") return &c->dzhi4;
    if (n == "rhoref") return &c->rhoref; if (n == "rhorefh") return &c->rhorefh;
    return nullptr;
}

int mhh_get_metric(const mhh_ctx* c, const char* name, double* out)
{
    CTX_CHECK(c);
    const std::string n(name);
    if (n == "scalars")
    {
        out[0] = c->gd.dx; out[1] = c->gd.dy; out[2] = c->gd.dxi; out[3] = c->gd.dyi;
        out[4] = c->gd.dzhi4bot; out[5] = c->gd.dzhi4top;
        return MHH_OK;
    }
    const std::vector<double>* v = metric_host(c, name);
    if (!v) return MHH_ERR_ARG;
    std::memcpy(out, v->data(), sizeof(double) * v->size());
    return MHH_OK;
}

const double* mhh_get_metric_device(const mhh_ctx* c, const char* name)
{
    if (!c) return nullptr;
    const std::string n(name);
    const GridD& gd = c->gd;
    if (n == "z") return gd.z; if (n == "zh") return gd.zh; if (n == "dz") return gd.dz;
    if (n == "dzh") return gd.dzh; if (n == "dzi") return gd.dzi; if (n == "dzhi") return gd.dzhi;
    if (n == "dzi4") return gd.dzi4; if (n == "dzhi4") return gd.dzhi4;
    if (n == "rhoref") return gd.rhoref; if (n == "rhorefh") return gd.rhorefh;
    return nullptr;
}

// ---------------------------------------------------------------- launch accounting / live kernel timing
unsigned long long mhh_launch_count(const mhh_ctx* c) { return c ? c->launches : 0ULL; }
int mhh_profile_enable(mhh_ctx* c, int on) { CTX_CHECK(c); c->prof_on = on != 0; return MHH_OK; }
int mhh_profile_report(mhh_ctx* c, char* names, int names_size, mhh_double* ms, long* counts, int max_entries)
{
    CTX_CHECK(c);
    std::vector<double> t; std::vector<long> n;
    const int k = prof_collect(c, t, n);
    std::string joined;
    int written = 0;
    for (int i = 0; i < k && i < max_entries; ++i)
    {
        if ((int)(joined.size() + c->prof_names[i].size() + 2) > names_size) break;
        joined += c->prof_names[i]; joined += ';';
        ms[i] = t[i]; counts[i] = n[i];
        ++written;
    }
    if (names_size > 0) { std::strncpy(names, joined.c_str(), names_size - 1); names[names_size - 1] = 0; }
    return written;   // number of entries (>= 0)
}

// ---------------------------------------------------------------- Advec
int mhh_advec_exec(mhh_ctx* c, int scheme, const double* u, const double* v, const double* w,
                   double* ut, double* vt, double* wt, int ns, const double* const* s, double* const* st,
                   const int* limit_flags)
{
    CTX_CHECK(c);
    MHH_TEND(ut); MHH_TEND(vt); MHH_TEND(wt); MHH_TENDS(st, ns);
    switch (scheme)
    {
        case MHH_ADVEC_2: return advec2_exec(c, u, v, w, ut, vt, wt, ns, s, st);
        case MHH_ADVEC_2I5: return advec2i5_exec(c, u, v, w, ut, vt, wt, ns, s, st, limit_flags);
        case MHH_ADVEC_4: return advec4_exec(c, u, v, w, ut, vt, wt, ns, s, st);
        default: c->last_error = "illegal value for swadvec"; return MHH_ERR_ARG;
    }
}

int mhh_advec_cfl(mhh_ctx* c, int scheme, const double* u, const double* v, const double* w, mhh_double dt, double* cfl)
{
    CTX_CHECK(c);
    int e;
    switch (scheme)
    {
        case MHH_ADVEC_2: e = advec2_cfl_bits(c, u, v, w, c->max_bits_d); break;
        case MHH_ADVEC_2I5: e = advec2i5_cfl_bits(c, u, v, w, c->max_bits_d); break;
        case MHH_ADVEC_4: e = advec4_cfl_bits(c, u, v, w, c->max_bits_d); break;
        default: return MHH_ERR_ARG;
    }
    if (e) return e;
    double m;
    if ((e = read_max(c, c->max_bits_d, &m))) return e;
    *cfl = m * dt;     // cfl = cfl*dt after the global max, src/advec_2.cxx:73-75
    return MHH_OK;
}

int mhh_advec_time_limit(mhh_ctx* c, int scheme, const double* u, const double* v, const double* w,
                         unsigned long idt, mhh_double dt, double cflmax, unsigned long* out)
{
    double cfl;
    if (int e = mhh_advec_cfl(c, scheme, u, v, w, dt, &cfl)) return e;
    const double cflmin = 1.e-5;                 // src/advec.cxx:42-60
    cfl = std::fmax(cflmin, cfl);
    *out = static_cast<unsigned long>(idt * cflmax / cfl);   // src/advec_2.cxx:293
    return MHH_OK;
}

// ---------------------------------------------------------------- Diff
int mhh_diff2_exec(mhh_ctx* c, const double* u, const double* v, const double* w, double* ut, double* vt, double* wt,
                   double visc, int ns, const double* const* s, double* const* st, const double* svisc)
{ CTX_CHECK(c); MHH_TEND(ut); MHH_TEND(vt); MHH_TEND(wt); MHH_TENDS(st, ns); return diff2_exec(c, u, v, w, ut, vt, wt, visc, ns, s, st, svisc); }

int mhh_diff4_exec(mhh_ctx* c, const double* u, const double* v, const double* w, double* ut, double* vt, double* wt,
                   double visc, int ns, const double* const* s, double* const* st, const double* svisc)
{ CTX_CHECK(c); MHH_TEND(ut); MHH_TEND(vt); MHH_TEND(wt); MHH_TENDS(st, ns); return diff4_exec(c, u, v, w, ut, vt, wt, visc, ns, s, st, svisc); }

int mhh_advec4_diff4_exec(mhh_ctx* c, const double* u, const double* v, double* w, double* ut, double* vt, double* wt,
                          double visc, int ns, const double* const* s, double* const* st, const double* svisc)
{
    CTX_CHECK(c);
    if (c->order != 4) { c->last_error = "mhh_advec4_diff4_exec needs swspatialorder=4"; return MHH_ERR_ARG; }
    MHH_TEND(ut); MHH_TEND(vt); MHH_TEND(wt); MHH_TENDS(st, ns);
    // src/model.cxx:387-392: [w ghost rows Conservation-type on entry] advec->exec; set_ghost_cells_w(Normal); diff->exec
    if (int e = advec4_exec_diff_uv(c, u, v, w, ut, vt, wt, ns, s, st, visc)) return e;
    if (int e = ghost_cells_w_exec(c, w, 0)) return e;
    return diff4_exec_scalars(c, ns, s, st, svisc);
}

int mhh_diff_const_dnmul(mhh_ctx* c, double viscmax, double* out)
{
    CTX_CHECK(c);
    const GridD& gd = c->gd;
    double dnmul = 0;
    if (c->order == 2)   // src/diff_2.cxx:147-150
        for (int k = gd.kstart; k < gd.kend; ++k)
            dnmul = std::fmax(dnmul, std::abs(viscmax * (1. / (gd.dx * gd.dx) + 1. / (gd.dy * gd.dy) + 1. / (c->dz[k] * c->dz[k]))));
    else                 // src/diff_4.cxx:230-242
    {
        const double dxidxi = 1. / (gd.dx * gd.dx), dyidyi = 1. / (gd.dy * gd.dy);
        for (int k = gd.kstart; k < gd.kend; ++k)
            dnmul = std::fmax(dnmul, std::abs(viscmax * (dxidxi + dyidyi + 1. / (c->dz[k] * c->dz[k]))));
    }
    *out = dnmul;
    return MHH_OK;
}

int mhh_smag2_exec_viscosity(mhh_ctx* c, int surface_model, const double* u, const double* v, const double* w,
                             const double* N2, const double* dudz, const double* dvdz, const double* dbdz,
                             const double* z0m, double cs, double tPr, double visc, double* evisc)
{
    CTX_CHECK(c);
    if (surface_model && (!dudz || !dvdz || !z0m || (N2 && !dbdz))) return MHH_ERR_ARG;
    return smag2_exec_viscosity(c, surface_model, u, v, w, N2, dudz, dvdz, dbdz, z0m, cs, tPr, visc, evisc, 0);
}

int mhh_smag2_exec_viscosity_dry(mhh_ctx* c, int surface_model, const double* u, const double* v, const double* w,
                                 const double* th, const double* dudz, const double* dvdz, const double* dbdz,
                                 const double* z0m, double cs, double tPr, double visc, double* evisc)
{
    CTX_CHECK(c);
    if (!th) { c->last_error = "th is required"; return MHH_ERR_ARG; }
    return smag2_exec_viscosity(c, surface_model, u, v, w, th, dudz, dvdz, dbdz, z0m, cs, tPr, visc, evisc, 2);
}

// exec_viscosity of the first substep together with the two maxima Model::set_time_step asks for right after it
// (Advec_2::get_time_limit and Diff_smag2::get_time_limit, src/model.cxx:354-357,731-751): the CFL sum and dnmul are
// evaluated in the evisc pass, which already holds u,v,w and the new evisc, instead of two more sweeps over the fields.
int mhh_smag2_exec_viscosity_dry_limits(mhh_ctx* c, int surface_model, const double* u, const double* v, const double* w,
                                        const double* th, const double* dudz, const double* dvdz, const double* dbdz,
                                        const double* z0m, double cs, double tPr, double visc, double* evisc,
                                        double* cfl_over_dt, double* dnmul)
{
    CTX_CHECK(c);
    if (!th || !cfl_over_dt || !dnmul) { c->last_error = "th and the two outputs are required"; return MHH_ERR_ARG; }
    unsigned long long* slots = c->max_bits_d + 4;
    if (int e = smag2_exec_viscosity(c, surface_model, u, v, w, th, dudz, dvdz, dbdz, z0m, cs, tPr, visc, evisc, 2, slots)) return e;
    unsigned long long* h = reinterpret_cast<unsigned long long*>(c->host_pinned);
    MHH_CUDA_CHECK(cudaMemcpyAsync(h, slots, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
    MHH_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    mhh_f64 m[2];                                // bit patterns of doubles in either build
    std::memcpy(m, h, 2 * sizeof(mhh_f64));
    *cfl_over_dt = m[0]; *dnmul = m[1];
    return MHH_OK;
}

int mhh_smag2_strain2(mhh_ctx* c, int surface_model, const double* u, const double* v, const double* w,
                      const double* dudz, const double* dvdz, double* strain2)
{
    CTX_CHECK(c);
    return smag2_exec_viscosity(c, surface_model, u, v, w, nullptr, dudz, dvdz, nullptr, nullptr, 0.1, 1., 0., strain2, 1);
}

int mhh_smag2_exec(mhh_ctx* c, int surface_model, const double* u, const double* v, const double* w, const double* evisc,
                   double* ut, double* vt, double* wt, const double* ufb, const double* uft, const double* vfb, const double* vft,
                   double visc, int ns, const double* const* s, double* const* st,
                   const double* const* sfb, const double* const* sft, const double* svisc, double tPr)
{
    CTX_CHECK(c);
    if (surface_model && (!ufb || !uft || !vfb || !vft)) return MHH_ERR_ARG;
    return smag2_exec(c, surface_model, u, v, w, evisc, ut, vt, wt, ufb, uft, vfb, vft, visc, ns, s, st, sfb, sft, svisc, tPr);
}

// Thermo_dry::exec (th_buoy != NULL) + Advec_2::exec + Diff_smag2::exec, one pass per prognostic field
int mhh_advec2_smag2_exec(mhh_ctx* c, int surface_model, const double* u, const double* v, const double* w, const double* evisc,
                          double* ut, double* vt, double* wt, const double* ufb, const double* uft, const double* vfb, const double* vft,
                          double visc, const double* th_buoy, int ns, const double* const* s, double* const* st,
                          const double* const* sfb, const double* const* sft, const double* svisc, double tPr)
{
    CTX_CHECK(c);
    if (c->order != 2) { c->last_error = "mhh_advec2_smag2_exec needs the 2nd-order grid"; return MHH_ERR_UNSUPPORTED; }
    if (surface_model && (!ufb || !uft || !vfb || !vft)) return MHH_ERR_ARG;
    return advec2_smag2_exec(c, surface_model, u, v, w, evisc, ut, vt, wt, ufb, uft, vfb, vft, visc, th_buoy, ns, s, st, sfb, sft, svisc, tPr);
}

int mhh_advec2_smag2_exec_rk(mhh_ctx* c, int surface_model, const double* u, const double* v, const double* w, const double* evisc,
                             double* ut, double* vt, double* wt, const double* ufb, const double* uft, const double* vfb, const double* vft,
                             double visc, const double* th_buoy, int ns, const double* const* s, double* const* st,
                             const double* const* sfb, const double* const* sft, const double* svisc, double tPr,
                             const double* const* sbuf, double* const* s_new, int rkorder, int substep, mhh_double dt)
{
    CTX_CHECK(c);
    if (c->order != 2) { c->last_error = "mhh_advec2_smag2_exec_rk needs the 2nd-order grid"; return MHH_ERR_UNSUPPORTED; }
    if (surface_model && (!ufb || !uft || !vfb || !vft)) return MHH_ERR_ARG;
    if (s_new && rkorder != 3 && rkorder != 4) { c->last_error = "rkorder must be 3 or 4"; return MHH_ERR_ARG; }
    if (s_new) for (int n = 0; n < ns; ++n) if (s_new[n] && s_new[n] == s[n])
    { c->last_error = "mhh_advec2_smag2_exec_rk: s_new must not alias s (neighbouring tiles still read s)"; return MHH_ERR_ARG; }
    return advec2_smag2_exec_rk(c, surface_model, u, v, w, evisc, ut, vt, wt, ufb, uft, vfb, vft, visc, th_buoy, ns, s, st, sfb, sft,
                                svisc, tPr, sbuf, s_new, rkorder, substep, dt);
}

// mhh_advec2_smag2_exec_rk in parts: the momentum pass (do_uvw) and / or the scalar passes (do_scalars) on the 64-level chunks
// [kc0, kc0 + nkc) of the column (nkc <= 0: to the top).  The scalar passes are independent of the pressure solver, so a
// decomposed caller issues them on a second stream in two halves that run while the copy engines move the solver's
// all-to-all chunks (multigpu.SlabModel).  The parts of one substep together equal one mhh_advec2_smag2_exec_rk call.
int mhh_advec2_smag2_exec_rk_parts(mhh_ctx* c, int surface_model, const double* u, const double* v, const double* w, const double* evisc,
                                   double* ut, double* vt, double* wt, const double* ufb, const double* uft, const double* vfb, const double* vft,
                                   double visc, const double* th_buoy, int ns, const double* const* s, double* const* st,
                                   const double* const* sfb, const double* const* sft, const double* svisc, double tPr,
                                   const double* const* sbuf, double* const* s_new, int rkorder, int substep, mhh_double dt,
                                   int do_uvw, int do_scalars, int kc0, int nkc)
{
    CTX_CHECK(c);
    if (c->order != 2) { c->last_error = "mhh_advec2_smag2_exec_rk_parts needs the 2nd-order grid"; return MHH_ERR_UNSUPPORTED; }
    if (ns < 0 || ns > 8 || (ns > 0 && (!s || !st || !svisc)) || kc0 < 0) return MHH_ERR_ARG;
    if (s_new && rkorder != 3 && rkorder != 4) { c->last_error = "rkorder must be 3 or 4"; return MHH_ERR_ARG; }
    if (s_new) for (int n = 0; n < ns; ++n) if (s_new[n] && s_new[n] == s[n])
    { c->last_error = "mhh_advec2_smag2_exec_rk_parts: s_new must not alias s"; return MHH_ERR_ARG; }
    return advec2_smag2_exec_rk_parts(c, surface_model, u, v, w, evisc, ut, vt, wt, ufb, uft, vfb, vft, visc, th_buoy, ns, s, st, sfb, sft,
                                      svisc, tPr, sbuf, s_new, rkorder, substep, dt, do_uvw, do_scalars, kc0, nkc);
}

// Lazy-tendency mode of the fused 2nd-order LES path.  Timeloop::exec leaves every tendency at zero after the last substep
// (at *= cA with cA = 0, src/timeloop.cxx:250-285) and the next step's first operator reads those zeros back.  With the mode
// on, the last-substep tails (mhh_pres_exec_rk for ut,vt,wt; the scalars of mhh_advec2_smag2_exec_rk) leave the arrays
// UNDEFINED instead, and the next mhh_advec2_smag2_exec[_rk] / mhh_smag2_exec call on the same arrays starts from 0 without
// reading them: 3 + 3 (+ 2 per scalar) field passes less per step, results bit-identical.  Contract: between the end of a step
// and that next call nothing else may read those tendency arrays.
int mhh_set_lazy_tendencies(mhh_ctx* c, int on)
{
    CTX_CHECK(c);
    c->lazy_tend = on != 0;
    if (!on)
    {
        if (c->fresh_uvw[0] != nullptr) { c->last_error = "lazy-tendency mode switched off while tendencies are undefined"; return MHH_ERR_STATE; }
        for (auto fp : c->fresh_s) if (fp != nullptr) { c->last_error = "lazy-tendency mode switched off while tendencies are undefined"; return MHH_ERR_STATE; }
    }
    return MHH_OK;
}

// Buffer::exec (src/buffer.cxx:163) and Force::exec's swlspres=geo branch (src/force.cxx:609-625) for u,v,w inside the fused
// momentum pass of mhh_advec2_smag2_exec[_rk]: armed here, applied by every later call until cleared (all NULL / 0).
int mhh_advec2_smag2_set_force_tail(mhh_ctx* c, const double* ubuf, const double* vbuf, const double* wbuf, int coriolis,
                                    double fc, const double* ug, const double* vg, double ugrid, double vgrid)
{
    CTX_CHECK(c);
    if (c->order != 2) { c->last_error = "mhh_advec2_smag2_set_force_tail needs swspatialorder=2"; return MHH_ERR_ARG; }
    const bool buf = ubuf || vbuf || wbuf;
    if (buf && !(ubuf && vbuf && wbuf)) { c->last_error = "mhh_advec2_smag2_set_force_tail: give all three buffer profiles or none"; return MHH_ERR_ARG; }
    if (buf && !c->buffer_sigma_d) { c->last_error = "buffer not planned (mhh_buffer_plan)"; return MHH_ERR_STATE; }
    if (coriolis && !(ug && vg)) { c->last_error = "mhh_advec2_smag2_set_force_tail: coriolis needs ug and vg"; return MHH_ERR_ARG; }
    auto& t = c->uvw_tail;
    t.buffer = buf; t.ubuf = ubuf; t.vbuf = vbuf; t.wbuf = wbuf;
    t.coriolis = coriolis != 0; t.fc = fc; t.ug = ug; t.vg = vg; t.ugrid = ugrid; t.vgrid = vgrid;
    return MHH_OK;
}

// 0: stage the tile planes with per-thread cp.async (LDGSTS) instead of TMA bulk tensor copies (default on where the layout allows)
int mhh_set_tma(mhh_ctx* c, int on) { CTX_CHECK(c); c->tma_off = on == 0; return MHH_OK; }
int mhh_tma_active(mhh_ctx* c) { CTX_CHECK(c); return tma_planes_supported(c, MHH_TILE_X_PUB + 2) ? 1 : 0; }

int mhh_smag2_dnmul(mhh_ctx* c, const double* evisc, double tPr, double* out)
{
    CTX_CHECK(c);
    if (int e = smag2_dnmul_bits(c, evisc, tPr, c->max_bits_d + 1)) return e;
    return read_max(c, c->max_bits_d + 1, out);
}

// ---------------------------------------------------------------- Thermo_dry (next row)
int mhh_thermo_set_ref(mhh_ctx* c, const double* thref, const double* threfh) { CTX_CHECK(c); return thermo_set_ref(c, thref, threfh); }
int mhh_thermo_dry_exec(mhh_ctx* c, double* wt, const double* th) { CTX_CHECK(c); MHH_TEND(wt); return thermo_buoyancy_tend(c, wt, th); }
int mhh_thermo_dry_N2(mhh_ctx* c, const double* th, double* N2) { CTX_CHECK(c); return thermo_n2(c, N2, th); }

// ---- Boundary_surface ----
int mhh_surface_init_solver(mhh_ctx* c, int mbcbot, int thermobc, double z0m, double z0h)
{ CTX_CHECK(c); return surface_lut(c, mbcbot, thermobc, z0m, z0h); }
int mhh_surface_get_lut(mhh_ctx* c, float* zL, float* f)
{
    CTX_CHECK(c);
    if (c->surf_lut_host.empty()) { c->last_error = "surface solver not prepared (mhh_surface_init_solver)"; return MHH_ERR_STATE; }
    const size_t n = c->surf_lut_host.size() / 2;
    std::memcpy(zL, c->surf_lut_host.data(), n * sizeof(float));
    std::memcpy(f, c->surf_lut_host.data() + n, n * sizeof(float));
    return MHH_OK;
}
int mhh_boundary_surface_exec_stages(mhh_ctx* c, const mhh_surface_args* a, int stages)
{
    CTX_CHECK(c);
    if (!a || !a->u || !a->v || !a->ustar || !a->obuk || !a->dutot || !a->z0m || (a->thermo && (!a->th || !a->nobuk)))
    { c->last_error = "mhh_boundary_surface_exec: null pointer"; return MHH_ERR_ARG; }
    if (stages < 1 || stages > 7) { c->last_error = "mhh_boundary_surface_exec_stages: stages is a mask of 1, 2, 4"; return MHH_ERR_ARG; }
    SurfaceArgs s{a->mbcbot, a->thermo, a->thermobc, a->u, a->v, a->ubot, a->vbot, a->ufluxbot, a->vfluxbot, a->ugradbot,
                  a->vgradbot, a->th, a->thbot, a->thgradbot, a->thfluxbot, a->ustar, a->obuk, a->nobuk, a->z0m, a->z0h,
                  a->dutot, a->dudz, a->dvdz, a->dbdz};
    s.constant_z0 = a->iterative_solver ? 0 : 1; s.charnock = a->charnock;
    s.alpha_m = a->alpha_m; s.alpha_ch = a->alpha_ch; s.alpha_h = a->alpha_h;
    return surface_exec(c, s, stages);
}
int mhh_boundary_surface_exec(mhh_ctx* c, const mhh_surface_args* a) { return mhh_boundary_surface_exec_stages(c, a, 7); }
int mhh_surface_scalar(mhh_ctx* c, int bcbot, const double* var, double* varbot, double* vargradbot, double* varfluxbot,
                       const double* ustar, const double* obuk, const double* z0h)
{ CTX_CHECK(c); return surface_scalar(c, bcbot, var, varbot, vargradbot, varfluxbot, ustar, obuk, z0h); }

// ---- Buffer / Force ----
int mhh_buffer_plan(mhh_ctx* c, double zstart, double sigma, double beta, int* bufferkstart, int* bufferkstarth)
{
    CTX_CHECK(c);
    if (int e = buffer_plan(c, zstart, sigma, beta)) return e;
    if (bufferkstart) *bufferkstart = c->buffer_kstart;
    if (bufferkstarth) *bufferkstarth = c->buffer_kstarth;
    return MHH_OK;
}
int mhh_buffer_exec(mhh_ctx* c, double* at, const double* a, const double* abuf, int is_w)
{ CTX_CHECK(c); if (!at || !a || !abuf) { c->last_error = "mhh_buffer_exec: null pointer"; return MHH_ERR_ARG; } MHH_TEND(at); return buffer_exec(c, at, a, abuf, is_w); }
int mhh_force_pressure(mhh_ctx* c, double* ut, double fbody)
{ CTX_CHECK(c); MHH_TEND(ut); return force_level_source(c, 0, ut, fbody, nullptr, nullptr, nullptr); }
int mhh_force_fixed_flux(mhh_ctx* c, double* ut, const double* u, double uflux, double utrans, mhh_double dt, double* fbody_out)
{
    CTX_CHECK(c);
    // enforce_fixed_flux, src/force.cxx:63-75, with the two calc_mean calls of Force::exec :593-594
    MHH_TEND(ut);
    double u_mean, ut_mean;
    if (int e = mhh_mean(c, u, &u_mean)) return e;
    if (int e = mhh_mean(c, ut, &ut_mean)) return e;
    const double fbody = (uflux - u_mean - utrans) / dt - ut_mean;
    if (fbody_out) *fbody_out = fbody;
    return force_level_source(c, 0, ut, fbody, nullptr, nullptr, nullptr);
}
int mhh_force_coriolis(mhh_ctx* c, double* ut, double* vt, const double* u, const double* v, const double* ug,
                       const double* vg, double fc, double utrans, double vtrans)
{ CTX_CHECK(c); MHH_TEND(ut); MHH_TEND(vt); return force_coriolis(c, ut, vt, u, v, ug, vg, fc, utrans, vtrans); }
int mhh_force_large_scale_source(mhh_ctx* c, double* st, const double* sls)
{ CTX_CHECK(c); MHH_TEND(st); return force_level_source(c, 1, st, 0., sls, nullptr, nullptr); }
int mhh_force_nudge(mhh_ctx* c, double* at, const double* fldmean, const double* ref, const double* factor)
{ CTX_CHECK(c); MHH_TEND(at); return force_level_source(c, 2, at, 0., fldmean, ref, factor); }
int mhh_force_subsidence(mhh_ctx* c, int mode, double* st, const double* s, const double* wls)
{ CTX_CHECK(c); MHH_TEND(st); return force_wls(c, mode, st, s, wls); }

// ---------------------------------------------------------------- Pres
int mhh_pres_set_values(mhh_ctx* c)
{
    CTX_CHECK(c);
    const GridD& gd = c->gd;
    int e = (c->order == 2) ? pres2_set_values(c) : pres4_set_values(c);
    if (e) return e;
    if (!c->plan_i.tw) { if ((e = fft_plan_create(&c->plan_i, gd.itot))) { c->last_error = "itot has a prime factor > 13"; return e; } }
    if (!c->plan_j.tw) { if ((e = fft_plan_create(&c->plan_j, gd.jtot))) { c->last_error = "jtot has a prime factor > 13"; return e; } }
    auto upload = [&](double** d, const std::vector<double>& h) -> int
    {
        if (h.empty()) return MHH_OK;
        if (!*d) MHH_CUDA_CHECK(cudaMalloc(d, sizeof(double) * h.size()));
        MHH_CUDA_CHECK(cudaMemcpy(*d, h.data(), sizeof(double) * h.size(), cudaMemcpyHostToDevice));
        return MHH_OK;
    };
    if ((e = upload(&c->bmati_d, c->bmati))) return e;
    if ((e = upload(&c->bmatj_d, c->bmatj))) return e;
    if ((e = upload(&c->pres_a_d, c->pres_a))) return e;
    if ((e = upload(&c->pres_c_d, c->pres_c))) return e;
    if ((e = upload(&c->bands_d, c->bands))) return e;
    const size_t n = (size_t)gd.imax * gd.jmax * gd.kmax;
    if (!c->work_p && c->npy == 1)
    {
        MHH_CUDA_CHECK(cudaMalloc(&c->work_p, sizeof(double) * n));
        MHH_CUDA_CHECK(cudaMalloc(&c->work_c, sizeof(double) * n));
        c->work_elems = n;
    }
    c->pres_ready = true;
    return MHH_OK;
}

double* mhh_pres_work(mhh_ctx* c) { return c ? c->work_p : nullptr; }

#define PRES_READY(c) do { CTX_CHECK(c); if (!(c)->pres_ready) { if (int e__ = mhh_pres_set_values(c)) return e__; } } while (0)

int mhh_pres_input(mhh_ctx* c, const double* u, const double* v, const double* w, double* ut, double* vt, double* wt, mhh_double dt)
{
    PRES_READY(c);
    MHH_TEND(ut); MHH_TEND(vt); MHH_TEND(wt);
    // cyclic BCs of the tendencies, src/pres_2.cxx:180-181 / src/pres_4.cxx:284-286
    if (c->npy == 1)
    {
        if (int e = cyclic_exec(c, ut, MHH_EDGE_EAST_WEST)) return e;
        if (int e = cyclic_exec(c, vt, MHH_EDGE_NORTH_SOUTH)) return e;
    }
    if (c->order == 2) return pres2_input(c, u, v, w, ut, vt, wt, dt, c->work_p);
    return pres4_input(c, u, v, w, ut, vt, wt, dt, c->work_p);
}

// Pres_2::input into a caller-supplied ghost-free array (y-slab driver); halos of ut/vt are the caller's.
int mhh_pres_input_to(mhh_ctx* c, const double* u, const double* v, const double* w, const double* ut, const double* vt,
                      const double* wt, mhh_double dt, double* dst)
{
    PRES_READY(c);
    MHH_TEND(ut); MHH_TEND(vt); MHH_TEND(wt);
    // Pres_4::input also mirrors wt's ghost rows (src/pres_4.cxx:290-303): wt is written there
    if (c->order == 4) return pres4_input(c, u, v, w, ut, vt, const_cast<double*>(wt), dt, dst);
    return pres2_input(c, u, v, w, ut, vt, wt, dt, dst);
}
// Off by default: on B200 the fused pass (1.70 ms at 512^3) only equals pres2_in (1.16) + fft_fwd_x (0.53); it moves 16 B/pt less.
int mhh_pres_set_fused_input(mhh_ctx* c, int on) { CTX_CHECK(c); c->fuse_pres_input = on != 0; return MHH_OK; }
int mhh_fft_force_generic(mhh_ctx* c, int on) { CTX_CHECK(c); c->fft_force_generic = on != 0; return MHH_OK; }
int mhh_fft_forward(mhh_ctx* c, double* data) { PRES_READY(c); return fft_forward_inplace(c, data); }
int mhh_fft_backward(mhh_ctx* c, double* data) { PRES_READY(c); return fft_backward_impl(c, data, nullptr); }

int mhh_pres_solve_spectral(mhh_ctx* c, double* data)
{
    PRES_READY(c);
    return c->order == 2 ? pres2_solve_spectral(c, data) : pres4_solve_spectral(c, data);
}

int mhh_pres_output(mhh_ctx* c, double* ut, double* vt, double* wt, double* p)
{
    PRES_READY(c);
    return c->order == 2 ? pres2_output(c, ut, vt, wt, p) : pres4_output(c, ut, vt, wt, p);
}

// ---- y-slab stages (multi-GPU driver): see microhh_b200/multigpu.py for the call order
int mhh_fft_x(mhh_ctx* c, int dir, double* data, double* p_or_null) { PRES_READY(c); return fft_x(c, dir, data, p_or_null); }
int mhh_fft_y(mhh_ctx* c, int dir, double* data) { PRES_READY(c); return fft_y(c, dir, data); }
int mhh_fft_x_chunks(mhh_ctx* c, int dir, const double* in, double* out, double* p_or_null, int k0, int nk)
{ PRES_READY(c); return fft_x_chunks(c, dir, in, out, p_or_null, k0, nk); }
int mhh_fft_y_chunks(mhh_ctx* c, int dir, const double* in, double* out, int k0, int nk)
{ PRES_READY(c); return fft_y_chunks(c, dir, in, out, k0, nk); }
// 1 when the chunk-layout FFT stages are available (both lengths on the register-resident power-of-two kernel and the
// per-rank blocks powers of two); otherwise the driver falls back to mhh_transpose_copy around mhh_fft_x / mhh_fft_y
int mhh_fft_chunks_supported(mhh_ctx* c)
{
    if (!c) return 0;
    if (!c->pres_ready) { if (mhh_pres_set_values(c)) return 0; }
    const int ib = c->iblock, jm = c->gd.jmax;
    return (c->plan_i.p2_nst > 0 && c->plan_j.p2_nst > 0 && ib > 0 && jm > 0 && (ib & (ib - 1)) == 0 && (jm & (jm - 1)) == 0) ? 1 : 0;
}
int mhh_transpose_copy(mhh_ctx* c, int mode, double* dst, const double* src) { CTX_CHECK(c); return transpose_copy(c, mode, dst, src); }
int mhh_halo_ns(mhh_ctx* c, double* fld, double* north, double* south, int unpack, int is2d)
{ CTX_CHECK(c); return halo_ns(c, fld, north, south, unpack, is2d ? 1 : c->gd.kcells); }
int mhh_pres_solve_spectral_ws(mhh_ctx* c, double* data, double* work3d)
{
    PRES_READY(c);
    if (c->order == 4) return pres4_solve_spectral(c, data);      // the 7-band factors live in the context's own scratch
    return pres2_solve_spectral_ws(c, data, work3d);
}
int mhh_slab_fft_x_fwd_sub(mhh_ctx* c, const double* in, double* const* segptrs, int nsub)
{ PRES_READY(c); if (!in || !segptrs) return MHH_ERR_ARG; if (int e = sub_check(c, nsub, 0)) return e; return fft_x_fwd_sub(c, in, segptrs, nsub); }
int mhh_slab_fft_y_fwd_sub(mhh_ctx* c, const double* in, double* out_Y, int b, int nsub)
{ PRES_READY(c); if (!in || !out_Y) return MHH_ERR_ARG; if (int e = sub_check(c, nsub, b)) return e; return fft_y_fwd_sub(c, in, out_Y, b, nsub); }
int mhh_slab_solve_sub(mhh_ctx* c, double* Y, double* work, int b, int nsub)
{
    PRES_READY(c);
    if (!Y || !work) return MHH_ERR_ARG;
    if (int e = sub_check(c, nsub, b)) return e;
    if (c->order != 2) { c->last_error = "mhh_slab_solve_sub: Pres_2 only"; return MHH_ERR_UNSUPPORTED; }
    return pres2_solve_spectral_sub(c, Y, work, b, nsub);
}
int mhh_slab_fft_y_bwd_sub(mhh_ctx* c, const double* in_Y, double* const* segptrs, int b, int nsub)
{ PRES_READY(c); if (!in_Y || !segptrs) return MHH_ERR_ARG; if (int e = sub_check(c, nsub, b)) return e; return fft_y_bwd_sub(c, in_Y, segptrs, b, nsub); }
int mhh_slab_fft_x_bwd_sub(mhh_ctx* c, const double* in, double* p, int nsub)
{ PRES_READY(c); if (!in || !p) return MHH_ERR_ARG; if (int e = sub_check(c, nsub, 0)) return e; return fft_x_bwd_sub(c, in, p, nsub); }
int mhh_pres_output_stage(mhh_ctx* c, int stage, double* ut, double* vt, double* wt, double* p)
{
    PRES_READY(c);
    if (c->order == 4) return stage == 0 ? pres4_output_ghosts(c, p) : pres4_output(c, ut, vt, wt, p, false);
    return pres2_output_stage(c, stage, ut, vt, wt, p);
}

int mhh_pres_scatter(mhh_ctx* c, double* p, const double* src) { CTX_CHECK(c); return pres_scatter(c, p, src); }

int mhh_pres_exec(mhh_ctx* c, const double* u, const double* v, const double* w,
                  double* ut, double* vt, double* wt, double* p, mhh_double dt)
{
    PRES_READY(c);
    if (c->npy != 1) { c->last_error = "mhh_pres_exec is the single-GPU pipeline; y-slab ranks drive the stages"; return MHH_ERR_STATE; }
    MHH_TEND(ut); MHH_TEND(vt); MHH_TEND(wt);
    int e;
    if ((e = pres_input_fft_forward(c, u, v, w, ut, vt, wt, dt))) return e;
    if ((e = mhh_pres_solve_spectral(c, c->work_p))) return e;
    if ((e = fft_backward_impl(c, c->work_p, p))) return e;
    return mhh_pres_output(c, ut, vt, wt, p);
}

// Pres_2::output's gradient stage fused with Timeloop::exec for u,v,w (the caller still runs mhh_rk_substep on the
// remaining prognostic fields with the same substep).  p must carry its ghost row and halos (output stage 0).
int mhh_pres_output_rk(mhh_ctx* c, double* ut, double* vt, double* wt, double* p, double* u, double* v, double* w,
                       int rkorder, int substep, mhh_double dt)
{
    PRES_READY(c);
    // 4th order: p's ghost rows and halos are stage 0 of mhh_pres_output_stage; this stage also puts w's ghost rows in
    // their Normal state (src/model.cxx:412) before the update
    if (c->order == 4) return pres4_output_rk(c, ut, vt, wt, p, u, v, w, rkorder, substep, dt, false);
    return pres2_output_rk(c, ut, vt, wt, p, u, v, w, rkorder, substep, dt);
}

// mhh_pres_exec with the fused last stage: on return u,v,w hold the substep's new velocities and ut,vt,wt the scaled
// tendencies the next substep starts from
int mhh_pres_exec_rk(mhh_ctx* c, double* u, double* v, double* w, double* ut, double* vt, double* wt, double* p,
                     mhh_double dt_sub, int rkorder, int substep, mhh_double dt)
{
    PRES_READY(c);
    if (c->npy != 1) { c->last_error = "mhh_pres_exec_rk is the single-GPU pipeline; y-slab ranks drive the stages"; return MHH_ERR_STATE; }
    MHH_TEND(ut); MHH_TEND(vt); MHH_TEND(wt);
    int e;
    if ((e = pres_input_fft_forward(c, u, v, w, ut, vt, wt, dt_sub))) return e;
    if ((e = mhh_pres_solve_spectral(c, c->work_p))) return e;
    if ((e = fft_backward_impl(c, c->work_p, p))) return e;
    if (c->out_wait)        // work on another stream still reads the old u,v,w (mhh_pres_set_output_wait)
    {
        MHH_CUDA_CHECK(cudaStreamWaitEvent(c->stream, c->out_wait, 0));
        c->out_wait = nullptr;
    }
    // 4th order: also leaves w's ghost rows in the Normal state of the not yet updated w (src/model.cxx:412)
    if (c->order == 4) return pres4_output_rk(c, ut, vt, wt, p, u, v, w, rkorder, substep, dt);
    if ((e = pres2_output_stage(c, 0, ut, vt, wt, p))) return e;
    return pres2_output_rk(c, ut, vt, wt, p, u, v, w, rkorder, substep, dt);
}

int mhh_pres_set_output_wait(mhh_ctx* c, void* cuda_event)
{ CTX_CHECK(c); c->out_wait = static_cast<cudaEvent_t>(cuda_event); return MHH_OK; }

int mhh_pres_check_divergence(mhh_ctx* c, const double* u, const double* v, const double* w, double* out)
{
    CTX_CHECK(c);
    int e = (c->order == 2) ? pres2_divergence_bits(c, u, v, w, c->max_bits_d + 2)
                            : pres4_divergence_bits(c, u, v, w, c->max_bits_d + 2);
    if (e) return e;
    return read_max(c, c->max_bits_d + 2, out);
}

int mhh_pres_get_coefs(const mhh_ctx* c, double* bmati, double* bmatj, double* a_or_bands, double* cc)
{
    CTX_CHECK(c);
    if (!c->pres_ready) return MHH_ERR_STATE;
    std::memcpy(bmati, c->bmati.data(), sizeof(double) * c->bmati.size());
    std::memcpy(bmatj, c->bmatj.data(), sizeof(double) * c->bmatj.size());
    if (c->order == 2)
    {
        std::memcpy(a_or_bands, c->pres_a.data(), sizeof(double) * c->pres_a.size());
        if (cc) std::memcpy(cc, c->pres_c.data(), sizeof(double) * c->pres_c.size());
    }
    else
        std::memcpy(a_or_bands, c->bands.data(), sizeof(double) * c->bands.size());
    return MHH_OK;
}

// ---------------------------------------------------------------- Boundary
int mhh_cyclic(mhh_ctx* c, double* fld, int edge) { CTX_CHECK(c); return cyclic_exec(c, fld, edge); }
int mhh_cyclic_2d(mhh_ctx* c, double* fld) { CTX_CHECK(c); return cyclic_exec_2d(c, fld); }
int mhh_ghost_cells(mhh_ctx* c, double* a, int bcbot, int bctop, double* abot, double* atop,
                    const double* agradbot, const double* agradtop)
{ CTX_CHECK(c); return ghost_cells_exec(c, a, bcbot, bctop, abot, atop, agradbot, agradtop); }
int mhh_ghost_cells_w(mhh_ctx* c, double* w, int type) { CTX_CHECK(c); return ghost_cells_w_exec(c, w, type); }
int mhh_set_bc(mhh_ctx* c, double* a, double* agrad, double* aflux, int sw, double aval, double visc, double offset)
{ CTX_CHECK(c); return set_bc_exec(c, a, agrad, aflux, sw, aval, visc, offset); }

// ---------------------------------------------------------------- Timeloop
int mhh_rk_substep(mhh_ctx* c, int rkorder, int substep, mhh_double dt, int nfields, double* const* a, double* const* at)
{ CTX_CHECK(c); MHH_TENDS(at, nfields); return rk_substep(c, rkorder, substep, dt, nfields, a, at); }
mhh_double mhh_rk_sub_dt(int rkorder, int substep, mhh_double dt) { return rk_sub_dt(rkorder, substep, dt); }

// ---------------------------------------------------------------- peer-memory data path of the y-slab decomposition
int mhh_peer_alloc(mhh_ctx* c, unsigned long long bytes, void** ptr)
{
    CTX_CHECK(c);
    if (!ptr || bytes == 0) return MHH_ERR_ARG;
    MHH_CUDA_CHECK(cudaMalloc(ptr, bytes));
    MHH_CUDA_CHECK(cudaMemset(*ptr, 0, bytes));
    return MHH_OK;
}
int mhh_peer_free(mhh_ctx* c, void* ptr) { CTX_CHECK(c); MHH_CUDA_CHECK(cudaFree(ptr)); return MHH_OK; }
int mhh_peer_export(mhh_ctx* c, void* ptr, unsigned char* handle64)
{
    CTX_CHECK(c);
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    cudaIpcMemHandle_t h;
    MHH_CUDA_CHECK(cudaIpcGetMemHandle(&h, ptr));
    std::memcpy(handle64, &h, 64);
    return MHH_OK;
}
int mhh_peer_import(mhh_ctx* c, const unsigned char* handle64, void** ptr)
{
    CTX_CHECK(c);
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle64, 64);
    MHH_CUDA_CHECK(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return MHH_OK;
}
int mhh_peer_unmap(mhh_ctx* c, void* ptr) { CTX_CHECK(c); MHH_CUDA_CHECK(cudaIpcCloseMemHandle(ptr)); return MHH_OK; }
int mhh_peer_register(mhh_ctx* c, int world, int me, void* my_base, unsigned long long bytes, void* const* bases,
                      unsigned long long flags_offset)
{
    CTX_CHECK(c);
    if (world < 1 || world > 8 || me < 0 || me >= world || !my_base || !bases || flags_offset + 32 * 8 > bytes || (flags_offset & 7))
    { c->last_error = "mhh_peer_register: bad arguments (world <= 8)"; return MHH_ERR_ARG; }
    if (world != c->npy || me != c->rank_y) { c->last_error = "mhh_peer_register: world / rank differ from the grid's npy / rank_y"; return MHH_ERR_ARG; }
    auto& P = c->peer;
    P.world = world; P.me = me; P.north = (me + 1) % world; P.south = (me + world - 1) % world;
    P.base = my_base; P.bytes = bytes;
    for (int r = 0; r < world; ++r) P.delta[r] = (long)(reinterpret_cast<const char*>(bases[r]) - reinterpret_cast<const char*>(my_base));
    if (P.delta[me] != 0) { c->last_error = "mhh_peer_register: bases[me] must be my_base"; return MHH_ERR_ARG; }
    P.flags = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(my_base) + flags_offset);
    P.halo_epoch = P.barrier_epoch = 0;
    return MHH_OK;
}
int mhh_peer_halo_push(mhh_ctx* c, int nfields, double* const* flds, int is2d) { CTX_CHECK(c); return peer_halo_push(c, nfields, flds, is2d); }
int mhh_peer_barrier(mhh_ctx* c) { CTX_CHECK(c); return peer_barrier(c); }
int mhh_peer_error(mhh_ctx* c, int* err) { CTX_CHECK(c); if (!err) return MHH_ERR_ARG; return peer_error(c, err); }
int mhh_fft_x_chunks_to_peers(mhh_ctx* c, const double* in, double* const* peer_out, int k0, int nk)
{
    PRES_READY(c);
    if (!peer_out) return MHH_ERR_ARG;
    return fft_x_chunks(c, -1, in, nullptr, nullptr, k0, nk, peer_out);
}
int mhh_fft_y_chunks_to_peers(mhh_ctx* c, const double* in, double* const* peer_out, int k0, int nk)
{
    PRES_READY(c);
    if (!peer_out) return MHH_ERR_ARG;
    return fft_y_chunks(c, +1, in, nullptr, k0, nk, peer_out);
}

// ---------------------------------------------------------------- Field3d_operators / diagnostics
int mhh_mean_profile(mhh_ctx* c, const double* fld, double* prof)
{
    CTX_CHECK(c);
    if (int e = level_sums(c, fld, prof)) return e;
    const double n = (double)c->gd.itot * c->gd.jtot;   // src/field3d_operators.cxx:49
    for (int k = 0; k < c->gd.kcells; ++k) prof[k] = prof[k] / n;
    return MHH_OK;
}
int mhh_level_sums(mhh_ctx* c, const double* fld, double* sums) { CTX_CHECK(c); return level_sums(c, fld, sums); }

// Stats profiles over the default mask, src/stats.cxx:264-398 (calc_mean, calc_moment, calc_cov): levels kstart .. kend
int mhh_stats_mean(mhh_ctx* c, const double* fld, double* prof)
{ CTX_CHECK(c); if (!fld || !prof) return MHH_ERR_ARG; return stats_profile(c, fld, nullptr, 0., 1, nullptr, nullptr, 0., 1, prof); }
int mhh_stats_moment(mhh_ctx* c, const double* fld, const double* fld_mean, double offset, int power, double* prof)
{ CTX_CHECK(c); if (!fld || !fld_mean || !prof) return MHH_ERR_ARG; return stats_profile(c, fld, fld_mean, offset, power, nullptr, nullptr, 0., 1, prof); }
int mhh_stats_cov(mhh_ctx* c, const double* fld1, const double* mean1, double offset1, int pow1,
                  const double* fld2, const double* mean2, double offset2, int pow2, double* prof)
{
    CTX_CHECK(c);
    if (!fld1 || !mean1 || !fld2 || !mean2 || !prof) return MHH_ERR_ARG;
    return stats_profile(c, fld1, mean1, offset1, pow1, fld2, mean2, offset2, pow2, prof);
}

int mhh_mean(mhh_ctx* c, const double* fld, double* out)
{
    CTX_CHECK(c);
    double s;
    if (int e = weighted_sum(c, 1, fld, nullptr, nullptr, &s)) return e;
    *out = s / (c->gd.itot * c->gd.jtot * c->gd.zsize);   // src/field3d_operators.cxx:152
    return MHH_OK;
}
int mhh_mean_2d(mhh_ctx* c, const double* fld, double* out)
{
    CTX_CHECK(c);
    double s;
    if (int e = plane_sum(c, fld, &s)) return e;
    *out = s / ((double)c->gd.itot * c->gd.jtot);
    return MHH_OK;
}
int mhh_max(mhh_ctx* c, const double* fld, double* out) { CTX_CHECK(c); return interior_max(c, fld, out); }

int mhh_check_momentum(mhh_ctx* c, const double* u, const double* v, const double* w, double* out)
{
    CTX_CHECK(c);
    double s;
    if (int e = weighted_sum(c, 2, u, v, w, &s)) return e;
    *out = s / (c->gd.itot * c->gd.jtot * c->gd.zsize);
    return MHH_OK;
}
int mhh_check_tke(mhh_ctx* c, const double* u, const double* v, const double* w, double* out)
{
    CTX_CHECK(c);
    double s;
    if (int e = weighted_sum(c, 3, u, v, w, &s)) return e;
    s /= (c->gd.itot * c->gd.jtot * c->gd.zsize);
    *out = s * 0.5;
    return MHH_OK;
}
int mhh_check_mass(mhh_ctx* c, const double* s_fld, double* out)
{
    CTX_CHECK(c);
    double s;
    if (int e = weighted_sum(c, 1, s_fld, nullptr, nullptr, &s)) return e;
    *out = s / (c->gd.itot * c->gd.jtot * c->gd.zsize);
    return MHH_OK;
}

}  // extern "C"
