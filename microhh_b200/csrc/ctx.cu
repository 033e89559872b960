// Context creation: grid dimensions (Grid::init, src/grid.cxx:106-189) and vertical metrics
// (Grid::calculate, src/grid.cxx:237-368) computed on the host exactly as the reference does,
// then uploaded once.  y-slab decomposition: npx = 1, jmax = jtot/npy (src/grid.cxx:142-150).
#include "sys.h"
#include <cstring>

#include "ctx.h"
#include "fft.h"

namespace mhh
{

int ctx_init_grid(Ctx* c, int itot, int jtot, int ktot, int igc, int jgc, int kgc,
                  double xsize, double ysize, double zsize, int order, int npy, int rank_y)
{
    if (itot < 1 || jtot < 1 || ktot < 1 || (order != 2 && order != 4) || npy < 1 || rank_y < 0 || rank_y >= npy)
        return MHH_ERR_ARG;
    // divisibility rules of src/grid.cxx:111-136 with npx = 1
    if (itot % npy != 0 || jtot % npy != 0)
    {
        c->last_error = "itot and jtot must be multiples of npy";
        return MHH_ERR_ARG;
    }
    GridD& gd = c->gd;
    std::memset(&gd, 0, sizeof(gd));
    c->order = order;
    c->npy = npy;
    c->rank_y = rank_y;
    gd.itot = itot; gd.jtot = jtot; gd.ktot = ktot;
    gd.imax = itot; gd.jmax = jtot / npy; gd.kmax = ktot;
    c->iblock = itot / npy;
    c->jblock = jtot;
    const int gmin = (order == 2) ? 1 : 3;
    gd.igc = igc > gmin ? igc : gmin;
    gd.jgc = jgc > gmin ? jgc : gmin;
    gd.kgc = kgc > gmin ? kgc : gmin;
    gd.icells = gd.imax + 2 * gd.igc;
    gd.jcells = gd.jmax + 2 * gd.jgc;
    gd.ijcells = gd.icells * gd.jcells;
    gd.kcells = gd.kmax + 2 * gd.kgc;
    const long long ncells = (long long)gd.ijcells * gd.kcells;
    if (ncells > 2000000000LL)
    {
        c->last_error = "grid too large for 32-bit cell indices on one GPU; use more y-slabs";
        return MHH_ERR_ARG;
    }
    gd.ncells = (int)ncells;
    gd.istart = gd.igc; gd.jstart = gd.jgc; gd.kstart = gd.kgc;
    gd.iend = gd.imax + gd.igc; gd.jend = gd.jmax + gd.jgc; gd.kend = gd.kmax + gd.kgc;
    // check_ghost_cells: a slab must be at least as thick as its halo (except 2-D runs)
    if (gd.imax < gd.igc || (gd.jmax < gd.jgc && gd.jtot > 1) || gd.kmax < gd.kgc)
    {
        c->last_error = "grid smaller than its ghost-cell depth";
        return MHH_ERR_ARG;
    }
    gd.xsize = xsize; gd.ysize = ysize; gd.zsize = zsize;
    gd.dx = xsize / itot;
    gd.dy = ysize / jtot;
    gd.dxi = 1. / gd.dx;
    gd.dyi = 1. / gd.dy;
    return MHH_OK;
}

int ctx_set_metrics(Ctx* c, const double* z_in)
{
    GridD& gd = c->gd;
    const int kcells = gd.kcells, kstart = gd.kstart, kend = gd.kend;
    auto& z = c->z; auto& zh = c->zh; auto& dz = c->dz; auto& dzh = c->dzh;
    auto& dzi = c->dzi; auto& dzhi = c->dzhi; auto& dzi4 = c->dzi4; auto& dzhi4 = c->dzhi4;
    for (auto* v : {&z, &zh, &dz, &dzh, &dzi, &dzhi, &dzi4, &dzhi4}) v->assign(kcells, 0.);
    for (int k = 0; k < gd.ktot; ++k) z[k + kstart] = z_in[k];
    if (z[kend - 1] > gd.zsize)
    {
        c->last_error = "Highest grid point is above prescribed zsize";
        return MHH_ERR_ARG;
    }
    const double zsize = gd.zsize;

    if (c->order == 2)
    {
        // src/grid.cxx:267-299
        z[kstart - 1] = -z[kstart];
        z[kend] = 2. * zsize - z[kend - 1];
        for (int k = kstart + 1; k < kend; ++k) zh[k] = 0.5 * (z[k - 1] + z[k]);
        zh[kstart] = 0.;
        zh[kend] = zsize;
        for (int k = 1; k < kcells; ++k) { dzh[k] = z[k] - z[k - 1]; dzhi[k] = 1. / dzh[k]; }
        dzh[kstart - 1] = dzh[kstart + 1];
        dzhi[kstart - 1] = dzhi[kstart + 1];
        for (int k = 1; k < kcells - 1; ++k) { dz[k] = zh[k + 1] - zh[k]; dzi[k] = 1. / dz[k]; }
        dz[kstart - 1] = dz[kstart];
        dzi[kstart - 1] = dzi[kstart];
        dz[kend] = dz[kend - 1];
        dzi[kend] = dzi[kend - 1];
    }
    else
    {
        // src/grid.cxx:301-367
        using namespace fd;
        z[kstart - 1] = -2. * z[kstart] + (1. / 3.) * z[kstart + 1];
        z[kstart - 2] = -9. * z[kstart] + 2. * z[kstart + 1];
        z[kend] = (8. / 3.) * zsize - 2. * z[kend - 1] + (1. / 3.) * z[kend - 2];
        z[kend + 1] = 8. * zsize - 9. * z[kend - 1] + 2. * z[kend - 2];
        z[kstart - 3] = kDhuge;
        z[kend + 2] = kDhuge;
        zh[kstart] = 0.;
        for (int k = kstart + 1; k < kend; ++k)
            zh[k] = ci0 * z[k - 2] + ci1 * z[k - 1] + ci2 * z[k] + ci3 * z[k + 1];
        zh[kend] = zsize;
        zh[kstart - 1] = bi0 * z[kstart - 2] + bi1 * z[kstart - 1] + bi2 * z[kstart] + bi3 * z[kstart + 1];
        zh[kend + 1] = ti0 * z[kend - 2] + ti1 * z[kend - 1] + ti2 * z[kend] + ti3 * z[kend + 1];
        for (int k = 1; k < kcells; ++k) { dzh[k] = z[k] - z[k - 1]; dzhi[k] = 1. / dzh[k]; }
        dzh[kstart - 3] = dzh[kstart + 3];
        dzhi[kstart - 3] = dzhi[kstart + 3];
        for (int k = 1; k < kcells - 1; ++k) { dz[k] = zh[k + 1] - zh[k]; dzi[k] = 1. / dz[k]; }
        dz[kstart - 3] = dz[kstart + 2];
        dzi[kstart - 3] = dzi[kstart + 2];
        dz[kend + 2] = dz[kend - 3];
        dzi[kend + 2] = dzi[kend - 3];
        for (int k = kstart; k < kend; ++k)
        {
            dzi4[k] = 1. / (cg0 * zh[k - 1] + cg1 * zh[k] + cg2 * zh[k + 1] + cg3 * zh[k + 2]);
            dzhi4[k] = 1. / (cg0 * z[k - 2] + cg1 * z[k - 1] + cg2 * z[k] + cg3 * z[k + 1]);
        }
        dzhi4[kend] = 1. / (cg0 * z[kend - 2] + cg1 * z[kend - 1] + cg2 * z[kend] + cg3 * z[kend + 1]);
        dzi4[kstart - 1] = 1. / (bg0 * zh[kstart - 1] + bg1 * zh[kstart] + bg2 * zh[kstart + 1] + bg3 * zh[kstart + 2]);
        dzhi4[kstart - 1] = 1. / (bg0 * z[kstart - 2] + bg1 * z[kstart - 1] + bg2 * z[kstart] + bg3 * z[kstart + 1]);
        dzi4[kend] = 1. / (tg0 * zh[kend - 2] + tg1 * zh[kend - 1] + tg2 * zh[kend] + tg3 * zh[kend + 1]);
        dzhi4[kend + 1] = 1. / (tg0 * z[kend - 2] + tg1 * z[kend - 1] + tg2 * z[kend] + tg3 * z[kend + 1]);
        gd.dzhi4bot = 1. / (bg0 * z[kstart - 1] + bg1 * z[kstart] + bg2 * z[kstart + 1] + bg3 * z[kstart + 2]);
        gd.dzhi4top = 1. / (tg0 * z[kend - 3] + tg1 * z[kend - 2] + tg2 * z[kend - 1] + tg3 * z[kend]);
        dzi4[kstart - 2] = kDhuge;
        dzi4[kstart - 3] = kDhuge;
        dzi4[kend + 1] = kDhuge;
        dzi4[kend + 2] = kDhuge;
    }
    return MHH_OK;
}

int ctx_upload_metrics(Ctx* c)
{
    GridD& gd = c->gd;
    const int kcells = gd.kcells;
    if (!c->metrics_d) MHH_CUDA_CHECK(cudaMalloc(&c->metrics_d, sizeof(double) * kcells * 12));
    // correctly rounded reciprocals of the density profiles: operands of div_by() (mhh_common.cuh)
    std::vector<double> rho_i(kcells), rhoh_i(kcells);
    for (int k = 0; k < kcells; ++k) { rho_i[k] = 1. / c->rhoref[k]; rhoh_i[k] = 1. / c->rhorefh[k]; }
    const std::vector<double>* src[12] = {&c->z, &c->zh, &c->dz, &c->dzh, &c->dzi, &c->dzhi, &c->dzi4, &c->dzhi4,
                                          &c->rhoref, &c->rhorefh, &rho_i, &rhoh_i};
    for (int n = 0; n < 12; ++n)
        MHH_CUDA_CHECK(cudaMemcpy(c->metrics_d + (size_t)n * kcells, src[n]->data(), sizeof(double) * kcells,
                                  cudaMemcpyHostToDevice));
    const double* m = c->metrics_d;
    gd.z = m; gd.zh = m + kcells; gd.dz = m + 2 * kcells; gd.dzh = m + 3 * kcells; gd.dzi = m + 4 * kcells;
    gd.dzhi = m + 5 * kcells; gd.dzi4 = m + 6 * kcells; gd.dzhi4 = m + 7 * kcells;
    gd.rhoref = m + 8 * kcells; gd.rhorefh = m + 9 * kcells;
    gd.rhoref_i = m + 10 * kcells; gd.rhorefh_i = m + 11 * kcells;
    return MHH_OK;
}

int ctx_alloc_scratch(Ctx* c)
{
    const GridD& gd = c->gd;
    const size_t nred = (size_t)gd.kcells + 64;
    MHH_CUDA_CHECK(cudaMalloc(&c->max_bits_d, sizeof(unsigned long long) * 8));
    MHH_CUDA_CHECK(cudaMemset(c->max_bits_d, 0, sizeof(unsigned long long) * 8));
    MHH_CUDA_CHECK(cudaMalloc(&c->sum_d, sizeof(mhh_f64) * nred));
    MHH_CUDA_CHECK(cudaMallocHost(&c->host_pinned, sizeof(mhh_f64) * nred));
    return MHH_OK;
}

static cudaEvent_t prof_get_event(Ctx* c)
{
    if (!c->prof_pool.empty())
    {
        cudaEvent_t e = c->prof_pool.back();
        c->prof_pool.pop_back();
        return e;
    }
    cudaEvent_t e;
    cudaEventCreate(&e);
    return e;
}

ProfScope::ProfScope(Ctx* ctx, const char* name) : c(ctx)
{
    ++c->launches;
    if (!c->prof_on) return;
    int id = -1;
    for (size_t n = 0; n < c->prof_names.size(); ++n)
        if (c->prof_names[n] == name) { id = (int)n; break; }
    if (id < 0) { id = (int)c->prof_names.size(); c->prof_names.push_back(name); }
    Ctx::ProfEntry pe;
    pe.id = id;
    pe.e0 = prof_get_event(c);
    pe.e1 = prof_get_event(c);
    cudaEventRecord(pe.e0, c->stream);
    e1 = pe.e1;
    c->prof_entries.push_back(pe);
}

ProfScope::~ProfScope()
{
    if (e1) cudaEventRecord(e1, c->stream);
}

// Sums the recorded intervals per kernel name; returns the number of names. ms/counts sized >= names.
int prof_collect(Ctx* c, std::vector<double>& ms, std::vector<long>& counts)
{
    cudaStreamSynchronize(c->stream);
    ms.assign(c->prof_names.size(), 0.);
    counts.assign(c->prof_names.size(), 0);
    for (auto& pe : c->prof_entries)
    {
        float t = 0.f;
        cudaEventElapsedTime(&t, pe.e0, pe.e1);
        ms[pe.id] += t;
        counts[pe.id] += 1;
        c->prof_pool.push_back(pe.e0);
        c->prof_pool.push_back(pe.e1);
    }
    c->prof_entries.clear();
    return (int)c->prof_names.size();
}

void ctx_free(Ctx* c)
{
    for (auto& pe : c->prof_entries) { cudaEventDestroy(pe.e0); cudaEventDestroy(pe.e1); }
    for (auto e : c->prof_pool) cudaEventDestroy(e);
    cudaFree(c->metrics_d);
    cudaFree(c->max_bits_d);
    cudaFree(c->sum_d); cudaFree(c->stats_d);
    if (c->host_pinned) cudaFreeHost(c->host_pinned);
    cudaFree(c->bmati_d); cudaFree(c->bmatj_d); cudaFree(c->pres_a_d); cudaFree(c->pres_c_d); cudaFree(c->bands_d);
    cudaFree(c->work_p); cudaFree(c->work_c); cudaFree(c->work_h);
    cudaFree(c->mlen0_d);
    cudaFree(c->thref_d);
    cudaFree(c->buffer_sigma_d);
    cudaFree(c->surf_lut_d);
    fft_plan_destroy(&c->plan_i);
    fft_plan_destroy(&c->plan_j);
}

}  // namespace mhh
