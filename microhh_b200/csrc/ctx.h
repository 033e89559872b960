// Host-side context behind the opaque `mhh_ctx*` of include/mhh_b200.h.
// Owns only what the reference's operators own (small coefficient arrays, FFT twiddles,
// solver scratch: src/pres_2.cu:217-250, src/diff_smag2.cu:521-548); all 3-D fields stay
// owned by the host model and arrive as device pointers.
#pragma once

#include "sys.h"
#include <string>
#include <vector>

#include <cuda.h>

#include "mhh_common.cuh"

namespace mhh
{

struct FftPlan
{
    int n = 0;
    int nfac = 0;
    int fac[20];
    double2* tw = nullptr;   // device, n entries exp(-2 pi i k/n)
    // register-resident power-of-two path (fft_pow2.cu): radices in {8,4,2}, per-stage twiddle tables
    // laid out [r-1][k] (k < ns) so that consecutive lanes read consecutive entries
    int p2_nst = 0;
    int p2_rad[8];
    int p2_off[8];
    double2* p2_tw = nullptr;
};

struct Ctx
{
    GridD gd;
    int order = 2;           // swspatialorder
    int npy = 1, rank_y = 0; // y-slab decomposition (npx = 1), SURVEY 8(e)
    int iblock = 0, jblock = 0;
    cudaStream_t stream = 0;

    // host copies of metric arrays (length kcells)
    std::vector<double> z, zh, dz, dzh, dzi, dzhi, dzi4, dzhi4, rhoref, rhorefh;
    // device storage backing gd.* pointers: 10 arrays of kcells
    double* metrics_d = nullptr;

    // reductions
    unsigned long long* max_bits_d = nullptr;   // 8 slots
    bool fuse_pres_input = false;   // mhh_pres_set_fused_input: Pres_2::input evaluated inside the forward x transform
    // lazy-tendency mode (mhh_set_lazy_tendencies): arrays the fused last-substep tails left undefined instead of zeroed
    bool lazy_tend = false;
    const double* fresh_uvw[3] = {nullptr, nullptr, nullptr};
    const double* fresh_s[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    mhh_f64* sum_d = nullptr;                   // kcells + 16 slots; sums are accumulated in double in either build
    double* stats_d = nullptr;                  // 2 * kcells: the mean profiles of mhh_stats_moment / mhh_stats_cov
    mhh_f64* host_pinned = nullptr;             // kcells + 16 doubles, pinned

    // pressure solver
    bool pres_ready = false;
    FftPlan plan_i, plan_j;
    bool fft_force_generic = false;   // tests: exercise the mixed-radix kernel on power-of-two sizes too
    double* bmati_d = nullptr;
    double* bmatj_d = nullptr;
    double* pres_a_d = nullptr;     // pres_2: a, c ; pres_4: m1..m7 (7*kmax)
    double* pres_c_d = nullptr;
    double* bands_d = nullptr;
    std::vector<double> bmati, bmatj, pres_a, pres_c, bands;
    double* work_p = nullptr;       // ghost-free imax*jmax*kmax (spectral work array)
    double* work_c = nullptr;       // tdma c' / transpose buffer
    double* work_h = nullptr;       // pres_4 hdma factors
    size_t work_elems = 0;

    // smag2: per-level Smagorinsky length table (rebuilt when cs changes)
    double* mlen0_d = nullptr;
    double mlen0_cs = -1.;

    // thermo_dry base state: thref (kcells) followed by threfh (kcells)
    double* thref_d = nullptr;
    double thref_ks = 0., threfh_ks = 0.;    // thref[kstart], threfh[kstart] (surface buoyancy, src/thermo_dry.cxx:129-160)

    // Boundary_surface: z/L lookup table of the surface solver (zL_sl | f_sl, nzL_lut floats each)
    float* surf_lut_d = nullptr;
    std::vector<float> surf_lut_host;

    // Buffer (sponge layer): sigma(z[k]) (kcells) followed by sigma(zh[k]) (kcells); first levels inside the layer
    double* buffer_sigma_d = nullptr;
    int buffer_kstart = 0, buffer_kstarth = 0;

    // Buffer / Coriolis terms of u,v,w applied inside the fused momentum pass (mhh_advec2_smag2_set_force_tail)
    struct UvwTail
    {
        bool buffer = false, coriolis = false;
        const double *ubuf = nullptr, *vbuf = nullptr, *wbuf = nullptr, *ug = nullptr, *vg = nullptr;
        double fc = 0., ugrid = 0., vgrid = 0.;
    } uvw_tail;

    // peer-memory data path of the y-slab decomposition (peer.cu): this rank's exchange arena and where the same arena of
    // every other rank is mapped in this process (delta[r] = rank r's base - my base, bytes)
    struct Peer
    {
        int world = 1, me = 0, north = 0, south = 0;
        void* base = nullptr;
        size_t bytes = 0;
        long delta[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        unsigned long long* flags = nullptr;        // 32 words inside the arena
        unsigned long long halo_epoch = 0, barrier_epoch = 0;
    } peer;

    cudaEvent_t out_wait = nullptr;  // mhh_pres_set_output_wait: the next fused pressure call waits for it before it updates u,v,w
    int sc_kc0 = 0, sc_nkc = 0;     // k-chunk window of the scalar tendency pass (mhh_advec2_smag2_exec_rk_parts); 0,0: whole column

    // TMA tensor maps of the fields the tile kernels stage (tma.cu); keyed by (pointer, box): the host model's arrays are stable
    struct TmaEntry { const double* fld; int boxx, boxy; CUtensorMap map; };
    std::vector<TmaEntry> tma_cache;
    bool tma_off = false;           // mhh_set_tma(ctx, 0): force the cp.async (LDGSTS) plane ring (tests compare the two)

    // launch accounting / live per-kernel timing (bench.py: gpu_launches and roofline.achieved)
    unsigned long long launches = 0;
    bool prof_on = false;
    std::vector<std::string> prof_names;
    struct ProfEntry { int id; cudaEvent_t e0, e1; };
    std::vector<ProfEntry> prof_entries;
    std::vector<cudaEvent_t> prof_pool;

    std::string last_error;
};

// Counts one kernel launch; when profiling is on, brackets it with CUDA events on the launch stream.
struct ProfScope
{
    Ctx* c;
    cudaEvent_t e1 = nullptr;
    ProfScope(Ctx* ctx, const char* name);
    ~ProfScope();
};

#define MHH_LAUNCH(c, name, ...) do { mhh::ProfScope ps__(c, name); __VA_ARGS__; } while (0)

int ctx_set_metrics(Ctx* c, const double* z_host);
int pres_set_values(Ctx* c);

}  // namespace mhh

struct mhh_ctx : public mhh::Ctx {};
