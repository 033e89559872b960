// Batched real <-> half-complex FFTs as hand-written shared-memory Stockham kernels (sm_100a).
//
// Semantics = FFTW r2r FFTW_R2HC / FFTW_HC2R as the reference plans them (src/fft.cxx:140-154):
// unnormalised, half-complex order r0,r1,...,r(n/2),i((n+1)/2-1),...,i1, in place in a real array
// (SURVEY F7).  The reference GPU path uses cuFFT D2Z + repack + transpose + 2 copies
// (src/pres.cu:288-387); here one kernel per direction does load -> FFT -> half-complex store.
//
// Method: two real lines are packed into one complex line z = a + i b (so an n-point complex FFT
// yields both spectra), transformed with a mixed-radix (4,2,3,5,7 + generic) autosort Stockham FFT
// that ping-pongs between two shared-memory buffers, and unpacked with the Hermitian split
//   A_k = (Z_k + conj(Z_{n-k}))/2,   B_k = (Z_k - conj(Z_{n-k}))/(2i).
// A CTA transforms `lines` consecutive lines; for strided (y) transforms consecutive threads
// touch consecutive lines so global accesses stay coalesced along i.
#include "sys.h"
#include <cmath>

#include "ctx.h"
#include "fft.h"

namespace mhh
{
namespace
{
    __device__ __forceinline__ double2 cmul(double2 a, double2 b)
    {
        return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
    }
    __device__ __forceinline__ double2 cadd(double2 a, double2 b) { return make_double2(a.x + b.x, a.y + b.y); }
    __device__ __forceinline__ double2 csub(double2 a, double2 b) { return make_double2(a.x - b.x, a.y - b.y); }
    // multiply by -i (forward) or +i (backward)
    template<int DIR> __device__ __forceinline__ double2 rot90(double2 a)
    {
        return DIR < 0 ? make_double2(a.y, -a.x) : make_double2(-a.y, a.x);
    }
    template<int DIR> __device__ __forceinline__ double2 twd(double2 t)
    {
        return DIR < 0 ? t : make_double2(t.x, -t.y);
    }

    // One radix-R butterfly of the Stockham stage: v[r] are the twiddled inputs.
    template<int R, int DIR>
    __device__ __forceinline__ void butterfly(double2* v, const double2* __restrict__ tw, int n)
    {
        if (R == 2)
        {
            const double2 a = v[0], b = v[1];
            v[0] = cadd(a, b); v[1] = csub(a, b);
        }
        else if (R == 4)
        {
            const double2 a = cadd(v[0], v[2]), b = csub(v[0], v[2]);
            const double2 c = cadd(v[1], v[3]), d = rot90<DIR>(csub(v[1], v[3]));
            v[0] = cadd(a, c); v[2] = csub(a, c);
            v[1] = cadd(b, d); v[3] = csub(b, d);
        }
        else
        {
            // generic small DFT with table twiddles exp(-+2 pi i q r / R) = tw[(q*r mod R) * n/R]
            double2 o[R];
            const int step = n / R;
            #pragma unroll
            for (int q = 0; q < R; ++q)
            {
                double2 s = v[0];
                #pragma unroll
                for (int r = 1; r < R; ++r)
                    s = cadd(s, cmul(v[r], twd<DIR>(tw[((q * r) % R) * step])));
                o[q] = s;
            }
            #pragma unroll
            for (int q = 0; q < R; ++q) v[q] = o[q];
        }
    }

    template<int R, int DIR>
    __device__ __forceinline__ void stage(const double2* __restrict__ x, double2* __restrict__ y,
                                          const double2* __restrict__ tw, int n, int npad, int ns, int npairs)
    {
        const int nb = n / R;                 // butterflies per line
        const int tstep = n / (ns * R);
        const int total = nb * npairs;
        for (int t = threadIdx.x; t < total; t += blockDim.x)
        {
            const int p = t / nb, j = t - p * nb;
            const int k = j % ns;
            const double2* xp = x + p * npad;
            double2 v[R];
            v[0] = xp[j];
            #pragma unroll
            for (int r = 1; r < R; ++r)
                v[r] = cmul(xp[j + r * nb], twd<DIR>(tw[k * r * tstep]));
            butterfly<R, DIR>(v, tw, n);
            double2* yp = y + p * npad + (j / ns) * ns * R + k;
            #pragma unroll
            for (int q = 0; q < R; ++q)
                yp[q * ns] = v[q];
        }
    }

    struct FftArgs
    {
        int n, nfac;
        int fac[20];
        const double2* tw;
        // line l (0 <= l < nlines): element m at  base + (l % n1)*s1 + (l / n1)*s2 + m*sn
        int nlines, n1;
        long s1, s2, sn;
        long base_in, base_out;
        int n1o; long s1o, s2o, sno;       // output mapping (may differ: scatter into the ghosted layout)
        int line_fast;                     // 1: consecutive threads -> consecutive lines (strided transforms)
        double scale;
    };

    template<int DIR>
    __device__ __forceinline__ double2* run_stages(double2* x, double2* y, const double2* tw, const FftArgs& A,
                                                   int npad, int npairs)
    {
        int ns = 1;
        for (int f = 0; f < A.nfac; ++f)
        {
            const int R = A.fac[f];
            switch (R)
            {
                case 4: stage<4, DIR>(x, y, tw, A.n, npad, ns, npairs); break;
                case 2: stage<2, DIR>(x, y, tw, A.n, npad, ns, npairs); break;
                case 3: stage<3, DIR>(x, y, tw, A.n, npad, ns, npairs); break;
                case 5: stage<5, DIR>(x, y, tw, A.n, npad, ns, npairs); break;
                case 7: stage<7, DIR>(x, y, tw, A.n, npad, ns, npairs); break;
                case 11: stage<11, DIR>(x, y, tw, A.n, npad, ns, npairs); break;
                default: stage<13, DIR>(x, y, tw, A.n, npad, ns, npairs); break;
            }
            __syncthreads();
            double2* t = x; x = y; y = t;
            ns *= R;
        }
        return x;   // buffer holding the result
    }

    // FWD: real lines -> half-complex.  LINES must be even.
    template<int DIR>
    __global__ void __launch_bounds__(256)
    fft_lines_kernel(const double* __restrict__ in, double* __restrict__ out, const FftArgs A, const int lines_per_cta)
    {
        extern __shared__ double2 smem[];
        const int n = A.n;
        const int npad = n + 1;
        const int npairs = lines_per_cta >> 1;
        double2* tw = smem;                       // n
        double2* x = smem + n;                    // npairs * npad
        double2* y = x + npairs * npad;           // npairs * npad
        double* xr = reinterpret_cast<double*>(x);
        double* yr = reinterpret_cast<double*>(y);

        for (int t = threadIdx.x; t < n; t += blockDim.x) tw[t] = A.tw[t];

        const int l0 = blockIdx.x * lines_per_cta;
        const int total = lines_per_cta * n;

        if (DIR < 0)
        {
            // load two real lines per complex line: x[p][m] = (a[m], b[m])
            for (int t = threadIdx.x; t < total; t += blockDim.x)
            {
                int ll, m;
                if (A.line_fast) { ll = t % lines_per_cta; m = t / lines_per_cta; }
                else             { m = t % n; ll = t / n; }
                const int l = l0 + ll;
                double val = 0.;
                if (l < A.nlines)
                    val = in[A.base_in + (long)(l % A.n1) * A.s1 + (long)(l / A.n1) * A.s2 + (long)m * A.sn];
                xr[2 * ((ll >> 1) * npad + m) + (ll & 1)] = val;
            }
            __syncthreads();
            double2* z = run_stages<DIR>(x, y, tw, A, npad, npairs);
            double* o = (z == x) ? yr : xr;       // the free buffer, viewed as lines_per_cta x (2*npad) doubles... use npad stride
            // Hermitian split into half-complex lines stored at o[ll*npad + m]
            const int nh = n / 2;
            for (int t = threadIdx.x; t < npairs * (nh + 1); t += blockDim.x)
            {
                const int p = t / (nh + 1), k = t - p * (nh + 1);
                const double2 zk = z[p * npad + k];
                const double2 zn = z[p * npad + (k == 0 ? 0 : n - k)];
                const double are = 0.5 * (zk.x + zn.x), aim = 0.5 * (zk.y - zn.y);
                const double bre = 0.5 * (zk.y + zn.y), bim = -0.5 * (zk.x - zn.x);
                double* oa = o + (2 * p) * npad;
                double* ob = o + (2 * p + 1) * npad;
                oa[k] = are; ob[k] = bre;
                if (k > 0 && 2 * k < n) { oa[n - k] = aim; ob[n - k] = bim; }
            }
            __syncthreads();
            for (int t = threadIdx.x; t < total; t += blockDim.x)
            {
                int ll, m;
                if (A.line_fast) { ll = t % lines_per_cta; m = t / lines_per_cta; }
                else             { m = t % n; ll = t / n; }
                const int l = l0 + ll;
                if (l < A.nlines)
                    out[A.base_out + (long)(l % A.n1o) * A.s1o + (long)(l / A.n1o) * A.s2o + (long)m * A.sno] = o[ll * npad + m] * A.scale;
            }
        }
        else
        {
            // load half-complex lines into y (as reals, stride npad)
            for (int t = threadIdx.x; t < total; t += blockDim.x)
            {
                int ll, m;
                if (A.line_fast) { ll = t % lines_per_cta; m = t / lines_per_cta; }
                else             { m = t % n; ll = t / n; }
                const int l = l0 + ll;
                double val = 0.;
                if (l < A.nlines)
                    val = in[A.base_in + (long)(l % A.n1) * A.s1 + (long)(l / A.n1) * A.s2 + (long)m * A.sn];
                yr[ll * npad + m] = val;
            }
            __syncthreads();
            // Z_k = A_k + i B_k with A, B rebuilt from the half-complex lines
            const int nh = n / 2;
            for (int t = threadIdx.x; t < npairs * (nh + 1); t += blockDim.x)
            {
                const int p = t / (nh + 1), k = t - p * (nh + 1);
                const double* ia = yr + (2 * p) * npad;
                const double* ib = yr + (2 * p + 1) * npad;
                const bool self = (k == 0) || (2 * k == n);
                const double are = ia[k], bre = ib[k];
                const double aim = self ? 0. : ia[n - k], bim = self ? 0. : ib[n - k];
                x[p * npad + k] = make_double2(are - bim, aim + bre);
                if (!self) x[p * npad + n - k] = make_double2(are + bim, -aim + bre);
            }
            __syncthreads();
            double2* z = run_stages<DIR>(x, y, tw, A, npad, npairs);
            const double* zr = reinterpret_cast<const double*>(z);
            for (int t = threadIdx.x; t < total; t += blockDim.x)
            {
                int ll, m;
                if (A.line_fast) { ll = t % lines_per_cta; m = t / lines_per_cta; }
                else             { m = t % n; ll = t / n; }
                const int l = l0 + ll;
                if (l < A.nlines)
                    out[A.base_out + (long)(l % A.n1o) * A.s1o + (long)(l / A.n1o) * A.s2o + (long)m * A.sno]
                        = zr[2 * ((ll >> 1) * npad + m) + (ll & 1)] * A.scale;
            }
        }
    }
}  // namespace

bool fft_pow2_supported(int n);
int fft_pow2_plan(FftPlan* p);
int fft_pow2_lines(Ctx* c, const FftPlan& plan, int dir, const double* in, double* out, const FftLayout& L);

int fft_plan_create(FftPlan* p, int n)
{
    p->n = n;
    p->nfac = 0;
    int m = n;
    const int radices[] = {4, 2, 3, 5, 7, 11, 13};
    for (int r : radices)
        while (m % r == 0 && p->nfac < 20) { p->fac[p->nfac++] = r; m /= r; }
    if (m != 1) return MHH_ERR_UNSUPPORTED;   // prime factor > 13
    std::vector<double2> tw(n);
    const mhh_f80 pi = std::acos(-1.0L);
    for (int k = 0; k < n; ++k)
    {
        const mhh_f80 ang = -2.0L * pi * (mhh_f80)k / (mhh_f80)n;
        tw[k] = make_double2((double)std::cos(ang), (double)std::sin(ang));
    }
    MHH_CUDA_CHECK(cudaMalloc(&p->tw, sizeof(double2) * n));
    MHH_CUDA_CHECK(cudaMemcpy(p->tw, tw.data(), sizeof(double2) * n, cudaMemcpyHostToDevice));
    return fft_pow2_plan(p);
}

void fft_plan_destroy(FftPlan* p)
{
    if (p->tw) cudaFree(p->tw);
    if (p->p2_tw) cudaFree(p->p2_tw);
    p->tw = nullptr;
    p->p2_tw = nullptr;
}

int fft_lines(Ctx* c, const FftPlan& plan, int dir, const double* in, double* out, const FftLayout& L)
{
    FftArgs A;
    A.n = plan.n; A.nfac = plan.nfac;
    for (int f = 0; f < plan.nfac; ++f) A.fac[f] = plan.fac[f];
    A.tw = plan.tw;
    A.nlines = L.nlines; A.n1 = L.n1; A.s1 = L.s1; A.s2 = L.s2; A.sn = L.sn;
    A.base_in = L.base_in; A.base_out = L.base_out;
    A.n1o = L.n1o; A.s1o = L.s1o; A.s2o = L.s2o; A.sno = L.sno;
    A.line_fast = (L.sn != 1);
    A.scale = L.scale;
    if (L.nlines == 0) return MHH_OK;
    const bool split = L.esplit_in > 0 || L.esplit_out > 0;
    if (plan.p2_nst > 0 && (!c->fft_force_generic || split)) return fft_pow2_lines(c, plan, dir, in, out, L);
    if (split) return MHH_ERR_UNSUPPORTED;
    if (plan.n == 1)
    {
        // length-1 transform = identity (2-D runs, jtot == 1): only scale / relayout
        // handled by the generic kernel as well (no stages), fall through
    }
    // lines per CTA: as many as fit ~96 KB (two complex buffers of lines/2 x (n+1))
    int lines = 16;
    const size_t budget = 96 * 1024;
    while (lines > 2 && (size_t)(2 * (lines / 2) * (plan.n + 1) + plan.n) * sizeof(double2) > budget) lines >>= 1;
    if (A.line_fast == 0) { while (lines > 2 && lines * plan.n > 8192) lines >>= 1; }
    const size_t smem = (size_t)(2 * (lines / 2) * (plan.n + 1) + plan.n) * sizeof(double2);
    if (smem > 200 * 1024) return MHH_ERR_UNSUPPORTED;
    const int nblocks = (L.nlines + lines - 1) / lines;
    if (dir < 0)
    {
        static bool attr_f = false;
        if (!attr_f) { cudaFuncSetAttribute(fft_lines_kernel<-1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024); attr_f = true; }
        MHH_LAUNCH(c, A.line_fast ? "fft_fwd_y" : "fft_fwd_x", fft_lines_kernel<-1><<<nblocks, 256, smem, c->stream>>>(in, out, A, lines));
    }
    else
    {
        static bool attr_b = false;
        if (!attr_b) { cudaFuncSetAttribute(fft_lines_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024); attr_b = true; }
        MHH_LAUNCH(c, A.line_fast ? "fft_bwd_y" : "fft_bwd_x", fft_lines_kernel<1><<<nblocks, 256, smem, c->stream>>>(in, out, A, lines));
    }
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

}  // namespace mhh
