// Register-resident power-of-two real <-> half-complex line FFTs (the fast path behind fft_lines).
//
// Same semantics as fft.cu (FFTW R2HC / HC2R, two real lines per complex transform), different
// execution shape, chosen after the first B200 measurement (profiles/r01_*: the generic kernel
// ran at 7 % of HBM peak, bound by index arithmetic, 5 radix-4/2 stages and bank conflicts):
//   * N = 8^a 4^b 2^c: radix-8 stages first, every thread owns 8 complex values in registers and
//     ALWAYS reads the same 8 shared-memory slots tf + m*N/8 (Stockham autosort, one buffer,
//     read -> sync -> write -> sync), so a 512-point transform takes 3 exchanges, not 5;
//   * all index arithmetic is shifts/masks; twiddles come from per-stage tables laid out [r][k]
//     so consecutive lanes read consecutive entries (L1-resident, __ldg);
//   * one slot of padding per 8 keeps the stride-8 writes of the first stage conflict-free;
//   * contiguous (x) transforms load their inputs straight into registers and store the result
//     straight from registers (the backward pass scatters into the ghost-celled pressure field);
//     strided (y) transforms stage a [n][line] tile through the same buffer so that global
//     accesses stay coalesced along i.
#include "sys.h"
#include <cmath>
#include <vector>

#include <cstdlib>

#include "ctx.h"
#include "fft.h"

namespace mhh
{
namespace
{
    constexpr int BLOCK = 256;
    constexpr double kS = 0.70710678118654752440;

    struct P2Args
    {
        int n, log2n, nst;
        int rad[8];
        int off[8];
        const double2* tw;
        int nlines, n1;
        long s1, s2, sn, base_in;
        int n1o;
        long s1o, s2o, sno, base_out;
        double scale;
        int esh_in, esh_out;          // log2 of the element split (31: none)
        long eseg_in, eseg_out;
        int use_tab;                  // output segments start at segtab[r] (peer receive buffers) instead of r * eseg_out
        long segtab[FFT_SEGTAB];
    };

    __device__ __forceinline__ double2 cmul(double2 a, double2 b)
    { return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
    __device__ __forceinline__ double2 cadd(double2 a, double2 b) { return make_double2(a.x + b.x, a.y + b.y); }
    __device__ __forceinline__ double2 csub(double2 a, double2 b) { return make_double2(a.x - b.x, a.y - b.y); }
    template<int DIR> __device__ __forceinline__ double2 rot(double2 a)      // * (-i) fwd, * (+i) bwd
    { return DIR < 0 ? make_double2(a.y, -a.x) : make_double2(-a.y, a.x); }
    template<int DIR> __device__ __forceinline__ double2 w8_1(double2 a)     // * W8^1
    { return DIR < 0 ? make_double2(kS * (a.x + a.y), kS * (a.y - a.x)) : make_double2(kS * (a.x - a.y), kS * (a.x + a.y)); }
    template<int DIR> __device__ __forceinline__ double2 w8_3(double2 a)     // * W8^3
    { return DIR < 0 ? make_double2(kS * (a.y - a.x), kS * (-a.x - a.y)) : make_double2(kS * (-a.x - a.y), kS * (a.x - a.y)); }
    template<int DIR> __device__ __forceinline__ double2 ldtw(const double2* __restrict__ t, int idx)
    {
        const double2 w = __ldg(t + idx);
        return DIR < 0 ? w : make_double2(w.x, -w.y);
    }

    __device__ __forceinline__ int pad(int n) { return n + (n >> 3); }

    // DFT of size R on v[b + NB*r] (r < R), NB = 8/R butterflies interleaved in the 8 registers.
    template<int DIR> __device__ __forceinline__ void dft8(double2* v)
    {
        const double2 b0 = cadd(v[0], v[4]), b4 = csub(v[0], v[4]);
        const double2 b1 = cadd(v[1], v[5]), b5 = w8_1<DIR>(csub(v[1], v[5]));
        const double2 b2 = cadd(v[2], v[6]), b6 = rot<DIR>(csub(v[2], v[6]));
        const double2 b3 = cadd(v[3], v[7]), b7 = w8_3<DIR>(csub(v[3], v[7]));
        const double2 c0 = cadd(b0, b2), c2 = csub(b0, b2), c1 = cadd(b1, b3), c3 = rot<DIR>(csub(b1, b3));
        const double2 c4 = cadd(b4, b6), c6 = csub(b4, b6), c5 = cadd(b5, b7), c7 = rot<DIR>(csub(b5, b7));
        v[0] = cadd(c0, c1); v[4] = csub(c0, c1); v[2] = cadd(c2, c3); v[6] = csub(c2, c3);
        v[1] = cadd(c4, c5); v[5] = csub(c4, c5); v[3] = cadd(c6, c7); v[7] = csub(c6, c7);
    }
    template<int DIR> __device__ __forceinline__ void dft4(double2& v0, double2& v1, double2& v2, double2& v3)
    {
        const double2 a = cadd(v0, v2), b = csub(v0, v2), c = cadd(v1, v3), d = rot<DIR>(csub(v1, v3));
        v0 = cadd(a, c); v2 = csub(a, c); v1 = cadd(b, d); v3 = csub(b, d);
    }

    // One Stockham stage of radix R at sub-transform length ns (log2 = lns): twiddle, butterfly,
    // and scatter of the 8 register values to their output slots in z (this pair's buffer).
    template<int R, int DIR, int ZSH = 0>
    __device__ __forceinline__ void stage(double2* v, double2* __restrict__ z, const double2* __restrict__ tw,
                                          const int tf, const int n8, const int lns, const bool first)
    {
        constexpr int NB = 8 / R;
        const int ns = 1 << lns;
        #pragma unroll
        for (int b = 0; b < NB; ++b)
        {
            const int j = tf + b * n8;
            const int k = j & (ns - 1);
            if (!first)
            {
                #pragma unroll
                for (int r = 1; r < R; ++r)
                    v[b + NB * r] = cmul(v[b + NB * r], ldtw<DIR>(tw, (r - 1) * ns + k));
            }
        }
        if (R == 8) dft8<DIR>(v);
        else if (R == 4) { dft4<DIR>(v[0], v[2], v[4], v[6]); dft4<DIR>(v[1], v[3], v[5], v[7]); }
        else
        {
            #pragma unroll
            for (int b = 0; b < 4; ++b)
            {
                const double2 a = v[b], c = v[b + 4];
                v[b] = cadd(a, c); v[b + 4] = csub(a, c);
            }
        }
        #pragma unroll
        for (int b = 0; b < NB; ++b)
        {
            const int j = tf + b * n8;
            const int k = j & (ns - 1);
            const int j0 = ((j >> lns) << lns) * R + k;
            #pragma unroll
            for (int q = 0; q < R; ++q)
                z[pad(j0 + (q << lns)) << ZSH] = v[b + NB * q];
        }
    }

    // Runs all stages. On return the transform is in registers in "position order": v[m] is element
    // tf + m*N/8 of the result (the last stage's output slots coincide with the read slots).
    template<int DIR, int ZSH = 0>
    __device__ __forceinline__ void run(double2* v, double2* __restrict__ z, const P2Args& A, const int tf)
    {
        const int n8 = A.n >> 3;
        int lns = 0;
        for (int s = 0; s < A.nst; ++s)
        {
            const int R = A.rad[s];
            const bool last = (s == A.nst - 1);
            if (last)
            {
                // twiddle + butterfly only; outputs stay in registers (slot of v[b+NB*q] is tf + (b+NB*q)*N/8)
                const int ns = 1 << lns;
                const double2* tw = A.tw + A.off[s];
                if (R == 8)
                {
                    if (s > 0) { const int k = tf & (ns - 1);
                        #pragma unroll
                        for (int r = 1; r < 8; ++r) v[r] = cmul(v[r], ldtw<DIR>(tw, (r - 1) * ns + k)); }
                    dft8<DIR>(v);
                }
                else if (R == 4)
                {
                    #pragma unroll
                    for (int b = 0; b < 2; ++b)
                    {
                        const int k = (tf + b * n8) & (ns - 1);
                        if (s > 0) {
                            #pragma unroll
                            for (int r = 1; r < 4; ++r) v[b + 2 * r] = cmul(v[b + 2 * r], ldtw<DIR>(tw, (r - 1) * ns + k)); }
                    }
                    dft4<DIR>(v[0], v[2], v[4], v[6]); dft4<DIR>(v[1], v[3], v[5], v[7]);
                }
                else
                {
                    #pragma unroll
                    for (int b = 0; b < 4; ++b)
                    {
                        const int k = (tf + b * n8) & (ns - 1);
                        if (s > 0) v[b + 4] = cmul(v[b + 4], ldtw<DIR>(tw, k));
                        const double2 a = v[b], c = v[b + 4];
                        v[b] = cadd(a, c); v[b + 4] = csub(a, c);
                    }
                }
            }
            else
            {
                if (s > 0) __syncthreads();      // everyone has read its inputs of this stage
                const double2* tw = A.tw + A.off[s];
                if (R == 8) stage<8, DIR, ZSH>(v, z, tw, tf, n8, lns, s == 0);
                else if (R == 4) stage<4, DIR, ZSH>(v, z, tw, tf, n8, lns, s == 0);
                else stage<2, DIR, ZSH>(v, z, tw, tf, n8, lns, s == 0);
                __syncthreads();
                #pragma unroll
                for (int m = 0; m < 8; ++m) v[m] = z[pad(tf + m * n8) << ZSH];
            }
            lns += (R == 8) ? 3 : (R == 4 ? 2 : 1);
        }
    }

    __device__ __forceinline__ long line_addr(int l, int n1, long s1, long s2)
    {
        return (long)(l % n1) * s1 + (long)(l / n1) * s2;
    }

    // element offset within a line; SPLIT: two-level index of the slab chunk layouts (FftLayout, fft.h)
    template<bool SPLIT> __device__ __forceinline__ long elem_off(int m, long sn, int esh, long eseg)
    {
        if (!SPLIT) return (long)m * sn;
        return (long)(m >> esh) * eseg + (long)(m & ((1 << esh) - 1)) * sn;
    }
    // output side: the segments may be the peers' receive buffers (P2Args::segtab)
    template<bool SPLIT> __device__ __forceinline__ long elem_off_out(int m, long sn, const P2Args& A)
    {
        if (!SPLIT) return (long)m * sn;
        const int r = m >> A.esh_out;
        const long seg = A.use_tab ? A.segtab[r & (FFT_SEGTAB - 1)] : (long)r * A.eseg_out;
        return seg + (long)(m & ((1 << A.esh_out) - 1)) * sn;
    }

    // DIR = -1: real lines -> half-complex; +1: half-complex -> real lines (unnormalised * scale).
    // STRIDED: element stride sn != 1 (lines adjacent in memory): tiles are staged through shared memory.
    // NG: pair groups per CTA.  The CTA stages 2*npairs*NG lines at once (so a strided row of the tile is
    //     16*NG contiguous doubles for n = 512) and transforms them group after group with the same threads.
    // The cooperative tile loops use only shifts and ONE line-address computation per thread: BLOCK is a
    // multiple of the (power-of-two) line count, so a thread always serves the same line (ncu r01: the y passes
    // sat at 16 % of DRAM throughput behind per-element integer div/mod and 64-byte rows).
    template<int DIR, bool STRIDED, int NG, bool SPLIT>
    __global__ void __launch_bounds__(BLOCK, 3)
    fft_pow2_kernel(const double* __restrict__ in, double* __restrict__ out, const P2Args A)
    {
        extern __shared__ double2 zbuf[];
        const int n = A.n, n8 = n >> 3, nh = n >> 1;
        const int tfn = n8;                              // threads per complex transform
        const int npairs = BLOCK / tfn;                  // complex transforms per group
        const int lines = 2 * npairs * NG;
        const int NP = pad(n) + 1;                       // slots per pair (odd skew between pairs)
        const int p = threadIdx.x / tfn;
        const int tf = threadIdx.x - p * tfn;
        double* zr = reinterpret_cast<double*>(zbuf);
        const int l0 = blockIdx.x * lines;
        double2 v[8];

        // ---------------- cooperative tile load (strided passes only) ----------------
        if (STRIDED)
        {
            // thread -> fixed line ll (lines <= BLOCK; for n = 8 a thread serves lines ll, ll + BLOCK),
            // elements m = m0, m0 + BLOCK/lines, ...
            const int lw = lines < BLOCK ? lines : BLOCK;
            const int llog = 31 - __clz(lw);
            const int mstep = BLOCK >> llog;
            for (int ll = threadIdx.x & (lw - 1); ll < lines; ll += BLOCK)
            {
                const int l = l0 + ll;
                double* dst = zr + 2 * ((ll >> 1) * NP) + (ll & 1);
                if (l < A.nlines)
                {
                    const double* src = in + A.base_in + line_addr(l, A.n1, A.s1, A.s2);
                    #pragma unroll 8
                    for (int m = threadIdx.x >> llog; m < n; m += mstep)
                        dst[2 * pad(m)] = src[elem_off<SPLIT>(m, A.sn, A.esh_in, A.eseg_in)];
                }
                else
                    for (int m = threadIdx.x >> llog; m < n; m += mstep) dst[2 * pad(m)] = 0.;
            }
            __syncthreads();
        }

        #pragma unroll 1
        for (int g = 0; g < NG; ++g)
        {
            double2* z = zbuf + (g * npairs + p) * NP;
            const int la = l0 + 2 * (g * npairs + p), lb = la + 1;
            const bool va = la < A.nlines, vb = lb < A.nlines;

            if (STRIDED)
            {
                if (DIR < 0)
                {
                    #pragma unroll
                    for (int m = 0; m < 8; ++m) v[m] = z[pad(tf + m * n8)];
                }
                else
                {
                    // Z_n = A_n + i B_n rebuilt from the two half-complex lines
                    #pragma unroll
                    for (int m = 0; m < 8; ++m)
                    {
                        const int pos = tf + m * n8;
                        const int np = pos <= nh ? pos : n - pos;
                        const bool self = (np == 0) || (np == nh);
                        const double sg = pos <= nh ? 1. : -1.;
                        const double2 re = z[pad(np)];
                        const double2 im = self ? make_double2(0., 0.) : z[pad(n - np)];
                        v[m] = make_double2(re.x - sg * im.y, sg * im.x + re.y);
                    }
                }
                __syncthreads();
            }
            else if (DIR < 0)
            {
                const double* pa = in + A.base_in + line_addr(va ? la : 0, A.n1, A.s1, A.s2);
                const double* pb = in + A.base_in + line_addr(vb ? lb : 0, A.n1, A.s1, A.s2);
                #pragma unroll
                for (int m = 0; m < 8; ++m)
                {
                    const int pos = tf + m * n8;
                    const long eo = elem_off<SPLIT>(pos, 1, A.esh_in, A.eseg_in);
                    v[m] = make_double2(va ? pa[eo] : 0., vb ? pb[eo] : 0.);
                }
            }
            else
            {
                // contiguous half-complex lines: Z_pos = A_pos + i B_pos rebuilt straight from global memory; the
                // mirrored element n - pos is a second (L1-resident) load instead of a shared-memory round trip
                const double* pa = in + A.base_in + line_addr(va ? la : 0, A.n1, A.s1, A.s2);
                const double* pb = in + A.base_in + line_addr(vb ? lb : 0, A.n1, A.s1, A.s2);
                double a1[8], a2[8], b1[8], b2[8];
                #pragma unroll
                for (int m = 0; m < 8; ++m)
                {
                    const int pos = tf + m * n8;
                    const int mir = (n - pos) & (n - 1);
                    const long ep = elem_off<SPLIT>(pos, 1, A.esh_in, A.eseg_in), em = elem_off<SPLIT>(mir, 1, A.esh_in, A.eseg_in);
                    a1[m] = va ? pa[ep] : 0.; a2[m] = va ? pa[em] : 0.;
                    b1[m] = vb ? pb[ep] : 0.; b2[m] = vb ? pb[em] : 0.;
                }
                #pragma unroll
                for (int m = 0; m < 8; ++m)
                {
                    const int pos = tf + m * n8;
                    if (pos == 0 || pos == nh) v[m] = make_double2(a1[m], b1[m]);
                    else if (pos < nh)         v[m] = make_double2(a1[m] - b2[m], a2[m] + b1[m]);
                    else                       v[m] = make_double2(a2[m] + b1[m], b2[m] - a1[m]);
                }
            }

            run<DIR>(v, z, A, tf);

            // ---------------- result back to shared memory / straight to global ----------------
            if (DIR > 0)
            {
                if (!STRIDED)
                {
                    double* qa = out + A.base_out + line_addr(va ? la : 0, A.n1o, A.s1o, A.s2o);
                    double* qb = out + A.base_out + line_addr(vb ? lb : 0, A.n1o, A.s1o, A.s2o);
                    #pragma unroll
                    for (int m = 0; m < 8; ++m)
                    {
                        const int pos = tf + m * n8;
                        const long eo = elem_off_out<SPLIT>(pos, 1, A);
                        if (va) qa[eo] = v[m].x * A.scale;
                        if (vb) qb[eo] = v[m].y * A.scale;
                    }
                    continue;
                }
                __syncthreads();
                #pragma unroll
                for (int m = 0; m < 8; ++m) z[pad(tf + m * n8)] = v[m];
                __syncthreads();
            }
            else
            {
                // result to natural order in shared memory, then the Hermitian split
                if (A.nst > 1) __syncthreads();
                #pragma unroll
                for (int m = 0; m < 8; ++m) z[pad(tf + m * n8)] = v[m];
                __syncthreads();
                double* qa = nullptr; double* qb = nullptr;
                if (!STRIDED)
                {
                    qa = out + A.base_out + line_addr(va ? la : 0, A.n1o, A.s1o, A.s2o);
                    qb = out + A.base_out + line_addr(vb ? lb : 0, A.n1o, A.s1o, A.s2o);
                }
                // k = tf + m*n8 for m < 4 covers 0..n/2-1; thread 0 also does k = n/2
                #pragma unroll
                for (int m = 0; m < 5; ++m)
                {
                    if (m == 4 && tf != 0) break;
                    const int k = (m == 4) ? nh : tf + m * n8;
                    const int kn = (k == 0) ? 0 : n - k;
                    const double2 zk = z[pad(k)], zn = z[pad(kn)];
                    const double are = 0.5 * (zk.x + zn.x), aim = 0.5 * (zk.y - zn.y);
                    const double bre = 0.5 * (zk.y + zn.y), bim = -0.5 * (zk.x - zn.x);
                    const bool has_im = (k > 0) && (k < nh);
                    if (!STRIDED)
                    {
                        const long ek = elem_off_out<SPLIT>(k, 1, A), en = elem_off_out<SPLIT>(n - k, 1, A);
                        if (va) { qa[ek] = are * A.scale; if (has_im) qa[en] = aim * A.scale; }
                        if (vb) { qb[ek] = bre * A.scale; if (has_im) qb[en] = bim * A.scale; }
                    }
                    else
                    {
                        z[pad(k)] = make_double2(are, bre);
                        if (has_im) z[pad(n - k)] = make_double2(aim, bim);
                    }
                }
                if (STRIDED) __syncthreads();
            }
        }
        if (!STRIDED) return;

        // ---------------- cooperative coalesced tile store (STRIDED only) ----------------
        {
            const int lw = lines < BLOCK ? lines : BLOCK;
            const int llog = 31 - __clz(lw);
            const int mstep = BLOCK >> llog;
            for (int ll = threadIdx.x & (lw - 1); ll < lines; ll += BLOCK)
            {
                const int l = l0 + ll;
                if (l >= A.nlines) continue;
                double* dst = out + A.base_out + line_addr(l, A.n1o, A.s1o, A.s2o);
                const double* src = zr + 2 * ((ll >> 1) * NP) + (ll & 1);
                #pragma unroll 8
                for (int m = threadIdx.x >> llog; m < n; m += mstep)
                    dst[elem_off_out<SPLIT>(m, A.sno, A)] = src[2 * pad(m)] * A.scale;
            }
        }
    }

    // Strided (y) passes with the LANES ALONG THE LINES.  Two adjacent lines (i, i+1) are one complex transform whose
    // elements are 16-byte (double2) words of the array itself, so a thread loads its eight elements straight into
    // registers and stores its results straight from them; 2^CPL such pairs ride side by side in the CTA (thread =
    // pair c + 2^CPL * tf), a row of the access is 2^CPL * 16 contiguous bytes, and the Stockham exchanges address
    // shared memory as [slot][pair]: consecutive lanes hit consecutive 16-byte words.  Against the staged-tile kernel
    // above this drops both tile passes through shared memory (5 -> 3 round trips forward, 4 -> 2 backward).
    // Needs: s1 = s1o = 1, n1 = n1o a multiple of 2^(CPL+1), even strides and bases (fft_pow2_lines checks).
    constexpr int CBLOCK = 512;
    template<int DIR, int CPL, bool SPLIT>
    __global__ void __launch_bounds__(CBLOCK, 2)
    fft_pow2_col_kernel(const double* __restrict__ in, double* __restrict__ out, const P2Args A)
    {
        extern __shared__ double2 zbuf[];
        constexpr int CP = 1 << CPL;
        const int n = A.n, n8 = n >> 3, nh = n >> 1;
        const int c = threadIdx.x & (CP - 1);
        const int tf = threadIdx.x >> CPL;
        const int l0 = blockIdx.x * (2 * CP);            // first line of this CTA; its 2 CP lines are adjacent in memory
        const double* src = in + A.base_in + line_addr(l0, A.n1, A.s1, A.s2) + 2 * c;
        double* dst = out + A.base_out + line_addr(l0, A.n1o, A.s1o, A.s2o) + 2 * c;
        double2* z = zbuf + c;
        double2 v[8];

        if (DIR < 0)
        {
            #pragma unroll
            for (int m = 0; m < 8; ++m)
                v[m] = *reinterpret_cast<const double2*>(src + elem_off<SPLIT>(tf + m * n8, A.sn, A.esh_in, A.eseg_in));
        }
        else
        {
            // Z_pos = A_pos + i B_pos rebuilt from the two half-complex lines: element np holds the real parts (a, b),
            // element n - np the imaginary parts; the mirrored word is a second (L1/L2-resident) load
            #pragma unroll
            for (int m = 0; m < 8; ++m)
            {
                const int pos = tf + m * n8;
                const int np = pos <= nh ? pos : n - pos;
                const bool self = (np == 0) || (np == nh);
                const double sg = pos <= nh ? 1. : -1.;
                const double2 re = *reinterpret_cast<const double2*>(src + elem_off<SPLIT>(np, A.sn, A.esh_in, A.eseg_in));
                double2 im = make_double2(0., 0.);
                if (!self) im = *reinterpret_cast<const double2*>(src + elem_off<SPLIT>(n - np, A.sn, A.esh_in, A.eseg_in));
                v[m] = make_double2(re.x - sg * im.y, sg * im.x + re.y);
            }
        }

        run<DIR, CPL>(v, z, A, tf);

        if (DIR > 0)
        {
            #pragma unroll
            for (int m = 0; m < 8; ++m)
                *reinterpret_cast<double2*>(dst + elem_off_out<SPLIT>(tf + m * n8, A.sno, A))
                    = make_double2(v[m].x * A.scale, v[m].y * A.scale);
        }
        else
        {
            // result to natural order in shared memory, then the Hermitian split; (a, b) leave as one 16-byte word
            if (A.nst > 1) __syncthreads();
            #pragma unroll
            for (int m = 0; m < 8; ++m) z[pad(tf + m * n8) << CPL] = v[m];
            __syncthreads();
            #pragma unroll
            for (int m = 0; m < 5; ++m)
            {
                if (m == 4 && tf != 0) break;
                const int k = (m == 4) ? nh : tf + m * n8;
                const int kn = (k == 0) ? 0 : n - k;
                const double2 zk = z[pad(k) << CPL], zn = z[pad(kn) << CPL];
                const double are = 0.5 * (zk.x + zn.x), aim = 0.5 * (zk.y - zn.y);
                const double bre = 0.5 * (zk.y + zn.y), bim = -0.5 * (zk.x - zn.x);
                *reinterpret_cast<double2*>(dst + elem_off_out<SPLIT>(k, A.sno, A))
                    = make_double2(are * A.scale, bre * A.scale);
                if (k > 0 && k < nh)
                    *reinterpret_cast<double2*>(dst + elem_off_out<SPLIT>(n - k, A.sno, A))
                        = make_double2(aim * A.scale, bim * A.scale);
            }
        }
    }

    // Pres_2::input (src/pres_2.cxx:183-195) fused with the forward x pass: the right-hand side of the Poisson equation
    //   p = rho [(ut + u/dt)_{i+1} - (ut + u/dt)_i] dxi + rho [(vt + v/dt)_{j+1} - (vt + v/dt)_j] dyi
    //       + [rhoh_{k+1} (wt + w/dt)_{k+1} - rhoh_k (wt + w/dt)_k] dzi
    // is evaluated straight into the registers of the transform that consumes it (two lines j, j+1 of one level per complex
    // transform), so the ghost-free RHS array is never written and read back (16 of the 72 B/pt of the two passes).  The
    // neighbour values (i+1: same line; j+1: the pair's other line or the next CTA's; k+1: the plane the CTAs of the next
    // level read a moment later) are L1 / L2 hits.  Same expression as pres2_in_kernel, same transform: bit-identical.
    struct PresInArgs
    {
        const double* u; const double* v; const double* w;
        const double* ut; const double* vt; const double* wt;
        double dti;
        int k0;              // first level of this call (y-slab k ranges)
        GridD gd;
    };

    template<bool SPLIT>
    __global__ void __launch_bounds__(BLOCK, 3)
    pres2_in_fft_x_kernel(const PresInArgs F, double* __restrict__ out, const P2Args A)
    {
        extern __shared__ double2 zbuf[];
        const GridD& gd = F.gd;
        const int n = A.n, n8 = n >> 3, nh = n >> 1;
        const int tfn = n8;
        const int npairs = BLOCK / tfn;
        const int NP = pad(n) + 1;
        const int p = threadIdx.x / tfn;
        const int tf = threadIdx.x - p * tfn;
        double2* z = zbuf + p * NP;
        const int la = blockIdx.x * (2 * npairs) + 2 * p, lb = la + 1;    // jmax is even: both lines lie in one level
        const bool va = la < A.nlines, vb = lb < A.nlines;
        double2 v[8];
        {
            const int ls = va ? la : 0;
            const int j = ls % gd.jmax, k = F.k0 + ls / gd.jmax;
            const int jj = gd.icells, kk = gd.ijcells;
            const long row = (long)gd.igc + (long)(j + gd.jgc) * jj + (long)(k + gd.kgc) * kk;
            const double rho = gd.rhoref[k + gd.kgc], rhoh_t = gd.rhorefh[k + gd.kgc + 1], rhoh_b = gd.rhorefh[k + gd.kgc];
            const double dzi = gd.dzi[k + gd.kgc];
            const double dxi = 1. / gd.dx, dyi = 1. / gd.dy, dti = F.dti;
            const double* __restrict__ u = F.u + row;   const double* __restrict__ ut = F.ut + row;
            const double* __restrict__ vv = F.v + row;  const double* __restrict__ vt = F.vt + row;
            const double* __restrict__ w = F.w + row;   const double* __restrict__ wt = F.wt + row;
            #pragma unroll
            for (int m = 0; m < 8; ++m)
            {
                const int i = tf + m * n8;
                // line a (row j) and line b (row j + 1) share the v row between them
                const double fv0 = vt[i] + vv[i] * dti, fv1 = vt[i + jj] + vv[i + jj] * dti, fv2 = vt[i + 2 * jj] + vv[i + 2 * jj] * dti;
                const double pa = rho * ((ut[i + 1] + u[i + 1] * dti) - (ut[i] + u[i] * dti)) * dxi
                                + rho * (fv1 - fv0) * dyi
                                + (rhoh_t * (wt[i + kk] + w[i + kk] * dti) - rhoh_b * (wt[i] + w[i] * dti)) * dzi;
                const double pb = rho * ((ut[i + jj + 1] + u[i + jj + 1] * dti) - (ut[i + jj] + u[i + jj] * dti)) * dxi
                                + rho * (fv2 - fv1) * dyi
                                + (rhoh_t * (wt[i + jj + kk] + w[i + jj + kk] * dti) - rhoh_b * (wt[i + jj] + w[i + jj] * dti)) * dzi;
                v[m] = make_double2(va ? pa : 0., vb ? pb : 0.);
            }
        }

        run<-1>(v, z, A, tf);

        // result to natural order in shared memory, then the Hermitian split (as fft_pow2_kernel<-1, false>)
        if (A.nst > 1) __syncthreads();
        #pragma unroll
        for (int m = 0; m < 8; ++m) z[pad(tf + m * n8)] = v[m];
        __syncthreads();
        double* qa = out + A.base_out + line_addr(va ? la : 0, A.n1o, A.s1o, A.s2o);
        double* qb = out + A.base_out + line_addr(vb ? lb : 0, A.n1o, A.s1o, A.s2o);
        #pragma unroll
        for (int m = 0; m < 5; ++m)
        {
            if (m == 4 && tf != 0) break;
            const int k = (m == 4) ? nh : tf + m * n8;
            const int kn = (k == 0) ? 0 : n - k;
            const double2 zk = z[pad(k)], zn = z[pad(kn)];
            const double are = 0.5 * (zk.x + zn.x), aim = 0.5 * (zk.y - zn.y);
            const double bre = 0.5 * (zk.y + zn.y), bim = -0.5 * (zk.x - zn.x);
            const bool has_im = (k > 0) && (k < nh);
            const long ek = elem_off_out<SPLIT>(k, 1, A), en = elem_off_out<SPLIT>(n - k, 1, A);
            if (va) { qa[ek] = are * A.scale; if (has_im) qa[en] = aim * A.scale; }
            if (vb) { qb[ek] = bre * A.scale; if (has_im) qb[en] = bim * A.scale; }
        }
    }
}  // namespace

bool fft_pow2_supported(int n) { return n >= 8 && n <= 8 * BLOCK && (n & (n - 1)) == 0; }

int fft_pow2_plan(FftPlan* p)
{
    const int n = p->n;
    if (!fft_pow2_supported(n)) return MHH_OK;
    int m = n, nst = 0;
    while (m % 8 == 0 && m > 1) { p->p2_rad[nst++] = 8; m /= 8; }
    if (m == 4) { p->p2_rad[nst++] = 4; m = 1; }
    if (m == 2) { p->p2_rad[nst++] = 2; m = 1; }
    p->p2_nst = nst;
    std::vector<double2> tw;
    const mhh_f80 pi = std::acos(-1.0L);
    int ns = 1;
    for (int s = 0; s < nst; ++s)
    {
        const int R = p->p2_rad[s];
        p->p2_off[s] = (int)tw.size();
        for (int r = 1; r < R; ++r)
            for (int k = 0; k < ns; ++k)
            {
                const mhh_f80 ang = -2.0L * pi * (mhh_f80)k * (mhh_f80)r / ((mhh_f80)ns * R);
                tw.push_back(make_double2((double)std::cos(ang), (double)std::sin(ang)));
            }
        ns *= R;
    }
    MHH_CUDA_CHECK(cudaMalloc(&p->p2_tw, sizeof(double2) * tw.size()));
    MHH_CUDA_CHECK(cudaMemcpy(p->p2_tw, tw.data(), sizeof(double2) * tw.size(), cudaMemcpyHostToDevice));
    return MHH_OK;
}

// Pres_2::input + forward x transform in one pass (pres2_in_fft_x_kernel).  L describes the OUTPUT side as for a forward
// fft_lines call on the ghost-free layout [k][j < jmax][i < itot] (or the slab chunk layout); k0: first level of the call.
// Returns MHH_ERR_UNSUPPORTED when the fused pass does not apply (the caller then runs the two separate stages).
int pres2_in_fft_x(Ctx* c, const FftPlan& plan, const double* u, const double* v, const double* w,
                   const double* ut, const double* vt, const double* wt, double dt, double* out, const FftLayout& L, int k0)
{
    const GridD& gd = c->gd;
    if (!c->fuse_pres_input || plan.p2_nst <= 0 || plan.n != gd.itot || (gd.jmax & 1) || L.sno != 1 || c->order != 2 || plan.n / 8 > BLOCK)
        return MHH_ERR_UNSUPPORTED;
    if (L.esplit_out > 0 && (L.esplit_out & (L.esplit_out - 1))) return MHH_ERR_UNSUPPORTED;
    P2Args A;
    A.n = plan.n; A.nst = plan.p2_nst;
    A.log2n = 0; while ((1 << A.log2n) < plan.n) ++A.log2n;
    for (int s = 0; s < plan.p2_nst; ++s) { A.rad[s] = plan.p2_rad[s]; A.off[s] = plan.p2_off[s]; }
    A.tw = plan.p2_tw;
    A.nlines = L.nlines; A.n1 = L.n1; A.s1 = L.s1; A.s2 = L.s2; A.sn = L.sn; A.base_in = L.base_in;
    A.n1o = L.n1o; A.s1o = L.s1o; A.s2o = L.s2o; A.sno = L.sno; A.base_out = L.base_out;
    A.scale = L.scale;
    auto lg = [](int x) { int l = 0; while ((1 << l) < x) ++l; return l; };
    A.esh_in = 31; A.eseg_in = 0;
    A.esh_out = L.esplit_out > 0 ? lg(L.esplit_out) : 31; A.eseg_out = L.eseg_out;
    A.use_tab = L.use_segtab_out ? 1 : 0;
    for (int r = 0; r < FFT_SEGTAB; ++r) A.segtab[r] = L.segtab_out[r];
    PresInArgs F{u, v, w, ut, vt, wt, 1. / dt, k0, gd};
    const int npairs = BLOCK / (plan.n / 8);
    const size_t smem = (size_t)npairs * ((plan.n + plan.n / 8) + 1) * sizeof(double2);
    const int nblocks = (L.nlines + 2 * npairs - 1) / (2 * npairs);
    if (L.esplit_out > 0)
    {
        static bool attr = false;
        if (!attr) { cudaFuncSetAttribute(pres2_in_fft_x_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 208 * 1024); attr = true; }
        MHH_LAUNCH(c, "pres2_in_fft_x", (pres2_in_fft_x_kernel<true><<<nblocks, BLOCK, smem, c->stream>>>(F, out, A)));
    }
    else
    {
        static bool attr = false;
        if (!attr) { cudaFuncSetAttribute(pres2_in_fft_x_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 208 * 1024); attr = true; }
        MHH_LAUNCH(c, "pres2_in_fft_x", (pres2_in_fft_x_kernel<false><<<nblocks, BLOCK, smem, c->stream>>>(F, out, A)));
    }
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

int fft_pow2_lines(Ctx* c, const FftPlan& plan, int dir, const double* in, double* out, const FftLayout& L)
{
    P2Args A;
    A.n = plan.n; A.nst = plan.p2_nst;
    A.log2n = 0; while ((1 << A.log2n) < plan.n) ++A.log2n;
    for (int s = 0; s < plan.p2_nst; ++s) { A.rad[s] = plan.p2_rad[s]; A.off[s] = plan.p2_off[s]; }
    A.tw = plan.p2_tw;
    A.nlines = L.nlines; A.n1 = L.n1; A.s1 = L.s1; A.s2 = L.s2; A.sn = L.sn; A.base_in = L.base_in;
    A.n1o = L.n1o; A.s1o = L.s1o; A.s2o = L.s2o; A.sno = L.sno; A.base_out = L.base_out;
    A.scale = L.scale;
    const bool strided = (L.sn != 1) || (L.sno != 1);
    const bool split = L.esplit_in > 0 || L.esplit_out > 0;
    auto lg = [](int v) { int l = 0; while ((1 << l) < v) ++l; return l; };
    if ((L.esplit_in > 0 && (L.esplit_in & (L.esplit_in - 1))) || (L.esplit_out > 0 && (L.esplit_out & (L.esplit_out - 1))))
        return MHH_ERR_UNSUPPORTED;
    A.esh_in = L.esplit_in > 0 ? lg(L.esplit_in) : 31;   A.eseg_in = L.eseg_in;
    A.esh_out = L.esplit_out > 0 ? lg(L.esplit_out) : 31; A.eseg_out = L.eseg_out;
    A.use_tab = L.use_segtab_out ? 1 : 0;
    for (int r = 0; r < FFT_SEGTAB; ++r) A.segtab[r] = L.segtab_out[r];
    if (L.use_segtab_out && (L.esplit_out <= 0 || plan.n / L.esplit_out > FFT_SEGTAB)) return MHH_ERR_UNSUPPORTED;
    // y passes whose lines are adjacent words: the lanes-along-the-lines kernel (n = 256, 512, 1024: 16, 8, 4 pairs per CTA)
    {
        static const bool col_off = []() { const char* e = getenv("MHH_FFT_NOCOL"); return e && atoi(e) != 0; }();
        const int tfn = plan.n / 8;
        const int cp = tfn > 0 ? CBLOCK / tfn : 0;
        auto even = [](long v) { return (v & 1L) == 0; };
        bool tab_even = true;
        if (L.use_segtab_out) for (int r = 0; r < FFT_SEGTAB; ++r) tab_even = tab_even && even(L.segtab_out[r]);
        const bool aligned = even(L.s2) && even(L.sn) && even(L.base_in) && even(L.s2o) && even(L.sno) && even(L.base_out)
                          && even(L.eseg_in) && even(L.eseg_out)
                          && tab_even
                          && (reinterpret_cast<uintptr_t>(in) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0;
        if (!col_off && strided && (cp == 4 || cp == 8 || cp == 16) && L.s1 == 1 && L.s1o == 1 && L.n1 == L.n1o
            && L.n1 % (2 * cp) == 0 && L.nlines % (2 * cp) == 0 && aligned)
        {
            const size_t smem_c = (size_t)(cp * (plan.n + plan.n / 8 + 1)) * sizeof(double2);
            const int nb = L.nlines / (2 * cp);
            const char* nm = dir < 0 ? "fft_fwd_y" : "fft_bwd_y";
            #define P2C_LAUNCH(D, CPL, SP) do { \
                static bool attr = false; \
                if (!attr) { cudaFuncSetAttribute(fft_pow2_col_kernel<D, CPL, SP>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024); attr = true; } \
                MHH_LAUNCH(c, nm, (fft_pow2_col_kernel<D, CPL, SP><<<nb, CBLOCK, smem_c, c->stream>>>(in, out, A))); } while (0)
            #define P2C_DISPATCH(D, SP) do { if (cp == 4) P2C_LAUNCH(D, 2, SP); else if (cp == 8) P2C_LAUNCH(D, 3, SP); else P2C_LAUNCH(D, 4, SP); } while (0)
            if (dir < 0) { if (split) P2C_DISPATCH(-1, true); else P2C_DISPATCH(-1, false); }
            else         { if (split) P2C_DISPATCH(1, true);  else P2C_DISPATCH(1, false); }
            #undef P2C_DISPATCH
            #undef P2C_LAUNCH
            MHH_LAUNCH_CHECK();
            return MHH_OK;
        }
    }
    const int npairs = BLOCK / (plan.n / 8);
    // strided (y) passes: several pair groups per CTA (wider rows of the staged tile) while the tile fits the shared-memory
    // budget and the grid still fills the machine; MHH_FFT_NG (test hook) caps the group count
    const size_t pair_bytes = (size_t)((plan.n + plan.n / 8) + 1) * sizeof(double2);
    int ng = 1;
    if (strided)
    {
        static const int ng_cap = []() { const char* e = getenv("MHH_FFT_NG"); return e ? atoi(e) : 2; }();
        for (int cand = 2; cand <= ng_cap && cand <= 4; cand *= 2)
            if ((size_t)cand * npairs * pair_bytes <= (cand == 2 ? 76 : 200) * 1024 && L.nlines >= 2 * npairs * cand * 296) ng = cand;
    }
    const int lines = 2 * npairs * ng;
    const size_t smem = (size_t)npairs * ng * pair_bytes;
    const int nblocks = (L.nlines + lines - 1) / lines;
    const char* name = dir < 0 ? (strided ? "fft_fwd_y" : "fft_fwd_x") : (strided ? "fft_bwd_y" : "fft_bwd_x");
    #define P2_LAUNCH(D, S, G, SP) do { \
        static bool attr = false; \
        if (!attr) { cudaFuncSetAttribute(fft_pow2_kernel<D, S, G, SP>, cudaFuncAttributeMaxDynamicSharedMemorySize, 208 * 1024); attr = true; } \
        MHH_LAUNCH(c, name, (fft_pow2_kernel<D, S, G, SP><<<nblocks, BLOCK, smem, c->stream>>>(in, out, A))); } while (0)
    #define P2_DISPATCH(SP) do { \
        if (dir < 0) { if (strided && ng == 4) P2_LAUNCH(-1, true, 4, SP); else if (strided && ng == 2) P2_LAUNCH(-1, true, 2, SP); else if (strided) P2_LAUNCH(-1, true, 1, SP); else P2_LAUNCH(-1, false, 1, SP); } \
        else         { if (strided && ng == 4) P2_LAUNCH(1, true, 4, SP);  else if (strided && ng == 2) P2_LAUNCH(1, true, 2, SP);  else if (strided) P2_LAUNCH(1, true, 1, SP);  else P2_LAUNCH(1, false, 1, SP); } } while (0)
    if (split) P2_DISPATCH(true); else P2_DISPATCH(false);
    #undef P2_DISPATCH
    #undef P2_LAUNCH
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

}  // namespace mhh
