// Batched real <-> half-complex line FFTs (fft.cu).
#pragma once
#include "sys.h"
#include "ctx.h"

namespace mhh
{
// Line l (0 <= l < nlines), element m (0 <= m < n):
//   input  address = base_in  + (l % n1 )*s1  + (l / n1 )*s2  + m*sn
//   output address = base_out + (l % n1o)*s1o + (l / n1o)*s2o + m*sno
// Optional two-level element index (the all-to-all chunk layouts of the y-slab solver, slab.cu): with esplit > 0 (a power
// of two) element m sits at (m / esplit)*eseg + (m % esplit)*sn instead of m*sn.  register-resident pow2 kernel only.
constexpr int FFT_SEGTAB = 32;    // output segments a line can be scattered to (ranks x kx sub-blocks)
struct FftLayout
{
    int nlines;
    int n1;  long s1, s2, sn;   long base_in;
    int n1o; long s1o, s2o, sno; long base_out;
    double scale;
    int esplit_in = 0;  long eseg_in = 0;
    int esplit_out = 0; long eseg_out = 0;
    // with use_segtab_out, output segment r = m / esplit_out starts at segtab_out[r] (elements, relative to `out`) instead of
    // r*eseg_out: the segments are the receive buffers of the peer ranks (peer.cu), the transform writes them directly
    bool use_segtab_out = false;
    long segtab_out[FFT_SEGTAB] = {};
};

int fft_plan_create(FftPlan* p, int n);
void fft_plan_destroy(FftPlan* p);
// dir = -1: real -> half-complex (FFTW_R2HC); +1: half-complex -> real (FFTW_HC2R); both unnormalised * scale.
int fft_lines(Ctx* c, const FftPlan& plan, int dir, const double* in, double* out, const FftLayout& L);
}
