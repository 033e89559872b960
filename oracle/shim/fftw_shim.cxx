/* TEST INFRASTRUCTURE ONLY (oracle). See fftw3.h in this directory.
 *
 * Implements the r2r R2HC / HC2R "many" interface the reference uses
 * (src/fft.cxx:140-154) with a recursive mixed-radix complex FFT in double
 * precision (twiddles from long-double sin/cos).  Definitions follow the FFTW
 * manual ("The 1d Real-data DFT", "The Halfcomplex-format DFT"):
 *   R2HC: Y_k = sum_j X_j exp(-2 pi i jk/n); out = r0..r(n/2), i((n+1)/2-1)..i1
 *   HC2R: the unnormalised inverse of the above.
 */
#include "fftw3.h"

#include <cmath>
#include <complex>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace
{
    using cplx = std::complex<double>;

    struct Engine
    {
        int n = 0;
        std::vector<cplx> tw;       // exp(-2 pi i k / n)
        std::vector<int> factors;
        std::vector<cplx> a, b;     // work buffers

        void setup(int nin)
        {
            n = nin;
            tw.resize(n);
            const long double pi = std::acos(-1.0L);
            for (int k = 0; k < n; ++k)
            {
                long double ang = -2.0L * pi * (long double)k / (long double)n;
                tw[k] = cplx((double)std::cos(ang), (double)std::sin(ang));
            }
            int m = n;
            for (int f : {4, 2, 3, 5, 7})
                while (m % f == 0) { factors.push_back(f); m /= f; }
            for (int f = 11; m > 1; f += 2)
                while (m % f == 0) { factors.push_back(f); m /= f; }
            a.resize(n);
            b.resize(n);
        }

        // Stockham autosort, forward sign (-). Input/outputs in a.
        void forward()
        {
            cplx* x = a.data();
            cplx* y = b.data();
            int ns = 1;
            std::vector<cplx> v(32);
            for (int R : factors)
            {
                if ((int)v.size() < R) v.resize(R);
                const int nb = n / R;
                for (int j = 0; j < nb; ++j)
                {
                    const int k = j % ns;
                    const int tstep = n / (ns * R);
                    for (int r = 0; r < R; ++r)
                        v[r] = x[j + r * nb] * tw[(size_t)((long)k * r * tstep) % n];
                    const int j0 = (j / ns) * ns * R + k;
                    const int rstep = n / R;
                    for (int q = 0; q < R; ++q)
                    {
                        cplx s = v[0];
                        for (int r = 1; r < R; ++r)
                            s += v[r] * tw[(size_t)((long)q * r * rstep) % n];
                        y[j0 + q * ns] = s;
                    }
                }
                std::swap(x, y);
                ns *= R;
            }
            if (x != a.data())
                std::memcpy(a.data(), x, sizeof(cplx) * n);
        }
    };
}

struct mhh_shim_plan
{
    int n, howmany;
    void* in;
    void* out;
    int istride, idist, ostride, odist;
    fftw_r2r_kind kind;
    bool single;
    Engine eng;
};

namespace
{
    template<typename T>
    void execute_plan(mhh_shim_plan* p)
    {
        const int n = p->n;
        Engine& e = p->eng;
        const T* in = static_cast<const T*>(p->in);
        T* out = static_cast<T*>(p->out);

        for (int b = 0; b < p->howmany; ++b)
        {
            const T* x = in + (size_t)b * p->idist;
            T* y = out + (size_t)b * p->odist;

            if (p->kind == FFTW_R2HC)
            {
                for (int j = 0; j < n; ++j)
                    e.a[j] = cplx((double)x[(size_t)j * p->istride], 0.);
                e.forward();
                for (int k = 0; k <= n / 2; ++k)
                    y[(size_t)k * p->ostride] = (T)e.a[k].real();
                for (int k = 1; k < (n + 1) / 2; ++k)
                    y[(size_t)(n - k) * p->ostride] = (T)e.a[k].imag();
            }
            else
            {
                // Build the Hermitian spectrum, conjugate, forward transform, take real part:
                // x_j = sum_k Y_k exp(+2 pi i jk/n) = conj( FFT_fwd( conj(Y) ) )_j
                e.a[0] = cplx((double)x[0], 0.);
                for (int k = 1; k < (n + 1) / 2; ++k)
                {
                    const double re = (double)x[(size_t)k * p->istride];
                    const double im = (double)x[(size_t)(n - k) * p->istride];
                    e.a[k] = cplx(re, -im);     // conj(Y_k)
                    e.a[n - k] = cplx(re, im);  // conj(Y_{n-k}) = Y_k
                }
                if (n % 2 == 0)
                    e.a[n / 2] = cplx((double)x[(size_t)(n / 2) * p->istride], 0.);
                e.forward();
                for (int j = 0; j < n; ++j)
                    y[(size_t)j * p->ostride] = (T)e.a[j].real();
            }
        }
    }

    mhh_shim_plan* make_plan(int rank, const int* n, int howmany, void* in, int istride, int idist,
                             void* out, int ostride, int odist, fftw_r2r_kind kind, bool single)
    {
        if (rank != 1) return nullptr;
        mhh_shim_plan* p = new mhh_shim_plan;
        p->n = n[0]; p->howmany = howmany; p->in = in; p->out = out;
        p->istride = istride; p->idist = idist; p->ostride = ostride; p->odist = odist;
        p->kind = kind; p->single = single;
        p->eng.setup(n[0]);
        return p;
    }
}

extern "C"
{
double* fftw_alloc_real(size_t n) { return (double*)std::malloc(n * sizeof(double)); }
float* fftwf_alloc_real(size_t n) { return (float*)std::malloc(n * sizeof(float)); }
void fftw_free(void* p) { std::free(p); }
void fftwf_free(void* p) { std::free(p); }

fftw_plan fftw_plan_many_r2r(int rank, const int* n, int howmany,
                             double* in, const int*, int istride, int idist,
                             double* out, const int*, int ostride, int odist,
                             const fftw_r2r_kind* kind, unsigned)
{ return make_plan(rank, n, howmany, in, istride, idist, out, ostride, odist, kind[0], false); }

fftwf_plan fftwf_plan_many_r2r(int rank, const int* n, int howmany,
                               float* in, const int*, int istride, int idist,
                               float* out, const int*, int ostride, int odist,
                               const fftwf_r2r_kind* kind, unsigned)
{ return make_plan(rank, n, howmany, in, istride, idist, out, ostride, odist, kind[0], true); }

void fftw_execute(const fftw_plan p) { execute_plan<double>(p); }
void fftwf_execute(const fftwf_plan p) { execute_plan<float>(p); }
void fftw_destroy_plan(fftw_plan p) { delete p; }
void fftwf_destroy_plan(fftwf_plan p) { delete p; }
void fftw_cleanup(void) {}
void fftwf_cleanup(void) {}
void fftw_forget_wisdom(void) {}
void fftwf_forget_wisdom(void) {}
int fftw_import_wisdom_from_filename(const char*) { return 1; }
int fftwf_import_wisdom_from_filename(const char*) { return 1; }
int fftw_export_wisdom_to_filename(const char*) { return 1; }
int fftwf_export_wisdom_to_filename(const char*) { return 1; }
}
