/* TEST INFRASTRUCTURE ONLY (oracle). Minimal stand-in for <fftw3.h>.
 *
 * FFTW 3.3.x is a third-party dependency of the reference that is NOT vendored
 * under /root/reference (config/default.cmake pins 3.3.10 on the author's box) and
 * is not installed in this image.  The reference only uses the r2r "many" planner
 * with FFTW_R2HC / FFTW_HC2R (src/fft.cxx:140-154), fftw_execute (src/fft.cxx:48-51),
 * alloc/free and wisdom import/export.  This header declares exactly that surface;
 * fftw_shim.cxx implements it with a plain double-precision mixed-radix DFT that
 * emits FFTW's half-complex order r0,r1,...,r(n/2),i((n+1)/2-1),...,i1 and is
 * unnormalised in both directions, as FFTW documents.
 * Bit-level parity with real FFTW is therefore UNPINNED; tolerance-level parity is
 * pinned by the DFT definition (tests compare against numpy.fft as well).
 */
#ifndef MHH_ORACLE_FFTW3_SHIM_H
#define MHH_ORACLE_FFTW3_SHIM_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef enum { FFTW_R2HC = 0, FFTW_HC2R = 1 } fftw_r2r_kind;
typedef fftw_r2r_kind fftwf_r2r_kind;
#define FFTW_ESTIMATE (1U << 6)
#define FFTW_MEASURE (0U)
#define FFTW_EXHAUSTIVE (1U << 3)
#define FFTW_PATIENT (1U << 5)

struct mhh_shim_plan;
typedef struct mhh_shim_plan* fftw_plan;
typedef struct mhh_shim_plan* fftwf_plan;

double* fftw_alloc_real(size_t n);
float* fftwf_alloc_real(size_t n);
void fftw_free(void* p);
void fftwf_free(void* p);

fftw_plan fftw_plan_many_r2r(int rank, const int* n, int howmany,
                             double* in, const int* inembed, int istride, int idist,
                             double* out, const int* onembed, int ostride, int odist,
                             const fftw_r2r_kind* kind, unsigned flags);
fftwf_plan fftwf_plan_many_r2r(int rank, const int* n, int howmany,
                               float* in, const int* inembed, int istride, int idist,
                               float* out, const int* onembed, int ostride, int odist,
                               const fftwf_r2r_kind* kind, unsigned flags);
void fftw_execute(const fftw_plan p);
void fftwf_execute(const fftwf_plan p);
void fftw_destroy_plan(fftw_plan p);
void fftwf_destroy_plan(fftwf_plan p);
void fftw_cleanup(void);
void fftwf_cleanup(void);
void fftw_forget_wisdom(void);
void fftwf_forget_wisdom(void);
int fftw_import_wisdom_from_filename(const char* filename);
int fftwf_import_wisdom_from_filename(const char* filename);
int fftw_export_wisdom_to_filename(const char* filename);
int fftwf_export_wisdom_to_filename(const char* filename);

#ifdef __cplusplus
}
#endif
#endif
