/* shim: the six FFTW3 symbols donut-classifier/classifier.c uses (lines 533-589), backed by a plain
 * double-precision r2c DFT in fftw3_shim.c.  FFTW itself is not in this image; any correct double
 * DFT of length 256 agrees with it to ~1e-13.  Test infrastructure only. */
#ifndef SJ_SHIM_FFTW3_H
#define SJ_SHIM_FFTW3_H
#include <stddef.h>
typedef double fftw_complex[2];
typedef struct sj_fftw_plan_s *fftw_plan;
#define FFTW_ESTIMATE (1U << 6)
void *fftw_malloc(size_t n);
void fftw_free(void *p);
fftw_plan fftw_plan_dft_r2c_1d(int n, double *in, fftw_complex *out, unsigned flags);
void fftw_execute(const fftw_plan p);
void fftw_destroy_plan(fftw_plan p);
#endif
