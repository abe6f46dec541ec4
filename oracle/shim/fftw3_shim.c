/* fftw3_shim.c -- see fftw3.h in this directory.  r2c DFT: X[k] = sum_n x[n] exp(-2 pi i k n / N),
 * k = 0..N/2, double precision, radix-2 for powers of two (long-double twiddles), naive otherwise. */
#include "fftw3.h"
#include <math.h>
#include <stdlib.h>

struct sj_fftw_plan_s { int n; double *in; fftw_complex *out; double *wr, *wi, *re, *im; };

void *fftw_malloc(size_t n) { return malloc(n); }
void fftw_free(void *p) { free(p); }

fftw_plan fftw_plan_dft_r2c_1d(int n, double *in, fftw_complex *out, unsigned flags)
{
    (void)flags;
    struct sj_fftw_plan_s *p = (struct sj_fftw_plan_s *)malloc(sizeof *p);
    p->n = n; p->in = in; p->out = out;
    p->wr = (double *)malloc(sizeof(double) * n); p->wi = (double *)malloc(sizeof(double) * n);
    p->re = (double *)malloc(sizeof(double) * n); p->im = (double *)malloc(sizeof(double) * n);
    for (int k = 0; k < n; k++) {
        long double a = -2.0L * 3.14159265358979323846264338327950288L * k / n;
        p->wr[k] = (double)cosl(a); p->wi[k] = (double)sinl(a);
    }
    return p;
}

void fftw_execute(const fftw_plan p)
{
    const int n = p->n;
    if ((n & (n - 1)) == 0) {
        for (int i = 0; i < n; i++) { p->re[i] = p->in[i]; p->im[i] = 0.0; }
        for (int i = 1, j = 0; i < n; i++) {
            int bit = n >> 1;
            for (; j & bit; bit >>= 1) j ^= bit;
            j ^= bit;
            if (i < j) { double t = p->re[i]; p->re[i] = p->re[j]; p->re[j] = t; }
        }
        for (int len = 2; len <= n; len <<= 1) {
            int half = len >> 1, step = n / len;
            for (int i = 0; i < n; i += len)
                for (int j = 0; j < half; j++) {
                    double wr = p->wr[j * step], wi = p->wi[j * step];
                    double vr = p->re[i + j + half] * wr - p->im[i + j + half] * wi;
                    double vi = p->re[i + j + half] * wi + p->im[i + j + half] * wr;
                    double ur = p->re[i + j], ui = p->im[i + j];
                    p->re[i + j] = ur + vr; p->im[i + j] = ui + vi;
                    p->re[i + j + half] = ur - vr; p->im[i + j + half] = ui - vi;
                }
        }
        for (int k = 0; k <= n / 2; k++) { p->out[k][0] = p->re[k]; p->out[k][1] = p->im[k]; }
    } else {
        for (int k = 0; k <= n / 2; k++) {
            double sr = 0, si = 0;
            for (int i = 0; i < n; i++) { int m = (int)(((long)k * i) % n); sr += p->in[i] * p->wr[m]; si += p->in[i] * p->wi[m]; }
            p->out[k][0] = sr; p->out[k][1] = si;
        }
    }
}

void fftw_destroy_plan(fftw_plan p) { free(p->wr); free(p->wi); free(p->re); free(p->im); free(p); }
