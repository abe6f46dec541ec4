/*
 * donut_oracle.c -- CPU restatement ("port") of the reference's donut classifier.
 *
 * TEST INFRASTRUCTURE ONLY (same rules as cepstrum_oracle.c): never linked, imported or called by
 * the product; used by tests/, smoke() and bench.py's CPU-baseline legs.
 *
 * Parity PINNED: unlike the cepstrum path the reference code for this path compiles here, so
 * oracle/Makefile builds the UNMODIFIED reference into oracle/_ref/ (libdonut_ref_f32.so from
 * sync/lib/classifier.cpp + PlainFFT.cpp; libdonut_ref_f64.so from donut-classifier/classifier.c with
 * the two shim headers in oracle/shim/) and tests/test_donut_oracle_cpu.py checks this port against
 * them function by function, and against the reference's own IIR artefact
 * donut-classifier/_postbutter.txt.
 *
 * One source, two precisions (-DSJ_REAL): float follows sync/lib/classifier.cpp (the API
 * `int classify(float*, int)`, sync/lib/classifier.h:19), double follows
 * donut-classifier/classifier.c (main, :30-212).  Variant constants come in through
 * sj_donut_params so every in-repo variant is expressible.
 */
#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#ifndef SJ_REAL
#define SJ_REAL double
#endif
typedef SJ_REAL dreal;

#define DN_WIN 256
#define DN_HOP 224
#define DN_BINS 129
#define DN_PI 3.14159265358979323846

typedef struct {
    double lo_norm, hi_norm;   /* normalised-dB band kept for the sums (classifier.c:141-142) */
    double mid_thr_db;         /* find_midpoints threshold (classifier.c:660) */
    double gap_tol, min_dur;   /* clustering (classifier.c:749-750) */
    double thr_mid, thr_above, thr_below; /* verdict: mid < thr_mid && above > thr_above && below > thr_below */
} sj_donut_params;

/* variant 0: donut-classifier/classifier.c (:141-142, :660, :749-750, :184)
 * variant 1: sync/lib/classifier.cpp       (:67-68,  :436, :523-524, :109)
 * variant 2: donut-classifier/revert.c (:713-714 gap tolerance 0.15, :169) = microphone/src/classifier.c (:164)
 * variant 3: microphone/src/classifier.cpp (:79-80, :448, :535-536, :123) */
void sj_donut_default_params(sj_donut_params *p, int variant)
{
    p->gap_tol = 0.05; p->min_dur = 0.15; p->lo_norm = 0.70; p->hi_norm = 0.85; p->mid_thr_db = 45.0;
    if (variant == 0)      { p->thr_mid = 75; p->thr_above = 300; p->thr_below = 100; }
    else if (variant == 1) { p->lo_norm = 0.65; p->hi_norm = 0.80; p->mid_thr_db = 70.0; p->thr_mid = 100; p->thr_above = 200; p->thr_below = 80; }
    else if (variant == 2) { p->gap_tol = 0.15; p->thr_mid = 50; p->thr_above = 200; p->thr_below = 200; }
    else                   { p->thr_mid = 100; p->thr_above = 200; p->thr_below = 150; }
}

/* butter_bandpass: hard-coded scipy butter(4, [lo,hi], 'band', fs=16000) coefficients, as printed in
 * classifier.c:342-360,383-401 (= classifier.cpp:142-182).  which: 0 = (1000,3000), 1 = (3000,7500) */
static const double DN_B[2][9] = {
    {0.01020948, 0., -0.04083792, 0., 0.06125688, 0., -0.04083792, 0., 0.01020948},
    {0.1362017, 0., -0.5448068, 0., 0.8172102, 0., -0.5448068, 0., 0.1362017}};
static const double DN_A[2][9] = {
    {1., -4.56803686, 9.95922498, -13.49912589, 12.43979269, -7.94997696, 3.43760562, -0.92305481, 0.1203896},
    {1., 2.60935592, 2.32553038, 1.20262614, 1.11690211, 0.76154474, 0.10005124, -0.0129829, 0.02236815}};

/* butter_bandpass_filter (classifier.c:420-446 / classifier.cpp:193-219): direct form II, zero state,
 * operations in exactly the reference's order (no fused multiply-add: build with -ffp-contract=off) */
void sj_oracle_donut_filter_coef(const dreal *x, int n, const dreal *b, const dreal *a, dreal *out)
{
    dreal w[9] = {0};
    for (int i = 0; i < n; i++) {
        dreal w0 = x[i];
        for (int j = 1; j < 9; j++) w0 -= a[j] * w[j - 1];
        dreal y = b[0] * w0;
        for (int j = 1; j < 9; j++) y += b[j] * w[j - 1];
        for (int j = 8; j > 0; j--) w[j] = w[j - 1];
        w[0] = w0;
        out[i] = y;
    }
}

void sj_oracle_donut_filter(const dreal *x, int n, int which, dreal *out)
{
    dreal b[9], a[9];
    for (int j = 0; j < 9; j++) { b[j] = (dreal)DN_B[which][j]; a[j] = (dreal)DN_A[which][j]; }
    sj_oracle_donut_filter_coef(x, n, b, a, out);
}

static void tukey_window(dreal *window)
{
    /* classifier.c:484-521 / classifier.cpp:263-297: first 256 points of the 257-point symmetric
     * Tukey(0.25); the reference's write to window[256] is out of bounds and is not reproduced */
    const dreal alpha = (dreal)0.25;
    const int M = DN_WIN + 1, width = 32;
    for (int n = 0; n <= width; n++) {
        if (sizeof(dreal) == 4)
            window[n] = 0.5f * (1.0f + cosf((float)DN_PI * (-1.0f + 2.0f * (float)n / ((float)alpha * ((float)M - 1.0f)))));
        else
            window[n] = 0.5 * (1 + cos(DN_PI * (-1 + 2.0 * n / ((double)alpha * (M - 1)))));
    }
    for (int n = width + 1; n <= M - width - 2; n++) window[n] = 1;
    for (int n = M - width - 1; n < DN_WIN; n++) {
        if (sizeof(dreal) == 4)
            window[n] = 0.5f * (1.0f + cosf((float)DN_PI * (-2.0f / (float)alpha + 1.0f + 2.0f * (float)n / ((float)alpha * ((float)M - 1.0f)))));
        else
            window[n] = 0.5 * (1 + cos(DN_PI * (-2.0 / (double)alpha + 1 + 2.0 * n / ((double)alpha * (M - 1)))));
    }
}

/* forward complex FFT, n = 256, in place, radix-2 (the reference uses FFTW r2c / PlainFFT::Compute) */
static void cfft256(dreal *re, dreal *im, const dreal *wr, const dreal *wi)
{
    const int n = DN_WIN;
    for (int i = 1, j = 0; i < n; i++) {
        int bit = n >> 1;
        for (; j & bit; bit >>= 1) j ^= bit;
        j ^= bit;
        if (i < j) { dreal t = re[i]; re[i] = re[j]; re[j] = t; t = im[i]; im[i] = im[j]; im[j] = t; }
    }
    for (int len = 2; len <= n; len <<= 1) {
        int half = len >> 1, step = n / len;
        for (int i = 0; i < n; i += len)
            for (int j = 0; j < half; j++) {
                dreal c = wr[j * step], s = wi[j * step];
                dreal vr = re[i + j + half] * c - im[i + j + half] * s;
                dreal vi = re[i + j + half] * s + im[i + j + half] * c;
                dreal ur = re[i + j], ui = im[i + j];
                re[i + j] = ur + vr; im[i + j] = ui + vi;
                re[i + j + half] = ur - vr; im[i + j + half] = ui - vi;
            }
    }
}

/* The float reference transforms with PlainFFT::Compute (sync/lib/PlainFFT.cpp:29-94): radix-2
 * decimation in time after a bit-reversal permutation, the twiddle (u1, u2) advanced by a float
 * recurrence per butterfly offset and the per-stage rotation (c1, c2) by half-angle square roots
 * evaluated in double.  Restated here operation for operation, because the spectrogram's global
 * minimum (classifier.cpp:34-55) sits on round-off-level cells and is therefore sensitive to the
 * exact rounding of the transform. */
static void cfft256_plain_f32(float *re, float *im)
{
    const int n = DN_WIN;
    for (int i = 1, j = 0; i < n; i++) {
        int bit = n >> 1;
        for (; j & bit; bit >>= 1) j ^= bit;
        j ^= bit;
        if (i < j) { float t = re[i]; re[i] = re[j]; re[j] = t; t = im[i]; im[i] = im[j]; im[j] = t; }
    }
    float c1 = -1.0f, c2 = 0.0f;
    for (int l = 0, l2 = 1; l < 8; l++) {
        const int l1 = l2;
        l2 <<= 1;
        float u1 = 1.0f, u2 = 0.0f;
        for (int m = 0; m < l1; m++) {
            for (int i = m; i < n; i += l2) {
                const int i1 = i + l1;
                const float t1 = u1 * re[i1] - u2 * im[i1];
                const float t2 = u1 * im[i1] + u2 * re[i1];
                re[i1] = re[i] - t1; im[i1] = im[i] - t2;
                re[i] += t1; im[i] += t2;
            }
            const float z = (u1 * c1) - (u2 * c2);
            u2 = (u1 * c2) + (u2 * c1);
            u1 = z;
        }
        c2 = (float)sqrt((1.0 - c1) / 2.0);
        c2 = -c2;
        c1 = (float)sqrt((1.0 + c1) / 2.0);
    }
}

/* compute_spectrogram (classifier.c:448-592 / classifier.cpp:221-368).  Sxx is dense [129][tb]
 * (row = frequency bin, like the reference's Sxx[f][t]); returns tb = (n-256)/224+1 (may be <= 0). */
int sj_oracle_donut_spectrogram(const dreal *sig, int n, int fs, dreal *Sxx, int tb_cap)
{
    int tb = (n - DN_WIN) / DN_HOP + 1;
    if (n < DN_WIN) tb = 0;
    if (tb > tb_cap) return -1;
    dreal window[DN_WIN], wr[DN_WIN], wi[DN_WIN], re[DN_WIN], im[DN_WIN];
    tukey_window(window);
    for (int k = 0; k < DN_WIN; k++) { wr[k] = (dreal)cos(2 * DN_PI * k / DN_WIN); wi[k] = (dreal)(-sin(2 * DN_PI * k / DN_WIN)); }
    dreal U = 0;
    for (int i = 0; i < DN_WIN; i++) U += window[i] * window[i];
    U *= (dreal)fs;
    for (int t = 0; t < tb; t++) {
        const int start = t * DN_HOP;
        dreal sum = 0;
        for (int i = 0; i < DN_WIN; i++) { re[i] = sig[start + i]; im[i] = 0; }
        for (int i = 0; i < DN_WIN; i++) sum += re[i];
        const dreal mean = sum / (dreal)DN_WIN;
        for (int i = 0; i < DN_WIN; i++) re[i] -= mean;
        for (int i = 0; i < DN_WIN; i++) re[i] *= window[i];
        if (sizeof(dreal) == 4) cfft256_plain_f32((float *)re, (float *)im);
        else cfft256(re, im, wr, wi);
        for (int f = 0; f < DN_BINS; f++) {
            dreal m2 = re[f] * re[f] + im[f] * im[f];
            Sxx[f * tb + t] = m2 / U;
        }
        for (int f = 1; f < DN_BINS - 1; f++) Sxx[f * tb + t] *= (dreal)2;
    }
    return tb;
}

/* dB conversion + min/max (classifier.c:104-123 / classifier.cpp:34-55): cells <= 0 become NaN */
static void to_db(dreal *S, int cells, dreal *mn, dreal *mx)
{
    dreal lo = (dreal)DBL_MAX, hi = (dreal)-DBL_MAX; /* float: +-inf, as in the reference */
    for (int i = 0; i < cells; i++) {
        if (S[i] > 0) {
            S[i] = (dreal)(10 * log10(S[i] / 1e-12));
            if (S[i] < lo) lo = S[i];
            if (S[i] > hi) hi = S[i];
        } else S[i] = (dreal)NAN;
    }
    *mn = lo; *mx = hi;
}

/* sum_intense (classifier.c:594-655 / classifier.cpp:370-431) on a dense [129][tb] array */
static dreal sum_intense(dreal lower, dreal upper, dreal half_range, const dreal *freqs, const dreal *times, int tb,
                         const dreal *I, dreal midpoint)
{
    int f0 = 0, f1 = DN_BINS - 1, t0 = 0, t1 = tb - 1, tmp;
    while (f0 < DN_BINS && freqs[f0] < lower) f0++;
    while (f1 >= 0 && freqs[f1] > upper) f1--;
    if (f0 >= DN_BINS) f0 = DN_BINS - 1;
    if (f1 < 0) f1 = 0;
    if (f0 > f1) { tmp = f0; f0 = f1; f1 = tmp; }
    while (t0 < tb && times[t0] < midpoint - half_range) t0++;
    while (t1 >= 0 && times[t1] > midpoint + half_range) t1--;
    if (t0 >= tb) t0 = tb - 1;
    if (t1 < 0) t1 = 0;
    if (t0 > t1) { tmp = t0; t0 = t1; t1 = tmp; }
    dreal total = 0;
    for (int i = f0; i <= f1; i++)
        for (int j = t0; j <= t1; j++)
            if (!isnan(I[i * tb + j])) total += I[i * tb + j];
    return total;
}

#define DN_MAX_MID 16

/* Whole classifier on one clip: classify (classifier.cpp:9-136) / the per-file body of main
 * (classifier.c:84-198).  Returns 0/1; -1 if the clip is shorter than one spectrogram frame.
 * n_mid: number of midpoints found; mids[<=16]; sums[3*k..] = (above, middle, below) for every
 * midpoint the reference evaluates (it stops at the first that passes). */
int sj_oracle_donut_classify(const float *x, int n, const sj_donut_params *P, int *n_mid, dreal *mids, dreal *sums)
{
    const int fs = 16000;
    int tb = (n - DN_WIN) / DN_HOP + 1;
    if (n < DN_WIN || tb < 1) { if (n_mid) *n_mid = 0; return -1; }
    dreal *data = (dreal *)malloc(sizeof(dreal) * n), *filt = (dreal *)malloc(sizeof(dreal) * n);
    dreal *A = (dreal *)malloc(sizeof(dreal) * DN_BINS * tb), *B = (dreal *)malloc(sizeof(dreal) * DN_BINS * tb);
    dreal freqs[DN_BINS], *times = (dreal *)malloc(sizeof(dreal) * tb);
    for (int i = 0; i < n; i++) data[i] = (dreal)x[i];
    for (int i = 0; i < DN_BINS; i++) freqs[i] = (dreal)i * (dreal)fs / (dreal)DN_WIN;
    for (int t = 0; t < tb; t++) times[t] = ((dreal)(t * DN_HOP + DN_WIN / 2)) / (dreal)fs;

    /* 3000-7500 Hz branch: spectrogram -> dB -> global normalise -> keep (lo, hi) */
    sj_oracle_donut_filter(data, n, 1, filt);
    sj_oracle_donut_spectrogram(filt, n, fs, A, tb);
    dreal mn, mx;
    to_db(A, DN_BINS * tb, &mn, &mx);
    const dreal lo = (dreal)P->lo_norm, hi = (dreal)P->hi_norm;
    for (int i = 0; i < DN_BINS * tb; i++) {
        if (!isnan(A[i])) A[i] = (A[i] - mn) / (mx - mn);
        if (!(A[i] > lo && A[i] < hi)) A[i] = (dreal)NAN;
    }
    /* find_midpoints (classifier.c:657-825 / classifier.cpp:433-598): 1000-3000 Hz branch on the RAW clip */
    sj_oracle_donut_filter(data, n, 0, filt);
    sj_oracle_donut_spectrogram(filt, n, fs, B, tb);
    dreal mn2, mx2;
    to_db(B, DN_BINS * tb, &mn2, &mx2);
    const dreal thr = (dreal)P->mid_thr_db;
    dreal *blob = (dreal *)malloc(sizeof(dreal) * tb);
    int nblob = 0;
    for (int j = 0; j < tb; j++) {
        int any = 0;
        for (int i = 0; i < DN_BINS; i++) if (B[i * tb + j] > thr) { any = 1; break; }
        if (any) blob[nblob++] = times[j];
    }
    const dreal tol = (dreal)P->gap_tol, mind = (dreal)P->min_dur;
    int count = 0, k = 0;
    while (k < nblob) {
        int s = k;
        while (k + 1 < nblob && blob[k + 1] - blob[k] <= tol) k++;
        if (blob[k] - blob[s] >= mind) {
            dreal st = 0;
            for (int j = s; j <= k; j++) st += blob[j];
            if (count < DN_MAX_MID) mids[count] = st / (dreal)(k - s + 1);
            count++;
        }
        k++;
    }
    if (n_mid) *n_mid = count;
    int label = 0;
    for (int m = 0; m < count && m < DN_MAX_MID; m++) {
        dreal above = sum_intense(5000, 7000, (dreal)0.18, freqs, times, tb, A, mids[m]);
        dreal middle = sum_intense(2500, 5000, (dreal)0.05, freqs, times, tb, A, mids[m]);
        dreal below = sum_intense(500, 2500, (dreal)0.18, freqs, times, tb, A, mids[m]);
        if (sums) { sums[3 * m] = above; sums[3 * m + 1] = middle; sums[3 * m + 2] = below; }
        if (middle < (dreal)P->thr_mid && above > (dreal)P->thr_above && below > (dreal)P->thr_below) { label = 1; break; }
    }
    free(data); free(filt); free(A); free(B); free(times); free(blob);
    return label;
}

/* batch driver: labels[n_clips]; detail (optional) [n_clips][4] = n_mid, above, middle, below of the
 * LAST midpoint evaluated (the passing one, if any) */
long sj_oracle_donut_batch(const float *clips, long n_clips, int n_samples, int variant, int *labels, float *detail)
{
    sj_donut_params P;
    sj_donut_default_params(&P, variant);
    for (long c = 0; c < n_clips; c++) {
        int nm = 0;
        dreal mids[DN_MAX_MID], sums[3 * DN_MAX_MID];
        memset(sums, 0, sizeof sums);
        int l = sj_oracle_donut_classify(clips + c * (long)n_samples, n_samples, &P, &nm, mids, sums);
        labels[c] = l < 0 ? 0 : l;
        if (detail) {
            /* index of the last midpoint the reference evaluated: the first that passes, else the last */
            int cap = nm < DN_MAX_MID ? nm : DN_MAX_MID, last = cap - 1;
            for (int m = 0; m < cap; m++)
                if (sums[3 * m + 1] < (dreal)P.thr_mid && sums[3 * m] > (dreal)P.thr_above && sums[3 * m + 2] > (dreal)P.thr_below) { last = m; break; }
            detail[4 * c] = (float)nm;
            detail[4 * c + 1] = nm ? (float)sums[3 * last] : 0.f;
            detail[4 * c + 2] = nm ? (float)sums[3 * last + 1] : 0.f;
            detail[4 * c + 3] = nm ? (float)sums[3 * last + 2] : 0.f;
        }
    }
    return n_clips;
}
