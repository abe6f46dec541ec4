/* keyword_oracle.c -- CPU restatement of the reference's keyword ("stop") detector.  TEST INFRASTRUCTURE ONLY:
 * used by tests/, __graft_entry__.smoke() and bench tools as the checker, never by the product.
 *
 * Follows 2fa/audio/word/c: compute_mfcc (mfcc.c:108-232) with its radix-2 FFT (fft_real_forward, mfcc.c:16-95:
 * bit reversal, then per stage a float twiddle recurrence w *= w_len), classify_signal's coefficient-major
 * re-layout padded to 500 frames (stop_detector.c:12-55) and audio_classifier_predict
 * (audio_classifier_inference.c:38-90).  All constants come from the "SJKW" blob (tools/export_keyword.py).
 * Pinned: tests/test_keyword_oracle_cpu.py compares it with the UNMODIFIED reference sources compiled into
 * oracle/_ref/libkeyword_ref.so (same compiler, same libm: agreement is exact) and with tests/golden/keyword_golden.npz. */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define NFFT 512
#define FRAME 400
#define HOP 160
#define NMEL 40
#define NMFCC 13
#define NBINS 257
#define MAXFR 500
#define NIN (NMFCC * MAXFR)

typedef struct {
    const float *hann, *mel, *dct, *mean, *scale, *w1, *b1, *w2, *b2, *w3, *b3, *w4, *b4;
} kw_model;

static int kw_bind(const void *blob, size_t bytes, kw_model *m)
{
    uint32_t h[16];
    if (!blob || bytes < sizeof h) return -1;
    memcpy(h, blob, sizeof h);
    if (h[0] != 0x574B4A53u || h[1] != 1 || h[3] != NFFT || h[4] != FRAME || h[5] != HOP || h[6] != NMEL ||
        h[7] != NMFCC || h[8] != NBINS || h[9] != MAXFR || h[10] != NIN)
        return -1;
    const float *p = (const float *)((const char *)blob + sizeof h);
    m->hann = p; p += FRAME;
    m->mel = p; p += NMEL * NBINS;
    m->dct = p; p += NMFCC * NMEL;
    m->mean = p; p += NIN;
    m->scale = p; p += NIN;
    m->w1 = p; p += NIN * 4;
    m->b1 = p; p += 4;
    m->w2 = p; p += 8;
    m->b2 = p; p += 2;
    m->w3 = p; p += 4;
    m->b3 = p; p += 2;
    m->w4 = p; p += 2;
    m->b4 = p;
    return 0;
}

/* in-place complex FFT of length 512, float throughout; twiddles by the reference's recurrence (mfcc.c:58-85) */
static void fft512(float *re, float *im)
{
    for (int i = 0, j = 0; i < NFFT; ++i) { /* bit-reversal permutation (mfcc.c:36-52) */
        if (i < j) {
            float t = re[i]; re[i] = re[j]; re[j] = t;
            t = im[i]; im[i] = im[j]; im[j] = t;
        }
        int m = NFFT >> 1;
        while (m >= 1 && j >= m) { j -= m; m >>= 1; }
        j += m;
    }
    for (int len = 2; len <= NFFT; len <<= 1) {
        const float ang = -2.0f * 3.14159265358979323846f / (float)len;
        const float sc = cosf(ang), ss = sinf(ang);
        const int half = len / 2;
        for (int base = 0; base < NFFT; base += len) {
            float wc = 1.0f, ws = 0.0f;
            for (int k = 0; k < half; ++k) {
                float *ur = &re[base + k], *ui = &im[base + k], *vr = ur + half, *vi = ui + half;
                const float tr = wc * *vr - ws * *vi, ti = wc * *vi + ws * *vr;
                const float ar = *ur, ai = *ui;
                *vr = ar - tr; *vi = ai - ti;
                *ur = ar + tr; *ui = ai + ti;
                const float nc = wc * sc - ws * ss;
                ws = wc * ss + ws * sc;
                wc = nc;
            }
        }
    }
}

/* compute_mfcc: frame-major out[frames][13]; returns the frame count */
int kw_oracle_compute_mfcc(const void *blob, size_t bytes, const float *signal, int n, float *out, int max_frames)
{
    kw_model m;
    if (kw_bind(blob, bytes, &m)) return -1;
    if (n < FRAME || max_frames <= 0) return 0;
    float re[NFFT], im[NFFT], pw[NBINS], e[NMEL], db[NMEL];
    int f = 0;
    for (; f < max_frames && f * HOP + FRAME <= n; ++f) {
        const float *x = signal + f * HOP;
        for (int i = 0; i < NFFT; ++i) { re[i] = i < FRAME ? x[i] * m.hann[i] : 0.0f; im[i] = 0.0f; }
        fft512(re, im);
        for (int k = 0; k < NBINS; ++k) pw[k] = re[k] * re[k] + im[k] * im[k];
        for (int b = 0; b < NMEL; ++b) {
            float s = 0.0f;
            for (int k = 0; k < NBINS; ++k) s += m.mel[b * NBINS + k] * pw[k];
            e[b] = s;
        }
        /* power_to_db(ref = max, amin = 1e-10, top_db = 80), mfcc.c:171-210 */
        const float amin = 1e-10f;
        float ref = 0.0f;
        for (int b = 0; b < NMEL; ++b) if (e[b] > ref) ref = e[b];
        if (ref < amin) ref = amin;
        const float log_ref = 10.0f * log10f(ref);
        for (int b = 0; b < NMEL; ++b) db[b] = 10.0f * log10f(e[b] < amin ? amin : e[b]) - log_ref;
        float top = db[0];
        for (int b = 1; b < NMEL; ++b) if (db[b] > top) top = db[b];
        for (int b = 0; b < NMEL; ++b) if (db[b] < top - 80.0f) db[b] = top - 80.0f;
        for (int c = 0; c < NMFCC; ++c) {
            float s = 0.0f;
            for (int b = 0; b < NMEL; ++b) s += m.dct[c * NMEL + b] * db[b];
            out[f * NMFCC + c] = s;
        }
    }
    return f;
}

static void dense(const float *x, const float *w, const float *b, int nin, int nout, float *y, int relu)
{
    for (int j = 0; j < nout; ++j) {
        float s = b[j];
        for (int i = 0; i < nin; ++i) s += w[i * nout + j] * x[i];
        y[j] = (relu && !(s > 0.0f)) ? 0.0f : s;
    }
}

/* audio_classifier_predict over a coefficient-major [13][500] feature vector */
float kw_oracle_predict(const void *blob, size_t bytes, const float *feat)
{
    kw_model m;
    if (kw_bind(blob, bytes, &m)) return -1.0f;
    float *x = (float *)malloc(sizeof(float) * NIN);
    for (int i = 0; i < NIN; ++i) {
        float sc = m.scale[i];
        if (sc == 0.0f) sc = 1.0f;
        x[i] = (feat[i] - m.mean[i]) / sc;
    }
    float h1[4], h2[2], h3[2], o[1];
    dense(x, m.w1, m.b1, NIN, 4, h1, 1);
    dense(h1, m.w2, m.b2, 4, 2, h2, 1);
    dense(h2, m.w3, m.b3, 2, 2, h3, 1);
    dense(h3, m.w4, m.b4, 2, 1, o, 0);
    free(x);
    return 1.0f / (1.0f + expf(-o[0]));
}

/* classify_signal; mfcc_out (may be NULL) [500][13] frame-major, zero beyond the frames; *n_frames (may be NULL) */
float kw_oracle_classify(const void *blob, size_t bytes, const float *signal, int n, float *mfcc_out, int *n_frames)
{
    float *fm = (float *)calloc(NIN, sizeof(float)), *feat = (float *)calloc(NIN, sizeof(float));
    int T = kw_oracle_compute_mfcc(blob, bytes, signal, n, fm, MAXFR);
    if (T < 0) { free(fm); free(feat); return -1.0f; }
    for (int c = 0; c < NMFCC; ++c)
        for (int t = 0; t < T; ++t) feat[c * MAXFR + t] = fm[t * NMFCC + c];
    const float p = kw_oracle_predict(blob, bytes, feat);
    if (mfcc_out) memcpy(mfcc_out, fm, sizeof(float) * NIN);
    if (n_frames) *n_frames = T;
    free(fm); free(feat);
    return p;
}

/* batch helper for the timed CPU baseline */
long kw_oracle_classify_batch(const void *blob, size_t bytes, const float *signals, long n_clips, int n, float *prob)
{
    for (long i = 0; i < n_clips; ++i) prob[i] = kw_oracle_classify(blob, bytes, signals + i * (long)n, n, NULL, NULL);
    return n_clips;
}
