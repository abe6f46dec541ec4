/*
 * cepstrum_oracle.c -- CPU restatement of the reference's cepstrum (MFCC mean/std + RBF-SVM) path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product (dsp_b200/, include/) may link, import or call
 * this file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs use it, as the checker or as the timed CPU baseline.
 *
 * PARITY UNPINNED for the feature arithmetic: the reference file cepstrum/scrubjay_infer.c is glue
 * over two third-party libraries that are NOT vendored under /root/reference and are absent from
 * this image: aubio (pkg-config "aubio", version unpinned, cepstrum/CMakeLists.txt:9) and
 * onnxruntime ($ORT_ROOT, version unpinned, cepstrum/CMakeLists.txt:7,24).  The reference holds no
 * golden vectors for this path.  What follows restates the published algorithms of aubio 0.4.x
 * (pvoc.c, fft.c, filterbank_mel.c, mfcc.c, fvec.c) and of onnxruntime's ai.onnx.ml Scaler /
 * SVMClassifier + libsvm's sigmoid_predict / multiclass_probability, anchored on the reference's
 * own call sites.  The SVM half IS pinned: tests check it against sklearn loading the reference's
 * scrubjay_svm.joblib (golden vectors in tests/golden/, generator tools/make_golden.py).
 *
 * Build: see oracle/Makefile.  -DSJ_REAL=double gives the parity build (FFT and filterbank sums in
 * double, rounded to float where aubio holds floats); -DSJ_REAL=float gives the "port" timed as the
 * CPU baseline (same precision as aubio built on fftw3f, cepstrum/CMakeLists.txt:12).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* the parameter block (plain data) is shared with the product's public header, so that every switch of
 * scrubjay_cepstrum_params has its checker here; nothing else of the product is used */
#include "../include/scrubjay_cuda.h"

#ifndef SJ_REAL
#define SJ_REAL double
#endif
typedef SJ_REAL real_t;

/* constants: cepstrum/scrubjay_infer.c:10-14 */
#define N_MFCC 20
#define FEAT_DIM (N_MFCC * 2)
#define WIN_SIZE 2048
#define HOP_SIZE (WIN_SIZE / 2)
#define N_FILTERS 40
#define N_BINS (WIN_SIZE / 2 + 1)

#define SJ_PI 3.14159265358979323846

typedef struct {
    unsigned sr;
    scrubjay_cepstrum_params prm;     /* the switchable aubio-isms / ORT-isms; defaults = restated aubio 0.4.x / ORT */
    float window[WIN_SIZE];           /* aubio "hanningz": 0.5*(1-cos(2*pi*i/N)), periodic (or the symmetric form) */
    float fbank[N_FILTERS][N_BINS];   /* Slaney triangles, dense like aubio's fmat */
    float dct[N_MFCC][N_FILTERS];     /* orthonormal DCT-II rows 0..19 */
    real_t tw_re[WIN_SIZE / 2], tw_im[WIN_SIZE / 2]; /* exp(-2*pi*i*k/2048), k<1024 */
} sj_oracle_t;

/* aubio filterbank_mel.c: aubio_filterbank_set_mel_coeffs_slaney + set_triangle_bands, all in
 * float (smpl_t).  Selected because n_filters == 40 at the call new_aubio_mfcc(WIN_SIZE, N_FILTERS,
 * N_MFCC, sr), cepstrum/scrubjay_infer.c:30. */
static void build_slaney(float fb[N_FILTERS][N_BINS], unsigned sr)
{
    float freqs[N_FILTERS + 2];
    const float lowest = 133.3333f, lin = 66.66666666f, logsp = 1.0711703f;
    int fn, bin;
    for (fn = 0; fn < 13; fn++) freqs[fn] = lowest + fn * lin;
    float last = freqs[12];
    for (fn = 0; fn < 27 + 2; fn++) freqs[fn + 13] = last * powf(logsp, (float)(fn + 1));
    float fft_freqs[N_BINS];
    for (bin = 0; bin < N_BINS; bin++) fft_freqs[bin] = ((float)sr / (float)WIN_SIZE) * (float)bin;
    memset(fb, 0, sizeof(float) * N_FILTERS * N_BINS);
    for (fn = 0; fn < N_FILTERS; fn++) {
        float lo = freqs[fn], ce = freqs[fn + 1], up = freqs[fn + 2];
        float h = 2.f / (up - lo);
        for (bin = 0; bin < N_BINS - 1; bin++) {
            if (fft_freqs[bin] <= lo && fft_freqs[bin + 1] > lo) { bin++; break; }
        }
        float rise = h / (ce - lo);
        for (; bin < N_BINS - 1; bin++) {
            fb[fn][bin] = (fft_freqs[bin] - lo) * rise;
            if (fft_freqs[bin + 1] >= ce) { bin++; break; }
        }
        float down = h / (up - ce);
        for (; bin < N_BINS - 1; bin++) {
            fb[fn][bin] += (up - fft_freqs[bin]) * down;
            if (fb[fn][bin] < 0.f) fb[fn][bin] = 0.f;
            if (fft_freqs[bin + 1] >= up) break;
        }
    }
}

static void default_params(scrubjay_cepstrum_params *p)
{
    memset(p, 0, sizeof *p);
    p->window = SJ_WINDOW_HANN_PERIODIC;   /* new_aubio_pvoc -> new_aubio_window("hanningz") */
    p->spectrum = SJ_SPECTRUM_MAGNITUDE;   /* aubio_filterbank power = 1 */
    p->log_floor = 2.e-42f;                /* VERY_SMALL_NUMBER */
    p->dct_norm = SJ_DCT_ORTHO;            /* aubio_dct */
    p->half_swap = 1;                      /* fvec_shift in aubio_pvoc_do */
    p->label_rule = SJ_LABEL_PREDICT;      /* cepstrum/run.py:46 */
}

sj_oracle_t *sj_oracle_new_ex(unsigned sr, const scrubjay_cepstrum_params *prm)
{
    sj_oracle_t *o = (sj_oracle_t *)calloc(1, sizeof(*o));
    int i, j;
    if (!o) return NULL;
    o->sr = sr;
    default_params(&o->prm);
    if (prm) o->prm = *prm;
    for (i = 0; i < WIN_SIZE; i++) {
        double den = o->prm.window == SJ_WINDOW_HANN_SYMMETRIC ? (double)(WIN_SIZE - 1) : (double)WIN_SIZE;
        o->window[i] = (float)(0.5 * (1.0 - cos(2.0 * SJ_PI * i / den)));
    }
    build_slaney(o->fbank, sr);
    /* aubio mfcc.c / dct_plain.c: out[0] scaled by sqrt(1/N), others sqrt(2/N); or the unnormalised 2 sum cos form */
    for (j = 0; j < N_MFCC; j++)
        for (i = 0; i < N_FILTERS; i++) {
            double s = (j == 0) ? sqrt(1.0 / N_FILTERS) : sqrt(2.0 / N_FILTERS);
            if (o->prm.dct_norm == SJ_DCT_UNNORMALISED) s = 2.0;
            o->dct[j][i] = (float)(s * cos(SJ_PI * j * (i + 0.5) / N_FILTERS));
        }
    for (i = 0; i < WIN_SIZE / 2; i++) {
        o->tw_re[i] = (real_t)cos(2.0 * SJ_PI * i / WIN_SIZE);
        o->tw_im[i] = (real_t)(-sin(2.0 * SJ_PI * i / WIN_SIZE));
    }
    return o;
}

sj_oracle_t *sj_oracle_new(unsigned sr) { return sj_oracle_new_ex(sr, NULL); }

void sj_oracle_free(sj_oracle_t *o) { free(o); }

/* expose the tables so tests can compare the device-side tables with them */
const float *sj_oracle_fbank(const sj_oracle_t *o) { return &o->fbank[0][0]; }
const float *sj_oracle_window(const sj_oracle_t *o) { return o->window; }
const float *sj_oracle_dct(const sj_oracle_t *o) { return &o->dct[0][0]; }

/* in-place iterative radix-2 complex FFT, n = 1024, forward, unnormalised (as fftw r2c / aubio fft.c) */
static void cfft1024(const sj_oracle_t *o, real_t *re, real_t *im)
{
    const int n = 1024;
    int i, j, len;
    for (i = 1, j = 0; i < n; i++) {
        int bit = n >> 1;
        for (; j & bit; bit >>= 1) j ^= bit;
        j ^= bit;
        if (i < j) { real_t t = re[i]; re[i] = re[j]; re[j] = t; t = im[i]; im[i] = im[j]; im[j] = t; }
    }
    for (len = 2; len <= n; len <<= 1) {
        int half = len >> 1, step = 2048 / len; /* twiddle table is for N=2048 */
        for (i = 0; i < n; i += len)
            for (j = 0; j < half; j++) {
                real_t wr = o->tw_re[j * step], wi = o->tw_im[j * step];
                real_t ur = re[i + j], ui = im[i + j];
                real_t vr = re[i + j + half] * wr - im[i + j + half] * wi;
                real_t vi = re[i + j + half] * wi + im[i + j + half] * wr;
                re[i + j] = ur + vr; im[i + j] = ui + vi;
                re[i + j + half] = ur - vr; im[i + j + half] = ui - vi;
            }
    }
}

/* |rfft(frame)|, 1025 bins: aubio fft.c aubio_fft_get_norm after the unnormalised r2c transform */
static void rfft_mag(const sj_oracle_t *o, const float *frame, float *norm)
{
    real_t re[1024], im[1024];
    int k;
    for (k = 0; k < 1024; k++) { re[k] = frame[2 * k]; im[k] = frame[2 * k + 1]; }
    cfft1024(o, re, im);
    for (k = 0; k <= 1024; k++) {
        int a = k & 1023, b = (1024 - k) & 1023;
        real_t er = (real_t)0.5 * (re[a] + re[b]), ei = (real_t)0.5 * (im[a] - im[b]);
        real_t orr = (real_t)0.5 * (im[a] + im[b]), oi = (real_t)-0.5 * (re[a] - re[b]);
        real_t wr, wi;
        if (k < 1024) { wr = o->tw_re[k]; wi = o->tw_im[k]; } else { wr = -1; wi = 0; }
        real_t xr = er + orr * wr - oi * wi, xi = ei + orr * wi + oi * wr;
        norm[k] = (float)sqrt((double)(xr * xr + xi * xi));
    }
}

/* aubio_priv.h: VERY_SMALL_NUMBER 2.e-42; SAFE_LOG10 = log10(max(x, 2e-42)) -- params.log_floor */

/* One aubio_pvoc_do + aubio_mfcc_do (cepstrum/scrubjay_infer.c:44-45) on a 2048-sample buffer
 * `data` = [previous hop | new hop].  Writes 20 coefficients. */
static void frame_mfcc(const sj_oracle_t *o, const float *data, float *mfcc)
{
    float frame[WIN_SIZE], norm[N_BINS], in_dct[N_FILTERS];
    int i, m, c;
    /* pvoc.c: fvec_weight(data, w); fvec_shift(data) (swap halves: phase only); fft; norm */
    for (i = 0; i < WIN_SIZE; i++) frame[o->prm.half_swap ? (i + WIN_SIZE / 2) % WIN_SIZE : i] = data[i] * o->window[i];
    rfft_mag(o, frame, norm);
    /* filterbank.c: aubio_filterbank_do raises norm to `power` when power != 1 (fvec_pow), then
     * fmat_vecmul(filters, norm) -- dense */
    if (o->prm.spectrum == SJ_SPECTRUM_POWER)
        for (i = 0; i < N_BINS; i++) norm[i] = norm[i] * norm[i];
    for (m = 0; m < N_FILTERS; m++) {
        real_t acc = 0;
        for (i = 0; i < N_BINS; i++) acc += (real_t)o->fbank[m][i] * (real_t)norm[i];
        float e = (float)acc;
        if (e < o->prm.log_floor) e = o->prm.log_floor;
        in_dct[m] = (float)log10((double)e); /* fvec_log10 */
    }
    for (c = 0; c < N_MFCC; c++) {
        real_t acc = 0;
        for (m = 0; m < N_FILTERS; m++) acc += (real_t)o->dct[c][m] * (real_t)in_dct[m];
        mfcc[c] = (float)acc;
    }
}

/* mfcc_stats (cepstrum/scrubjay_infer.c:19-77) over an in-memory mono buffer instead of a path
 * (aubio_source decode is outside the hot path).  Frame count follows the do/while at :39-53 with
 * aubio_source_do zero-filling the last partial hop.  Returns the number of frames, or -1 for an
 * empty input (:55-59).  `frames_out` (optional) receives [frames][20] per-frame coefficients. */
long sj_oracle_mfcc_stats(const sj_oracle_t *o, const float *x, long n, float out[FEAT_DIM], float *frames_out)
{
    double sum[N_MFCC] = {0.0}, sq[N_MFCC] = {0.0};
    float data[WIN_SIZE], mfcc[N_MFCC];
    long frames = 0, pos = 0;
    int i;
    memset(data, 0, sizeof data);
    for (;;) {
        long rd = n - pos; if (rd > HOP_SIZE) rd = HOP_SIZE;
        if (rd <= 0) break;
        /* pvoc.c aubio_pvoc_swapbuffers: data = [dataold | new]; dataold = data[hop:] */
        memmove(data, data + HOP_SIZE, sizeof(float) * HOP_SIZE);
        memcpy(data + HOP_SIZE, x + pos, sizeof(float) * (size_t)rd);
        if (rd < HOP_SIZE) memset(data + HOP_SIZE + rd, 0, sizeof(float) * (size_t)(HOP_SIZE - rd));
        pos += rd;
        frame_mfcc(o, data, mfcc);
        for (i = 0; i < N_MFCC; i++) { double v = mfcc[i]; sum[i] += v; sq[i] += v * v; }
        if (frames_out) memcpy(frames_out + frames * N_MFCC, mfcc, sizeof mfcc);
        ++frames;
        if (rd != HOP_SIZE) break;
    }
    if (frames == 0) return -1;
    for (i = 0; i < N_MFCC; i++) {
        double mean = sum[i] / frames;
        double var = sq[i] / frames - mean * mean;
        out[i] = (float)mean;
        out[i + N_MFCC] = (float)sqrtf(var > 0 ? var : 0);
    }
    return frames;
}

/* ---- SVM: "SJSV" v1 blob written by tools/export_svm.py from cepstrum/scrubjay_svm.onnx ---- */
typedef struct {
    uint32_t magic, version, dim, n_sv, nsv0, nsv1;
    float gamma, rho, prob_a, prob_b;
} sjsv_hdr_t;

/* libsvm svm.cpp sigmoid_predict */
static double sigmoid_predict(double dv, double A, double B)
{
    double fApB = dv * A + B;
    if (fApB >= 0) return exp(-fApB) / (1.0 + exp(-fApB));
    return 1.0 / (1.0 + exp(fApB));
}

/* libsvm svm.cpp multiclass_probability for k = 2 (Wu, Lin, Weng 2004, second method) */
static void multiclass_probability2(double r01, double p[2])
{
    const int k = 2;
    double r[2][2] = {{0, r01}, {1 - r01, 0}}, Q[2][2], Qp[2], pQp, eps = 0.005 / k;
    int t, j, iter, max_iter = 100;
    for (t = 0; t < k; t++) {
        p[t] = 1.0 / k;
        Q[t][t] = 0;
        for (j = 0; j < t; j++) { Q[t][t] += r[j][t] * r[j][t]; Q[t][j] = Q[j][t]; }
        for (j = t + 1; j < k; j++) { Q[t][t] += r[j][t] * r[j][t]; Q[t][j] = -r[j][t] * r[t][j]; }
    }
    for (iter = 0; iter < max_iter; iter++) {
        pQp = 0;
        for (t = 0; t < k; t++) {
            Qp[t] = 0;
            for (j = 0; j < k; j++) Qp[t] += Q[t][j] * p[j];
            pQp += p[t] * Qp[t];
        }
        double max_error = 0;
        for (t = 0; t < k; t++) { double e = fabs(Qp[t] - pQp); if (e > max_error) max_error = e; }
        if (max_error < eps) break;
        for (t = 0; t < k; t++) {
            double diff = (-Qp[t] + pQp) / Q[t][t];
            p[t] += diff;
            pQp = (pQp + diff * (diff * Q[t][t] + 2 * Qp[t])) / (1 + diff) / (1 + diff);
            for (j = 0; j < k; j++) { Qp[j] = (Qp[j] + diff * Q[t][j]) / (1 + diff); p[j] /= (1 + diff); }
        }
    }
}

/* ort_predict (cepstrum/scrubjay_infer.c:105-141): Scaler -> SVMClassifier(RBF) -> label, conf.
 * `score` receives d = -(sum_j coef_j K_j + rho) = sklearn decision_function (positive => class 1).
 * Label rule: SURVEY.md Appendix B.3 (ii), the rule of cepstrum/run.py:46 (model.predict):
 * label = 1 iff d > 0.  conf = probability of the predicted class (:134). */
int sj_oracle_predict_ex(const void *blob, const float feat[FEAT_DIM], int label_rule, float *conf, float *score);
int sj_oracle_predict(const void *blob, const float feat[FEAT_DIM], float *conf, float *score)
{
    return sj_oracle_predict_ex(blob, feat, SJ_LABEL_PREDICT, conf, score);
}

/* label_rule: the three candidate rules of SURVEY.md B.3 -- (ii) run.py's model.predict d > 0 (default),
 * (i) libsvm's vote s <= 0, (iii) argmax of the probability pair with ties to class 0 */
int sj_oracle_predict_ex(const void *blob, const float feat[FEAT_DIM], int label_rule, float *conf, float *score)
{
    const sjsv_hdr_t *h = (const sjsv_hdr_t *)blob;
    const float *offset = (const float *)(h + 1), *scale = offset + h->dim;
    const float *coef = scale + h->dim, *sv = coef + h->n_sv;
    float z[64];
    uint32_t i, j;
    for (i = 0; i < h->dim; i++) z[i] = (feat[i] - offset[i]) * scale[i];
    double s = 0;
    for (j = 0; j < h->n_sv; j++) {
        double d2 = 0;
        for (i = 0; i < h->dim; i++) { double t = (double)z[i] - (double)sv[j * h->dim + i]; d2 += t * t; }
        s += (double)coef[j] * exp(-(double)h->gamma * d2);
    }
    s += h->rho;
    double d = -s;
    double r01 = sigmoid_predict(s, h->prob_a, h->prob_b), p[2];
    if (r01 < 1e-7) r01 = 1e-7;
    if (r01 > 1 - 1e-7) r01 = 1 - 1e-7;
    multiclass_probability2(r01, p);
    /* the GPU forms the score in float; the sign tests are taken on the float-rounded values there as here */
    int label = label_rule == SJ_LABEL_VOTE ? ((float)s <= 0.f ? 1 : 0)
              : label_rule == SJ_LABEL_PROB_ARGMAX ? (p[1] > p[0] ? 1 : 0) : ((float)d > 0.f ? 1 : 0);
    if (conf) *conf = (float)p[label];
    if (score) *score = (float)d;
    return label;
}

/* batch driver used by tests and by the timed CPU baseline: clips are [n_clips][n_samples] */
long sj_oracle_infer_batch(const sj_oracle_t *o, const void *blob, const float *clips, long n_clips,
                           long n_samples, int *labels, float *conf, float *score, float *feats)
{
    long c;
    for (c = 0; c < n_clips; c++) {
        float f[FEAT_DIM], cf = 0, sc = 0;
        if (sj_oracle_mfcc_stats(o, clips + c * n_samples, n_samples, f, NULL) < 0) return -1;
        int l = sj_oracle_predict_ex(blob, f, o->prm.label_rule, &cf, &sc);
        if (labels) labels[c] = l;
        if (conf) conf[c] = cf;
        if (score) score[c] = sc;
        if (feats) memcpy(feats + c * FEAT_DIM, f, sizeof f);
    }
    return n_clips;
}
