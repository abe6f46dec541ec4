// sj_tables.cpp -- host-side construction of the kernel tables (window, twiddles, sparse Slaney
// filterbank, DCT, SVM weights).  Runs once in scrubjay_create.
#include <math.h>
#include <stdio.h>
#include <string.h>
#include <vector>

#include "sj_tables.h"

extern const unsigned int sj_default_model_words[];
extern const size_t sj_default_model_bytes;
const unsigned int sj_default_model_words[] = {
#include "scrubjay_svm_default.inc"
};
const size_t sj_default_model_bytes = sizeof(sj_default_model_words);

// Slaney mel bank exactly as aubio builds it when n_filters == 40 (the case selected by
// new_aubio_mfcc(WIN_SIZE, 40, 20, sr) at cepstrum/scrubjay_infer.c:30): 13 linearly spaced edges
// from 133.3333 Hz in steps of 66.66666666 Hz, then 29 edges growing by 1.0711703 per step;
// triangle m spans edges m, m+1, m+2 with unit area; bin k sits at k*sr/2048 Hz.  All float, as
// aubio's smpl_t.  The walk over bins mirrors aubio_filterbank_set_triangle_bands: the rising side
// starts at the bin after the last one <= lower, the falling side takes over at the first bin whose
// successor test crosses the centre, and negative values are clamped.
void sj_dense_slaney(float *fb, unsigned samplerate)
{
    float edge[SJ_NFILT + 2];
    for (int i = 0; i < 13; ++i) edge[i] = 133.3333f + i * 66.66666666f;
    for (int i = 0; i < 29; ++i) edge[13 + i] = edge[12] * powf(1.0711703f, (float)(i + 1));
    std::vector<float> hz(SJ_NBINS);
    const float per_bin = (float)samplerate / (float)SJ_WIN;
    for (int k = 0; k < SJ_NBINS; ++k) hz[k] = per_bin * (float)k;
    memset(fb, 0, sizeof(float) * SJ_NFILT * SJ_NBINS);
    const int last = SJ_NBINS - 1;
    for (int m = 0; m < SJ_NFILT; ++m) {
        float *row = fb + (size_t)m * SJ_NBINS;
        const float lo = edge[m], mid = edge[m + 1], hi = edge[m + 2];
        const float peak = 2.f / (hi - lo);
        int k = 0;
        while (k < last) {
            bool hit = hz[k] <= lo && hz[k + 1] > lo;
            ++k;
            if (hit) break;
        }
        const float up = peak / (mid - lo);
        while (k < last) {
            row[k] = (hz[k] - lo) * up;
            bool crossed = hz[k + 1] >= mid;
            ++k;
            if (crossed) break;
        }
        const float down = peak / (hi - mid);
        while (k < last) {
            row[k] += (hi - hz[k]) * down;
            if (row[k] < 0.f) row[k] = 0.f;
            if (hz[k + 1] >= hi) break;
            ++k;
        }
    }
}

static int fail(char *err, const char *msg)
{
    snprintf(err, 256, "%s", msg);
    return -1;
}

static int build_mel(SjTables *t, unsigned samplerate, char *err)
{
    std::vector<float> fb((size_t)SJ_NFILT * SJ_NBINS);
    sj_dense_slaney(fb.data(), samplerate);
    // per bin: lowest touching filter f, weights for f and f+1
    std::vector<int> f0(SJ_NBINS, -1);
    std::vector<float> wlo(SJ_NBINS, 0.f), whi(SJ_NBINS, 0.f);
    int lo = -1, hi = -1;
    for (int k = 0; k < SJ_NBINS; ++k) {
        int first = -1, cnt = 0, lastm = -1;
        for (int m = 0; m < SJ_NFILT; ++m)
            if (fb[(size_t)m * SJ_NBINS + k] != 0.f) { if (first < 0) first = m; lastm = m; ++cnt; }
        if (!cnt) continue;
        if (cnt > 2 || lastm - first > 1)
            return fail(err, "filterbank at this samplerate is not two-adjacent-filters-per-bin sparse");
        f0[k] = first;
        wlo[k] = fb[(size_t)first * SJ_NBINS + k];
        if (cnt == 2) whi[k] = fb[(size_t)lastm * SJ_NBINS + k];
        if (lo < 0) lo = k;
        hi = k;
    }
    if (lo < 0) return fail(err, "empty filterbank");
    if (lo == 0 || hi == SJ_NBINS - 1) {
        // bins 0 and 1024 are produced only as a by-product of the untangling step; keep it simple
        if ((lo == 0 && wlo[0] != 0.f) || hi == SJ_NBINS - 1)
            return fail(err, "filterbank touches the DC/Nyquist bin (unsupported samplerate)");
    }
    int ch = (hi - lo + 1 + 31) / 32;
    if ((ch & 1) == 0) ++ch; // odd stride => conflict-free shared-memory reads across lanes
    if (ch > SJ_MAX_CH) return fail(err, "filterbank support too wide");
    if (lo + 32 * ch > 1056) lo = 1056 - 32 * ch; // stay inside the per-warp magnitude buffer
    t->mel_lo = lo;
    t->mel_ch = ch;
    // Per lane: walk the chunk; f (lowest filter touching the bin) never decreases and steps by one at
    // a triangle edge.  The lane keeps two running sums (filter f, filter f+1); at a step the sum of
    // filter f is complete for this lane and is stored, the f+1 sum becomes the new "lower" sum.
    // Slot s of lane L therefore holds the lane's partial energy of filter first_f[L] + s.
    std::vector<int> first_f(32, -1), last_f(32, -1);
    int maxslot = 2;
    for (int L = 0; L < 32; ++L) {
        int cur = -1;
        for (int i = 0; i < ch; ++i) {
            int k = lo + L * ch + i;
            float a = 0.f, b = 0.f;
            bool step = false;
            if (k < SJ_NBINS && f0[k] >= 0) {
                a = 0.5f * wlo[k];
                b = 0.5f * whi[k];
                if (cur < 0) { cur = f0[k]; first_f[L] = cur; }
                else if (f0[k] != cur) {
                    if (f0[k] != cur + 1) return fail(err, "filterbank edges skip a filter inside a bin chunk (unsupported samplerate)");
                    cur = f0[k];
                    step = true;
                }
            }
            // step flag = sign bit of w_lo (weights are >= 0); -0.0f works for a zero weight
            t->melw[i][L].x = step ? -a : a;
            t->melw[i][L].y = b;
            if (step && a == 0.f) { uint32_t bits = 0x80000000u; memcpy(&t->melw[i][L].x, &bits, 4); }
        }
        last_f[L] = cur;
        if (cur >= 0 && cur - first_f[L] + 2 > maxslot) maxslot = cur - first_f[L] + 2;
    }
    t->mel_maxseg = maxslot;
    // combine lists: partial (L, s) lives at float index L*maxslot + s; the zero slot is 32*maxslot
    const int zero_slot = 32 * maxslot;
    std::vector<std::vector<int>> lists(64);
    for (int L = 0; L < 32; ++L) {
        if (first_f[L] < 0) continue;
        for (int f = first_f[L]; f <= last_f[L] + 1 && f < SJ_NFILT; ++f) lists[f].push_back(L * maxslot + (f - first_f[L]));
    }
    size_t ncomb = 1;
    for (int m = 0; m < SJ_NFILT; ++m) if (lists[m].size() > ncomb) ncomb = lists[m].size();
    if (ncomb > SJ_MAX_COMB) return fail(err, "mel combine list too long");
    t->mel_ncomb = (int)ncomb;
    for (size_t e = 0; e < SJ_MAX_COMB; ++e)
        for (int m = 0; m < 64; ++m)
            t->comb[e][m] = (uint16_t)((m < SJ_NFILT && e < lists[m].size()) ? lists[m][e] : zero_slot);
    return 0;
}

struct SjsvHeader {
    uint32_t magic, version, dim, n_sv, nsv0, nsv1;
    float gamma, rho, prob_a, prob_b;
};

int sj_build_tables(SjTables *t, const void *blob, size_t blob_bytes, unsigned samplerate, char *err)
{
    memset(t, 0, sizeof(*t));
    if (samplerate == 0) return fail(err, "samplerate must be > 0");
    const double two_pi = 6.283185307179586476925286766559;
    for (int i = 0; i < SJ_WIN; ++i) t->window[i] = (float)(0.5 * (1.0 - cos(two_pi * i / SJ_WIN)));
    for (int a = 0; a < 32; ++a)
        for (int b = 0; b < 32; ++b) {
            double ang = two_pi * (double)(a * b) / 1024.0;
            t->tw32[a][b].x = (float)cos(ang);
            t->tw32[a][b].y = (float)(-sin(ang));
        }
    for (int k = 0; k < 512; ++k) {
        double ang = two_pi * k / 2048.0;
        t->twpost[k].x = (float)cos(ang);
        t->twpost[k].y = (float)(-sin(ang));
    }
    if (build_mel(t, samplerate, err)) return -1;
    // aubio mfcc.c: orthonormal DCT-II over the 40 log energies, first 20 rows kept
    const double pi = two_pi / 2;
    for (int m = 0; m < SJ_NFILT; ++m)
        for (int c = 0; c < SJ_NCOEF; ++c) {
            double s = c == 0 ? sqrt(1.0 / SJ_NFILT) : sqrt(2.0 / SJ_NFILT);
            t->dct[m][c] = (float)(s * cos(pi * c * (m + 0.5) / SJ_NFILT));
        }
    if (!blob) { blob = sj_default_model_words; blob_bytes = sj_default_model_bytes; }
    if (blob_bytes < sizeof(SjsvHeader)) return fail(err, "model blob too small");
    SjsvHeader h;
    memcpy(&h, blob, sizeof h);
    if (h.magic != 0x56534A53u || h.version != 1) return fail(err, "bad model blob magic/version");
    if (h.dim != SJ_FEAT) return fail(err, "model feature dimension must be 40");
    if (h.n_sv == 0 || h.n_sv > SJ_MAX_SV) return fail(err, "unsupported support-vector count");
    size_t need = sizeof h + sizeof(float) * (2 * h.dim + h.n_sv + (size_t)h.n_sv * h.dim);
    if (blob_bytes < need) return fail(err, "model blob truncated");
    const float *p = (const float *)((const char *)blob + sizeof h);
    memcpy(t->sv_off, p, sizeof(float) * SJ_FEAT); p += SJ_FEAT;
    memcpy(t->sv_scale, p, sizeof(float) * SJ_FEAT); p += SJ_FEAT;
    memcpy(t->sv_coef, p, sizeof(float) * h.n_sv); p += h.n_sv;
    for (uint32_t j = 0; j < h.n_sv; ++j)
        for (int i = 0; i < SJ_FEAT; ++i) t->sv[i][j] = p[(size_t)j * SJ_FEAT + i];
    t->gamma = h.gamma; t->rho = h.rho; t->prob_a = h.prob_a; t->prob_b = h.prob_b;
    t->n_sv = (int)h.n_sv;
    return 0;
}
