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

void sj_default_cepstrum_params(scrubjay_cepstrum_params *p)
{
    memset(p, 0, sizeof *p);
    p->window = SJ_WINDOW_HANN_PERIODIC;
    p->spectrum = SJ_SPECTRUM_MAGNITUDE;
    p->log_floor = 2.e-42f;
    p->dct_norm = SJ_DCT_ORTHO;
    p->half_swap = 1;
    p->label_rule = SJ_LABEL_PREDICT;
}

// `wscale`: factor folded into the weights -- the untangling step leaves |X| doubled (0.5), |X|^2 quadrupled (0.25)
// `allow_fast`: false forces the generic partial-sum layout (the specialised 16 kHz instantiation hard-codes the
// in-register periodic Hann, |X| and the 2e-42 floor, so non-default parameters must not get its layout)
static int build_mel(SjTables *t, unsigned samplerate, float wscale, bool allow_fast, char *err)
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
    // the untangling step can skip the mirror bin 1024-k of the first `pskip` rows (k = lane + 32 j,
    // j < pskip) when those mirrors lie above the bank's support
    int pskip = 0;
    while (pskip < 16 && 1024 - 32 * pskip - 31 > hi) ++pskip;
    t->mel_pskip = pskip;
    // chunk layout: lane L owns bins base + L*ch + i; base and ch multiples of 4 (128-bit loads)
    const int base = lo & ~3;
    int ch = (hi - base + 1 + 31) / 32;
    ch = (ch + 3) & ~3;
    if (ch > SJ_MAX_CH) return fail(err, "filterbank support too wide");
    bool fast = allow_fast && ch <= SJ_MEL_FAST_CH;
    if (fast) ch = SJ_MEL_FAST_CH;
    int lo_al = base;
    if (lo_al + 32 * ch > 1152) lo_al = (1152 - 32 * ch) & ~3; // stay inside the per-warp magnitude buffer
    // Per lane: walk the chunk; f (lowest filter touching the bin) never decreases and steps by one at
    // a triangle edge.  The lane keeps two running sums (filter f, filter f+1); at a step the sum of
    // filter f is complete for this lane and is flushed, the f+1 sum becomes the new "lower" sum.
    std::vector<int> first_f(32, -1), last_f(32, -1);
    std::vector<std::vector<int>> step_at(32); // chunk positions of the steps of each lane
    int maxslot = 2;
    memset(t->melw4, 0, sizeof t->melw4);
    for (int L = 0; L < 32; ++L) {
        int cur = -1;
        for (int i = 0; i < ch; ++i) {
            int k = lo_al + L * ch + i;
            float a = 0.f, b = 0.f;
            bool step = false;
            if (k < SJ_NBINS && f0[k] >= 0) {
                a = wscale * wlo[k];
                b = wscale * whi[k];
                if (cur < 0) { cur = f0[k]; first_f[L] = cur; }
                else if (f0[k] != cur) {
                    if (f0[k] != cur + 1) return fail(err, "filterbank edges skip a filter inside a bin chunk (unsupported samplerate)");
                    cur = f0[k];
                    step = true;
                    step_at[L].push_back(i);
                }
            }
            // step flag = sign bit of w_lo (weights are >= 0); -0.0f works for a zero weight
            float ax = step ? -a : a;
            if (step && a == 0.f) { uint32_t bits = 0x80000000u; memcpy(&ax, &bits, 4); }
            SjQuad &q = t->melw4[i >> 1][L];
            if (i & 1) { q.z = ax; q.w = b; } else { q.x = ax; q.y = b; }
        }
        last_f[L] = cur;
        if (cur >= 0 && cur - first_f[L] + 2 > maxslot) maxslot = cur - first_f[L] + 2;
        for (size_t s = 1; s < step_at[L].size(); ++s)
            if (step_at[L][s] / 8 == step_at[L][s - 1] / 8) fast = false; // two steps in one group of 8 bins
    }
    t->mel_lo = lo_al;
    t->mel_ch = ch;
    t->mel_maxseg = maxslot;
    t->mel_fast = fast ? 1 : 0;
    // combine lists.  Slot s (in flush order) of lane L holds the lane's partial energy of filter
    // first_f[L] + s; its position in P depends on the layout.
    const int zero_slot = fast ? 32 * SJ_MEL_FAST_SLOTS : 32 * maxslot;
    std::vector<std::vector<int>> lists(64);
    for (int L = 0; L < 32; ++L) {
        if (first_f[L] < 0) continue;
        const int nstep = (int)step_at[L].size();
        for (int s = 0; s <= nstep + 1; ++s) {
            const int f = first_f[L] + s;
            if (f >= SJ_NFILT) continue;
            int idx;
            if (!fast) idx = L * maxslot + s;
            else if (s < nstep) idx = (step_at[L][s] / 8) * 32 + L;
            else idx = (SJ_MEL_FAST_SLOTS - 2 + (s - nstep)) * 32 + L;
            lists[f].push_back(idx);
        }
    }
    size_t ncomb = 1;
    for (int m = 0; m < SJ_NFILT; ++m) if (lists[m].size() > ncomb) ncomb = lists[m].size();
    if (ncomb > SJ_MAX_COMB) return fail(err, "mel combine list too long");
    t->mel_ncomb = (int)ncomb;
    for (size_t e = 0; e < SJ_MAX_COMB; ++e)
        for (int m = 0; m < 32; ++m) {
            const uint32_t i0 = (uint32_t)((m < SJ_NFILT && e < lists[m].size()) ? lists[m][e] : zero_slot);
            const uint32_t i1 = (uint32_t)((m + 32 < SJ_NFILT && e < lists[m + 32].size()) ? lists[m + 32][e] : zero_slot);
            t->comb2[e][m] = i0 | (i1 << 16);
        }
    return 0;
}

// Replays the kernel's mel walk symbolically and expands the chunked sparse tables back to a dense
// [40][1025] bank (diagnostic; tests compare it with an independent construction).
void sj_replay_mel_dense(const SjTables *t, float *out)
{
    memset(out, 0, sizeof(float) * SJ_NFILT * SJ_NBINS);
    const bool fast = t->mel_fast != 0;
    const int nP = (fast ? 32 * SJ_MEL_FAST_SLOTS : 32 * t->mel_maxseg) + 1;
    typedef std::vector<std::pair<int, float>> Terms; // (bin, weight)
    std::vector<Terms> P(nP);
    for (int L = 0; L < 32; ++L) {
        Terms lo, hi;
        int slot = 0;
        for (int i = 0; i < t->mel_ch; ++i) {
            const SjQuad &q = t->melw4[i >> 1][L];
            const float wx = (i & 1) ? q.z : q.x, wy = (i & 1) ? q.w : q.y;
            uint32_t bits;
            memcpy(&bits, &wx, 4);
            if (bits & 0x80000000u) {
                const int idx = fast ? (i / 8) * 32 + L : L * t->mel_maxseg + slot;
                P[idx] = lo;
                ++slot;
                lo = hi;
                hi.clear();
            }
            const int k = t->mel_lo + L * t->mel_ch + i;
            lo.push_back({k, fabsf(wx)});
            hi.push_back({k, wy});
        }
        if (fast) {
            P[(SJ_MEL_FAST_SLOTS - 2) * 32 + L] = lo;
            P[(SJ_MEL_FAST_SLOTS - 1) * 32 + L] = hi;
        } else {
            P[L * t->mel_maxseg + slot] = lo;
            if (slot + 1 < t->mel_maxseg) P[L * t->mel_maxseg + slot + 1] = hi;
        }
    }
    for (int m = 0; m < SJ_NFILT; ++m)
        for (int e = 0; e < t->mel_ncomb; ++e) {
            const uint32_t c = t->comb2[e][m & 31];
            const int idx = (m < 32) ? (int)(c & 0xffffu) : (int)(c >> 16);
            if (idx >= nP - 1) continue;
            for (auto &bw : P[idx])
                if (bw.first < SJ_NBINS && bw.second != 0.f) out[(size_t)m * SJ_NBINS + bw.first] += bw.second;
        }
}

struct SjsvHeader {
    uint32_t magic, version, dim, n_sv, nsv0, nsv1;
    float gamma, rho, prob_a, prob_b;
};

int sj_build_tables(SjTables *t, const void *blob, size_t blob_bytes, unsigned samplerate,
                    const scrubjay_cepstrum_params *params, char *err)
{
    memset(t, 0, sizeof(*t));
    if (samplerate == 0) return fail(err, "samplerate must be > 0");
    scrubjay_cepstrum_params prm;
    sj_default_cepstrum_params(&prm);
    if (params) prm = *params;
    if (prm.window != SJ_WINDOW_HANN_PERIODIC && prm.window != SJ_WINDOW_HANN_SYMMETRIC) return fail(err, "params: unknown window");
    if (prm.spectrum != SJ_SPECTRUM_MAGNITUDE && prm.spectrum != SJ_SPECTRUM_POWER) return fail(err, "params: unknown spectrum kind");
    if (prm.dct_norm != SJ_DCT_ORTHO && prm.dct_norm != SJ_DCT_UNNORMALISED) return fail(err, "params: unknown DCT normalisation");
    if (prm.label_rule < SJ_LABEL_PREDICT || prm.label_rule > SJ_LABEL_PROB_ARGMAX) return fail(err, "params: unknown label rule");
    if (!(prm.log_floor > 0.f)) return fail(err, "params: log_floor must be > 0");
    if (prm.reserved[0] || prm.reserved[1]) return fail(err, "params: reserved fields must be 0");
    const double two_pi = 6.283185307179586476925286766559;
    for (int a = 0; a < 16; ++a)
        for (int b = 0; b < 32; ++b) {
            double a0 = two_pi * (double)(a * b) / 1024.0, a1 = two_pi * (double)((a + 16) * b) / 1024.0;
            t->tw4[a][b].x = (float)cos(a0);
            t->tw4[a][b].y = (float)(-sin(a0));
            t->tw4[a][b].z = (float)cos(a1);
            t->tw4[a][b].w = (float)(-sin(a1));
        }
    for (int b = 0; b < 32; ++b) {
        const double a0 = two_pi * b / 2048.0;
        t->tp4[0][b].x = (float)cos(a0);
        t->tp4[0][b].y = (float)(-sin(a0));
        t->tp4[0][b].z = 0.f;
        t->tp4[0][b].w = 0.f;
    }
    const bool power = prm.spectrum == SJ_SPECTRUM_POWER;
    const int gen_flags = (power ? SJ_GEN_POWER : 0) | (prm.window != SJ_WINDOW_HANN_PERIODIC ? SJ_GEN_WINDOW_TABLE : 0);
    if (build_mel(t, samplerate, power ? 0.25f : 0.5f, !gen_flags && prm.log_floor == 2.e-42f, err)) return -1;
    t->gen_flags = gen_flags;
    t->log_floor = prm.log_floor;
    t->label_rule = prm.label_rule;
    for (int i = 0; i < SJ_WIN; ++i) {
        const double den = prm.window == SJ_WINDOW_HANN_SYMMETRIC ? (double)(SJ_WIN - 1) : (double)SJ_WIN;
        t->window[i] = (float)(0.5 * (1.0 - cos(two_pi * (double)i / den)));
    }
    // aubio mfcc.c: orthonormal DCT-II over the 40 log energies, first 20 rows kept (or the unnormalised 2 sum cos form)
    const double pi = two_pi / 2;
    for (int m = 0; m < SJ_NFILT; ++m)
        for (int c = 0; c < SJ_NCOEF; ++c) {
            double s = c == 0 ? sqrt(1.0 / SJ_NFILT) : sqrt(2.0 / SJ_NFILT);
            if (prm.dct_norm == SJ_DCT_UNNORMALISED) s = 2.0;
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
