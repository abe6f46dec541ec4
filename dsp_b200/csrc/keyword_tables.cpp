// keyword_tables.cpp -- host-side construction of the keyword kernel's tables from an "SJKW" blob
// (the constants of 2fa/audio/word/c/mfcc_params.h and model_params.h, exported by tools/export_keyword.py).
#include <math.h>
#include <stdio.h>
#include <string.h>
#include <vector>

#include "keyword_kernel.h"

static int kw_fail(char *err, const char *msg)
{
    snprintf(err, 256, "%s", msg);
    return -1;
}

int kw_build_tables(KwTables *t, const void *blob, size_t bytes, char *err)
{
    memset(t, 0, sizeof *t);
    uint32_t h[16];
    if (!blob || bytes < sizeof h) return kw_fail(err, "keyword model blob too small");
    memcpy(h, blob, sizeof h);
    if (h[0] != 0x574B4A53u || h[1] != 1) return kw_fail(err, "bad keyword model magic/version");
    if (h[2] != 16000 || h[3] != KW_NFFT || h[4] != KW_FRAME || h[5] != KW_HOP || h[6] != KW_NMEL || h[7] != KW_NMFCC ||
        h[8] != KW_NBINS || h[9] != KW_MAX_FRAMES || h[10] != KW_INPUT || h[11] != 4 || h[12] != 2 || h[13] != 2 || h[14] != 1)
        return kw_fail(err, "keyword model shape differs from the compiled kernel (512/400/160, 40 mels, 13 mfcc, 500 frames, 4-2-2-1)");
    const size_t nfl = KW_FRAME + KW_NMEL * KW_NBINS + KW_NMFCC * KW_NMEL + 2 * KW_INPUT + KW_INPUT * 4 + 4 + 8 + 2 + 4 + 2 + 2 + 1;
    if (bytes < sizeof h + nfl * 4) return kw_fail(err, "keyword model blob truncated");
    const float *p = (const float *)((const char *)blob + sizeof h);
    const float *hann = p; p += KW_FRAME;
    const float *mel = p; p += KW_NMEL * KW_NBINS;
    const float *dct = p; p += KW_NMFCC * KW_NMEL;
    memcpy(t->mean, p, sizeof t->mean); p += KW_INPUT;
    memcpy(t->scale, p, sizeof t->scale); p += KW_INPUT;
    memcpy(t->w1, p, sizeof t->w1); p += KW_INPUT * 4;
    memcpy(t->b1, p, 16); p += 4;
    memcpy(t->w2, p, 32); p += 8;
    memcpy(t->b2, p, 8); p += 2;
    memcpy(t->w3, p, 16); p += 4;
    memcpy(t->b3, p, 8); p += 2;
    memcpy(t->w4, p, 8); p += 2;
    memcpy(t->b4, p, 4);
    for (int m = 0; m < KW_FRAME / 2; ++m) { t->hann2[m].x = hann[2 * m]; t->hann2[m].y = hann[2 * m + 1]; }
    const double two_pi = 6.283185307179586476925286766559;
    for (int a = 0; a < 16; ++a)
        for (int b = 0; b < 16; ++b) {
            const double ang = two_pi * (a * b) / 256.0;
            t->tw16[a][b].x = (float)cos(ang);
            t->tw16[a][b].y = (float)(-sin(ang));
        }
    for (int k = 0; k <= 128; ++k) {
        const double ang = two_pi * k / 512.0;
        t->tp[k].x = (float)cos(ang);
        t->tp[k].y = (float)(-sin(ang));
    }
    for (int l = 0; l < 26; ++l)
        for (int j = 0; j < 20; ++j) t->dct[j][l] = dct[(l % 13) * KW_NMEL + 20 * (l / 13) + j];
    // sparse chunked mel bank (same walk as the cepstrum kernel's generic layout, sj_tables.cpp build_mel)
    std::vector<int> f0(KW_NBINS, -1);
    std::vector<float> wlo(KW_NBINS, 0.f), whi(KW_NBINS, 0.f);
    for (int k = 0; k < KW_NBINS; ++k) {
        int first = -1, last = -1, cnt = 0;
        for (int m = 0; m < KW_NMEL; ++m)
            if (mel[m * KW_NBINS + k] != 0.f) { if (first < 0) first = m; last = m; ++cnt; }
        if (!cnt) continue;
        if (cnt > 2 || last - first > 1) return kw_fail(err, "mel bank is not two-adjacent-filters-per-bin sparse");
        f0[k] = first;
        wlo[k] = mel[first * KW_NBINS + k];
        if (cnt == 2) whi[k] = mel[last * KW_NBINS + k];
    }
    // per lane: f (lowest filter touching the bin) never decreases and steps by one at a triangle edge
    std::vector<int> first_f(32, -1), nstep(32, 0);
    int maxslot = 2;
    for (int L = 0; L < 32; ++L) {
        int cur = -1;
        for (int i = 0; i < KW_MEL_CH; ++i) {
            const int k = L * KW_MEL_CH + i;
            float a = 0.f, b = 0.f;
            bool step = false;
            if (k < KW_NBINS && f0[k] >= 0) {
                a = 0.25f * wlo[k];
                b = 0.25f * whi[k];
                if (cur < 0) { cur = f0[k]; first_f[L] = cur; }
                else if (f0[k] != cur) {
                    if (f0[k] != cur + 1) return kw_fail(err, "mel bank edges skip a filter inside a lane chunk");
                    cur = f0[k];
                    step = true;
                    ++nstep[L];
                }
            }
            float ax = step ? -a : a;
            if (step && a == 0.f) { uint32_t bits = 0x80000000u; memcpy(&ax, &bits, 4); }
            t->melw[i][L].x = ax;
            t->melw[i][L].y = b;
        }
        if (cur >= 0 && nstep[L] + 2 > maxslot) maxslot = nstep[L] + 2;
    }
    t->mel_maxseg = maxslot;
    const int zero_slot = 32 * maxslot;
    std::vector<std::vector<int>> lists(64);
    for (int L = 0; L < 32; ++L) {
        if (first_f[L] < 0) continue;
        for (int s = 0; s <= nstep[L] + 1; ++s) {
            const int f = first_f[L] + s;
            if (f < KW_NMEL) lists[f].push_back(L * maxslot + s);
        }
    }
    size_t ncomb = 1;
    for (int m = 0; m < KW_NMEL; ++m) if (lists[m].size() > ncomb) ncomb = lists[m].size();
    if (ncomb > KW_MAX_COMB) return kw_fail(err, "mel combine list too long");
    t->mel_ncomb = (int)ncomb;
    for (size_t e = 0; e < KW_MAX_COMB; ++e)
        for (int m = 0; m < 32; ++m) {
            const uint32_t i0 = (uint32_t)((e < lists[m].size()) ? lists[m][e] : zero_slot);
            const uint32_t i1 = (uint32_t)((m + 32 < KW_NMEL && e < lists[m + 32].size()) ? lists[m + 32][e] : zero_slot);
            t->comb2[e][m] = i0 | (i1 << 16);
        }
    return 0;
}
