// sj_tables.h -- constant tables of the cepstrum kernel, built once per context on the host
// (sj_tables.cpp) and copied to device global memory; every CTA stages them into shared memory.
#pragma once
#include <stddef.h>
#include <stdint.h>

#include "../../include/scrubjay_cuda.h"

#define SJ_WIN 2048
#define SJ_HOP 1024
#define SJ_NBINS 1025
#define SJ_NFILT 40
#define SJ_NCOEF 20
#define SJ_FEAT 40
#define SJ_MAX_SV 64  /* support vectors are padded to 64 (the shipped model has 55) */
#define SJ_MAX_CH 32  /* max bins per lane chunk in the sparse mel stage */
#define SJ_MAX_COMB 16

struct SjPair { float x, y; };
struct SjQuad { float x, y, z, w; };

#define SJ_MEL_FAST_CH 28   /* bins per lane of the fast mel path (multiple of 4) */
#define SJ_MEL_FAST_SLOTS 6 /* fast path: one flush slot per aligned group of 8 bins (4) + the two final sums */

// Everything the kernel reads besides the clips.  Plain POD, identical layout on host and device.
// The per-frame ("hot") tables come first; the kernel stages everything before sv_off into shared memory.
struct SjTables {
    // step-2 twiddles of the 32x32 split of the 1024-point complex FFT, paired for the merged first
    // butterfly stage: tw4[n][lane] = (W^(n*lane), W^((n+16)*lane)), W = exp(-2 pi i/1024)
    SjQuad tw4[16][32];
    // sparse Slaney bank, chunked: lane L owns bins mel_lo + L*mel_ch + i, i < mel_ch (mel_lo and
    // mel_ch are multiples of 4).  melw4[g][L] = (w_lo, w_hi) of bins 2g and 2g+1 of the chunk: weights
    // for filters f and f+1 where f is the lowest filter touching the bin; they carry the 0.5 that the
    // untangling step leaves out.  The sign bit of w_lo is set on the first bin where f has stepped to
    // f+1: the running sum of the old f is flushed and the f+1 sum becomes the lower sum.
    //   fast layout (mel_fast = 1, mel_ch = 28): the flush of a step at chunk bin i goes to slot i/8 (a
    //     chunk never holds two steps in one aligned group of 8 bins), the two final sums to slots 4, 5;
    //     partial (slot, L) lives at P[slot*32 + L].
    //   generic layout: slots are filled in order, partial (L, s) at P[L*mel_maxseg + s].
    SjQuad melw4[SJ_MAX_CH / 2][32];
    int32_t mel_lo, mel_ch, mel_maxseg, mel_ncomb;
    int32_t mel_fast, mel_pskip;
    // table-driven switches of the generic instantiation (scrubjay_cepstrum_params): SJ_GEN_WINDOW_TABLE = the window
    // comes from `window` below instead of the in-register periodic Hann; SJ_GEN_POWER = the filterbank works on
    // |X|^2 (the mel weights then carry 0.25 instead of 0.5)
    int32_t gen_flags;
    float log_floor; // energies are raised to this before log10 (aubio VERY_SMALL_NUMBER = 2e-42)
    // combine lists: energy of filter m = sum_e P[comb[e][m]] over e < mel_ncomb; unused entries point
    // at a zero slot.  comb2[e][lane] packs the entries of filters lane (low 16 bits) and lane+32 (high).
    uint32_t comb2[SJ_MAX_COMB][32];
    // orthonormal DCT-II, transposed: dct[m][c], c < 20 (padded to 32)
    float dct[SJ_NFILT][32];
    // SVM (ONNX ai.onnx.ml Scaler + SVMClassifier attributes)
    float sv_off[SJ_FEAT], sv_scale[SJ_FEAT];
    float sv[SJ_FEAT][SJ_MAX_SV]; // support vectors transposed: sv[i][j] = SV_j[i]
    float sv_coef[SJ_MAX_SV];     // zero beyond n_sv
    float gamma, rho, prob_a, prob_b;
    int32_t n_sv, label_rule /* SJ_LABEL_* */, pad1, pad2;
    // real-FFT untangling base twiddles W_2048^lane = (x, y) (the kernel reads its own entry once, from global memory)
    SjQuad tp4[1][32];
    // analysis window of the generic instantiation when SJ_GEN_WINDOW_TABLE is set (read through L1 from global memory)
    float window[SJ_WIN];
};
#define SJ_GEN_WINDOW_TABLE 1
#define SJ_GEN_POWER 2

// Builds the tables.  Returns 0, or a negative value and a message in `err` (size >= 256).
// `params` NULL = the defaults (restated aubio 0.4.x / onnxruntime behaviour).
int sj_build_tables(SjTables *t, const void *blob, size_t blob_bytes, unsigned samplerate,
                    const scrubjay_cepstrum_params *params, char *err);
void sj_default_cepstrum_params(scrubjay_cepstrum_params *p);
// dense [40][1025] Slaney bank (aubio filterbank_mel.c semantics), used by the builder and by tests
void sj_dense_slaney(float *fb /*[40][1025]*/, unsigned samplerate);
// dense [40][1025] bank recovered by replaying the kernel's sparse walk over the built tables (weights carry 0.5)
void sj_replay_mel_dense(const SjTables *t, float *out);
