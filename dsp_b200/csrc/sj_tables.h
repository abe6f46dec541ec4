// sj_tables.h -- constant tables of the cepstrum kernel, built once per context on the host
// (sj_tables.cpp) and copied to device global memory; every CTA stages them into shared memory.
#pragma once
#include <stdint.h>

#define SJ_WIN 2048
#define SJ_HOP 1024
#define SJ_NBINS 1025
#define SJ_NFILT 40
#define SJ_NCOEF 20
#define SJ_FEAT 40
#define SJ_MAX_SV 64  /* support vectors are padded to 64 (the shipped model has 55) */
#define SJ_MAX_CH 33  /* max bins per lane chunk in the sparse mel stage */
#define SJ_MAX_COMB 16

struct SjPair { float x, y; };

// Everything the kernel reads besides the clips.  Plain POD, identical layout on host and device.
struct SjTables {
    // aubio "hanningz" window, w[i] = 0.5(1-cos(2 pi i/2048))  (pvoc.c / mathutils.c new_aubio_window)
    float window[SJ_WIN];
    // step-2 twiddles of the 32x32 split of the 1024-point complex FFT: W_1024^(k1*m2) at [k1][m2]
    SjPair tw32[32][32];
    // real-FFT untangling twiddles W_2048^k, k = 0..511
    SjPair twpost[512];
    // sparse Slaney bank, chunked: lane L owns bins mel_lo + L*mel_ch + i, i < mel_ch.
    // melw[i][L] = (w_lo, w_hi): weights of that bin for filters f and f+1 where f is the lowest
    // filter touching the bin; weights already carry the 0.5 that the untangling step leaves out.
    // The sign bit of w_lo is set on the first bin where f has stepped to f+1: the running sum of the
    // old f is stored to the lane's next slot and the f+1 sum becomes the lower sum.
    // mel_maxseg = slots per lane (filters a lane's chunk can touch).
    SjPair melw[SJ_MAX_CH][32];
    int32_t mel_lo, mel_ch, mel_maxseg, mel_ncomb;
    // combine lists: energy of filter m = sum_e P[comb[e][m]] over e < mel_ncomb, P = flat float
    // array [lane][slot] of per-lane partial energies; unused entries point at a zero slot.
    uint16_t comb[SJ_MAX_COMB][64];
    // orthonormal DCT-II, transposed: dct[m][c], c < 20 (padded to 32)
    float dct[SJ_NFILT][32];
    // SVM (ONNX ai.onnx.ml Scaler + SVMClassifier attributes)
    float sv_off[SJ_FEAT], sv_scale[SJ_FEAT];
    float sv[SJ_FEAT][SJ_MAX_SV]; // support vectors transposed: sv[i][j] = SV_j[i]
    float sv_coef[SJ_MAX_SV];     // zero beyond n_sv
    float gamma, rho, prob_a, prob_b;
    int32_t n_sv, pad0, pad1, pad2;
};

// Builds the tables.  Returns 0, or a negative value and a message in `err` (size >= 256).
int sj_build_tables(SjTables *t, const void *blob, size_t blob_bytes, unsigned samplerate, char *err);
// dense [40][1025] Slaney bank (aubio filterbank_mel.c semantics), used by the builder and by tests
void sj_dense_slaney(float *fb /*[40][1025]*/, unsigned samplerate);
