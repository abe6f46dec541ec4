// keyword_kernel.h -- tables and launch interface of the fused keyword ("stop") detector kernel:
// compute_mfcc (2fa/audio/word/c/mfcc.c:108-232) + classify_signal's re-layout (stop_detector.c:12-55) +
// audio_classifier_predict (audio_classifier_inference.c:38-90), one CTA per clip.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#define KW_NFFT 512
#define KW_FRAME 400
#define KW_HOP 160
#define KW_NMEL 40
#define KW_NMFCC 13
#define KW_NBINS 257
#define KW_MAX_FRAMES 500
#define KW_INPUT (KW_NMFCC * KW_MAX_FRAMES)
#define KW_MEL_CH 9      /* power-spectrum bins per lane in the sparse mel stage (32 * 9 >= 257) */
#define KW_MAX_COMB 8

struct KwPair { float x, y; };

// per-frame tables (staged into shared memory by every CTA) followed by the network (read once per clip)
struct KwTables {
    KwPair hann2[208];         // window as pairs (w[2m], w[2m+1]), m < 200, zero beyond (frame is zero-padded to 512)
    KwPair tw16[16][16];       // step-2 twiddles of the 16 x 16 split of the 256-point complex FFT: W_256^(k1*m2)
    KwPair tp[144];            // real-FFT untangling twiddles W_512^k, k = 0..128
    // sparse mel bank over the POWER spectrum, chunked like the cepstrum kernel's generic layout: lane L owns bins
    // L*9 + i; melw[i][L] = weights for filters f, f+1 (f = lowest filter touching the bin), already times 0.25
    // (the untangling step yields 2 X); sign bit of .x = "f stepped": flush the finished sum to the next slot.
    KwPair melw[KW_MEL_CH][32];
    int32_t mel_maxseg, mel_ncomb, pad0, pad1;
    uint32_t comb2[KW_MAX_COMB][32]; // packed P indices for filters lane (low 16) and lane + 32 (high 16)
    float dct[20][32];         // lane l < 26: coefficient l % 13, mel rows 20*(l / 13) + j
    // ---- network (global memory only)
    float mean[KW_INPUT], scale[KW_INPUT];
    float w1[KW_INPUT][4];
    float b1[4], w2[4][2], b2[2], w3[2][2], b3[2], w4[2], b4[1], pad2[3];
};

// parses an "SJKW" v1 blob (tools/export_keyword.py); returns 0 or -1 with a message in err (>= 256 bytes)
int kw_build_tables(KwTables *t, const void *blob, size_t bytes, char *err);

struct KwArgs {
    const float *signals;  // [n_clips][n_samples] device
    long n_clips;
    int n_samples;
    const KwTables *tables; // device
    float *prob;           // [n_clips]
    float *mfcc;           // [n_clips][500][13] frame-major like compute_mfcc's out_mfcc (zero beyond the frames), or NULL
    int32_t *n_frames;     // [n_clips] or NULL
};
cudaError_t kw_setup();
cudaError_t kw_launch(const KwArgs &a, int mel_maxseg, int num_sms, cudaStream_t stream);
