// donut_kernel.h -- launch interface and tables of the fused donut-classifier kernel
// (Butterworth IIR x2 -> spectrogram x2 -> band energies -> verdict), float and double flavours.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define DN_WIN 256
#define DN_HOP 224
#define DN_BINS 129
#define DN_MAX_TB 96      /* frames per clip the kernel supports: n_samples <= 256 + 224*95 */
#define DN_MAX_MID 8
#define DN_CLIPS_PER_WARP 16

// Variant constants; mirrors the reference's literals.
struct SjDonutParams {
    double lo_norm, hi_norm;   // classifier.c:141-142 / classifier.cpp:67-68
    double mid_thr_db;         // classifier.c:660 / classifier.cpp:436
    double gap_tol, min_dur;   // classifier.c:749-750 / classifier.cpp:523-524
    double thr_mid, thr_above, thr_below; // classifier.c:184 / classifier.cpp:109
    double band[3][2], half[3];           // sum_intense boxes: above, middle, below (classifier.c:175-177 / classifier.ino:123-125)
    int nperseg;                          // 256, or 64 for classifier.ino:249
};

template <typename T> struct SjDonutTables {
    T window[DN_WIN];          // periodic Tukey(0.25), built on the host with the reference's formula
    T tw128[32][4][2];         // W_128^(g*k1): [k1][g][re,im]
    T tw256[DN_BINS + 3][2];   // W_256^k, k = 0..128 (padded)
    // twiddles of the reference's float FFT (PlainFFT::Compute, sync/lib/PlainFFT.cpp:29-94: radix-2 DIT
    // whose (u1, u2) come from a float recurrence); stage l, butterfly offset m at index (2^l - 1) + m
    T bourke[256][2];
    T a[2][9], b[2][9];        // IIR coefficients: [0] = 1000-3000 Hz, [1] = 3000-7500 Hz
    T U;                       // fs * sum(window^2)
    T fs;
    T s_thr;                   // smallest Sxx whose dB value exceeds mid_thr_db (exact, by bisection)
    // The kernel works on a = re^2 + im^2 and applies the reference's "/ U" and "* 2" only where a value is
    // needed; every test on Sxx is moved to a through exact thresholds (Sxx is a monotone function of a):
    T a_thr[2];                // smallest a with Sxx(a) >= s_thr: [0] edge bins (0, 128), [1] inner bins (doubled)
    T a_pos;                   // smallest a with a / U > 0 (the reference's "Sxx > 0" test)
    int32_t box_f[3][2];       // sum_intense's inclusive frequency-bin range of the above / middle / below band
    int32_t pad_i[2];
    T lo_norm, hi_norm, gap_tol, min_dur, thr_mid, thr_above, thr_below;
    T pad;
};

struct SjDonutArgs {
    const float *clips;   // [n_clips][n_samples] device
    long n_clips;
    int clips_per_cta;    // set by sj_launch_donut: clips a CTA owns at a time (1 .. DN_CLIPS_PER_WARP)
    int n_samples;
    int tb;               // frames per clip = (n_samples - 256) / 224 + 1
    const void *tables;   // SjDonutTables<T> on the device
    void *scratch;        // [grid][16][tb][129] of T: the 3000-7500 Hz PSD of the clips in flight
    int32_t *labels;      // [n_clips]
    float *detail;        // [n_clips][4] = n_midpoints, above, middle, below (last midpoint evaluated) or NULL
    float *mids;          // [n_clips][DN_MAX_MID] midpoint times in seconds (NaN-padded) or NULL
    int8_t *track;        // [n_clips] label + 1 (the firmware's per-second classification byte) or NULL
    long row_stride;      // stream form: clip c = second (c % clips_per_row) of channel (c / clips_per_row),
    int clips_per_row;    // channels row_stride samples apart; 0 = plain contiguous clips
};

// builds the tables for precision T on the host (window/twiddles/threshold exactly as the reference computes them)
void sj_donut_build_tables_f32(SjDonutTables<float> *t, const SjDonutParams &p, int fs);
void sj_donut_build_tables_f64(SjDonutTables<double> *t, const SjDonutParams &p, int fs);
void sj_donut_default_params(SjDonutParams *p, int variant); // 0: classifier.c (f64 row), 1: sync/lib (f32 row), 2, 3, 4: classifier.ino

// PSD scratch bytes for a launch on `num_sms` SMs (whichever kernel it takes); the launch picks kernel and grid itself:
// SCRUBJAY_DONUT_LEGACY = 1 | f32 | f64 forces the round-1 kernels, = 0 the warp-specialised ones (donut_kernel.cu: use_round1)
size_t sj_donut_scratch_bytes(int num_sms, int tb, bool f64);
cudaError_t sj_donut_setup();
cudaError_t sj_launch_donut(const SjDonutArgs &a, bool f64, int num_sms, cudaStream_t stream);
