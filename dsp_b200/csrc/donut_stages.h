// donut_stages.h -- the donut classifier's stage functions as stand-alone batched kernels:
// butter_bandpass_filter (classifier.c:420-446 / sync/lib/classifier.cpp:193-219) and compute_spectrogram
// (classifier.c:448-592 / classifier.cpp:221-368; classifier.ino:245-395 for the 64-sample window).
// The fused kernel (donut_kernel.cu) is the hot path; these serve the reference's public helper symbols
// (sync/lib/classifier.h:14-18), arbitrary coefficient sets and signals of any length.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define SJ_STAGE_MAX_WIN 256
#define SJ_SCAN_BLOCK 256 /* samples per block of the state-space scan */

template <typename T> struct SjStageTab {
    T window[SJ_STAGE_MAX_WIN]; // first nperseg points of the (nperseg + 1)-point symmetric Tukey(0.25)
    T tw[SJ_STAGE_MAX_WIN][2];  // radix-2 DIT twiddles: stage l, butterfly offset m at index (2^l - 1) + m
    T U;                        // fs * sum(window^2)
    int32_t nperseg, hop, bins, pad;
};

// float: the twiddle sequence of PlainFFT::Compute (bit-exact replay); double: cos/sin evaluated in double
void sj_stage_build_tab_f32(SjStageTab<float> *t, int nperseg, int fs);
void sj_stage_build_tab_f64(SjStageTab<double> *t, int nperseg, int fs);

// 8 x 8 transition matrix of the zero-input recurrence over `steps` samples (state k = w0 at time n-1-k), row-major
void sj_stage_transition(const double a[9], int steps, double M[64]);

struct SjIirArgs {
    const void *in;   // [n_signals] rows of T, `in_stride` elements apart
    void *out;        // same shape, `out_stride`
    long n_signals, n_samples, in_stride, out_stride;
    const void *coef; // device: T a[9], T b[9]
    void *state;      // device [n_signals][8] of T, read at entry and written at exit; NULL = zero state, not returned
    int in_f32;       // exact mode: the input rows are float even when T is double (converted on load)
    bool vec_in, vec_out; // set by the launchers: rows are 16-byte aligned (128-bit accesses)
    // scan mode only
    void *blk_final;  // [n_signals][n_blocks][8] of T
    const void *M;    // device: T M[64]
    long n_blocks;
};

cudaError_t sj_launch_iir_exact(const SjIirArgs &a, bool f64, cudaStream_t stream);
cudaError_t sj_launch_iir_scan(const SjIirArgs &a, bool f64, cudaStream_t stream); // 3 launches

struct SjSpecArgs {
    const void *signals; // [n_signals] rows of T, `stride` elements apart
    long n_signals, n_samples, stride;
    int frames;          // (n_samples - nperseg) / hop + 1
    const void *tab;     // SjStageTab<T> on the device
    void *sxx;           // [n_signals][bins][frames] of T (the reference's Sxx[f][t] per signal)
};
cudaError_t sj_stage_setup();
cudaError_t sj_launch_spectrogram(const SjSpecArgs &a, bool f64, int nperseg, cudaStream_t stream);

// sum_intense (classifier.cpp:370-431) on a dense [bins][frames] float plane, one thread, the reference's order
cudaError_t sj_launch_sum_intense(const float *d_plane, int bins, int frames, int f0, int f1, int t0, int t1,
                                  float *d_out, cudaStream_t stream);

// ---- generic epilogue of the classifier on dense spectrograms (any length, any window): dB, global min/max,
// level band, find_midpoints' clustering, sum_intense boxes, verdict (classifier.c:104-198 / classifier.cpp:34-135 /
// classifier.ino:52-139)
template <typename T> struct SjEpilogueParams {
    T fs;
    int32_t nperseg, hop, bins, max_mid;
    T s_thr;              // smallest Sxx whose dB value exceeds the midpoint threshold (exact, by bisection on the host)
    T lo_norm, hi_norm, gap_tol, min_dur, thr_mid, thr_above, thr_below;
    T half[3];            // sum_intense half ranges: above, middle, below
    int32_t box_f[3][2];  // sum_intense's inclusive frequency-bin ranges (index searches done on the host)
    int32_t pad[2];
};
struct SjEpilogueArgs {
    const void *sxx_band; // [n][bins][frames] of T: spectrogram of the 3000-7500 Hz filtered signal
    const void *sxx_mid;  // [n][bins][frames] of T: spectrogram of the 1000-3000 Hz filtered signal
    long n_signals;
    int frames;
    const void *params;   // SjEpilogueParams<T> on the device
    unsigned char *flags; // scratch [n][frames]
    void *mids_t;         // scratch [n][max_mid] of T: the midpoints in working precision
    float *mids;          // [n][max_mid] midpoint times as float (NaN-padded) or NULL
    int32_t *labels;      // [n]
    float *detail;        // [n][4] or NULL
};
cudaError_t sj_launch_donut_epilogue(const SjEpilogueArgs &a, bool f64, cudaStream_t stream);
// constants of reference variant `variant` (0-3: the 256-sample programs, 4: classifier.ino) for sampling rate fs
bool sj_epilogue_variant_f64(int variant);
int sj_epilogue_variant_nperseg(int variant);
void sj_epilogue_build_params_f32(SjEpilogueParams<float> *p, int variant, int fs, int max_mid);
void sj_epilogue_build_params_f64(SjEpilogueParams<double> *p, int variant, int fs, int max_mid);
