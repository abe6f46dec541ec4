// cepstrum_kernel.h -- launch interface of the fused cepstrum (MFCC stats + SVM) kernel.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "sj_tables.h"

// gather targets of the fused multi-GPU result gather (device memory, written by scrubjay_gather_bind): label /
// score of output index i go to labels[p][out_base + i], score[p][out_base + i], p < n_peers
#define SJ_KMAX_PEERS 8
struct SjPeerTable {
    int32_t *labels[SJ_KMAX_PEERS];
    float *score[SJ_KMAX_PEERS];
    long out_base;
    int n_peers, pad;
};

struct SjCepstrumArgs {
    const float *clips;   // [n_clips][n_samples], device (int16_t samples when pcm16 != 0)
    int pcm16;            // clips are int16 PCM, converted as s / 32768 in the kernel (clip mode only)
    long n_clips;
    int n_samples;
    const SjTables *tables; // device copy
    int32_t *labels;      // [n_clips] or NULL
    float *conf;          // [n_clips] or NULL
    float *score;         // [n_clips] or NULL
    float *feats;         // [n_clips][40] or NULL
    int do_svm;           // 0: features only
    const SjPeerTable *peers; // non-NULL: labels / score go to every target of the table instead of labels / score
    // stream mode (n_windows > 0): `clips` holds n_clips channels of n_samples samples, chan_stride apart; the
    // pvoc runs on continuously; window j covers frames [j*step_frames, j*step_frames + win_frames) and the
    // outputs are [channel][window].  prev_hop ([n_clips][1024] or NULL) is the hop before sample 0 (carried
    // state of a chunked stream; NULL = stream start, zeros like a fresh aubio_pvoc).
    long chan_stride;
    const float *prev_hop;
    long n_windows;       // windows per channel; 0 = clip mode
    int win_frames, step_frames;
    int chunk_windows;    // windows per work item
    long chunks_per_chan;
};
#define SJ_STREAM_MAX_WIN 16 /* frames per stream-mode window (ring of MFCC vectors in shared memory) */

// Enqueues the kernel on `stream`.  `host_tables` is the host copy of a.tables (selects the mel layout),
// `num_sms` the SM count of the device.  Returns the cudaError_t of the launch.
cudaError_t sj_launch_cepstrum(const SjCepstrumArgs &a, const SjTables *host_tables, int num_sms, cudaStream_t stream);
// frame-parallel clip mode, second kernel: per-clip statistics over frames[clip][n_frames][40] (+ Scaler + SVM, outputs
// as named by `a`; a.clips / stream fields are not used)
cudaError_t sj_launch_clip_stats(const SjCepstrumArgs &a, const float *frames, int n_frames, cudaStream_t stream);
// allocation-free variant: the frame scratch is a static device array of SJ_NOALLOC_MAX_FRAMES x 40 floats (2.5 MiB)
#define SJ_NOALLOC_MAX_FRAMES 16384
cudaError_t sj_noalloc_frame_scratch(int device, float **ptr); // device must be current; looked up once per device
cudaError_t sj_launch_clip_stats_static(const SjCepstrumArgs &a, int n_frames, cudaStream_t stream);
// warps the persistent grid keeps resident (work-item sizing of the stream mode)
int sj_cepstrum_resident_warps(int num_sms);
// one-time attribute set-up (max dynamic shared memory); returns cudaError_t
cudaError_t sj_cepstrum_setup(int mel_maxseg);
// batched standalone SVM: feats [n][40] -> labels/conf/score
cudaError_t sj_launch_svm(const float *feats, long n, const SjTables *tables, int32_t *labels, float *conf,
                          float *score, cudaStream_t stream);
