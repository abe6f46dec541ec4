/*
 * scrubjay_cuda.h -- C ABI of libscrubjay_cuda: the batched, B200-native (sm_100a) drop-in for the
 * audio-classification hot path of cornell-c2s2/dsp.
 *
 * Plain C: pointers and sizes only.  No torch types, no C++ in the signatures.  Every entry point
 * names the reference interface it replaces (file:line under the reference tree).
 *
 * Two families of entry points:
 *   *_dev   : pointers are DEVICE pointers on the context's GPU; work is enqueued on the context's
 *             stream (scrubjay_set_stream) and is asynchronous with respect to the host.
 *   others  : pointers are HOST pointers; the call stages through the context's pinned ring,
 *             copies, computes, copies back and returns when the results are in the caller's arrays.
 *
 * Error handling: every call returns SJ_OK (0) or a negative SJ_E* code; scrubjay_last_error()
 * returns a message.  The library never prints and never calls exit() (the reference does both:
 * cepstrum/scrubjay_infer.c:22-26,84-91).  There is NO CPU fallback: without a usable sm_100 GPU
 * scrubjay_create fails with SJ_ENODEVICE.
 *
 * Threading: one context per GPU; calls on one context must not overlap between host threads.
 */
#ifndef SCRUBJAY_CUDA_H
#define SCRUBJAY_CUDA_H

#include <stddef.h>
#include <stdbool.h>
#include <stdint.h>

#if defined(__GNUC__)
#define SJ_API __attribute__((visibility("default")))
#else
#define SJ_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

#define SJ_OK 0
#define SJ_EINVAL (-1)    /* bad argument (NULL, negative size, empty clip: scrubjay_infer.c:55-59) */
#define SJ_ENODEVICE (-2) /* no CUDA device / not sm_100 */
#define SJ_ECUDA (-3)     /* CUDA runtime error, see scrubjay_last_error */
#define SJ_EMODEL (-4)    /* malformed model blob */
#define SJ_ENOMEM (-5)

/* constants of the cepstrum path: cepstrum/scrubjay_infer.c:10-14 */
#define SJ_N_MFCC 20
#define SJ_FEAT_DIM 40
#define SJ_WIN_SIZE 2048
#define SJ_HOP_SIZE 1024
#define SJ_N_FILTERS 40

typedef struct scrubjay_ctx scrubjay_ctx;

/* Replaces ort_init (cepstrum/scrubjay_infer.c:93-102) and the per-file new_aubio_pvoc /
 * new_aubio_mfcc set-up (:29-30).  `model_blob` is an "SJSV" v1 blob (tools/export_svm.py) or NULL
 * for the built-in export of cepstrum/scrubjay_svm.onnx.  `samplerate` is what aubio_source reports
 * for the material (:28); it selects the Slaney filterbank. */
SJ_API int scrubjay_create(scrubjay_ctx **out, int device, const void *model_blob, size_t blob_bytes,
                    unsigned samplerate);

/* The literals of the cepstrum path that live in the two UN-VENDORED libraries behind
 * cepstrum/scrubjay_infer.c (aubio: :21,29-30,44-45; onnxruntime: :119-134) as ONE plain-data parameter
 * block, so that each "aubio-ism" / "ORT-ism" this library restates from the published algorithms can be
 * switched without a code change the day a real aubio / onnxruntime vector disagrees with the restatement
 * (SURVEY 5 "Config / flags", 7 "Hard parts").  The defaults (scrubjay_cepstrum_params_default, also what
 * scrubjay_create uses) are the restated aubio 0.4.x / ORT behaviour.  The same struct drives the CPU oracle
 * (oracle/cepstrum_oracle.c), so every switch has a checker. */
#define SJ_WINDOW_HANN_PERIODIC 0  /* aubio "hanningz": 0.5 (1 - cos(2 pi i / N)) -- new_aubio_pvoc's window */
#define SJ_WINDOW_HANN_SYMMETRIC 1 /* 0.5 (1 - cos(2 pi i / (N - 1))) (numpy.hanning, aubio "hanning") */
#define SJ_SPECTRUM_MAGNITUDE 0    /* filterbank on |X| (aubio_filterbank power = 1, the 0.4.9 default) */
#define SJ_SPECTRUM_POWER 1        /* filterbank on |X|^2 (aubio_filterbank_set_power(fb, 2); librosa's convention) */
#define SJ_DCT_ORTHO 0             /* aubio_dct: row 0 scaled sqrt(1/N), others sqrt(2/N) */
#define SJ_DCT_UNNORMALISED 1      /* 2 sum_m e[m] cos(pi c (m + 1/2) / N) (scipy dct(type=2, norm=None), FFTW REDFT10) */
#define SJ_LABEL_PREDICT 0         /* label 1 iff decision_function > 0  (cepstrum/run.py:46 model.predict; SURVEY B.3 rule ii) */
#define SJ_LABEL_VOTE 1            /* label 1 iff ORT score <= 0, i.e. decision_function >= 0 (libsvm vote; rule i) */
#define SJ_LABEL_PROB_ARGMAX 2     /* argmax of the Platt probability pair, ties -> class 0 (rule iii) */
typedef struct scrubjay_cepstrum_params {
    int32_t window;     /* SJ_WINDOW_* */
    int32_t spectrum;   /* SJ_SPECTRUM_* */
    float log_floor;    /* aubio VERY_SMALL_NUMBER: energies below it are raised to it before log10 (2e-42) */
    int32_t dct_norm;   /* SJ_DCT_* */
    int32_t half_swap;  /* aubio fvec_shift before the FFT (1 = aubio).  Multiplies bin k by (-1)^k: only the phase
                           changes and the path uses |X| alone, so the kernel ignores it; the oracle honours it */
    int32_t label_rule; /* SJ_LABEL_* */
    int32_t reserved[2]; /* must be 0 */
} scrubjay_cepstrum_params;
SJ_API void scrubjay_cepstrum_params_default(scrubjay_cepstrum_params *p);
/* scrubjay_create with an explicit parameter block (NULL = defaults).  Non-default window / spectrum / log_floor
 * select the kernel's table-driven (generic) instantiation instead of the specialised 16 kHz one. */
SJ_API int scrubjay_create_ex(scrubjay_ctx **out, int device, const void *model_blob, size_t blob_bytes,
                       unsigned samplerate, const scrubjay_cepstrum_params *params);
SJ_API void scrubjay_destroy(scrubjay_ctx *ctx);
/* ctx may be NULL: returns the message of the last failed scrubjay_create on this thread */
SJ_API const char *scrubjay_last_error(const scrubjay_ctx *ctx);
/* cudaStream_t as void*; NULL = the legacy default stream.  Device entry points enqueue here. */
SJ_API int scrubjay_set_stream(scrubjay_ctx *ctx, void *cuda_stream);
/* number of kernels this context has launched so far (for bench.py's gpu_launches) */
SJ_API long scrubjay_launch_count(const scrubjay_ctx *ctx);
/* milliseconds between the start and the end of the kernels of the most recent *_dev call,
 * measured with CUDA events on the context's stream; synchronises the stream. */
SJ_API int scrubjay_last_kernel_ms(scrubjay_ctx *ctx, float *ms);

/* ---- cepstrum path ------------------------------------------------------------------------- */

/* Batched mfcc_stats (cepstrum/scrubjay_infer.c:19-77) over in-memory mono clips instead of paths:
 * d_clips is [n_clips][n_samples] float32 in [-1,1); d_feats is [n_clips][40] (mean[20] | std[20]).
 * Frame count per clip = ceil(n_samples/1024) (the do/while at :39-53). */
SJ_API int scrubjay_mfcc_stats_batch_dev(scrubjay_ctx *ctx, const float *d_clips, long n_clips,
                                  int n_samples, float *d_feats);

/* Batched ort_predict (cepstrum/scrubjay_infer.c:105-141): Scaler -> SVMClassifier(RBF) -> label.
 * d_labels[n] int32 0/1; d_conf[n] probability of the predicted class (:134); d_score[n] the
 * decision value d (positive => class 1, = sklearn decision_function = -ORT score).  d_conf and
 * d_score may be NULL. */
SJ_API int scrubjay_predict_batch_dev(scrubjay_ctx *ctx, const float *d_feats, long n, int32_t *d_labels,
                               float *d_conf, float *d_score);

/* Fused mfcc_stats + ort_predict for a batch: the loop body of main (cepstrum/scrubjay_infer.c:
 * 166-176) for n_clips clips in one kernel.  d_conf, d_score, d_feats may be NULL. */
SJ_API int scrubjay_infer_batch_dev(scrubjay_ctx *ctx, const float *d_clips, long n_clips, int n_samples,
                             int32_t *d_labels, float *d_conf, float *d_score, float *d_feats);

/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  scrubjay_infer_batch_dev
 * for SMALL batches that allocates nothing: its frame scratch ([clip][frame][40] floats) is a static device array
 * inside the library -- SJ_NOALLOC_MAX_TOTAL_FRAMES frames = 2.5 MiB that every GPU which loads the library holds
 * until exit.  Limit: n_clips * ceil(n_samples / 1024) <= SJ_NOALLOC_MAX_TOTAL_FRAMES (1024 one-second 16 kHz clips);
 * beyond it the call returns SJ_EINVAL with a message and does nothing.  Inputs and outputs are the caller's device
 * arrays as in scrubjay_infer_batch_dev (d_conf, d_score, d_feats may be NULL); the outputs are bit-identical to that
 * call.  Work is enqueued on the context's stream (two kernels); the call itself makes no cudaMalloc / cudaFree and
 * creates no stream or event.  ALL calls on a GPU share the one array, which keeps the last call's contents (nothing
 * in it needs to start at zero: every entry read is written first): calls that overlap on the GPU -- two contexts or
 * two streams on one device -- race on it; serialise them. */
#define SJ_NOALLOC_MAX_TOTAL_FRAMES 16384
SJ_API int scrubjay_infer_batch_noalloc_dev(scrubjay_ctx *ctx, const float *d_clips, long n_clips, int n_samples,
                                            int32_t *d_labels, float *d_conf, float *d_score, float *d_feats);

/* ---- multi-GPU result gather fused into the kernel (clips are sharded by contiguous ranges, the path's
 * only exchange is the final gather of label int32 + score fp32 = 8 B per clip).  The reference has no counterpart:
 * its main loop prints every file's label as it goes (cepstrum/scrubjay_infer.c:166-176).  Instead of a collective
 * after the kernel, the kernel itself stores every clip's label and score into the result arrays of ALL ranks
 * (its own and, through NVLink/NVSwitch peer mappings, the other GPUs'), so that when the kernels of all ranks have
 * finished every rank holds the complete arrays.  One process per GPU: each rank allocates its result arrays
 * with scrubjay_peer_alloc, hands the 64-byte handles to the other ranks over any transport it likes, and maps theirs
 * with scrubjay_peer_open (CUDA IPC).  In a single process the arrays of other devices can be passed as they are
 * (after cudaDeviceEnablePeerAccess). */
#define SJ_MAX_PEERS 8
#define SJ_PEER_HANDLE_BYTES 64
/* device allocation on the context's GPU that other processes can map; `handle` may be NULL */
SJ_API int scrubjay_peer_alloc(scrubjay_ctx *ctx, size_t bytes, void **d_ptr, unsigned char handle[SJ_PEER_HANDLE_BYTES]);
SJ_API int scrubjay_peer_free(scrubjay_ctx *ctx, void *d_ptr);
/* map / unmap an allocation exported by another process (peer access is enabled as needed) */
SJ_API int scrubjay_peer_open(scrubjay_ctx *ctx, const unsigned char handle[SJ_PEER_HANDLE_BYTES], void **d_ptr);
SJ_API int scrubjay_peer_close(scrubjay_ctx *ctx, void *d_ptr);
/* Bind the gather targets: result i of a later scrubjay_infer_batch_gather_dev goes to
 * peer_labels[p][out_base + i] and peer_score[p][out_base + i] for every p < n_peers (the caller's own arrays are
 * one of the entries; peer_score may be NULL or hold NULL entries).  n_peers = 0 unbinds. */
SJ_API int scrubjay_gather_bind(scrubjay_ctx *ctx, int n_peers, int32_t *const *peer_labels, float *const *peer_score,
                         long out_base);
/* scrubjay_infer_batch_dev with the labels and scores stored to every bound target.  Stream-ordered like every
 * *_dev call; the arrays of a rank are complete once the calls of all ranks have completed (the caller's own
 * barrier).  d_conf, d_feats (local, [n_clips]...) may be NULL. */
SJ_API int scrubjay_infer_batch_gather_dev(scrubjay_ctx *ctx, const float *d_clips, long n_clips, int n_samples,
                                    float *d_conf, float *d_feats);

/* int16 PCM input: the format the reference actually reads (WAV decoded by aubio_source_do,
 * cepstrum/scrubjay_infer.c:21,41; `data[i] / 32768` at donut-classifier/classifier.c:56-60).  Clips are
 * [n_clips][n_samples] int16 mono; the conversion x = s / 32768 is exact and done inside the kernel, so the
 * outputs are bit-identical to scrubjay_infer_batch* on the converted floats, at half the bytes per clip
 * (32 008 B algorithmic).  The host form pipelines H2D -> kernel -> D2H like scrubjay_infer_batch. */
SJ_API int scrubjay_infer_batch_pcm16_dev(scrubjay_ctx *ctx, const int16_t *d_clips, long n_clips, int n_samples,
                                   int32_t *d_labels, float *d_conf, float *d_score, float *d_feats);
SJ_API int scrubjay_infer_batch_pcm16(scrubjay_ctx *ctx, const int16_t *clips, long n_clips, int n_samples,
                               int32_t *labels, float *conf, float *score, float *feats);

/* WAV ingest.  scrubjay_wav_parse is host-only (no GPU): it walks a RIFF/WAVE image in memory with the
 * acceptance rules of read_wav_file (donut-classifier/classifier.c:214-306): "RIFF"/"WAVE", the "fmt " chunk
 * first, AudioFormat 1 (PCM), then any chunks are skipped up to "data".  frames = data_bytes /
 * (bits_per_sample / 8) / channels.  Returns SJ_OK or SJ_EINVAL (the reference returns 0 and prints). */
typedef struct scrubjay_wav_info {
    int32_t audio_format, channels, sample_rate, bits_per_sample;
    int64_t data_offset, data_bytes, frames;
} scrubjay_wav_info;
SJ_API int scrubjay_wav_parse(const void *bytes, size_t n_bytes, scrubjay_wav_info *info);
/* strict = 1: the rules above (read_wav_file, the donut path's reader).  strict = 0: chunks in front of "fmt " (JUNK,
 * bext, LIST ...) are skipped, the way libsndfile -- the reader behind aubio_source (cepstrum/scrubjay_infer.c:21) --
 * walks a RIFF file; cepstrum/testing/P_1363v2-sj-short.WAV starts with a JUNK chunk. */
SJ_API int scrubjay_wav_parse_ex(const void *bytes, size_t n_bytes, scrubjay_wav_info *info, int strict);
/* Device decode of interleaved int16 frames to mono float: policy SJ_MIX_FIRST = channel 0 / 32768
 * (classifier.c:293-297, 56-60), SJ_MIX_MEAN = mean over channels of s / 32768 (aubio_source_do's down-mix,
 * the input of mfcc_stats).  d_out[frames]. */
#define SJ_MIX_FIRST 0
#define SJ_MIX_MEAN 1
SJ_API int scrubjay_decode_pcm16_dev(scrubjay_ctx *ctx, const int16_t *d_interleaved, long frames, int channels,
                              int policy, float *d_out);
/* The firmware's ADC path in front of classify() (sync/sync.cpp:236-241): 12-bit counts -> int16 pcm = (adc - 2048) * 16
 * -> pcm / 32768.0 as float.  d_adc[n] uint16, d_out[n] float. */
SJ_API int scrubjay_decode_adc12_dev(scrubjay_ctx *ctx, const uint16_t *d_adc, long n, float *d_out);
/* upsampleLinear (microphone/src/microphone.cpp:55-70) for n_signals buffers [n_signals][old_size] ->
 * [n_signals][new_size], float arithmetic in the reference's order (no fused multiply-add). */
SJ_API int scrubjay_upsample_linear_dev(scrubjay_ctx *ctx, const float *d_in, long n_signals, int old_size, int new_size,
                                        float *d_out);

/* Stream mode (BASELINE config 5: long continuous channels, sliding-frame cepstrum + SVM).  The reference
 * has no streaming entry point: main (cepstrum/scrubjay_infer.c:158-177) reduces a whole file to one
 * feature vector.  Here the loop body of mfcc_stats (:39-53) runs on without restarting aubio_pvoc, and the
 * statistics of :61-67 followed by ort_predict are taken over every window of `win_frames` consecutive
 * frames, advancing by `step_frames` (1 <= step <= win <= 16): window j of a channel covers frames
 * [j*step, j*step + win), frame f = [hop f-1 | hop f].  d_streams holds n_channels channels of n_samples
 * float32 samples, chan_stride samples apart; outputs are [channel][window] with
 * scrubjay_stream_windows(n_samples, win, step) windows per channel.  d_prev_hop ([n_channels][1024] or
 * NULL) is the hop that precedes sample 0 of every channel: the state a caller carries between the time
 * chunks of an endless stream (overlap consecutive chunks by win - step hops to keep every window);
 * NULL = stream start (zeros, like a fresh aubio_pvoc).  d_conf / d_score / d_feats may be NULL. */
SJ_API long scrubjay_stream_windows(long n_samples, int win_frames, int step_frames);
SJ_API int scrubjay_stream_infer_dev(scrubjay_ctx *ctx, const float *d_streams, long n_channels, long n_samples,
                              long chan_stride, const float *d_prev_hop, int win_frames, int step_frames,
                              int32_t *d_labels, float *d_conf, float *d_score, float *d_feats);
/* The same on int16 PCM streams (d_streams / d_prev_hop hold samples s, the kernel works on s / 32768): outputs
 * bit-identical to the float entry point on the converted samples, at half the resident bytes per channel. */
SJ_API int scrubjay_stream_infer_pcm16_dev(scrubjay_ctx *ctx, const int16_t *d_streams, long n_channels, long n_samples,
                                           long chan_stride, const int16_t *d_prev_hop, int win_frames, int step_frames,
                                           int32_t *d_labels, float *d_conf, float *d_score, float *d_feats);
/* Stream mode with the fused multi-GPU gather (scrubjay_gather_bind): window j of channel c of this call is result
 * c * n_windows + j, stored at out_base + that index into every bound target (channels are sharded over ranks, so
 * out_base = first channel of the rank * n_windows). */
SJ_API int scrubjay_stream_infer_gather_dev(scrubjay_ctx *ctx, const float *d_streams, long n_channels, long n_samples,
                                     long chan_stride, const float *d_prev_hop, int win_frames, int step_frames,
                                     float *d_conf, float *d_feats);
SJ_API int scrubjay_stream_infer_pcm16_gather_dev(scrubjay_ctx *ctx, const int16_t *d_streams, long n_channels,
                                           long n_samples, long chan_stride, const int16_t *d_prev_hop, int win_frames,
                                           int step_frames, float *d_conf, float *d_feats);
/* Small batches (fewer clips than half the warps the GPU keeps resident, i.e. < 888 on a B200) run frame-parallel by
 * default: the frames of a clip are spread over warps and SMs and a second small kernel forms the per-clip statistics in
 * frame order, so a single clip costs about two frames of latency instead of ceil(n/1024) (the reference's use: one
 * file / one second at a time, cepstrum/scrubjay_infer.c:156-177).  Outputs are bit-identical to the one-kernel
 * path.  on = 0 forces the one-kernel path. */
SJ_API int scrubjay_set_latency_mode(scrubjay_ctx *ctx, int on);
/* windows per work item of the stream mode (0 = automatic); a tuning / test knob */
SJ_API int scrubjay_set_stream_chunk(scrubjay_ctx *ctx, int windows_per_item);

/* Host-pointer forms.  scrubjay_mfcc_stats keeps the reference's contract: 0 on success, -1
 * (SJ_EINVAL) on an empty clip. */
SJ_API int scrubjay_mfcc_stats(scrubjay_ctx *ctx, const float *samples, long n_samples, float out[SJ_FEAT_DIM]);
/* returns the label (0/1) like ort_predict, or a negative error */
SJ_API int scrubjay_predict(scrubjay_ctx *ctx, const float feat[SJ_FEAT_DIM], float *conf);
/* host clips [n_clips][n_samples] (pageable or pinned); chunks are pipelined H2D -> kernel -> D2H
 * on internal streams.  conf / score / feats may be NULL.  One to three clips in PINNED memory (scrubjay_host_alloc,
 * cudaHostAlloc, cudaHostRegister) are read by the kernels in place and answered through mapped memory: no copy
 * operation, 33 us per single-clip call instead of 46 us; the call then runs on the context's stream. */
SJ_API int scrubjay_infer_batch(scrubjay_ctx *ctx, const float *clips, long n_clips, int n_samples,
                         int32_t *labels, float *conf, float *score, float *feats);

/* ---- donut path ---------------------------------------------------------------------------- */

/* variants: which reference binary's constants and precision to follow */
#define SJ_DONUT_DESKTOP_F64 0 /* donut-classifier/classifier.c: double, band 0.70-0.85, 45 dB, mid<75 above>300 below>100 */
#define SJ_DONUT_SYNC_F32 1    /* sync/lib/classifier.cpp: float, band 0.65-0.80, 70 dB, mid<100 above>200 below>80 */
#define SJ_DONUT_REVERT_F64 2  /* donut-classifier/revert.c = microphone/src/classifier.c: double, 0.70-0.85, 45 dB, gap tolerance 0.15 s, mid<50 above>200 below>200 */
#define SJ_DONUT_MIC_F32 3     /* microphone/src/classifier.cpp: float, 0.70-0.85, 45 dB, mid<100 above>200 below>150 */
#define SJ_DONUT_N_VARIANTS 4  /* variants of the fused 256-sample-window kernel */
#define SJ_DONUT_INO_F32 4     /* donut-classifier/classifier.ino: float, 64-sample windows, 0.70-0.85, 45 dB, boxes 4500-7500 /
                                  3500-4000 / 500-3000 Hz, mid<75 above>300 below>100 (staged path only) */

/* The reference's own entry point, same name and signature: sync/lib/classifier.h:19, called as
 * classify(buffer, 16000) at sync/sync.cpp:202.  Host pointer in, 1/0 out (0 also on failure, as the
 * reference does at classifier.cpp:87-91).  Runs on a process-wide context (device $SCRUBJAY_DEVICE). */
SJ_API int classify(float *data, int data_size);

/* Batched classify: the per-file body of main (classifier.c:84-198) / classify (classifier.cpp:9-136)
 * for n_clips clips [n_clips][n_samples] (256 <= n_samples <= 21536).  d_labels[n] int32 0/1;
 * d_detail (may be NULL) [n][4] = number of midpoints (find_midpoints), then the above / middle / below
 * band sums (sum_intense) of the last midpoint the reference would have evaluated. */
SJ_API int scrubjay_donut_classify_batch_dev(scrubjay_ctx *ctx, const float *d_clips, long n_clips,
                                             int n_samples, int variant, int32_t *d_labels, float *d_detail);
SJ_API int scrubjay_donut_classify_batch(scrubjay_ctx *ctx, const float *clips, long n_clips, int n_samples,
                                         int variant, int32_t *labels, float *detail);

/* As above, also returning find_midpoints' array (classifier.cpp:433-598 / classifier.c:657-825): d_mids
 * [n][SJ_DONUT_MAX_MID] midpoint times in seconds, NaN-padded (d_detail[n][0] holds the count). */
#define SJ_DONUT_MAX_MID 8
SJ_API int scrubjay_donut_midpoints_batch_dev(scrubjay_ctx *ctx, const float *d_clips, long n_clips, int n_samples,
                                              int variant, int32_t *d_labels, float *d_detail, float *d_mids);

/* Signals of ANY length (the desktop program classifies whole recordings, classifier.c:37-210) and the
 * 64-sample-window variant: the same classifier as a staged pipeline (filter -> spectrogram, twice, -> epilogue)
 * instead of the fused kernel; scrubjay_donut_classify_batch(_dev) switches to it by itself for n_samples > 21536
 * or variant SJ_DONUT_INO_F32.  d_mids (may be NULL) [n][max_mid] float midpoints, NaN-padded; at most max_mid
 * (1..4096) midpoints are evaluated per signal.  d_detail as above. */
SJ_API int scrubjay_donut_classify_long_dev(scrubjay_ctx *ctx, const float *d_signals, long n_signals, long n_samples,
                                            int variant, int32_t *d_labels, float *d_detail, float *d_mids, int max_mid);

/* Decision tracks of continuous streams, the firmware's loop (sync/sync.cpp:190-212): every full block of
 * clip_samples samples (16000 there) of every channel is classified from zero state and the result stored as the
 * classification byte of IMUClassificationFrame (sync.cpp:49-63, :202-203): 1 = no bird, 2 = scrub jay.
 * d_streams [n_channels] rows chan_stride samples apart; d_track [n_channels][seconds] int8, seconds =
 * scrubjay_donut_stream_seconds(n_samples, clip_samples) (a trailing partial block is not classified, as in the
 * firmware, which waits for a full buffer); d_detail (may be NULL) [n_channels][seconds][4] as in classify_batch. */
SJ_API long scrubjay_donut_stream_seconds(long n_samples, int clip_samples);
SJ_API int scrubjay_donut_classify_streams_dev(scrubjay_ctx *ctx, const float *d_streams, long n_channels, long n_samples,
                                               long chan_stride, int clip_samples, int variant, int8_t *d_track,
                                               float *d_detail);

/* ---- donut stage functions: the reference's public helpers (sync/lib/classifier.h:14-18; the double forms
 * donut-classifier/classifier.c:22-28) as batched calls, plus the helpers themselves under their own names.
 * The reference's C++ symbols are exported here with C linkage, like classify. ---------------------------- */
#define SJ_F32 0
#define SJ_F64 1
#define SJ_IIR_EXACT 0 /* one recurrence per lane in the reference's operation order: bit-identical outputs */
#define SJ_IIR_SCAN 1  /* block-wise state-space scan over time, for few long signals: rounding differs (~1e-15 rel, f64) */
#define SJ_BUTTER_16K 0        /* the active coefficient sets, classifier.c:342-360 and :383-401 (= classifier.cpp:142-182) */
#define SJ_BUTTER_48K_DESIGN 1 /* the commented-out sets, classifier.c:323-341 and :364-382 (they produced _postbutter.txt) */

/* butter_bandpass (classifier.h:14 / classifier.cpp:138-191): table lookup for (1000,3000) and (3000,7500),
 * false otherwise (the reference also prints "invalid bandpass range"). */
SJ_API bool butter_bandpass(float lowcut, float highcut, float *b, float *a);
/* double form (classifier.c:319-409) with the choice of coefficient set; SJ_OK or SJ_EINVAL */
SJ_API int scrubjay_butter_bandpass_f64(double lowcut, double highcut, int design, double *b, double *a);

/* butter_bandpass_filter (classifier.h:15 / classifier.cpp:193-219 / classifier.c:420-446) for n_signals rows of
 * n_samples elements (rows `stride` elements apart in both arrays; element type float or double by `precision`),
 * coefficients b[9], a[9] of the same type (host pointers).  d_state (may be NULL = zero initial state as in the
 * reference) is [n_signals][8] elements: the delay line w[0..7] (classifier.c:422) read at entry and written back
 * at exit, so a stream can be filtered in pieces with results identical to one call.  In place (d_out == d_in)
 * is allowed in SJ_IIR_EXACT mode. */
SJ_API int scrubjay_butter_filter_batch_dev(scrubjay_ctx *ctx, const void *d_in, void *d_out, long n_signals,
                                            long n_samples, long stride, int precision, const void *b, const void *a,
                                            void *d_state, int mode);
SJ_API int scrubjay_butter_filter_batch(scrubjay_ctx *ctx, const void *in, void *out, long n_signals, long n_samples,
                                        int precision, const void *b, const void *a, int mode);
/* the reference's symbol (host pointers, one signal, zero state), run on the process-wide context */
SJ_API void butter_bandpass_filter(float *data, int n, float *b, float *a, float *output);

/* compute_spectrogram (classifier.h:16 / classifier.cpp:221-368 / classifier.c:448-592; classifier.ino:245-395 with
 * nperseg 64): frames of nperseg samples every nperseg - nperseg/8, mean removed, periodic Tukey(0.25), PSD.
 * d_sxx is dense [n_signals][nperseg/2 + 1][frames] (the reference's Sxx[f][t] per signal), frames =
 * scrubjay_spectrogram_frames(n_samples, nperseg).  nperseg 256 or 64.  float: bit-identical to the reference
 * (PlainFFT replayed operation for operation); double: ~1e-15 relative (FFTW's order is not reproducible). */
SJ_API long scrubjay_spectrogram_frames(long n_samples, int nperseg);
SJ_API int scrubjay_spectrogram_batch_dev(scrubjay_ctx *ctx, const void *d_signals, long n_signals, long n_samples,
                                          long stride, int precision, int nperseg, int fs, void *d_sxx);
SJ_API int scrubjay_spectrogram_batch(scrubjay_ctx *ctx, const void *signals, long n_signals, long n_samples,
                                      int precision, int nperseg, int fs, void *sxx);
/* the reference's symbol: allocates *frequencies, *times, *Sxx and each (*Sxx)[f] with malloc; the caller frees
 * them (classifier.cpp:239-245 vs :127-133).  On failure *freq_bins = *time_bins = 0 and nothing is allocated. */
SJ_API void compute_spectrogram(float *signal, int signal_length, int fs, float **frequencies, float **times,
                                float ***Sxx, int *freq_bins, int *time_bins);
/* find_midpoints (classifier.h:18 / classifier.cpp:433-598): malloc'd array of *num_midpoints times, NULL on failure */
SJ_API float *find_midpoints(float *data, int num_frames, int samplingFreq, int *num_midpoints);
/* sum_intense (classifier.h:17 / classifier.cpp:370-431): the index searches run on the host, the NaN-skipping
 * sum over the box on the device in the reference's order */
SJ_API float sum_intense(float lower, float upper, float half_range, float *frequencies, int freq_bins, float *times,
                         int time_bins, float **intensity_dB_filtered, float midpoint);

/* ---- keyword ("stop") detector: 2fa/audio/word/c ------------------------------------------------ */

/* Loads the detector's constants (mfcc_params.h + model_params.h as an "SJKW" blob, tools/export_keyword.py;
 * shipped as dsp_b200/data/keyword_model.bin).  Replaces the compiled-in tables of the reference. */
SJ_API int scrubjay_keyword_load(scrubjay_ctx *ctx, const void *blob, size_t bytes);
/* Batched classify_signal (stop_detector.c:12-55 = compute_mfcc, mfcc.c:108-232, + audio_classifier_predict,
 * audio_classifier_inference.c:38-90) for n_clips signals [n_clips][n_samples] of 16 kHz float audio:
 * d_prob[n] = probability of "stop".  d_mfcc (may be NULL) [n][500][13] receives compute_mfcc's frame-major
 * output, zero beyond the frames computed; d_n_frames (may be NULL) [n] its return value. */
SJ_API int scrubjay_keyword_classify_batch_dev(scrubjay_ctx *ctx, const float *d_signals, long n_clips, int n_samples,
                                        float *d_prob, float *d_mfcc, int32_t *d_n_frames);
SJ_API int scrubjay_keyword_classify_batch(scrubjay_ctx *ctx, const float *signals, long n_clips, int n_samples,
                                    float *prob);
/* The reference's own entry points, same names and signatures (stop_detector.h:9-15); process-wide context,
 * model blob from $SCRUBJAY_KEYWORD_MODEL. */
SJ_API float classify_signal(const float *signal, int num_samples);
SJ_API int classify_signal_binary(const float *signal, int num_samples, float threshold);

/* Synthetic test/bench audio generated on the device: d_out[n_clips][n_samples], a pure function of
 * (seed, first_clip + clip, sample) -- white noise floor + up to three gated chirps, ~5 % of clips
 * silent or with a silent tail.  Not a reference function; the bench's "data": "synthetic". */
SJ_API int scrubjay_synth_clips_dev(scrubjay_ctx *ctx, float *d_out, long first_clip, long n_clips,
                                    int n_samples, unsigned seed);
/* The same with a choice of flavour: SJ_SYNTH_MIXED = the generator above; SJ_SYNTH_CALLS = "donut"-shaped calls (a
 * faint noise floor and three frequency-modulated bursts at 600-2400 / 2600-4900 / 5100-6900 Hz around a common
 * centre) on which every donut variant's verdict -- the double-precision desktop one included -- says 1 on a
 * sizeable fraction of the clips. */
#define SJ_SYNTH_MIXED 0
#define SJ_SYNTH_CALLS 1
SJ_API int scrubjay_synth_clips_ex_dev(scrubjay_ctx *ctx, float *d_out, long first_clip, long n_clips,
                                       int n_samples, unsigned seed, int flavour);

/* Diagnostic, host only (no GPU needed): the kernel's sparse filterbank tables for `samplerate`,
 * expanded to a dense [40][1025] float array (weights carry a factor 0.5). */
SJ_API int scrubjay_debug_mel_dense(unsigned samplerate, float *out);

/* Test hook (fault injection): the next host-pointer batch call (scrubjay_infer_batch / _pcm16) returns SJ_ECUDA right
 * before it would enqueue its chunk number `chunk` (0-based; chunks are 256 MiB of clips), with the earlier chunks in
 * flight.  The call after a failed one must not see anything of it (tests/test_cepstrum_gpu.py).  -1 disarms. */
SJ_API int scrubjay_debug_inject_failure(scrubjay_ctx *ctx, long chunk);

/* Pinned (page-locked) host memory: clips handed to the host-pointer entry points from such a
 * buffer move at full PCIe rate; pageable memory works too, slower. */
SJ_API int scrubjay_host_alloc(void **p, size_t bytes);
SJ_API void scrubjay_host_free(void *p);

#ifdef __cplusplus
}
#endif
#endif /* SCRUBJAY_CUDA_H */
