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
/* Device decode of interleaved int16 frames to mono float: policy SJ_MIX_FIRST = channel 0 / 32768
 * (classifier.c:293-297, 56-60), SJ_MIX_MEAN = mean over channels of s / 32768 (aubio_source_do's down-mix,
 * the input of mfcc_stats).  d_out[frames]. */
#define SJ_MIX_FIRST 0
#define SJ_MIX_MEAN 1
SJ_API int scrubjay_decode_pcm16_dev(scrubjay_ctx *ctx, const int16_t *d_interleaved, long frames, int channels,
                              int policy, float *d_out);

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
/* windows per work item of the stream mode (0 = automatic); a tuning / test knob */
SJ_API int scrubjay_set_stream_chunk(scrubjay_ctx *ctx, int windows_per_item);

/* Host-pointer forms.  scrubjay_mfcc_stats keeps the reference's contract: 0 on success, -1
 * (SJ_EINVAL) on an empty clip. */
SJ_API int scrubjay_mfcc_stats(scrubjay_ctx *ctx, const float *samples, long n_samples, float out[SJ_FEAT_DIM]);
/* returns the label (0/1) like ort_predict, or a negative error */
SJ_API int scrubjay_predict(scrubjay_ctx *ctx, const float feat[SJ_FEAT_DIM], float *conf);
/* host clips [n_clips][n_samples] (pageable or pinned); chunks are pipelined H2D -> kernel -> D2H
 * on internal streams.  conf / score / feats may be NULL. */
SJ_API int scrubjay_infer_batch(scrubjay_ctx *ctx, const float *clips, long n_clips, int n_samples,
                         int32_t *labels, float *conf, float *score, float *feats);

/* ---- donut path ---------------------------------------------------------------------------- */

/* variants: which reference binary's constants and precision to follow */
#define SJ_DONUT_DESKTOP_F64 0 /* donut-classifier/classifier.c: double, band 0.70-0.85, 45 dB, mid<75 above>300 below>100 */
#define SJ_DONUT_SYNC_F32 1    /* sync/lib/classifier.cpp: float, band 0.65-0.80, 70 dB, mid<100 above>200 below>80 */
#define SJ_DONUT_REVERT_F64 2  /* donut-classifier/revert.c = microphone/src/classifier.c: double, 0.70-0.85, 45 dB, gap tolerance 0.15 s, mid<50 above>200 below>200 */
#define SJ_DONUT_MIC_F32 3     /* microphone/src/classifier.cpp: float, 0.70-0.85, 45 dB, mid<100 above>200 below>150 */
#define SJ_DONUT_N_VARIANTS 4

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

/* Diagnostic, host only (no GPU needed): the kernel's sparse filterbank tables for `samplerate`,
 * expanded to a dense [40][1025] float array (weights carry a factor 0.5). */
SJ_API int scrubjay_debug_mel_dense(unsigned samplerate, float *out);

/* Pinned (page-locked) host memory: clips handed to the host-pointer entry points from such a
 * buffer move at full PCIe rate; pageable memory works too, slower. */
SJ_API int scrubjay_host_alloc(void **p, size_t bytes);
SJ_API void scrubjay_host_free(void *p);

#ifdef __cplusplus
}
#endif
#endif /* SCRUBJAY_CUDA_H */
