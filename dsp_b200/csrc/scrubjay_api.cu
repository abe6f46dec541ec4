// scrubjay_api.cu -- the C ABI of libscrubjay_cuda (include/scrubjay_cuda.h): context management,
// device-pointer batch entry points and the host-pointer pipeline (H2D -> kernel -> D2H).
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>

#include <new>

#include "../../include/scrubjay_cuda.h"
#include "cepstrum_kernel.h"
#include "donut_kernel.h"
#include "donut_stages.h"
#include "ingest_kernel.h"
#include "keyword_kernel.h"
#include "sj_tables.h"

cudaError_t sj_launch_synth(float *out, long first_clip, long n_clips, int n_samples, uint32_t seed, float fs,
                            int flavour, cudaStream_t stream);

namespace {

constexpr int kStages = 3;          // host pipeline depth
constexpr long kChunkBytes = 256l << 20; // clips per chunk = kChunkBytes / clip bytes

struct Stage {
    cudaStream_t stream = nullptr;
    cudaEvent_t done = nullptr;
    float *d_clips = nullptr;
    int32_t *d_labels = nullptr, *h_labels = nullptr;
    float *d_conf = nullptr, *d_score = nullptr, *d_feats = nullptr;
    float *h_conf = nullptr, *h_score = nullptr, *h_feats = nullptr;
    long pending_first = -1, pending_count = 0; // chunk whose outputs still sit in the h_* buffers
};

thread_local char g_create_err[512];

} // namespace

struct scrubjay_ctx {
    int device = 0, num_sms = 0;
    unsigned samplerate = 0;
    cudaStream_t stream = nullptr;
    SjTables *h_tables = nullptr, *d_tables = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    bool have_timing = false;
    long launches = 0;
    char err[512] = {0};
    // host pipeline (allocated on first use)
    Stage st[kStages];
    long chunk_clips = 0;
    int chunk_n_samples = 0;
    // single-item scratch for scrubjay_predict / scrubjay_mfcc_stats
    float *d_small = nullptr;
    // mapped pinned results of the single-clip zero-copy path (run_host_batch_impl): labels, conf, score, features
    float *h_zc = nullptr;
    // grow-only device buffer of scrubjay_mfcc_stats (one clip): allocated on first use, never on the steady-state path
    float *d_one = nullptr;
    size_t one_bytes = 0;
    // fault injection for tests (scrubjay_debug_inject_failure): the host pipeline fails before chunk number `fail_at_chunk`
    long fail_at_chunk = -1;
    // keyword path (scrubjay_keyword_load): tables on the device, staging for the host form
    KwTables *d_kw_tables = nullptr;
    int kw_maxseg = 0;
    float *d_kw_signals = nullptr, *d_kw_prob = nullptr;
    size_t kw_stage_bytes = 0;
    long kw_stage_clips = 0;
    int stream_chunk_windows = 0; // 0 = automatic work-item size of the stream mode
    // frame-parallel clip mode for small batches (scrubjay_set_latency_mode; on by default)
    int latency_mode = 1;
    float *d_frame_scratch = nullptr;
    size_t frame_scratch_bytes = 0;
    cudaEvent_t frame_scratch_free = nullptr;
    // fused multi-GPU gather (scrubjay_gather_bind): the target table on the device
    SjPeerTable *d_peers = nullptr;
    int n_peers = 0;
    int peer_access_on[SJ_MAX_PEERS * 2] = {0}; // devices this context has enabled peer access to (index = device id)
    // donut path (allocated on first use): tables per variant, PSD scratch for the clips in flight
    void *d_donut_tables[SJ_DONUT_N_VARIANTS] = {nullptr, nullptr, nullptr, nullptr};
    void *d_donut_scratch = nullptr;
    size_t donut_scratch_bytes = 0;
    float *d_donut_clips = nullptr; // host-path staging
    int32_t *d_donut_labels = nullptr;
    float *d_donut_detail = nullptr;
    long donut_stage_clips = 0;
    int donut_stage_samples = 0;
    // donut stage functions (donut_stage_api.inc)
    bool stage_setup_done = false;
    void *d_stage_coef = nullptr, *d_stage_M = nullptr;
    void *d_stage_tab[2][2] = {{nullptr, nullptr}, {nullptr, nullptr}}; // [precision][nperseg == 64]
    int stage_tab_fs[2][2] = {{0, 0}, {0, 0}};
    void *d_stage_blk = nullptr, *d_stage_in = nullptr, *d_stage_out = nullptr;
    size_t stage_blk_bytes = 0, stage_in_bytes = 0, stage_out_bytes = 0;
    void *d_ep_params[SJ_DONUT_INO_F32 + 1] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    int ep_max_mid[SJ_DONUT_INO_F32 + 1] = {0, 0, 0, 0, 0};
    void *d_long_ws = nullptr;
    size_t long_ws_bytes = 0;
};

namespace {

int set_err(scrubjay_ctx *c, int code, const char *what, cudaError_t e = cudaSuccess)
{
    if (c) {
        if (e != cudaSuccess) snprintf(c->err, sizeof c->err, "%s: %s", what, cudaGetErrorString(e));
        else snprintf(c->err, sizeof c->err, "%s", what);
    }
    return code;
}

#define SJ_CUDA(c, call)                                                \
    do {                                                                \
        cudaError_t e_ = (call);                                        \
        if (e_ != cudaSuccess) return set_err((c), SJ_ECUDA, #call, e_); \
    } while (0)

void free_stages(scrubjay_ctx *c)
{
    for (Stage &s : c->st) {
        if (s.stream) cudaStreamDestroy(s.stream);
        if (s.done) cudaEventDestroy(s.done);
        cudaFree(s.d_clips); cudaFree(s.d_labels); cudaFree(s.d_conf); cudaFree(s.d_score); cudaFree(s.d_feats);
        cudaFreeHost(s.h_labels); cudaFreeHost(s.h_conf); cudaFreeHost(s.h_score); cudaFreeHost(s.h_feats);
        s = Stage();
    }
    c->chunk_clips = 0;
    c->chunk_n_samples = 0;
}

int ensure_stages(scrubjay_ctx *c, int n_samples)
{
    if (c->chunk_clips && c->chunk_n_samples == n_samples) return SJ_OK;
    free_stages(c);
    long clips = kChunkBytes / ((long)n_samples * 4);
    if (clips < 1) clips = 1;
    for (Stage &s : c->st) {
        SJ_CUDA(c, cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
        SJ_CUDA(c, cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming));
        SJ_CUDA(c, cudaMalloc(&s.d_clips, (size_t)clips * n_samples * 4));
        SJ_CUDA(c, cudaMalloc(&s.d_labels, (size_t)clips * 4));
        SJ_CUDA(c, cudaMalloc(&s.d_conf, (size_t)clips * 4));
        SJ_CUDA(c, cudaMalloc(&s.d_score, (size_t)clips * 4));
        SJ_CUDA(c, cudaMalloc(&s.d_feats, (size_t)clips * SJ_FEAT * 4));
        SJ_CUDA(c, cudaMallocHost(&s.h_labels, (size_t)clips * 4));
        SJ_CUDA(c, cudaMallocHost(&s.h_conf, (size_t)clips * 4));
        SJ_CUDA(c, cudaMallocHost(&s.h_score, (size_t)clips * 4));
        SJ_CUDA(c, cudaMallocHost(&s.h_feats, (size_t)clips * SJ_FEAT * 4));
    }
    c->chunk_clips = clips;
    c->chunk_n_samples = n_samples;
    return SJ_OK;
}

int launch_cepstrum(scrubjay_ctx *c, cudaStream_t stream, const float *d_clips, long n_clips, int n_samples,
                    int32_t *labels, float *conf, float *score, float *feats, int do_svm, bool timed, int pcm16 = 0,
                    const SjPeerTable *peers = nullptr)
{
    SjCepstrumArgs a;
    a.pcm16 = pcm16;
    a.peers = peers;
    a.clips = d_clips; a.n_clips = n_clips; a.n_samples = n_samples; a.tables = c->d_tables;
    a.labels = labels; a.conf = conf; a.score = score; a.feats = feats; a.do_svm = do_svm;
    a.chan_stride = 0; a.prev_hop = nullptr; a.n_windows = 0; a.win_frames = a.step_frames = 0;
    a.chunk_windows = 0; a.chunks_per_chan = 0;
    // Small batches (fewer clips than half the resident warps) run frame-parallel: the stream instantiation with
    // one-frame windows writes every frame's MFCC vector to a scratch ([clip][frame][40]), work items of k frames, and
    // a second small kernel adds them up per clip in frame order -- the same per-frame arithmetic and the same order
    // of the double sums as the per-clip loop, so the outputs are bit-identical; one clip takes 2 frames of latency
    // instead of 16.
    const long F = ((long)n_samples + SJ_HOP - 1) / SJ_HOP;
    const long resident = sj_cepstrum_resident_warps(c->num_sms);
    const bool frame_parallel = c->latency_mode && n_clips > 0 && F >= 2 && n_clips * 2 <= resident &&
                                (size_t)n_clips * F * SJ_FEAT * sizeof(float) <= (64u << 20);
    if (frame_parallel) {
        const size_t need = (size_t)n_clips * F * SJ_FEAT * sizeof(float);
        // one scratch per context: its users on different streams (the host pipeline's stages) are chained by an event
        if (!c->frame_scratch_free) SJ_CUDA(c, cudaEventCreateWithFlags(&c->frame_scratch_free, cudaEventDisableTiming));
        else SJ_CUDA(c, cudaStreamWaitEvent(stream, c->frame_scratch_free, 0));
        if (need > c->frame_scratch_bytes) {
            SJ_CUDA(c, cudaEventSynchronize(c->frame_scratch_free)); // unrecorded event: returns at once
            cudaFree(c->d_frame_scratch);
            c->d_frame_scratch = nullptr;
            c->frame_scratch_bytes = 0;
            SJ_CUDA(c, cudaMalloc(&c->d_frame_scratch, need));
            c->frame_scratch_bytes = need;
        }
        SjCepstrumArgs f = a;
        f.labels = nullptr; f.conf = nullptr; f.score = nullptr; f.peers = nullptr; f.do_svm = 0;
        f.feats = c->d_frame_scratch;
        f.chan_stride = n_samples; f.prev_hop = nullptr; f.n_windows = F; f.win_frames = f.step_frames = 1;
        long k = (n_clips * F + resident - 1) / resident; // frames per work item: about one item per resident warp
        if (k < 1) k = 1;
        if (k > F) k = F;
        f.chunk_windows = (int)k;
        f.chunks_per_chan = (F + k - 1) / k;
        if (timed) SJ_CUDA(c, cudaEventRecord(c->ev0, stream));
        SJ_CUDA(c, sj_launch_cepstrum(f, c->h_tables, c->num_sms, stream));
        SJ_CUDA(c, sj_launch_clip_stats(a, c->d_frame_scratch, (int)F, stream));
        SJ_CUDA(c, cudaEventRecord(c->frame_scratch_free, stream));
        if (timed) { SJ_CUDA(c, cudaEventRecord(c->ev1, stream)); c->have_timing = true; }
        c->launches += 2;
        return SJ_OK;
    }
    if (timed) SJ_CUDA(c, cudaEventRecord(c->ev0, stream));
    SJ_CUDA(c, sj_launch_cepstrum(a, c->h_tables, c->num_sms, stream));
    if (timed) { SJ_CUDA(c, cudaEventRecord(c->ev1, stream)); c->have_timing = true; }
    c->launches += 1;
    return SJ_OK;
}

} // namespace

extern "C" {

void scrubjay_cepstrum_params_default(scrubjay_cepstrum_params *p)
{
    if (p) sj_default_cepstrum_params(p);
}

int scrubjay_create(scrubjay_ctx **out, int device, const void *model_blob, size_t blob_bytes, unsigned samplerate)
{
    return scrubjay_create_ex(out, device, model_blob, blob_bytes, samplerate, nullptr);
}

int scrubjay_create_ex(scrubjay_ctx **out, int device, const void *model_blob, size_t blob_bytes, unsigned samplerate,
                       const scrubjay_cepstrum_params *params)
{
    g_create_err[0] = 0;
    if (!out) { snprintf(g_create_err, sizeof g_create_err, "out is NULL"); return SJ_EINVAL; }
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || device < 0 || device >= count) {
        snprintf(g_create_err, sizeof g_create_err, "no usable CUDA device %d (%s); libscrubjay_cuda has no CPU fallback",
                 device, e == cudaSuccess ? "index out of range" : cudaGetErrorString(e));
        return SJ_ENODEVICE;
    }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess || prop.major != 10) {
        snprintf(g_create_err, sizeof g_create_err, "device %d is not sm_100 (libscrubjay_cuda is built for sm_100a only)", device);
        return SJ_ENODEVICE;
    }
    scrubjay_ctx *c = new (std::nothrow) scrubjay_ctx();
    if (!c) return SJ_ENOMEM;
    c->device = device;
    c->samplerate = samplerate;
    c->num_sms = prop.multiProcessorCount;
    c->h_tables = new (std::nothrow) SjTables;
    char terr[256] = {0};
    int rc = SJ_OK;
    if (!c->h_tables) rc = SJ_ENOMEM;
    else if (sj_build_tables(c->h_tables, model_blob, blob_bytes, samplerate, params, terr)) {
        snprintf(g_create_err, sizeof g_create_err, "%s", terr);
        rc = strstr(terr, "model") || strstr(terr, "support-vector") ? SJ_EMODEL : SJ_EINVAL;
    }
    if (rc == SJ_OK) {
        cudaError_t ce = cudaSetDevice(device);
        if (ce == cudaSuccess) ce = cudaMalloc(&c->d_tables, sizeof(SjTables));
        if (ce == cudaSuccess) ce = cudaMemcpy(c->d_tables, c->h_tables, sizeof(SjTables), cudaMemcpyHostToDevice);
        if (ce == cudaSuccess) ce = cudaMalloc(&c->d_small, 4096);
        if (ce == cudaSuccess) ce = cudaHostAlloc(&c->h_zc, 4096, cudaHostAllocMapped);
        if (ce == cudaSuccess) ce = cudaEventCreate(&c->ev0);
        if (ce == cudaSuccess) ce = cudaEventCreate(&c->ev1);
        if (ce == cudaSuccess) ce = sj_cepstrum_setup(c->h_tables->mel_maxseg);
        if (ce != cudaSuccess) {
            snprintf(g_create_err, sizeof g_create_err, "CUDA set-up failed: %s", cudaGetErrorString(ce));
            rc = SJ_ECUDA;
        }
    }
    if (rc != SJ_OK) { scrubjay_destroy(c); return rc; }
    *out = c;
    return SJ_OK;
}

void scrubjay_destroy(scrubjay_ctx *c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    free_stages(c);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    cudaFree(c->d_tables);
    cudaFree(c->d_small);
    cudaFreeHost(c->h_zc);
    cudaFree(c->d_one);
    cudaFree(c->d_peers);
    cudaFree(c->d_frame_scratch);
    if (c->frame_scratch_free) cudaEventDestroy(c->frame_scratch_free);
    for (void *t : c->d_donut_tables) cudaFree(t);
    cudaFree(c->d_donut_scratch);
    cudaFree(c->d_donut_clips); cudaFree(c->d_donut_labels); cudaFree(c->d_donut_detail);
    cudaFree(c->d_stage_coef); cudaFree(c->d_stage_M); cudaFree(c->d_stage_blk); cudaFree(c->d_stage_in); cudaFree(c->d_stage_out);
    for (auto &r : c->d_stage_tab) for (void *t : r) cudaFree(t);
    for (void *t : c->d_ep_params) cudaFree(t);
    cudaFree(c->d_long_ws);
    cudaFree(c->d_kw_tables); cudaFree(c->d_kw_signals); cudaFree(c->d_kw_prob);
    delete c->h_tables;
    delete c;
}

const char *scrubjay_last_error(const scrubjay_ctx *c) { return c ? c->err : g_create_err; }

int scrubjay_set_stream(scrubjay_ctx *c, void *s)
{
    if (!c) return SJ_EINVAL;
    c->stream = (cudaStream_t)s;
    return SJ_OK;
}

long scrubjay_launch_count(const scrubjay_ctx *c) { return c ? c->launches : 0; }

int scrubjay_last_kernel_ms(scrubjay_ctx *c, float *ms)
{
    if (!c || !ms) return SJ_EINVAL;
    if (!c->have_timing) return set_err(c, SJ_EINVAL, "no timed launch yet");
    SJ_CUDA(c, cudaEventSynchronize(c->ev1));
    SJ_CUDA(c, cudaEventElapsedTime(ms, c->ev0, c->ev1));
    return SJ_OK;
}

int scrubjay_mfcc_stats_batch_dev(scrubjay_ctx *c, const float *d_clips, long n_clips, int n_samples, float *d_feats)
{
    if (!c) return SJ_EINVAL;
    if (!d_clips || !d_feats || n_clips < 0) return set_err(c, SJ_EINVAL, "NULL pointer or negative count");
    if (n_samples <= 0) return set_err(c, SJ_EINVAL, "empty clip"); // scrubjay_infer.c:55-59
    SJ_CUDA(c, cudaSetDevice(c->device));
    return launch_cepstrum(c, c->stream, d_clips, n_clips, n_samples, nullptr, nullptr, nullptr, d_feats, 0, true);
}

int scrubjay_predict_batch_dev(scrubjay_ctx *c, const float *d_feats, long n, int32_t *d_labels, float *d_conf,
                               float *d_score)
{
    if (!c) return SJ_EINVAL;
    if (!d_feats || !d_labels || n < 0) return set_err(c, SJ_EINVAL, "NULL pointer or negative count");
    SJ_CUDA(c, cudaSetDevice(c->device));
    SJ_CUDA(c, cudaEventRecord(c->ev0, c->stream));
    SJ_CUDA(c, sj_launch_svm(d_feats, n, c->d_tables, d_labels, d_conf, d_score, c->stream));
    SJ_CUDA(c, cudaEventRecord(c->ev1, c->stream));
    c->have_timing = true;
    if (n > 0) c->launches += 1;
    return SJ_OK;
}

int scrubjay_infer_batch_dev(scrubjay_ctx *c, const float *d_clips, long n_clips, int n_samples, int32_t *d_labels,
                             float *d_conf, float *d_score, float *d_feats)
{
    if (!c) return SJ_EINVAL;
    if (!d_clips || !d_labels || n_clips < 0) return set_err(c, SJ_EINVAL, "NULL pointer or negative count");
    if (n_samples <= 0) return set_err(c, SJ_EINVAL, "empty clip");
    SJ_CUDA(c, cudaSetDevice(c->device));
    return launch_cepstrum(c, c->stream, d_clips, n_clips, n_samples, d_labels, d_conf, d_score, d_feats, 1, true);
}

int scrubjay_infer_batch_pcm16_dev(scrubjay_ctx *c, const int16_t *d_clips, long n_clips, int n_samples,
                                   int32_t *d_labels, float *d_conf, float *d_score, float *d_feats)
{
    if (!c) return SJ_EINVAL;
    if (!d_clips || !d_labels || n_clips < 0) return set_err(c, SJ_EINVAL, "NULL pointer or negative count");
    if (n_samples <= 0) return set_err(c, SJ_EINVAL, "empty clip");
    SJ_CUDA(c, cudaSetDevice(c->device));
    return launch_cepstrum(c, c->stream, reinterpret_cast<const float *>(d_clips), n_clips, n_samples, d_labels, d_conf,
                           d_score, d_feats, 1, true, 1);
}

/* Allocation-free fused mfcc_stats + ort_predict for small batches (no counterpart in the reference).
 * Always the frame-parallel form: the STREAM instantiation with one-frame windows
 * writes every frame's MFCC vector into the library's static device array (address from cudaGetSymbolAddress, looked
 * up once per device in cepstrum_kernel.cu), then clip_stats_static_kernel forms the per-clip statistics in frame
 * order -- the arithmetic and the order of the double sums of scrubjay_infer_batch_dev, so the outputs are the same
 * bits.  Two kernel launches; no cudaMalloc / cudaFree / event or stream creation, nothing else that allocates. */
int scrubjay_infer_batch_noalloc_dev(scrubjay_ctx *c, const float *d_clips, long n_clips, int n_samples, int32_t *d_labels,
                                     float *d_conf, float *d_score, float *d_feats)
{
    static_assert(SJ_NOALLOC_MAX_FRAMES == SJ_NOALLOC_MAX_TOTAL_FRAMES, "public and internal limits");
    if (!c) return SJ_EINVAL;
    if (!d_clips || !d_labels || n_clips < 0) return set_err(c, SJ_EINVAL, "NULL pointer or negative count");
    if (n_samples <= 0) return set_err(c, SJ_EINVAL, "empty clip");
    if (n_clips == 0) return SJ_OK;
    const long F = ((long)n_samples + SJ_HOP - 1) / SJ_HOP;
    if (n_clips > SJ_NOALLOC_MAX_FRAMES || n_clips * F > SJ_NOALLOC_MAX_FRAMES) {
        snprintf(c->err, sizeof c->err,
                 "scrubjay_infer_batch_noalloc_dev: %ld clips x %ld frames = %ld frames exceed the static scratch of %d frames "
                 "(SJ_NOALLOC_MAX_TOTAL_FRAMES); use scrubjay_infer_batch_dev", n_clips, F, n_clips * F, SJ_NOALLOC_MAX_FRAMES);
        return SJ_EINVAL;
    }
    SJ_CUDA(c, cudaSetDevice(c->device));
    float *scratch = nullptr;
    SJ_CUDA(c, sj_noalloc_frame_scratch(c->device, &scratch));
    SjCepstrumArgs a;
    a.pcm16 = 0;
    a.peers = nullptr;
    a.clips = d_clips; a.n_clips = n_clips; a.n_samples = n_samples; a.tables = c->d_tables;
    a.labels = d_labels; a.conf = d_conf; a.score = d_score; a.feats = d_feats; a.do_svm = 1;
    a.chan_stride = 0; a.prev_hop = nullptr; a.n_windows = 0; a.win_frames = a.step_frames = 0;
    a.chunk_windows = 0; a.chunks_per_chan = 0;
    SjCepstrumArgs f = a; // first kernel: every frame's MFCC vector -> scratch[clip][frame][40]
    f.labels = nullptr; f.conf = nullptr; f.score = nullptr; f.do_svm = 0;
    f.feats = scratch;
    f.chan_stride = n_samples; f.n_windows = F; f.win_frames = f.step_frames = 1;
    const long resident = sj_cepstrum_resident_warps(c->num_sms);
    long k = (n_clips * F + resident - 1) / resident; // frames per work item: about one item per resident warp
    if (k < 1) k = 1;
    if (k > F) k = F;
    f.chunk_windows = (int)k;
    f.chunks_per_chan = (F + k - 1) / k;
    SJ_CUDA(c, sj_launch_cepstrum(f, c->h_tables, c->num_sms, c->stream));
    SJ_CUDA(c, sj_launch_clip_stats_static(a, (int)F, c->stream));
    c->launches += 2;
    return SJ_OK;
}

int scrubjay_peer_alloc(scrubjay_ctx *c, size_t bytes, void **d_ptr, unsigned char handle[SJ_PEER_HANDLE_BYTES])
{
    static_assert(sizeof(cudaIpcMemHandle_t) == SJ_PEER_HANDLE_BYTES, "handle size");
    if (!c) return SJ_EINVAL;
    if (!d_ptr || bytes == 0) return set_err(c, SJ_EINVAL, "NULL pointer or empty allocation");
    *d_ptr = nullptr;
    SJ_CUDA(c, cudaSetDevice(c->device));
    void *p = nullptr;
    SJ_CUDA(c, cudaMalloc(&p, bytes)); // a whole allocation: IPC handles name allocations, not sub-ranges
    if (handle) {
        cudaIpcMemHandle_t h;
        const cudaError_t e = cudaIpcGetMemHandle(&h, p);
        if (e != cudaSuccess) { cudaFree(p); return set_err(c, SJ_ECUDA, "cudaIpcGetMemHandle", e); }
        memcpy(handle, &h, sizeof h);
    }
    *d_ptr = p;
    return SJ_OK;
}

int scrubjay_peer_free(scrubjay_ctx *c, void *d_ptr)
{
    if (!c) return SJ_EINVAL;
    SJ_CUDA(c, cudaSetDevice(c->device));
    SJ_CUDA(c, cudaFree(d_ptr));
    return SJ_OK;
}

int scrubjay_peer_open(scrubjay_ctx *c, const unsigned char handle[SJ_PEER_HANDLE_BYTES], void **d_ptr)
{
    if (!c) return SJ_EINVAL;
    if (!handle || !d_ptr) return set_err(c, SJ_EINVAL, "NULL pointer");
    *d_ptr = nullptr;
    SJ_CUDA(c, cudaSetDevice(c->device));
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof h);
    SJ_CUDA(c, cudaIpcOpenMemHandle(d_ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return SJ_OK;
}

int scrubjay_peer_close(scrubjay_ctx *c, void *d_ptr)
{
    if (!c) return SJ_EINVAL;
    SJ_CUDA(c, cudaSetDevice(c->device));
    SJ_CUDA(c, cudaIpcCloseMemHandle(d_ptr));
    return SJ_OK;
}

int scrubjay_gather_bind(scrubjay_ctx *c, int n_peers, int32_t *const *peer_labels, float *const *peer_score, long out_base)
{
    if (!c) return SJ_EINVAL;
    if (n_peers < 0 || n_peers > SJ_MAX_PEERS || out_base < 0) return set_err(c, SJ_EINVAL, "n_peers out of range or negative base");
    if (n_peers > 0 && !peer_labels) return set_err(c, SJ_EINVAL, "NULL target list");
    static_assert(SJ_MAX_PEERS == SJ_KMAX_PEERS, "peer table size");
    SJ_CUDA(c, cudaSetDevice(c->device));
    SjPeerTable t;
    memset(&t, 0, sizeof t);
    for (int p = 0; p < n_peers; ++p) {
        if (!peer_labels[p]) return set_err(c, SJ_EINVAL, "NULL label target");
        t.labels[p] = peer_labels[p];
        t.score[p] = peer_score ? peer_score[p] : nullptr;
        // same-process targets on another device need peer access; targets mapped with scrubjay_peer_open already have it
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, peer_labels[p]) == cudaSuccess && at.type == cudaMemoryTypeDevice &&
            at.device != c->device && at.device >= 0 && at.device < SJ_MAX_PEERS * 2 && !c->peer_access_on[at.device]) {
            const cudaError_t e = cudaDeviceEnablePeerAccess(at.device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return set_err(c, SJ_ECUDA, "cudaDeviceEnablePeerAccess", e);
            (void)cudaGetLastError();
            c->peer_access_on[at.device] = 1;
        } else (void)cudaGetLastError();
    }
    t.out_base = out_base;
    t.n_peers = n_peers;
    if (!c->d_peers) SJ_CUDA(c, cudaMalloc(&c->d_peers, sizeof(SjPeerTable)));
    // stream-ordered with respect to the launches that use the table
    SJ_CUDA(c, cudaMemcpyAsync(c->d_peers, &t, sizeof t, cudaMemcpyHostToDevice, c->stream));
    SJ_CUDA(c, cudaStreamSynchronize(c->stream)); // `t` is on the stack
    c->n_peers = n_peers;
    return SJ_OK;
}

int scrubjay_infer_batch_gather_dev(scrubjay_ctx *c, const float *d_clips, long n_clips, int n_samples, float *d_conf,
                                    float *d_feats)
{
    if (!c) return SJ_EINVAL;
    if (!d_clips || n_clips < 0) return set_err(c, SJ_EINVAL, "NULL pointer or negative count");
    if (n_samples <= 0) return set_err(c, SJ_EINVAL, "empty clip");
    if (c->n_peers <= 0) return set_err(c, SJ_EINVAL, "no gather targets bound (scrubjay_gather_bind)");
    SJ_CUDA(c, cudaSetDevice(c->device));
    return launch_cepstrum(c, c->stream, d_clips, n_clips, n_samples, nullptr, d_conf, nullptr, d_feats, 1, true, 0,
                           c->d_peers);
}

int scrubjay_wav_parse(const void *bytes, size_t n, scrubjay_wav_info *info) { return scrubjay_wav_parse_ex(bytes, n, info, 1); }

int scrubjay_wav_parse_ex(const void *bytes, size_t n, scrubjay_wav_info *info, int strict)
{
    if (!bytes || !info) return SJ_EINVAL;
    const unsigned char *p = static_cast<const unsigned char *>(bytes);
    auto u16 = [&](size_t o) { return (uint32_t)p[o] | ((uint32_t)p[o + 1] << 8); };
    auto u32 = [&](size_t o) { return u16(o) | (u16(o + 2) << 16); };
    memset(info, 0, sizeof *info);
    if (n < 36 || memcmp(p, "RIFF", 4) || memcmp(p + 8, "WAVE", 4)) return SJ_EINVAL;
    // strict: "fmt " must be the first chunk, as read_wav_file demands (classifier.c:248-253); otherwise chunks in
    // front of it (JUNK, bext, LIST ...) are skipped the way libsndfile / aubio_source walk a RIFF file
    size_t f = 12;
    if (strict) { if (memcmp(p + f, "fmt ", 4)) return SJ_EINVAL; }
    else {
        while (f + 8 <= n && memcmp(p + f, "fmt ", 4)) f += 8 + (size_t)u32(f + 4) + (u32(f + 4) & 1);
        if (f + 24 > n) return SJ_EINVAL;
    }
    const size_t fmt_size = u32(f + 4);
    info->audio_format = (int32_t)u16(f + 8);
    info->channels = (int32_t)u16(f + 10);
    info->sample_rate = (int32_t)u32(f + 12);
    info->bits_per_sample = (int32_t)u16(f + 22);
    if (info->audio_format != 1 || info->channels < 1 || info->bits_per_sample < 8) return SJ_EINVAL;
    // the reference reads 16 bytes of fmt and expects the next chunk header right after them; a longer fmt
    // chunk (fmt_size > 16) would make it misparse -- here the declared size is honoured
    size_t o = f + 8 + (fmt_size < 16 ? 16 : fmt_size);
    o += fmt_size & 1;
    while (o + 8 <= n && memcmp(p + o, "data", 4)) o += 8 + (size_t)u32(o + 4) + (u32(o + 4) & 1);
    if (o + 8 > n) return SJ_EINVAL;
    size_t data_bytes = u32(o + 4);
    if (data_bytes > n - (o + 8)) data_bytes = n - (o + 8); // truncated file: what is there
    info->data_offset = (int64_t)(o + 8);
    info->data_bytes = (int64_t)data_bytes;
    info->frames = (int64_t)(data_bytes / (size_t)(info->bits_per_sample / 8) / (size_t)info->channels);
    return SJ_OK;
}

int scrubjay_decode_pcm16_dev(scrubjay_ctx *c, const int16_t *d_in, long frames, int channels, int policy, float *d_out)
{
    if (!c) return SJ_EINVAL;
    if (!d_in || !d_out || frames < 0) return set_err(c, SJ_EINVAL, "NULL pointer or negative count");
    if (channels < 1 || (policy != SJ_MIX_FIRST && policy != SJ_MIX_MEAN)) return set_err(c, SJ_EINVAL, "bad channel count or policy");
    SJ_CUDA(c, cudaSetDevice(c->device));
    SJ_CUDA(c, sj_launch_decode_pcm16(d_in, frames, channels, policy, d_out, c->num_sms, c->stream));
    if (frames > 0) c->launches += 1;
    return SJ_OK;
}

int scrubjay_decode_adc12_dev(scrubjay_ctx *c, const uint16_t *d_adc, long n, float *d_out)
{
    if (!c) return SJ_EINVAL;
    if (!d_adc || !d_out || n < 0) return set_err(c, SJ_EINVAL, "NULL pointer or negative count");
    SJ_CUDA(c, cudaSetDevice(c->device));
    SJ_CUDA(c, sj_launch_decode_adc12(d_adc, n, d_out, c->num_sms, c->stream));
    if (n > 0) c->launches += 1;
    return SJ_OK;
}

int scrubjay_upsample_linear_dev(scrubjay_ctx *c, const float *d_in, long n_signals, int old_size, int new_size, float *d_out)
{
    if (!c) return SJ_EINVAL;
    if (!d_in || !d_out || n_signals < 0) return set_err(c, SJ_EINVAL, "NULL pointer or negative count");
    if (old_size < 1 || new_size < 2) return set_err(c, SJ_EINVAL, "upsample: old_size >= 1 and new_size >= 2 (the reference divides by new_size - 1)");
    SJ_CUDA(c, cudaSetDevice(c->device));
    SJ_CUDA(c, sj_launch_upsample_linear(d_in, n_signals, old_size, new_size, d_out, c->num_sms, c->stream));
    if (n_signals > 0) c->launches += 1;
    return SJ_OK;
}

long scrubjay_stream_windows(long n_samples, int win_frames, int step_frames)
{
    if (n_samples <= 0 || win_frames < 1 || step_frames < 1) return 0;
    const long hops = (n_samples + SJ_HOP - 1) / SJ_HOP;
    return hops < win_frames ? 0 : (hops - win_frames) / step_frames + 1;
}

static int stream_infer_any(scrubjay_ctx *c, const void *d_streams, long n_channels, long n_samples, long chan_stride,
                            const void *d_prev_hop, int win_frames, int step_frames, int32_t *d_labels,
                            float *d_conf, float *d_score, float *d_feats, int pcm16, bool gather = false);

int scrubjay_stream_infer_dev(scrubjay_ctx *c, const float *d_streams, long n_channels, long n_samples, long chan_stride,
                              const float *d_prev_hop, int win_frames, int step_frames, int32_t *d_labels,
                              float *d_conf, float *d_score, float *d_feats)
{
    return stream_infer_any(c, d_streams, n_channels, n_samples, chan_stride, d_prev_hop, win_frames, step_frames, d_labels,
                            d_conf, d_score, d_feats, 0);
}

int scrubjay_stream_infer_pcm16_dev(scrubjay_ctx *c, const int16_t *d_streams, long n_channels, long n_samples,
                                    long chan_stride, const int16_t *d_prev_hop, int win_frames, int step_frames,
                                    int32_t *d_labels, float *d_conf, float *d_score, float *d_feats)
{
    return stream_infer_any(c, d_streams, n_channels, n_samples, chan_stride, d_prev_hop, win_frames, step_frames, d_labels,
                            d_conf, d_score, d_feats, 1);
}

int scrubjay_stream_infer_gather_dev(scrubjay_ctx *c, const float *d_streams, long n_channels, long n_samples,
                                     long chan_stride, const float *d_prev_hop, int win_frames, int step_frames,
                                     float *d_conf, float *d_feats)
{
    return stream_infer_any(c, d_streams, n_channels, n_samples, chan_stride, d_prev_hop, win_frames, step_frames, nullptr,
                            d_conf, nullptr, d_feats, 0, true);
}

int scrubjay_stream_infer_pcm16_gather_dev(scrubjay_ctx *c, const int16_t *d_streams, long n_channels, long n_samples,
                                           long chan_stride, const int16_t *d_prev_hop, int win_frames, int step_frames,
                                           float *d_conf, float *d_feats)
{
    return stream_infer_any(c, d_streams, n_channels, n_samples, chan_stride, d_prev_hop, win_frames, step_frames, nullptr,
                            d_conf, nullptr, d_feats, 1, true);
}

static int stream_infer_any(scrubjay_ctx *c, const void *d_streams, long n_channels, long n_samples, long chan_stride,
                            const void *d_prev_hop, int win_frames, int step_frames, int32_t *d_labels,
                            float *d_conf, float *d_score, float *d_feats, int pcm16, bool gather)
{
    if (!c) return SJ_EINVAL;
    if (!d_streams || (!d_labels && !gather) || n_channels < 0) return set_err(c, SJ_EINVAL, "NULL pointer or negative count");
    if (gather && c->n_peers <= 0) return set_err(c, SJ_EINVAL, "no gather targets bound (scrubjay_gather_bind)");
    if (n_samples <= 0 || n_samples > 0x7fffffffL) return set_err(c, SJ_EINVAL, "samples per channel must be in 1 .. 2^31-1");
    if (chan_stride < n_samples) return set_err(c, SJ_EINVAL, "channel stride shorter than the channel");
    if (win_frames < 1 || win_frames > SJ_STREAM_MAX_WIN || step_frames < 1 || step_frames > win_frames)
        return set_err(c, SJ_EINVAL, "need 1 <= step_frames <= win_frames <= 16");
    const long nw = scrubjay_stream_windows(n_samples, win_frames, step_frames);
    if (nw == 0 || n_channels == 0) return SJ_OK;
    SJ_CUDA(c, cudaSetDevice(c->device));
    SjCepstrumArgs a;
    a.pcm16 = pcm16;
    a.peers = gather ? c->d_peers : nullptr;
    a.clips = reinterpret_cast<const float *>(d_streams); a.n_clips = n_channels; a.n_samples = (int)n_samples; a.tables = c->d_tables;
    a.labels = d_labels; a.conf = d_conf; a.score = d_score; a.feats = d_feats; a.do_svm = 1;
    a.chan_stride = chan_stride; a.prev_hop = reinterpret_cast<const float *>(d_prev_hop); a.n_windows = nw;
    a.win_frames = win_frames; a.step_frames = step_frames;
    // work items: enough of them that the static round-robin over the resident warps ends evenly (>= 16 per
    // warp), but long enough that the win - step frames recomputed at the head of each item stay below ~20 %
    // of its work
    const long want = 16L * sj_cepstrum_resident_warps(c->num_sms);
    long per_chan = (want + n_channels - 1) / n_channels;
    long k = (nw + per_chan - 1) / per_chan;
    const long k_min = 4L * (win_frames - step_frames) / step_frames;
    if (k < k_min) k = k_min;
    if (k < 1) k = 1;
    if (c->stream_chunk_windows > 0) k = c->stream_chunk_windows;
    if (k > nw) k = nw;
    a.chunk_windows = (int)k;
    a.chunks_per_chan = (nw + k - 1) / k;
    SJ_CUDA(c, cudaEventRecord(c->ev0, c->stream));
    SJ_CUDA(c, sj_launch_cepstrum(a, c->h_tables, c->num_sms, c->stream));
    SJ_CUDA(c, cudaEventRecord(c->ev1, c->stream));
    c->have_timing = true;
    c->launches += 1;
    return SJ_OK;
}

int scrubjay_set_latency_mode(scrubjay_ctx *c, int on)
{
    if (!c) return SJ_EINVAL;
    c->latency_mode = on ? 1 : 0;
    return SJ_OK;
}

int scrubjay_set_stream_chunk(scrubjay_ctx *c, int windows_per_item)
{
    if (!c || windows_per_item < 0) return SJ_EINVAL;
    c->stream_chunk_windows = windows_per_item;
    return SJ_OK;
}

} // extern "C"

namespace {

// drain the outputs of a finished chunk from the stage's pinned buffers into the caller's arrays
int drain(scrubjay_ctx *c, Stage &s, int32_t *labels, float *conf, float *score, float *feats)
{
    if (s.pending_first < 0) return SJ_OK;
    SJ_CUDA(c, cudaEventSynchronize(s.done));
    const long f = s.pending_first, n = s.pending_count;
    if (labels) memcpy(labels + f, s.h_labels, (size_t)n * 4);
    if (conf) memcpy(conf + f, s.h_conf, (size_t)n * 4);
    if (score) memcpy(score + f, s.h_score, (size_t)n * 4);
    if (feats) memcpy(feats + f * SJ_FEAT, s.h_feats, (size_t)n * SJ_FEAT * 4);
    s.pending_first = -1;
    return SJ_OK;
}

// After a failed call nothing of it may survive into the next one: wait for whatever the stages still have in flight
// and forget their pending outputs (they belong to the failed call's arrays).
void reset_stages(scrubjay_ctx *c)
{
    for (Stage &s : c->st) {
        if (s.stream && s.pending_first >= 0) cudaStreamSynchronize(s.stream);
        s.pending_first = -1;
        s.pending_count = 0;
    }
    (void)cudaGetLastError();
}

int run_host_batch_impl(scrubjay_ctx *c, const void *clips_v, long n_clips, int n_samples, int32_t *labels, float *conf,
                        float *score, float *feats, int do_svm, int pcm16);

// host clips -> chunked 3-stage pipeline.  do_svm = 0 computes features only.
int run_host_batch(scrubjay_ctx *c, const void *clips_v, long n_clips, int n_samples, int32_t *labels, float *conf,
                   float *score, float *feats, int do_svm, int pcm16 = 0)
{
    reset_stages(c); // a previous call that failed half-way must not drain into this call's arrays
    const int rc = run_host_batch_impl(c, clips_v, n_clips, n_samples, labels, conf, score, feats, do_svm, pcm16);
    if (rc != SJ_OK) reset_stages(c);
    return rc;
}

int run_host_batch_impl(scrubjay_ctx *c, const void *clips_v, long n_clips, int n_samples, int32_t *labels, float *conf,
                        float *score, float *feats, int do_svm, int pcm16)
{
    const unsigned char *clips = static_cast<const unsigned char *>(clips_v);
    const size_t es = pcm16 ? 2 : 4; // bytes per sample; the staging buffers are sized for float
    SJ_CUDA(c, cudaSetDevice(c->device));
    // One to three clips in pinned host memory (BASELINE config 1, the latency case): no copy operations at all.  The
    // kernels read the clips through their device alias (bulk copies over PCIe: 64 KB per clip) and write the results into
    // mapped pinned memory of the context -- two launches and one synchronisation instead of a copy in, two launches,
    // three copies out: 32 us against 45 us per call for one clip (tools/zero_copy_probe.py).  From four clips on the
    // copy engine is as fast or faster.
    constexpr long kZeroCopyClips = 3;
    if (n_clips >= 1 && n_clips <= kZeroCopyClips && c->h_zc && c->fail_at_chunk < 0) {
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, clips_v) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer) {
            float *zc = nullptr; // [0, 8) labels, [8, 16) confidence, [16, 24) score, from 64 on [clip][40] features
            SJ_CUDA(c, cudaHostGetDevicePointer((void **)&zc, c->h_zc, 0));
            int rc = launch_cepstrum(c, c->stream, static_cast<const float *>(at.devicePointer), n_clips, n_samples, (int32_t *)zc,
                                     zc + 8, zc + 16, feats ? zc + 64 : nullptr, do_svm, false, pcm16);
            if (rc != SJ_OK) return rc;
            SJ_CUDA(c, cudaStreamSynchronize(c->stream));
            if (do_svm) {
                memcpy(labels, c->h_zc, (size_t)n_clips * 4);
                if (conf) memcpy(conf, c->h_zc + 8, (size_t)n_clips * 4);
                if (score) memcpy(score, c->h_zc + 16, (size_t)n_clips * 4);
            }
            if (feats) memcpy(feats, c->h_zc + 64, (size_t)n_clips * SJ_FEAT * 4);
            return SJ_OK;
        } else cudaGetLastError(); // a pageable pointer is not an error here
    }
    int rc = ensure_stages(c, n_samples);
    if (rc != SJ_OK) return rc;
    long done = 0;
    for (int i = 0; done < n_clips; ++i) {
        Stage &s = c->st[i % kStages];
        rc = drain(c, s, labels, conf, score, feats);
        if (rc != SJ_OK) return rc;
        const long n = (n_clips - done < c->chunk_clips) ? n_clips - done : c->chunk_clips;
        if (c->fail_at_chunk == i) { c->fail_at_chunk = -1; return set_err(c, SJ_ECUDA, "injected failure (scrubjay_debug_inject_failure)"); }
        SJ_CUDA(c, cudaMemcpyAsync(s.d_clips, clips + (size_t)done * n_samples * es, (size_t)n * n_samples * es,
                                   cudaMemcpyHostToDevice, s.stream));
        rc = launch_cepstrum(c, s.stream, s.d_clips, n, n_samples, s.d_labels, s.d_conf, s.d_score,
                             feats ? s.d_feats : nullptr, do_svm, false, pcm16);
        if (rc != SJ_OK) return rc;
        if (do_svm) {
            SJ_CUDA(c, cudaMemcpyAsync(s.h_labels, s.d_labels, (size_t)n * 4, cudaMemcpyDeviceToHost, s.stream));
            if (conf) SJ_CUDA(c, cudaMemcpyAsync(s.h_conf, s.d_conf, (size_t)n * 4, cudaMemcpyDeviceToHost, s.stream));
            if (score) SJ_CUDA(c, cudaMemcpyAsync(s.h_score, s.d_score, (size_t)n * 4, cudaMemcpyDeviceToHost, s.stream));
        }
        if (feats)
            SJ_CUDA(c, cudaMemcpyAsync(s.h_feats, s.d_feats, (size_t)n * SJ_FEAT * 4, cudaMemcpyDeviceToHost, s.stream));
        SJ_CUDA(c, cudaEventRecord(s.done, s.stream));
        s.pending_first = done;
        s.pending_count = n;
        done += n;
    }
    for (Stage &s : c->st) {
        rc = drain(c, s, do_svm ? labels : nullptr, do_svm ? conf : nullptr, do_svm ? score : nullptr, feats);
        if (rc != SJ_OK) return rc;
    }
    return SJ_OK;
}

} // namespace

extern "C" {

int scrubjay_infer_batch(scrubjay_ctx *c, const float *clips, long n_clips, int n_samples, int32_t *labels,
                         float *conf, float *score, float *feats)
{
    if (!c) return SJ_EINVAL;
    if (!clips || !labels || n_clips < 0) return set_err(c, SJ_EINVAL, "NULL pointer or negative count");
    if (n_samples <= 0) return set_err(c, SJ_EINVAL, "empty clip");
    return run_host_batch(c, clips, n_clips, n_samples, labels, conf, score, feats, 1);
}

int scrubjay_infer_batch_pcm16(scrubjay_ctx *c, const int16_t *clips, long n_clips, int n_samples, int32_t *labels,
                               float *conf, float *score, float *feats)
{
    if (!c) return SJ_EINVAL;
    if (!clips || !labels || n_clips < 0) return set_err(c, SJ_EINVAL, "NULL pointer or negative count");
    if (n_samples <= 0) return set_err(c, SJ_EINVAL, "empty clip");
    return run_host_batch(c, clips, n_clips, n_samples, labels, conf, score, feats, 1, 1);
}

int scrubjay_mfcc_stats(scrubjay_ctx *c, const float *samples, long n_samples, float out[SJ_FEAT_DIM])
{
    if (!c) return SJ_EINVAL;
    if (!samples || !out) return set_err(c, SJ_EINVAL, "NULL pointer");
    if (n_samples <= 0 || n_samples > 0x7fffffffL) return set_err(c, SJ_EINVAL, "empty clip");
    SJ_CUDA(c, cudaSetDevice(c->device));
    {   // a clip in pinned host memory is read in place and answered through mapped memory (see run_host_batch_impl)
        cudaPointerAttributes at;
        if (c->h_zc && cudaPointerGetAttributes(&at, samples) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer) {
            float *zc = nullptr;
            SJ_CUDA(c, cudaHostGetDevicePointer((void **)&zc, c->h_zc, 0));
            int rc = launch_cepstrum(c, c->stream, static_cast<const float *>(at.devicePointer), 1, (int)n_samples, nullptr, nullptr,
                                     nullptr, zc + 64, 0, false);
            if (rc != SJ_OK) return rc;
            SJ_CUDA(c, cudaStreamSynchronize(c->stream));
            memcpy(out, c->h_zc + 64, SJ_FEAT * 4);
            return SJ_OK;
        } else cudaGetLastError();
    }
    // the clip goes through a grow-only buffer owned by the context (sized for 4 s at 96 kHz at first use): no
    // allocation on the steady-state path
    const size_t need = (size_t)n_samples * 4;
    if (need > c->one_bytes) {
        SJ_CUDA(c, cudaStreamSynchronize(c->stream));
        cudaFree(c->d_one);
        c->d_one = nullptr;
        c->one_bytes = 0;
        const size_t want = need > (1536u << 10) ? need : (size_t)(1536u << 10);
        SJ_CUDA(c, cudaMalloc(&c->d_one, want));
        c->one_bytes = want;
    }
    float *d = c->d_one;
    cudaError_t e = cudaMemcpyAsync(d, samples, need, cudaMemcpyHostToDevice, c->stream);
    int rc = SJ_OK;
    if (e == cudaSuccess)
        rc = launch_cepstrum(c, c->stream, d, 1, (int)n_samples, nullptr, nullptr, nullptr, c->d_small, 0, false);
    if (e == cudaSuccess && rc == SJ_OK)
        e = cudaMemcpyAsync(out, c->d_small, SJ_FEAT * 4, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess && rc == SJ_OK) e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) return set_err(c, SJ_ECUDA, "scrubjay_mfcc_stats", e);
    return rc;
}

int scrubjay_predict(scrubjay_ctx *c, const float feat[SJ_FEAT_DIM], float *conf)
{
    if (!c) return SJ_EINVAL;
    if (!feat) return set_err(c, SJ_EINVAL, "NULL pointer");
    SJ_CUDA(c, cudaSetDevice(c->device));
    float *d_feat = c->d_small;            // [40]
    int32_t *d_label = (int32_t *)(c->d_small + 64);
    float *d_conf = c->d_small + 65, *d_score = c->d_small + 66;
    SJ_CUDA(c, cudaMemcpyAsync(d_feat, feat, SJ_FEAT * 4, cudaMemcpyHostToDevice, c->stream));
    SJ_CUDA(c, sj_launch_svm(d_feat, 1, c->d_tables, d_label, d_conf, d_score, c->stream));
    c->launches += 1;
    float host[3];
    SJ_CUDA(c, cudaMemcpyAsync(host, c->d_small + 64, 12, cudaMemcpyDeviceToHost, c->stream));
    SJ_CUDA(c, cudaStreamSynchronize(c->stream));
    int32_t label;
    memcpy(&label, &host[0], 4);
    if (conf) *conf = host[1];
    return (int)label;
}

int scrubjay_synth_clips_dev(scrubjay_ctx *c, float *d_out, long first_clip, long n_clips, int n_samples,
                             unsigned seed)
{
    return scrubjay_synth_clips_ex_dev(c, d_out, first_clip, n_clips, n_samples, seed, SJ_SYNTH_MIXED);
}

int scrubjay_synth_clips_ex_dev(scrubjay_ctx *c, float *d_out, long first_clip, long n_clips, int n_samples,
                                unsigned seed, int flavour)
{
    if (!c) return SJ_EINVAL;
    if (!d_out || n_clips < 0 || n_samples <= 0) return set_err(c, SJ_EINVAL, "bad argument");
    if (flavour != SJ_SYNTH_MIXED && flavour != SJ_SYNTH_CALLS) return set_err(c, SJ_EINVAL, "unknown synth flavour");
    SJ_CUDA(c, cudaSetDevice(c->device));
    SJ_CUDA(c, sj_launch_synth(d_out, first_clip, n_clips, n_samples, seed, (float)c->samplerate, flavour, c->stream));
    if (n_clips > 0) c->launches += 1;
    return SJ_OK;
}

/* diagnostic (host only, no GPU): expand the kernel's chunked sparse mel tables back to a dense
 * [40][1025] bank so tests can compare them with an independent construction */
int scrubjay_debug_mel_dense(unsigned samplerate, float *out)
{
    if (!out) return SJ_EINVAL;
    SjTables *t = new (std::nothrow) SjTables;
    if (!t) return SJ_ENOMEM;
    char err[256];
    int rc = sj_build_tables(t, nullptr, 0, samplerate, nullptr, err);
    if (rc) { snprintf(g_create_err, sizeof g_create_err, "%s", err); delete t; return SJ_EINVAL; }
    sj_replay_mel_dense(t, out);
    delete t;
    return SJ_OK;
}

/* test hook: the next host-pointer batch call fails (SJ_ECUDA) right before it would enqueue chunk number `chunk`
 * (0-based), after the earlier chunks are in flight; -1 disarms */
int scrubjay_debug_inject_failure(scrubjay_ctx *c, long chunk)
{
    if (!c) return SJ_EINVAL;
    c->fail_at_chunk = chunk;
    return SJ_OK;
}

/* pinned host memory for callers that want full PCIe rate on the host-pointer entry points */
int scrubjay_host_alloc(void **p, size_t bytes) { return cudaMallocHost(p, bytes) == cudaSuccess ? SJ_OK : SJ_ENOMEM; }
void scrubjay_host_free(void *p) { cudaFreeHost(p); }

} // extern "C"

// ------------------------------------------------------------------------------------------------
// donut path: int classify(float*, int) (sync/lib/classifier.h:19) and its batched forms
// ------------------------------------------------------------------------------------------------
namespace {

inline bool sj_donut_variant_is_f64(int variant) { return variant == SJ_DONUT_DESKTOP_F64 || variant == SJ_DONUT_REVERT_F64; }

int donut_prepare(scrubjay_ctx *c, int variant, int n_samples, int *tb_out)
{
    if (variant < 0 || variant >= SJ_DONUT_N_VARIANTS) return set_err(c, SJ_EINVAL, "unknown donut variant");
    if (c->samplerate != 16000)
        return set_err(c, SJ_EINVAL, "donut path: only 16 kHz has filter coefficients (classifier.c:403-407)");
    const int tb = n_samples >= DN_WIN ? (n_samples - DN_WIN) / DN_HOP + 1 : 0;
    if (tb < 1) return set_err(c, SJ_EINVAL, "donut path: clip shorter than one 256-sample frame");
    if (tb > DN_MAX_TB) return set_err(c, SJ_EINVAL, "donut path: clip longer than 21536 samples");
    const bool f64 = sj_donut_variant_is_f64(variant);
    if (!c->d_donut_tables[variant]) {
        SJ_CUDA(c, sj_donut_setup());
        SjDonutParams p;
        sj_donut_default_params(&p, variant);
        if (f64) {
            SjDonutTables<double> t;
            sj_donut_build_tables_f64(&t, p, 16000);
            SJ_CUDA(c, cudaMalloc(&c->d_donut_tables[variant], sizeof t));
            SJ_CUDA(c, cudaMemcpy(c->d_donut_tables[variant], &t, sizeof t, cudaMemcpyHostToDevice));
        } else {
            SjDonutTables<float> t;
            sj_donut_build_tables_f32(&t, p, 16000);
            SJ_CUDA(c, cudaMalloc(&c->d_donut_tables[variant], sizeof t));
            SJ_CUDA(c, cudaMemcpy(c->d_donut_tables[variant], &t, sizeof t, cudaMemcpyHostToDevice));
        }
    }
    const size_t need = sj_donut_scratch_bytes(c->num_sms, tb, f64);
    if (need > c->donut_scratch_bytes) {
        SJ_CUDA(c, cudaStreamSynchronize(c->stream));
        cudaFree(c->d_donut_scratch);
        c->d_donut_scratch = nullptr;
        c->donut_scratch_bytes = 0;
        SJ_CUDA(c, cudaMalloc(&c->d_donut_scratch, need));
        c->donut_scratch_bytes = need;
    }
    *tb_out = tb;
    return SJ_OK;
}

int donut_launch(scrubjay_ctx *c, const float *d_clips, long n_clips, int n_samples, int variant, int tb,
                 int32_t *d_labels, float *d_detail, float *d_mids = nullptr, int8_t *d_track = nullptr,
                 long row_stride = 0, int clips_per_row = 0)
{
    SjDonutArgs a;
    a.clips = d_clips; a.n_clips = n_clips; a.n_samples = n_samples; a.tb = tb;
    a.tables = c->d_donut_tables[variant]; a.scratch = c->d_donut_scratch; a.labels = d_labels; a.detail = d_detail; a.mids = d_mids;
    a.track = d_track; a.row_stride = row_stride; a.clips_per_row = clips_per_row;
    SJ_CUDA(c, cudaEventRecord(c->ev0, c->stream));
    SJ_CUDA(c, sj_launch_donut(a, sj_donut_variant_is_f64(variant), c->num_sms, c->stream));
    SJ_CUDA(c, cudaEventRecord(c->ev1, c->stream));
    c->have_timing = true;
    if (n_clips > 0) c->launches += 1;
    return SJ_OK;
}

scrubjay_ctx *g_default_ctx = nullptr;

} // namespace

extern "C" {

// signals longer than the fused kernel's 96 frames, and the 64-sample-window variant, take the staged pipeline
static bool donut_needs_staged(int variant, long n_samples) { return variant == SJ_DONUT_INO_F32 || n_samples > DN_WIN + (long)DN_HOP * (DN_MAX_TB - 1); }
// Midpoints the staged classifier keeps per signal.  The reference evaluates EVERY midpoint (classifier.cpp:92-115), so the
// capacity follows the signal: a cluster spans >= 0.15 s (>= 11 frames of hop 224, >= 43 of hop 56) and is followed by a
// gap > 0.05 s, hence fewer than frames / 12 + 1 clusters; 4096 (the entry point's limit) covers 3 hours of audio.
static int staged_max_mid(int variant, long n_samples)
{
    const long win = variant == SJ_DONUT_INO_F32 ? 64 : DN_WIN, hop = win - win / 8;
    const long frames = n_samples >= win ? (n_samples - win) / hop + 1 : 1;
    long mm = frames / 12 + 1;
    if (mm < 8) mm = 8;
    if (mm > 4096) mm = 4096;
    return (int)mm;
}

int scrubjay_donut_classify_batch_dev(scrubjay_ctx *c, const float *d_clips, long n_clips, int n_samples, int variant,
                                      int32_t *d_labels, float *d_detail)
{
    if (!c) return SJ_EINVAL;
    if (!d_clips || !d_labels || n_clips < 0) return set_err(c, SJ_EINVAL, "NULL pointer or negative count");
    SJ_CUDA(c, cudaSetDevice(c->device));
    if (donut_needs_staged(variant, n_samples))
        return scrubjay_donut_classify_long_dev(c, d_clips, n_clips, n_samples, variant, d_labels, d_detail, nullptr, staged_max_mid(variant, n_samples));
    int tb = 0, rc = donut_prepare(c, variant, n_samples, &tb);
    if (rc != SJ_OK) return rc;
    return donut_launch(c, d_clips, n_clips, n_samples, variant, tb, d_labels, d_detail);
}

int scrubjay_donut_classify_batch(scrubjay_ctx *c, const float *clips, long n_clips, int n_samples, int variant,
                                  int32_t *labels, float *detail)
{
    if (!c) return SJ_EINVAL;
    if (!clips || !labels || n_clips < 0) return set_err(c, SJ_EINVAL, "NULL pointer or negative count");
    SJ_CUDA(c, cudaSetDevice(c->device));
    const bool staged = donut_needs_staged(variant, n_samples);
    int tb = 0, rc = staged ? SJ_OK : donut_prepare(c, variant, n_samples, &tb);
    if (rc != SJ_OK) return rc;
    if (n_samples < 1) return set_err(c, SJ_EINVAL, "donut path: empty clip");
    long chunk = kChunkBytes / ((long)n_samples * 4);
    if (chunk < 1) chunk = 1;
    if (chunk > n_clips && n_clips > 0) chunk = n_clips;
    if (c->donut_stage_clips < chunk || c->donut_stage_samples != n_samples) {
        SJ_CUDA(c, cudaStreamSynchronize(c->stream));
        cudaFree(c->d_donut_clips); cudaFree(c->d_donut_labels); cudaFree(c->d_donut_detail);
        c->d_donut_clips = nullptr; c->d_donut_labels = nullptr; c->d_donut_detail = nullptr;
        c->donut_stage_clips = 0;
        SJ_CUDA(c, cudaMalloc(&c->d_donut_clips, (size_t)chunk * n_samples * 4));
        SJ_CUDA(c, cudaMalloc(&c->d_donut_labels, (size_t)chunk * 4));
        SJ_CUDA(c, cudaMalloc(&c->d_donut_detail, (size_t)chunk * 16));
        c->donut_stage_clips = chunk;
        c->donut_stage_samples = n_samples;
    }
    for (long done = 0; done < n_clips; done += chunk) {
        const long n = n_clips - done < chunk ? n_clips - done : chunk;
        SJ_CUDA(c, cudaMemcpyAsync(c->d_donut_clips, clips + done * (long)n_samples, (size_t)n * n_samples * 4,
                                   cudaMemcpyHostToDevice, c->stream));
        rc = staged ? scrubjay_donut_classify_long_dev(c, c->d_donut_clips, n, n_samples, variant, c->d_donut_labels,
                                                       c->d_donut_detail, nullptr, staged_max_mid(variant, n_samples))
                    : donut_launch(c, c->d_donut_clips, n, n_samples, variant, tb, c->d_donut_labels, c->d_donut_detail);
        if (rc != SJ_OK) return rc;
        SJ_CUDA(c, cudaMemcpyAsync(labels + done, c->d_donut_labels, (size_t)n * 4, cudaMemcpyDeviceToHost, c->stream));
        if (detail)
            SJ_CUDA(c, cudaMemcpyAsync(detail + done * 4, c->d_donut_detail, (size_t)n * 16, cudaMemcpyDeviceToHost, c->stream));
        SJ_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return SJ_OK;
}

/* The reference's own entry point, same name and signature (sync/lib/classifier.h:19; called as
 * classify(buffer, 16000) from sync/sync.cpp:202).  Uses a process-wide context on device
 * $SCRUBJAY_DEVICE (default 0), created on first use.  Returns 1/0 like the reference, and 0 on any
 * failure (the reference returns 0 when find_midpoints fails, classifier.cpp:87-91). */
int classify(float *data, int data_size)
{
    if (!g_default_ctx) {
        const char *env = getenv("SCRUBJAY_DEVICE");
        if (scrubjay_create(&g_default_ctx, env ? atoi(env) : 0, nullptr, 0, 16000) != SJ_OK) return 0;
    }
    int32_t label = 0;
    if (scrubjay_donut_classify_batch(g_default_ctx, data, 1, data_size, SJ_DONUT_SYNC_F32, &label, nullptr) != SJ_OK)
        return 0;
    return (int)label;
}

/* ---- keyword ("stop") detector: 2fa/audio/word/c ------------------------------------------------ */

int scrubjay_keyword_load(scrubjay_ctx *c, const void *blob, size_t bytes)
{
    if (!c) return SJ_EINVAL;
    KwTables *t = new (std::nothrow) KwTables;
    if (!t) return set_err(c, SJ_ENOMEM, "out of memory");
    char err[256];
    if (kw_build_tables(t, blob, bytes, err)) { delete t; return set_err(c, SJ_EMODEL, err); }
    cudaError_t e = cudaSetDevice(c->device);
    if (e == cudaSuccess && !c->d_kw_tables) e = cudaMalloc(&c->d_kw_tables, sizeof(KwTables));
    if (e == cudaSuccess) e = cudaMemcpy(c->d_kw_tables, t, sizeof(KwTables), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = kw_setup();
    c->kw_maxseg = t->mel_maxseg;
    delete t;
    if (e != cudaSuccess) return set_err(c, SJ_ECUDA, "scrubjay_keyword_load", e);
    return SJ_OK;
}

int scrubjay_keyword_classify_batch_dev(scrubjay_ctx *c, const float *d_signals, long n_clips, int n_samples,
                                        float *d_prob, float *d_mfcc, int32_t *d_n_frames)
{
    if (!c) return SJ_EINVAL;
    if (!c->d_kw_tables) return set_err(c, SJ_EMODEL, "no keyword model loaded (scrubjay_keyword_load)");
    if (!d_signals || !d_prob || n_clips < 0 || n_samples < 0) return set_err(c, SJ_EINVAL, "NULL pointer or negative count");
    SJ_CUDA(c, cudaSetDevice(c->device));
    KwArgs a;
    a.signals = d_signals; a.n_clips = n_clips; a.n_samples = n_samples; a.tables = c->d_kw_tables;
    a.prob = d_prob; a.mfcc = d_mfcc; a.n_frames = d_n_frames;
    SJ_CUDA(c, cudaEventRecord(c->ev0, c->stream));
    SJ_CUDA(c, kw_launch(a, c->kw_maxseg, c->num_sms, c->stream));
    SJ_CUDA(c, cudaEventRecord(c->ev1, c->stream));
    c->have_timing = true;
    if (n_clips > 0) c->launches += 1;
    return SJ_OK;
}

int scrubjay_keyword_classify_batch(scrubjay_ctx *c, const float *signals, long n_clips, int n_samples, float *prob)
{
    if (!c) return SJ_EINVAL;
    if (!c->d_kw_tables) return set_err(c, SJ_EMODEL, "no keyword model loaded (scrubjay_keyword_load)");
    if (!signals || !prob || n_clips < 0 || n_samples < 0) return set_err(c, SJ_EINVAL, "NULL pointer or negative count");
    SJ_CUDA(c, cudaSetDevice(c->device));
    const size_t per = (size_t)(n_samples > 0 ? n_samples : 1) * 4;
    long chunk = (long)((256u << 20) / per);
    if (chunk < 1) chunk = 1;
    if (chunk > n_clips) chunk = n_clips;
    if (chunk * per > c->kw_stage_bytes || chunk > c->kw_stage_clips) {
        cudaFree(c->d_kw_signals); cudaFree(c->d_kw_prob);
        c->d_kw_signals = c->d_kw_prob = nullptr; c->kw_stage_bytes = 0; c->kw_stage_clips = 0;
        SJ_CUDA(c, cudaMalloc(&c->d_kw_signals, chunk * per));
        SJ_CUDA(c, cudaMalloc(&c->d_kw_prob, (size_t)chunk * 4));
        c->kw_stage_bytes = chunk * per; c->kw_stage_clips = chunk;
    }
    for (long done = 0; done < n_clips; done += chunk) {
        const long n = n_clips - done < chunk ? n_clips - done : chunk;
        SJ_CUDA(c, cudaMemcpyAsync(c->d_kw_signals, signals + done * (long)n_samples, (size_t)n * n_samples * 4,
                                   cudaMemcpyHostToDevice, c->stream));
        int rc = scrubjay_keyword_classify_batch_dev(c, c->d_kw_signals, n, n_samples, c->d_kw_prob, nullptr, nullptr);
        if (rc != SJ_OK) return rc;
        SJ_CUDA(c, cudaMemcpyAsync(prob + done, c->d_kw_prob, (size_t)n * 4, cudaMemcpyDeviceToHost, c->stream));
        SJ_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return SJ_OK;
}

/* The reference's own entry points, same names and signatures (2fa/audio/word/c/stop_detector.h:9-15).  They run
 * on the process-wide context; the model blob is read once from $SCRUBJAY_KEYWORD_MODEL.  On any failure the
 * probability is 0 (the reference has no failure path). */
float classify_signal(const float *signal, int num_samples)
{
    if (!g_default_ctx) {
        const char *env = getenv("SCRUBJAY_DEVICE");
        if (scrubjay_create(&g_default_ctx, env ? atoi(env) : 0, nullptr, 0, 16000) != SJ_OK) return 0.f;
    }
    if (!g_default_ctx->d_kw_tables) {
        const char *path = getenv("SCRUBJAY_KEYWORD_MODEL");
        FILE *f = path ? fopen(path, "rb") : nullptr;
        if (!f) return 0.f;
        std::vector<char> buf(1 << 20);
        const size_t n = fread(buf.data(), 1, buf.size(), f);
        fclose(f);
        if (scrubjay_keyword_load(g_default_ctx, buf.data(), n) != SJ_OK) return 0.f;
    }
    float p = 0.f;
    if (scrubjay_keyword_classify_batch(g_default_ctx, signal, 1, num_samples, &p) != SJ_OK) return 0.f;
    return p;
}

int classify_signal_binary(const float *signal, int num_samples, float threshold)
{
    return classify_signal(signal, num_samples) > threshold ? 1 : 0;
}

} // extern "C"

#include "donut_stage_api.inc"
