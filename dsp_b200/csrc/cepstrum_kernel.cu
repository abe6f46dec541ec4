// cepstrum_kernel.cu -- fused MFCC-statistics + RBF-SVM kernel for sm_100a.
//
// Replaces, for a whole batch of clips in one launch, the per-file hot loop of the reference:
// mfcc_stats (cepstrum/scrubjay_infer.c:19-77: aubio_pvoc_do + aubio_mfcc_do per hop, mean/std over
// hops) followed by ort_predict (:105-141: ONNX Scaler + SVMClassifier).
//
// Mapping: persistent grid, one CTA per SM, each WARP owns whole clips (clip = CTA + grid * warp + k*stride)
// and walks their hops in order, so no block-level synchronisation exists after the prologue.
//   * HBM -> shared: every hop (1024 samples, 4 KB) is fetched exactly once by a bulk async copy
//     (cp.async.bulk, the TMA engine) into a 2-slot per-warp ring, completion on an mbarrier; hop
//     h+1 is in flight while frame h is computed.  Each sample is used by two frames (50 % overlap)
//     but read from HBM once.
//   * frame = [hop h-1 | hop h] * periodic Hann -> 2048-point real FFT done as a 1024-point complex
//     FFT split 32 x 32: a 32-point FFT in registers per lane, ONE transpose through shared memory
//     (scalar stores, 128-bit loads), a second 32-point register FFT whose first butterfly stage
//     absorbs the step-2 twiddles, then the real-FFT untangling with the partner bin fetched by one
//     warp shuffle.  The upper window half is 1 - w, so only 1024 window values are read.  aubio's fvec_shift (half swap) only flips the sign of odd
//     bins and is dropped because only |X| is used.
//   * |X| (bins mel_lo..) -> sparse Slaney mel sums (<= 2 filters per bin, lane-contiguous chunks
//     with an odd stride = conflict-free) -> log10 with aubio's 2e-42 floor -> 20x40 DCT -> the
//     reference's double-precision sum / sum of squares per coefficient.
//   * after the last hop: mean/std -> Scaler -> 55-SV RBF -> Platt probability (libsvm's
//     multiclass_probability for k = 2) -> label, confidence, decision value.
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>

#define SJ_PACKED_F32 1
#include "sj_pk.h"
#include "fft_gen.inc"
#include "sj_tables.h"
#include "cepstrum_kernel.h"

namespace {

#ifndef SJ_WIDE_TRANSPOSE
#define SJ_WIDE_TRANSPOSE 1
#endif
constexpr int kTStride = 36;                 // row stride of the transpose plane: 144 B => conflict-free 128-bit row reads
constexpr int kTFloats = 32 * kTStride;      // 1152 floats: transpose plane, then |X| bins 0..1024
// plane size per warp: the interleaved transpose needs 32 x 68 = 2176 floats INCLUDING the slot of the previous hop:
// 1024 + 1152 with float hops; int16 hops are 512 floats, so their plane grows to 1664
__host__ __device__ constexpr int t_floats(bool pcm16) { return (SJ_WIDE_TRANSPOSE && pcm16) ? 2176 - SJ_HOP / 2 : kTFloats; }
constexpr int kLogFloats = 96;               // folded log-energies of two frames (2 x 40, +pad)

// mel partial sums: generic layout 32*maxseg, fast layout 32*SJ_MEL_FAST_SLOTS, + the zero slot
__host__ __device__ constexpr int p_floats(int maxseg)
{
    return (32 * (maxseg > SJ_MEL_FAST_SLOTS ? maxseg : SJ_MEL_FAST_SLOTS) + 4) & ~3;
}
// per-warp shared memory in floats: the 2-slot hop ring (int16 hops need half a slot), the transpose plane, logs, partials
__host__ __device__ constexpr int warp_floats(int maxseg, bool pcm16)
{
    return (pcm16 ? SJ_HOP : 2 * SJ_HOP) + t_floats(pcm16) + kLogFloats + p_floats(maxseg);
}

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity)
        : "memory");
}
// bulk async copy global -> shared (TMA engine, no tensor map), completion counted on `bar`
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void fence_proxy_async()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
// log10(max(x, 2e-42)) (aubio fvec_log10 / SAFE_LOG10) on one MUFU.LG2: values below 1e-30 (the floor
// itself is subnormal) are scaled by 2^64 first.  Absolute error <= 2 ulp of log2(x) * 0.30103.
__device__ __forceinline__ float floor_log10(float x, float floor_v = 2.e-42f)
{
    x = fmaxf(x, floor_v);
    const bool tiny = x < 1e-30f;
    const float y = tiny ? x * 18446744073709551616.f : x;
    float l;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l) : "f"(y));
    return fmaf(l, 0.30102999566398120f, tiny ? -19.265919722494797f : 0.f);
}
__device__ __forceinline__ float fast_sqrt(float x)
{
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); // one MUFU.SQRT; |X| below 1e-19 flushes to 0
    return r;
}

// libsvm sigmoid_predict + multiclass_probability (k = 2) in double, as onnxruntime's SVMClassifier
// applies them when prob_a/prob_b are present.  s = ORT score (positive => class 0).
__device__ void platt_pair(double s, double A, double B, double p[2])
{
    double fApB = s * A + B;
    double r01 = fApB >= 0 ? exp(-fApB) / (1.0 + exp(-fApB)) : 1.0 / (1.0 + exp(fApB));
    r01 = fmin(fmax(r01, 1e-7), 1.0 - 1e-7);
    const double r10 = 1.0 - r01;
    double Q[2][2] = {{r10 * r10, -r10 * r01}, {-r10 * r01, r01 * r01}};
    p[0] = p[1] = 0.5;
    const double eps = 0.005 / 2;
    for (int iter = 0; iter < 100; ++iter) {
        double Qp[2] = {Q[0][0] * p[0] + Q[0][1] * p[1], Q[1][0] * p[0] + Q[1][1] * p[1]};
        double pQp = p[0] * Qp[0] + p[1] * Qp[1];
        double err = fmax(fabs(Qp[0] - pQp), fabs(Qp[1] - pQp));
        if (err < eps) break;
#pragma unroll
        for (int t = 0; t < 2; ++t) {
            double diff = (-Qp[t] + pQp) / Q[t][t];
            p[t] += diff;
            pQp = (pQp + diff * (diff * Q[t][t] + 2 * Qp[t])) / (1 + diff) / (1 + diff);
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                Qp[j] = (Qp[j] + diff * Q[t][j]) / (1 + diff);
                p[j] /= (1 + diff);
            }
        }
    }
}

} // namespace

namespace {

// only the per-frame tables are staged in shared memory; the SVM weights (used once per clip) are read
// through L1/L2 from the global copy, which leaves room for more resident warps
constexpr size_t kHotBytes = offsetof(SjTables, sv_off);
static_assert(kHotBytes % 16 == 0, "tables are staged with 16-byte copies");
constexpr unsigned kFull = 0xffffffffu;

struct SvmOut { int label; float conf, score; };

// 0.5 cos(2 pi m1/32) and 0.5 sin(2 pi m1/32), m1 < 16 (window rebuilt in registers, see the kernel)
__device__ constexpr float kWinC[16] = {0.5f, 0.49039264020161522f, 0.46193976625564337f, 0.41573480615127262f,
                                        0.35355339059327379f, 0.27778511650980114f, 0.19134171618254492f, 0.09754516100806417f,
                                        0.f, -0.09754516100806417f, -0.19134171618254492f, -0.27778511650980114f,
                                        -0.35355339059327379f, -0.41573480615127262f, -0.46193976625564337f, -0.49039264020161522f};
__device__ constexpr float kWinS[16] = {0.f, 0.09754516100806417f, 0.19134171618254492f, 0.27778511650980114f,
                                        0.35355339059327379f, 0.41573480615127262f, 0.46193976625564337f, 0.49039264020161522f,
                                        0.5f, 0.49039264020161522f, 0.46193976625564337f, 0.41573480615127262f,
                                        0.35355339059327379f, 0.27778511650980114f, 0.19134171618254492f, 0.09754516100806417f};

// Scaler + RBF SVM + Platt for one 40-d feature vector held as z[40] (already standardised) in
// shared memory.  All 32 lanes participate; the result is valid on every lane.
// W_64^j = (kUntC[j], -kUntS[j]), j < 16
__device__ constexpr float kUntC[16] = {1.f, 0.99518472667219693f, 0.98078528040323043f, 0.95694033573220882f,
                                        0.92387953251128674f, 0.88192126434835505f, 0.83146961230254524f, 0.77301045336273699f,
                                        0.70710678118654757f, 0.63439328416364549f, 0.55557023301960229f, 0.47139673682599781f,
                                        0.38268343236508984f, 0.29028467725446233f, 0.19509032201612833f, 0.09801714032956077f};
__device__ constexpr float kUntS[16] = {0.f, 0.09801714032956060f, 0.19509032201612825f, 0.29028467725446233f,
                                        0.38268343236508978f, 0.47139673682599764f, 0.55557023301960218f, 0.63439328416364549f,
                                        0.70710678118654746f, 0.77301045336273699f, 0.83146961230254524f, 0.88192126434835494f,
                                        0.92387953251128674f, 0.95694033573220894f, 0.98078528040323043f, 0.99518472667219682f};
__device__ __forceinline__ SvmOut svm_eval(const SjTables *T, const float *z, int lane)
{
    float d0 = 0.f, d1 = 0.f;
#pragma unroll 8
    for (int i = 0; i < SJ_FEAT; ++i) {
        float zi = z[i];
        float t0 = zi - T->sv[i][lane], t1 = zi - T->sv[i][lane + 32];
        d0 = fmaf(t0, t0, d0);
        d1 = fmaf(t1, t1, d1);
    }
    float part = T->sv_coef[lane] * expf(-T->gamma * d0) + T->sv_coef[lane + 32] * expf(-T->gamma * d1);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(kFull, part, o);
    const float s = part + T->rho; // onnxruntime score: positive => class 0
    SvmOut r;
    r.score = -s;                  // sklearn decision_function
    double p[2];
    platt_pair((double)s, (double)T->prob_a, (double)T->prob_b, p);
    // label rule at the onnxruntime boundary (scrubjay_cepstrum_params.label_rule); default = the rule
    // of cepstrum/run.py:46 (model.predict): decision_function > 0
    const int rule = T->label_rule;
    r.label = rule == SJ_LABEL_VOTE ? (s <= 0.f ? 1 : 0) : rule == SJ_LABEL_PROB_ARGMAX ? (p[1] > p[0] ? 1 : 0) : (r.score > 0.f ? 1 : 0);
    r.conf = (float)p[r.label];
    return r;
}

// window statistics (cepstrum/scrubjay_infer.c:61-67) -> Scaler -> SVM -> outputs at index `out`; `lg` is a
// per-warp shared scratch of >= 40 floats; all 32 lanes call, lanes < 20 hold the sums of their coefficient
__device__ __forceinline__ void sj_emit(const SjCepstrumArgs &a, float *lg, int lane, long out, double sum, double sq,
                                        int nframes)
{
    __syncwarp();
    if (lane < SJ_NCOEF) {
        const double mean = sum / (double)nframes;
        const double var = sq / (double)nframes - mean * mean;
        const float fm = (float)mean;
        const float fs = sqrtf((float)(var > 0 ? var : 0));
        if (a.feats) {
            a.feats[out * SJ_FEAT + lane] = fm;
            a.feats[out * SJ_FEAT + SJ_NCOEF + lane] = fs;
        }
        const SjTables *G = a.tables;
        lg[lane] = (fm - __ldg(&G->sv_off[lane])) * __ldg(&G->sv_scale[lane]);
        lg[lane + SJ_NCOEF] = (fs - __ldg(&G->sv_off[lane + SJ_NCOEF])) * __ldg(&G->sv_scale[lane + SJ_NCOEF]);
    }
    __syncwarp();
    if (a.do_svm) {
        const SvmOut r = svm_eval(a.tables, lg, lane);
        if (a.peers) { // fused gather: lane p stores the result into rank p's arrays (its own, or a peer's over NVLink)
            if (lane < a.peers->n_peers) {
                const long at = a.peers->out_base + out;
                a.peers->labels[lane][at] = r.label;
                float *ps = a.peers->score[lane];
                if (ps) ps[at] = r.score;
            }
            if (lane == 0 && a.conf) a.conf[out] = r.conf;
        } else if (lane == 0) {
            if (a.labels) a.labels[out] = r.label;
            if (a.conf) a.conf[out] = r.conf;
            if (a.score) a.score[out] = r.score;
        }
    }
    __syncwarp();
}

// FAST: the 16 kHz layout (28-bin lane chunks, static flush slots, <= 5 combine entries, mirror bins of
// the first 4 untangling rows skipped); otherwise every mel parameter is read from the tables.
// STREAM: the batch is a set of long channels cut into work items of `chunk_windows` analysis windows; the
// pvoc state (previous hop) runs on across windows and a ring of the last win_frames MFCC vectors feeds the
// per-window statistics.  Otherwise one work item = one clip = one window over all of its frames.
// PCM16: the clips are int16 PCM (the reference's on-disk format); a hop is 2 KB and the conversion
// x = s / 32768 (aubio_source / classifier.c:56-60, exact in float) happens in the load phase, so the
// results are bit-identical to the float path fed with the converted samples.
template <int WARPS, bool TMA, bool FAST, bool STREAM, bool PCM16>
__global__ void __launch_bounds__(WARPS * 32, 1) cepstrum_kernel(const SjCepstrumArgs a, const int maxseg)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    SjTables *T = reinterpret_cast<SjTables *>(smem_raw);
    size_t off = (kHotBytes + 127) & ~size_t(127);
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem_raw + off) + 2 * warp;
    off += (size_t(WARPS) * 2 * 8 + 127) & ~size_t(127);
    constexpr int kRingFloats = STREAM ? SJ_STREAM_MAX_WIN * SJ_NCOEF : 0;
    float *wbase = reinterpret_cast<float *>(smem_raw + off) + size_t(warp) * (warp_floats(maxseg, PCM16) + kRingFloats);
    constexpr int kSlot = PCM16 ? SJ_HOP / 2 : SJ_HOP; // floats per ring slot
    // [slot 0 | plane | slot 1], so that the plane and whichever slot holds the PREVIOUS hop are contiguous and together
    // hold the 32 x 68 floats of the interleaved transpose (see below)
    constexpr bool kWide = SJ_WIDE_TRANSPOSE;
    constexpr int kPlane = t_floats(PCM16);
    float *tb = kWide ? wbase + kSlot : wbase + 2 * kSlot; // transpose plane, later |X|
    float *lg = wbase + 2 * kSlot + kPlane;                 // log energies / standardised features
    auto slot_ptr = [&](int slot) -> float * { return kWide ? (slot ? wbase + kSlot + kPlane : wbase) : wbase + slot * kSlot; };
    float *P = lg + kLogFloats;       // mel partial sums
    float *mring = P + p_floats(maxseg); // STREAM: the last win_frames MFCC vectors, [frame % W][coef]

    {   // stage the tables
        const int4 *src = reinterpret_cast<const int4 *>(a.tables);
        int4 *dst = reinterpret_cast<int4 *>(T);
        for (int i = tid; i < int(kHotBytes / 16); i += WARPS * 32) dst[i] = __ldg(src + i);
    }
    if (lane == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = lane; i < p_floats(maxseg); i += 32) P[i] = 0.f; // includes the combine lists' zero slot
    __syncthreads();

    const int n_samples = a.n_samples;
    const int F = (n_samples + SJ_HOP - 1) / SJ_HOP; // hops of one clip / one channel
    const int last_valid = n_samples - (F - 1) * SJ_HOP;
    const long stride = long(gridDim.x) * WARPS;
    const long n_items = STREAM ? a.n_clips * a.chunks_per_chan : a.n_clips;
    // items are dealt to CTAs first and to the warps of a CTA second, so that a batch smaller than the resident
    // warp count spreads over the SMs instead of filling a few of them (latency of small batches: one warp alone on
    // an SM walks a clip in 42 us, twelve sharing one SM take 72 us); for large batches this is only a permutation
    long item = long(warp) * gridDim.x + blockIdx.x;
    const int W = a.win_frames, S = a.step_frames;
    // hann[64 m1 + 2 lane + d] = 0.5 - 0.5 cos(a + b), a = 2 pi m1/32, b = pi (2 lane + d)/1024
    float2 sb, cb;
    sincospif((float)(2 * lane) * (1.f / 1024.f), &sb.x, &cb.x);
    sincospif((float)(2 * lane + 1) * (1.f / 1024.f), &sb.y, &cb.y);
    const float4 *tw4 = reinterpret_cast<const float4 *>(T->tw4) + lane;
    // untangling base u = -i W_2048^lane with its quarter turn ur = i u: W_2048^(lane + 32 j) = W_2048^lane W_64^j, and the
    // W_64^j are immediates (this replaced a 4 KB table: 32 shared-memory wavefronts per frame)
    const float2 ut = make_float2(__ldg(&a.tables->tp4[0][lane].y), -__ldg(&a.tables->tp4[0][lane].x)), ur = make_float2(-ut.y, ut.x);
    const float4 *mw4 = reinterpret_cast<const float4 *>(T->melw4) + lane;
    const int mel_ch = FAST ? SJ_MEL_FAST_CH : T->mel_ch, ncomb = FAST ? 5 : T->mel_ncomb;
    const float4 *X4 = reinterpret_cast<const float4 *>(tb + T->mel_lo + lane * mel_ch);
    const int partner = (32 - lane) & 31;
    constexpr int PSKIP = FAST ? 4 : 0;
    // generic instantiation only: the switches of scrubjay_cepstrum_params (warp-uniform)
    const bool win_tab = !FAST && (T->gen_flags & SJ_GEN_WINDOW_TABLE), pw = !FAST && (T->gen_flags & SJ_GEN_POWER);
    const float log_floor = FAST ? 2.e-42f : T->log_floor;
    const float2 *wtab = reinterpret_cast<const float2 *>(a.tables->window) + lane; // pair (2m, 2m+1), m = 32 m1 + lane
    uint32_t cmb[5]; // fast layout: this lane's combine entries live in registers
#pragma unroll
    for (int e = 0; e < 5; ++e) cmb[e] = T->comb2[e][lane];

    // A work item walks `nh` consecutive hops of one clip / channel starting at hop `hop0` (-1 = the caller's
    // carried-over hop); its first `warm` hops are only loaded (they are the previous hop of the first frame).
    constexpr int ES = PCM16 ? 2 : 4; // bytes per sample
    struct Seg { const unsigned char *base, *prev; long out0; int hop0, nh, warm, f0; };
    const unsigned char *clips_b = reinterpret_cast<const unsigned char *>(a.clips);
    auto seg_of = [&](long it) -> Seg {
        Seg g;
        if (!STREAM) {
            g.base = clips_b + it * (long)n_samples * ES; g.prev = nullptr; g.out0 = it; g.hop0 = 0; g.nh = F; g.warm = 0; g.f0 = 0;
        } else {
            const long c = it / a.chunks_per_chan;
            const long j0 = (it - c * a.chunks_per_chan) * a.chunk_windows;
            const long j1 = j0 + a.chunk_windows < a.n_windows ? j0 + a.chunk_windows : a.n_windows;
            g.base = clips_b + c * a.chan_stride * ES;
            g.prev = a.prev_hop ? reinterpret_cast<const unsigned char *>(a.prev_hop) + c * SJ_HOP * ES : nullptr;
            g.out0 = c * a.n_windows + j0;
            g.f0 = (int)(j0 * S);
            g.warm = (g.f0 > 0 || g.prev) ? 1 : 0;
            g.hop0 = g.f0 - g.warm;
            g.nh = (int)((j1 - 1) * S) + W - g.hop0;
        }
        return g;
    };
    // fetch hop `h` of the item into ring slot `slot`
    auto fetch = [&](const Seg &g, int h, int slot) {
        const unsigned char *src = h < 0 ? g.prev : g.base + (long)h * SJ_HOP * ES;
        const int cnt = (h == F - 1) ? last_valid : SJ_HOP;
        float *dst = slot_ptr(slot);
        if (TMA) {
            if (lane == 0) {
                fence_proxy_async();
                mbar_expect_tx(&bars[slot], (uint32_t)cnt * ES);
                bulk_g2s(dst, src, (uint32_t)cnt * ES, &bars[slot]);
            }
        } else {
            if (PCM16)
                for (int i = lane; i < cnt; i += 32)
                    reinterpret_cast<int16_t *>(dst)[i] = __ldg(reinterpret_cast<const int16_t *>(src) + i);
            else
                for (int i = lane; i < cnt; i += 32) dst[i] = __ldg(reinterpret_cast<const float *>(src) + i);
            __syncwarp();
        }
    };
    auto emit = [&](long out, double sum, double sq, int nframes) { sj_emit(a, lg, lane, out, sum, sq, nframes); };

    uint32_t q = 0; // running hop number of this warp: ring slot q&1, mbarrier parity (q>>1)&1
    Seg sg = seg_of(item < n_items ? item : 0);
    if (item < n_items) fetch(sg, sg.hop0, 0);

    for (; item < n_items; item += stride) {
        double sum = 0.0, sq = 0.0;
        const int nh = sg.nh;
        Seg nxt = sg;
        int ring_at = 0, until_emit = W - 1, win_at = 0; // STREAM bookkeeping, see below
        for (int hr = 0; hr < nh; ++hr, ++q) {
            const int slot = q & 1, ha = sg.hop0 + hr;
            if (TMA) mbar_wait(&bars[slot], (q >> 1) & 1);
            float *curf = slot_ptr(slot);
            if (ha == F - 1 && last_valid < SJ_HOP) { // zero-padded last read (aubio_source_do)
                if (PCM16)
                    for (int i = last_valid + lane; i < SJ_HOP; i += 32) reinterpret_cast<int16_t *>(curf)[i] = 0;
                else
                    for (int i = last_valid + lane; i < SJ_HOP; i += 32) curf[i] = 0.f;
                __syncwarp();
            }
            if (STREAM && hr < sg.warm) { // previous hop of the item's first frame: nothing to compute
                fetch(sg, ha + 1, slot ^ 1);
                continue;
            }
            // sample pair (2m, 2m+1), m = 32*m1 + lane, of the current / previous hop as floats
            const float *prvf = slot_ptr(slot ^ 1);
            auto pair_at = [&](const float *hop, int m1) -> float2 {
                if (PCM16) {
                    // int16 -> float without the conversion pipe (I2F.S16 runs at a quarter of the FMA rate): the
                    // offset-binary sample s + 32768 becomes the mantissa of 2^23, and one packed FMA removes the
                    // bias and scales: (2^23 + 32768 + s) * 2^-15 - 257 = s / 32768, every step exact
                    const uint32_t w = reinterpret_cast<const uint32_t *>(hop)[32 * m1 + lane] ^ 0x80008000u;
                    const float2 f = make_float2(__uint_as_float(__byte_perm(w, 0x4B000000u, 0x7610)),
                                                 __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7632)));
                    return pk_fma(f, pk_bc(1.f / 32768.f), pk_bc(-257.f));
                }
                return reinterpret_cast<const float2 *>(hop)[32 * m1 + lane];
            };
            float2 v[32];
            // The window is rebuilt from the angle sum 2 pi (64 m1 + 2 lane + d)/2048: 0.5 cos/sin of the m1 part
            // are immediates, cos/sin of the lane part sit in four registers -- two FMAs per pair of samples
            // instead of a table read (shared-memory bandwidth binds this kernel, not issue).
            if (win_tab) { // any window, from the table: lower half on the previous hop, upper half on the current one
#pragma unroll
                for (int m1 = 0; m1 < 16; ++m1) {
                    const float2 c = pair_at(curf, m1), s = hr == 0 ? make_float2(0.f, 0.f) : pair_at(prvf, m1);
                    v[m1] = pk_mul(s, __ldg(wtab + 32 * m1));
                    v[m1 + 16] = pk_mul(c, __ldg(wtab + 512 + 32 * m1));
                }
            } else if (hr == 0) { // no previous hop: aubio_pvoc starts from zeros
#pragma unroll
                for (int m1 = 0; m1 < 16; ++m1) {
                    const float2 c = pair_at(curf, m1);
                    const float2 w = pk_fma(pk_bc(kWinS[m1]), sb, pk_fma(pk_bc(-kWinC[m1]), cb, pk_bc(0.5f)));
                    v[m1] = make_float2(0.f, 0.f);
                    v[m1 + 16] = pk_fma(pk_neg(c), w, c);
                }
            } else {
#pragma unroll
                for (int m1 = 0; m1 < 16; ++m1) {
                    const float2 c = pair_at(curf, m1), s = pair_at(prvf, m1);
                    const float2 w = pk_fma(pk_bc(kWinS[m1]), sb, pk_fma(pk_bc(-kWinC[m1]), cb, pk_bc(0.5f)));
                    v[m1] = pk_mul(s, w);
                    v[m1 + 16] = pk_fma(pk_neg(c), w, c);
                }
            }
            __syncwarp();
            // prefetch the next hop (of this item, else the first hop of the warp's next item) into the slot that
            // just became free
            auto prefetch_next = [&]() {
                if (hr + 1 < nh) fetch(sg, ha + 1, slot ^ 1);
                else if (item + stride < n_items) {
                    nxt = seg_of(item + stride);
                    fetch(nxt, nxt.hop0, slot ^ 1);
                }
            };
            if (!kWide) prefetch_next();
            // ---- 1024-point complex FFT, 32 x 32 split
            sj_fft32(v); // over m1: Y[k1] in v[SJ_BREV32(k1)]
            if (kWide) {
                // transpose (lane = m2, reg = k1) -> (lane = k1, reg = m2) with (re, im) interleaved: 32 64-bit stores
                // (256 contiguous bytes each) and 16 conflict-free 128-bit row loads (row stride 68 floats = 17 x 16 B)
                // that deliver the register pairs of the packed arithmetic as they are -- the same 128 wavefronts as two
                // float planes but half the store instructions and one round trip instead of two.  The 32 x 68 floats
                // span the plane and the slot of the previous hop, which is dead by now; the next hop is requested
                // into that slot afterwards.
                float *pl = (slot ^ 1) ? tb : wbase;
                float2 *pl2 = reinterpret_cast<float2 *>(pl);
#pragma unroll
                for (int k1 = 0; k1 < 32; ++k1) pl2[k1 * 34 + lane] = v[SJ_BREV32(k1)];
                __syncwarp();
                const float4 *row = reinterpret_cast<const float4 *>(pl + lane * 68);
#pragma unroll
                for (int g = 0; g < 16; ++g) {
                    const float4 q4 = row[g];
                    v[2 * g] = make_float2(q4.x, q4.y);
                    v[2 * g + 1] = make_float2(q4.z, q4.w);
                }
                __syncwarp();
                prefetch_next();
            } else
            // transpose (lane = m2, reg = k1) -> (lane = k1, reg = m2), one float plane at a time
            {
                const float4 *row = reinterpret_cast<const float4 *>(tb + lane * kTStride);
#pragma unroll
                for (int k1 = 0; k1 < 32; ++k1) tb[k1 * kTStride + lane] = v[SJ_BREV32(k1)].x;
                __syncwarp();
                float4 r[8];
#pragma unroll
                for (int g = 0; g < 8; ++g) r[g] = row[g];
                __syncwarp();
#pragma unroll
                for (int k1 = 0; k1 < 32; ++k1) tb[k1 * kTStride + lane] = v[SJ_BREV32(k1)].y;
                __syncwarp();
#pragma unroll
                for (int g = 0; g < 8; ++g) {
                    const float4 i4 = row[g];
                    v[4 * g + 0] = make_float2(r[g].x, i4.x);
                    v[4 * g + 1] = make_float2(r[g].y, i4.y);
                    v[4 * g + 2] = make_float2(r[g].z, i4.z);
                    v[4 * g + 3] = make_float2(r[g].w, i4.w);
                }
                __syncwarp();
            }
            sj_fft32tw(v, tw4); // twiddles W^(lane*m2), then over m2: Z[lane + 32*k2] in v[SJ_BREV32(k2)]

            // ---- untangle the real FFT: bins k = lane + 32 j and 1024 - k, j < 16 (values are 2|X|)
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const float2 A = v[SJ_BREV32(j)];
                const float2 S = v[SJ_BREV32(31 - j)];
                float bx = __shfl_sync(kFull, S.x, partner), by = __shfl_sync(kFull, S.y, partner);
                if (lane == 0) { bx = v[SJ_BREV32((32 - j) & 31)].x; by = v[SJ_BREV32((32 - j) & 31)].y; }
                const float2 cbar = make_float2(bx, -by);
                const float2 e = pk_add(A, cbar);
                float2 d = pk_sub(A, cbar);
                // t = -i W_2048^(lane + 32 j) d = u (W_64^j d)
                if (j) d = pk_fma(pk_bc(kUntS[j]), make_float2(d.y, -d.x), pk_mul(pk_bc(kUntC[j]), d));
                const float2 t = pk_fma(pk_bc(d.y), ur, pk_mul(pk_bc(d.x), ut));
                if (j >= PSKIP) { // (bin k, mirror bin 1024 - k) side by side: e +- t
                    const float2 X = pk_fma(pk_bc(t.x), make_float2(1.f, -1.f), pk_bc(e.x));
                    const float2 Y = pk_fma(pk_bc(t.y), make_float2(1.f, -1.f), pk_bc(e.y));
                    const float2 m = pk_fma(Y, Y, pk_mul(X, X));
                    tb[lane + 32 * j] = pw ? m.x : fast_sqrt(m.x); // (2|X|)^2 for the power filterbank
                    tb[1024 - lane - 32 * j] = pw ? m.y : fast_sqrt(m.y);
                } else { // the mirrors of the first rows lie above the filterbank
                    const float2 xa = pk_add(e, t);
                    tb[lane + 32 * j] = fast_sqrt(fmaf(xa.x, xa.x, xa.y * xa.y));
                }
            }
            if (lane == 0) {
                const float2 c = v[SJ_BREV32(16)];
                const float m = fmaf(c.x, c.x, c.y * c.y);
                tb[512] = pw ? 4.f * m : 2.f * fast_sqrt(m);
            }
            __syncwarp();

#ifdef SJ_EXP_SKIP_MEL /* timing experiment only (results are wrong): the FFT half of a frame alone */
            sum += (double)(tb[lane] * 1e-9f); sq += 1.0;
            if (false)
#endif
            // ---- sparse Slaney mel sums over this lane's bin chunk: two running sums (filter f, f+1);
            // where f steps, the finished sum is flushed and the f+1 sum moves down
            {
                float lo = 0.f, hi = 0.f;
                if (FAST) {
                    float *Pl = P + lane;
                    // a chunk has at most one step per aligned group of 8 bins: the flushed sum is kept in a
                    // register and stored once per group (a predicated store per bin costs a shared-memory
                    // wavefront whenever any lane steps there; stale values in groups without a step are never read)
                    float fl = 0.f;
#pragma unroll
                    for (int b = 0; b < SJ_MEL_FAST_CH / 4; ++b) {
                        const float4 x = X4[b], wa = mw4[64 * b], wb = mw4[64 * b + 32];
                        auto bin = [&](float av, float wx, float wy) {
                            const bool step = __float_as_int(wx) < 0;
                            fl = step ? lo : fl;
                            lo = fmaf(fabsf(wx), av, step ? hi : lo);
                            hi = fmaf(wy, av, step ? 0.f : hi);
                        };
                        bin(x.x, wa.x, wa.y);
                        bin(x.y, wa.z, wa.w);
                        bin(x.z, wb.x, wb.y);
                        bin(x.w, wb.z, wb.w);
                        if ((b & 1) || b == SJ_MEL_FAST_CH / 4 - 1) Pl[(b >> 1) * 32] = fl;
                    }
                    Pl[(SJ_MEL_FAST_SLOTS - 2) * 32] = lo;
                    Pl[(SJ_MEL_FAST_SLOTS - 1) * 32] = hi;
                } else {
                    float *Pw = P + lane * maxseg;
                    for (int b = 0; b < mel_ch / 4; ++b) {
                        const float4 x = X4[b], wa = mw4[64 * b], wb = mw4[64 * b + 32];
                        auto bin = [&](float av, float wx, float wy) {
                            const bool step = __float_as_int(wx) < 0;
                            if (step) *Pw = lo;
                            Pw += step;
                            lo = fmaf(fabsf(wx), av, step ? hi : lo);
                            hi = fmaf(wy, av, step ? 0.f : hi);
                        };
                        bin(x.x, wa.x, wa.y);
                        bin(x.y, wa.z, wa.w);
                        bin(x.z, wb.x, wb.y);
                        bin(x.w, wb.z, wb.w);
                    }
                    Pw[0] = lo;
                    Pw[1] = hi;
                }
            }
            __syncwarp();
#ifdef SJ_EXP_SKIP_MEL
            if (false)
#endif
            {
                float e0 = 0.f, e1 = 0.f;
                if (FAST) {
#pragma unroll
                    for (int e = 0; e < 5; ++e) {
                        e0 += P[cmb[e] & 0xffffu];
                        e1 += P[cmb[e] >> 16];
                    }
                } else {
                    for (int e = 0; e < ncomb; ++e) {
                        const uint32_t c = T->comb2[e][lane];
                        e0 += P[c & 0xffffu];
                        e1 += P[c >> 16];
                    }
                }
                // aubio fvec_log10: log10(max(x, 2e-42)); lane m holds band m, lanes 0..7 also band m+32
                e0 = floor_log10(e0, log_floor);
                e1 = floor_log10(e1, log_floor);
                // DCT-II symmetry D[c][39-m] = (-1)^c D[c][m]: fold the 40 log energies into 20 sums
                // (even c) and 20 differences (odd c)
                const float far0 = __shfl_sync(kFull, e0, (39 - lane) & 31); // band 39-m for m = 8..19
                const float far1 = __shfl_sync(kFull, e1, (7 - lane) & 31);  // band 39-m for m = 0..7
                const float far = lane < 8 ? far1 : far0;
                // clip mode runs the DCT on two frames at a time (one read of the matrix): odd frames use row 1
                float *lgw = lg + ((!STREAM && (hr & 1)) ? SJ_NFILT : 0);
                if (lane < SJ_NCOEF) {
                    lgw[lane] = e0 + far;
                    lgw[SJ_NCOEF + lane] = e0 - far;
                }
            }
            __syncwarp();
            // ---- DCT-II (first 20 rows) and the reference's double accumulators
            if (!STREAM) {
#ifdef SJ_EXP_SKIP_MEL
                if (false)
#endif
                if ((hr & 1) || hr == nh - 1) {
                    float acc0 = 0.f, acc1 = 0.f;
                    // folded log energies as 128-bit loads (two addresses per warp: even / odd coefficient lanes)
                    const float4 *eo = reinterpret_cast<const float4 *>(lg + (lane & 1) * SJ_NCOEF);
#pragma unroll
                    for (int q = 0; q < SJ_NFILT / 8; ++q) {
                        const float4 ea = eo[q], eb = eo[SJ_NFILT / 4 + q];
                        const float d0 = T->dct[4 * q][lane], d1 = T->dct[4 * q + 1][lane];
                        const float d2 = T->dct[4 * q + 2][lane], d3 = T->dct[4 * q + 3][lane];
                        acc0 = fmaf(d0, ea.x, acc0); acc1 = fmaf(d0, eb.x, acc1);
                        acc0 = fmaf(d1, ea.y, acc0); acc1 = fmaf(d1, eb.y, acc1);
                        acc0 = fmaf(d2, ea.z, acc0); acc1 = fmaf(d2, eb.z, acc1);
                        acc0 = fmaf(d3, ea.w, acc0); acc1 = fmaf(d3, eb.w, acc1);
                    }
                    const double d0 = (double)acc0; // frame hr-1 (or the unpaired last frame), then frame hr
                    sum += d0;
                    sq += d0 * d0;
                    if (hr & 1) {
                        const double d1 = (double)acc1;
                        sum += d1;
                        sq += d1 * d1;
                    }
                }
            } else {
                float acc = 0.f;
                const float4 *eo = reinterpret_cast<const float4 *>(lg + (lane & 1) * SJ_NCOEF);
#pragma unroll
                for (int q = 0; q < SJ_NFILT / 8; ++q) {
                    const float4 ea = eo[q];
                    acc = fmaf(T->dct[4 * q][lane], ea.x, acc);
                    acc = fmaf(T->dct[4 * q + 1][lane], ea.y, acc);
                    acc = fmaf(T->dct[4 * q + 2][lane], ea.z, acc);
                    acc = fmaf(T->dct[4 * q + 3][lane], ea.w, acc);
                }
                {
                    // frame `fr` of the item goes to ring row fr % W; window j ends at frame j*S + W-1 (both kept
                    // as running counters: ring_at = row to write, until_emit = frames before the next window end)
                    if (lane < SJ_NCOEF) mring[ring_at * SJ_NCOEF + lane] = acc;
                    ring_at = ring_at + 1 == W ? 0 : ring_at + 1; // now the oldest row
                    __syncwarp();
                    if (until_emit == 0) {
                        double ws = 0.0, wq = 0.0;
                        if (lane < SJ_NCOEF) {
                            int row = ring_at;
                            for (int i = 0; i < W; ++i) { // oldest frame first, like the reference's running sums
                                const double dv = (double)mring[row * SJ_NCOEF + lane];
                                ws += dv;
                                wq += dv * dv;
                                row = row + 1 == W ? 0 : row + 1;
                            }
                        }
                        emit(sg.out0 + win_at, ws, wq, W);
                        ++win_at;
                        until_emit = S;
                    }
                    --until_emit;
                }
            }
        }
        if (!STREAM) emit(sg.out0, sum, sq, F);
        sg = nxt;
    }
}

// standalone batched ort_predict: one warp per feature vector
__global__ void __launch_bounds__(256) svm_kernel(const float *feats, long n, const SjTables *tables,
                                                  int32_t *labels, float *conf, float *score)
{
    __shared__ float z[8][SJ_FEAT];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (long i = (long)blockIdx.x * 8 + warp; i < n; i += (long)gridDim.x * 8) {
        for (int c = lane; c < SJ_FEAT; c += 32)
            z[warp][c] = (feats[i * SJ_FEAT + c] - tables->sv_off[c]) * tables->sv_scale[c];
        __syncwarp();
        const SvmOut r = svm_eval(tables, z[warp], lane);
        if (lane == 0) {
            labels[i] = r.label;
            if (conf) conf[i] = r.conf;
            if (score) score[i] = r.score;
        }
        __syncwarp();
    }
}

// Second half of the frame-parallel clip mode (small batches): frames[clip][f][40] holds the MFCC vector of every
// frame in its first 20 floats (written by the STREAM instantiation run with one-frame windows); one warp per clip adds
// them up in frame order in double, exactly like the per-clip loop, and finishes with the same sj_emit.
__global__ void __launch_bounds__(128) clip_stats_kernel(const SjCepstrumArgs a, const float *frames, int F)
{
    __shared__ float z[4][SJ_FEAT];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (long i = (long)blockIdx.x * 4 + warp; i < a.n_clips; i += (long)gridDim.x * 4) {
        double sum = 0.0, sq = 0.0;
        if (lane < SJ_NCOEF) {
            const float *p = frames + i * F * SJ_FEAT + lane;
            for (int f = 0; f < F; ++f) {
                const double dv = (double)p[(long)f * SJ_FEAT];
                sum += dv;
                sq += dv * dv;
            }
        }
        sj_emit(a, z[warp], lane, i, sum, sq, F);
    }
}

// ---- allocation-free variant (scrubjay_infer_batch_noalloc_dev; no counterpart in the reference).
// The frame scratch [clip][frame][40] of the frame-parallel clip mode lives in a file-scope
// static array instead of a cudaMalloc'd buffer: SJ_NOALLOC_MAX_FRAMES frames x 40 floats = 2.5 MiB that every GPU
// which loads the library holds until exit.  Every call on a device shares the array, and it keeps the last call's
// contents; nothing in it has to start at zero (the first kernel writes all 40 floats of every (clip, frame) entry the
// second kernel reads), but two calls that overlap on one GPU race on it.
static __device__ __align__(256) float g_noalloc_frames[size_t(SJ_NOALLOC_MAX_FRAMES) * SJ_FEAT];

// clip_stats_kernel with the frames read from the static array, named directly by device code
__global__ void __launch_bounds__(128) clip_stats_static_kernel(const SjCepstrumArgs a, int F)
{
    __shared__ float z[4][SJ_FEAT];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (long i = (long)blockIdx.x * 4 + warp; i < a.n_clips; i += (long)gridDim.x * 4) {
        double sum = 0.0, sq = 0.0;
        if (lane < SJ_NCOEF) {
            const float *p = g_noalloc_frames + i * F * SJ_FEAT + lane;
            for (int f = 0; f < F; ++f) {
                const double dv = (double)p[(long)f * SJ_FEAT];
                sum += dv;
                sq += dv * dv;
            }
        }
        sj_emit(a, z[warp], lane, i, sum, sq, F);
    }
}

template <int WARPS> size_t smem_bytes(int maxseg, bool stream, bool pcm16)
{
    size_t off = (kHotBytes + 127) & ~size_t(127);
    off += (size_t(WARPS) * 2 * 8 + 127) & ~size_t(127);
    return off + size_t(WARPS) * (warp_floats(maxseg, pcm16) + (stream ? SJ_STREAM_MAX_WIN * SJ_NCOEF : 0)) * sizeof(float);
}

// warps per CTA (one CTA per SM): what 227 KB of shared memory and 64 K registers admit per variant
#ifndef SJ_WARPS
#define SJ_WARPS 12
#endif
#ifndef SJ_WARPS_PCM16
#define SJ_WARPS_PCM16 SJ_WARPS
#endif
#ifndef SJ_WARPS_STREAM
#define SJ_WARPS_STREAM 12
#endif
template <bool STREAM, bool PCM16> constexpr int warps_of() { return STREAM ? SJ_WARPS_STREAM : PCM16 ? SJ_WARPS_PCM16 : SJ_WARPS; }

template <bool TMA, bool FAST, bool STREAM, bool PCM16> cudaError_t setup_one(int maxseg)
{
    constexpr int W = warps_of<STREAM, PCM16>();
    return cudaFuncSetAttribute(cepstrum_kernel<W, TMA, FAST, STREAM, PCM16>,
                                cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes<W>(maxseg, STREAM, PCM16));
}

template <bool TMA, bool FAST, bool STREAM, bool PCM16>
cudaError_t launch_one(const SjCepstrumArgs &a, int maxseg, int num_sms, cudaStream_t stream)
{
    constexpr int W = warps_of<STREAM, PCM16>();
    const long items = STREAM ? a.n_clips * a.chunks_per_chan : a.n_clips;
    long ctas = items < num_sms ? items : num_sms;
    cepstrum_kernel<W, TMA, FAST, STREAM, PCM16>
        <<<(unsigned)ctas, W * 32, smem_bytes<W>(maxseg, STREAM, PCM16), stream>>>(a, maxseg);
    return cudaGetLastError();
}

template <bool STREAM, bool PCM16>
cudaError_t launch_any(const SjCepstrumArgs &a, const SjTables *h, int num_sms, cudaStream_t stream)
{
    // the bulk-copy path needs 16-byte aligned hop addresses and sizes
    const int per16 = PCM16 ? 8 : 4; // samples per 16 bytes
    bool tma = (a.n_samples % per16 == 0) && ((reinterpret_cast<uintptr_t>(a.clips) & 15) == 0);
    if (STREAM) tma = tma && (a.chan_stride % per16 == 0) && ((reinterpret_cast<uintptr_t>(a.prev_hop) & 15) == 0);
    // the fast instantiation hard-codes the 16 kHz layout: 28-bin chunks, static flush slots, <= 5
    // combine entries, the mirrors of the first 4 untangling rows above the bank
    const bool fast = h->mel_fast && h->mel_ch == SJ_MEL_FAST_CH && h->mel_ncomb <= 5 && h->mel_pskip >= 4 &&
                      !h->gen_flags && h->log_floor == 2.e-42f;
    const int maxseg = h->mel_maxseg;
    if (fast) return tma ? launch_one<true, true, STREAM, PCM16>(a, maxseg, num_sms, stream)
                         : launch_one<false, true, STREAM, PCM16>(a, maxseg, num_sms, stream);
    return tma ? launch_one<true, false, STREAM, PCM16>(a, maxseg, num_sms, stream)
               : launch_one<false, false, STREAM, PCM16>(a, maxseg, num_sms, stream);
}

template <bool STREAM, bool PCM16> cudaError_t setup_four(int maxseg)
{
    cudaError_t e = setup_one<true, true, STREAM, PCM16>(maxseg);
    if (e == cudaSuccess) e = setup_one<false, true, STREAM, PCM16>(maxseg);
    if (e == cudaSuccess) e = setup_one<true, false, STREAM, PCM16>(maxseg);
    if (e == cudaSuccess) e = setup_one<false, false, STREAM, PCM16>(maxseg);
    return e;
}

} // namespace

cudaError_t sj_cepstrum_setup(int maxseg)
{
    cudaError_t e = setup_four<false, false>(maxseg);
    if (e == cudaSuccess) e = setup_four<true, false>(maxseg);
    if (e == cudaSuccess) e = setup_four<false, true>(maxseg);
    if (e == cudaSuccess) e = setup_four<true, true>(maxseg);
    return e;
}

cudaError_t sj_launch_cepstrum(const SjCepstrumArgs &a, const SjTables *h, int num_sms, cudaStream_t stream)
{
    if (a.n_clips <= 0) return cudaSuccess;
    if (a.pcm16) return a.n_windows > 0 ? launch_any<true, true>(a, h, num_sms, stream) : launch_any<false, true>(a, h, num_sms, stream);
    if (a.n_windows > 0) return launch_any<true, false>(a, h, num_sms, stream);
    return launch_any<false, false>(a, h, num_sms, stream);
}

cudaError_t sj_launch_clip_stats(const SjCepstrumArgs &a, const float *frames, int n_frames, cudaStream_t stream)
{
    if (a.n_clips <= 0) return cudaSuccess;
    long blocks = (a.n_clips + 3) / 4;
    if (blocks > 148 * 8) blocks = 148 * 8;
    clip_stats_kernel<<<(unsigned)blocks, 128, 0, stream>>>(a, frames, n_frames);
    return cudaGetLastError();
}

// Device address of the static frame scratch on `device`, which the caller has made current.  In host code the array's
// name is only a host-side stand-in, not a device address: the address comes from cudaGetSymbolAddress, looked up once
// per device (set-up) and cached.  No allocation.
cudaError_t sj_noalloc_frame_scratch(int device, float **ptr)
{
    constexpr int kMaxDev = 64;
    static float *cache[kMaxDev] = {nullptr};
    if (!ptr || device < 0 || device >= kMaxDev) return cudaErrorInvalidDevice;
    if (!cache[device]) {
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_frames);
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        cache[device] = static_cast<float *>(p);
    }
    *ptr = cache[device];
    return cudaSuccess;
}

cudaError_t sj_launch_clip_stats_static(const SjCepstrumArgs &a, int n_frames, cudaStream_t stream)
{
    if (a.n_clips <= 0) return cudaSuccess;
    long blocks = (a.n_clips + 3) / 4;
    if (blocks > 148 * 8) blocks = 148 * 8;
    clip_stats_static_kernel<<<(unsigned)blocks, 128, 0, stream>>>(a, n_frames);
    return cudaGetLastError();
}

int sj_cepstrum_resident_warps(int num_sms) { return num_sms * warps_of<true, false>(); }

cudaError_t sj_launch_svm(const float *feats, long n, const SjTables *tables, int32_t *labels, float *conf,
                          float *score, cudaStream_t stream)
{
    if (n <= 0) return cudaSuccess;
    long blocks = (n + 7) / 8;
    if (blocks > 148 * 8) blocks = 148 * 8;
    svm_kernel<<<(unsigned)blocks, 256, 0, stream>>>(feats, n, tables, labels, conf, score);
    return cudaGetLastError();
}
