// cepstrum_kernel.cu -- fused MFCC-statistics + RBF-SVM kernel for sm_100a.
//
// Replaces, for a whole batch of clips in one launch, the per-file hot loop of the reference:
// mfcc_stats (cepstrum/scrubjay_infer.c:19-77: aubio_pvoc_do + aubio_mfcc_do per hop, mean/std over
// hops) followed by ort_predict (:105-141: ONNX Scaler + SVMClassifier).
//
// Mapping: persistent grid, one CTA per SM, each WARP owns whole clips (clip = warp id + k*stride)
// and walks their hops in order, so no block-level synchronisation exists after the prologue.
//   * HBM -> shared: every hop (1024 samples, 4 KB) is fetched exactly once by a bulk async copy
//     (cp.async.bulk, the TMA engine) into a 2-slot per-warp ring, completion on an mbarrier; hop
//     h+1 is in flight while frame h is computed.  Each sample is used by two frames (50 % overlap)
//     but read from HBM once.
//   * frame = [hop h-1 | hop h] * periodic Hann -> 2048-point real FFT done as a 1024-point complex
//     FFT split 32 x 32: a 32-point FFT in registers per lane, twiddle, ONE transpose through
//     shared memory, a second 32-point register FFT, then the real-FFT untangling with the partner
//     bin fetched by one warp shuffle.  aubio's fvec_shift (half swap) only flips the sign of odd
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

#include "fft_gen.inc"
#include "sj_tables.h"
#include "cepstrum_kernel.h"

namespace {

constexpr int kTStride = 33;                 // padded row stride of the transpose buffer
constexpr int kTFloats = 32 * kTStride;      // 1056 floats: transpose plane, then |X| bins 0..1024
constexpr int kLogFloats = 64;               // 40 log-energies (+pad)

__host__ __device__ constexpr int p_floats(int maxseg) { return ((32 * maxseg + 2) + 3) & ~3; }
__host__ __device__ constexpr int warp_floats(int maxseg)
{
    return 2 * SJ_HOP + kTFloats + kLogFloats + p_floats(maxseg);
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

// Scaler + RBF SVM + Platt for one 40-d feature vector held as z[40] (already standardised) in
// shared memory.  All 32 lanes participate; the result is valid on every lane.
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
    r.label = r.score > 0.f ? 1 : 0; // rule of cepstrum/run.py:46 (model.predict)
    double p[2];
    platt_pair((double)s, (double)T->prob_a, (double)T->prob_b, p);
    r.conf = (float)p[r.label];
    return r;
}

template <int WARPS, bool TMA, int MELCH>
__global__ void __launch_bounds__(WARPS * 32, 1) cepstrum_kernel(const SjCepstrumArgs a, const int maxseg)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    SjTables *T = reinterpret_cast<SjTables *>(smem_raw);
    size_t off = (kHotBytes + 127) & ~size_t(127);
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem_raw + off) + 2 * warp;
    off += (size_t(WARPS) * 2 * 8 + 127) & ~size_t(127);
    float *wbase = reinterpret_cast<float *>(smem_raw + off) + size_t(warp) * warp_floats(maxseg);
    float *tb = wbase + 2 * SJ_HOP;   // transpose plane, later |X|
    float *lg = tb + kTFloats;        // log energies / standardised features
    float *P = lg + kLogFloats;       // mel partial sums

    {   // stage the tables
        const int4 *src = reinterpret_cast<const int4 *>(a.tables);
        int4 *dst = reinterpret_cast<int4 *>(T);
        for (int i = tid; i < int(kHotBytes / 16); i += WARPS * 32) dst[i] = __ldg(src + i);
    }
    if (lane == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        P[32 * maxseg] = 0.f; // the combine lists' zero slot
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const int n_samples = a.n_samples;
    const int F = (n_samples + SJ_HOP - 1) / SJ_HOP;
    const int last_valid = n_samples - (F - 1) * SJ_HOP;
    const long stride = long(gridDim.x) * WARPS;
    long clip = long(blockIdx.x) * WARPS + warp;
    const float2 *win2 = reinterpret_cast<const float2 *>(T->window);
    const int mel_ch = T->mel_ch, mel_base = T->mel_lo + lane * T->mel_ch, ncomb = T->mel_ncomb;
    const int partner = (32 - lane) & 31;

    // fetch hop `h` of clip `c` into ring slot `slot`
    auto fetch = [&](long c, int h, int slot) {
        const float *src = a.clips + c * (long)n_samples + (long)h * SJ_HOP;
        const int cnt = (h == F - 1) ? last_valid : SJ_HOP;
        float *dst = wbase + slot * SJ_HOP;
        if (TMA) {
            if (lane == 0) {
                fence_proxy_async();
                mbar_expect_tx(&bars[slot], (uint32_t)cnt * 4u);
                bulk_g2s(dst, src, (uint32_t)cnt * 4u, &bars[slot]);
            }
        } else {
            for (int i = lane; i < cnt; i += 32) dst[i] = __ldg(src + i);
            __syncwarp();
        }
    };

    uint32_t q = 0; // running hop number of this warp: ring slot q&1, mbarrier parity (q>>1)&1
    if (clip < a.n_clips) fetch(clip, 0, 0);

    for (; clip < a.n_clips; clip += stride) {
        double sum = 0.0, sq = 0.0;
        for (int t = 0; t < F; ++t, ++q) {
            const int slot = q & 1;
            if (TMA) mbar_wait(&bars[slot], (q >> 1) & 1);
            const float2 *cur = reinterpret_cast<const float2 *>(wbase + slot * SJ_HOP);
            const float2 *prv = reinterpret_cast<const float2 *>(wbase + (slot ^ 1) * SJ_HOP);
            float2 v[32];
            // ---- frame = [hop t-1 | hop t] * hann, packed z[m] = (y[2m], y[2m+1]), m = 32*m1 + lane
            if (t == 0) {
#pragma unroll
                for (int m1 = 0; m1 < 16; ++m1) v[m1] = make_float2(0.f, 0.f);
            } else {
#pragma unroll
                for (int m1 = 0; m1 < 16; ++m1) {
                    float2 s = prv[32 * m1 + lane], w = win2[32 * m1 + lane];
                    v[m1] = make_float2(s.x * w.x, s.y * w.y);
                }
            }
            const int valid = (t == F - 1) ? last_valid : SJ_HOP;
            if (valid == SJ_HOP) {
#pragma unroll
                for (int m1 = 16; m1 < 32; ++m1) {
                    float2 s = cur[32 * (m1 - 16) + lane], w = win2[32 * m1 + lane];
                    v[m1] = make_float2(s.x * w.x, s.y * w.y);
                }
            } else {
#pragma unroll
                for (int m1 = 16; m1 < 32; ++m1) {
                    const int i0 = 64 * (m1 - 16) + 2 * lane;
                    float2 s = cur[32 * (m1 - 16) + lane], w = win2[32 * m1 + lane];
                    v[m1] = make_float2(i0 < valid ? s.x * w.x : 0.f, i0 + 1 < valid ? s.y * w.y : 0.f);
                }
            }
            __syncwarp();
            {   // prefetch the next hop of the stream into the slot that just became free
                int nt = t + 1;
                long nc = clip;
                if (nt == F) { nt = 0; nc = clip + stride; }
                if (nc < a.n_clips) fetch(nc, nt, slot ^ 1);
            }
            // ---- 1024-point complex FFT, 32 x 32 split
            sj_fft32(v); // over m1: Y[k1] in v[SJ_BREV32(k1)]
#pragma unroll
            for (int k1 = 1; k1 < 32; ++k1) {
                const SjPair w = T->tw32[k1][lane];
                float2 &y = v[SJ_BREV32(k1)];
                const float yr = y.x * w.x - y.y * w.y;
                y.y = fmaf(y.x, w.y, y.y * w.x);
                y.x = yr;
            }
            // transpose (lane = m2, reg = k1) -> (lane = k1, reg = m2), one float plane at a time
#pragma unroll
            for (int k1 = 0; k1 < 32; ++k1) tb[k1 * kTStride + lane] = v[SJ_BREV32(k1)].x;
            __syncwarp();
#pragma unroll
            for (int m2 = 0; m2 < 32; ++m2) v[m2].x = tb[lane * kTStride + m2];
            __syncwarp();
#pragma unroll
            for (int k1 = 0; k1 < 32; ++k1) tb[k1 * kTStride + lane] = v[SJ_BREV32(k1)].y;
            __syncwarp();
#pragma unroll
            for (int m2 = 0; m2 < 32; ++m2) v[m2].y = tb[lane * kTStride + m2];
            __syncwarp();
            sj_fft32(v); // over m2: Z[lane + 32*k2] in v[SJ_BREV32(k2)]

            // ---- untangle the real FFT: bins k = lane + 32 j and 1024 - k, j < 16 (values are 2|X|)
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const float2 A = v[SJ_BREV32(j)];
                const float2 S = v[SJ_BREV32(31 - j)];
                float bx = __shfl_sync(kFull, S.x, partner), by = __shfl_sync(kFull, S.y, partner);
                if (lane == 0) { bx = v[SJ_BREV32((32 - j) & 31)].x; by = v[SJ_BREV32((32 - j) & 31)].y; }
                const float ex = A.x + bx, ey = A.y - by, dx = A.x - bx, dy = A.y + by;
                const SjPair w = T->twpost[lane + 32 * j];
                const float tx = fmaf(w.x, dy, w.y * dx), ty = fmaf(w.y, dy, -w.x * dx);
                const float xa = ex + tx, ya = ey + ty, xb = ex - tx, yb = ey - ty;
                tb[lane + 32 * j] = fast_sqrt(fmaf(xa, xa, ya * ya));
                tb[1024 - lane - 32 * j] = fast_sqrt(fmaf(xb, xb, yb * yb));
            }
            if (lane == 0) {
                const float2 c = v[SJ_BREV32(16)];
                tb[512] = 2.f * fast_sqrt(fmaf(c.x, c.x, c.y * c.y));
            }
            __syncwarp();

            // ---- sparse Slaney mel sums over this lane's bin chunk: two running sums (filter f, f+1);
            // where f steps, the finished sum goes to the lane's next slot and the f+1 sum moves down
            {
                float lo = 0.f, hi = 0.f;
                float *Pw = P + lane * maxseg;
                auto bin = [&](int i) {
                    const float av = tb[mel_base + i];
                    const SjPair w = T->melw[i][lane];
                    const bool step = __float_as_int(w.x) < 0;
                    if (step) *Pw = lo;
                    Pw += step;
                    lo = fmaf(fabsf(w.x), av, step ? hi : lo);
                    hi = fmaf(w.y, av, step ? 0.f : hi);
                };
                if (MELCH > 0) {
#pragma unroll
                    for (int i = 0; i < MELCH; ++i) bin(i);
                } else {
                    for (int i = 0; i < mel_ch; ++i) bin(i);
                }
                Pw[0] = lo;
                Pw[1] = hi;
            }
            __syncwarp();
            {
                float e0 = 0.f, e1 = 0.f;
                for (int e = 0; e < ncomb; ++e) {
                    e0 += P[T->comb[e][lane]];
                    e1 += P[T->comb[e][lane + 32]];
                }
                // aubio fvec_log10: log10(max(x, 2e-42)); lane m holds band m, lanes 0..7 also band m+32
                e0 = log10f(fmaxf(e0, 2.e-42f));
                e1 = log10f(fmaxf(e1, 2.e-42f));
                // DCT-II symmetry D[c][39-m] = (-1)^c D[c][m]: fold the 40 log energies into 20 sums
                // (even c) and 20 differences (odd c)
                const float far0 = __shfl_sync(kFull, e0, (39 - lane) & 31); // band 39-m for m = 8..19
                const float far1 = __shfl_sync(kFull, e1, (7 - lane) & 31);  // band 39-m for m = 0..7
                const float far = lane < 8 ? far1 : far0;
                if (lane < SJ_NCOEF) {
                    lg[lane] = e0 + far;
                    lg[SJ_NCOEF + lane] = e0 - far;
                }
            }
            __syncwarp();
            // ---- DCT-II (first 20 rows) and the reference's double accumulators
            {
                float acc = 0.f;
                const float *eo = lg + (lane & 1) * SJ_NCOEF;
#pragma unroll
                for (int m = 0; m < SJ_NFILT / 2; ++m) acc = fmaf(T->dct[m][lane], eo[m], acc);
                const double dv = (double)acc;
                sum += dv;
                sq += dv * dv;
            }
        }
        // ---- clip statistics (cepstrum/scrubjay_infer.c:61-67) and the SVM
        __syncwarp();
        if (lane < SJ_NCOEF) {
            const double mean = sum / (double)F;
            const double var = sq / (double)F - mean * mean;
            const float fm = (float)mean;
            const float fs = sqrtf((float)(var > 0 ? var : 0));
            if (a.feats) {
                a.feats[clip * SJ_FEAT + lane] = fm;
                a.feats[clip * SJ_FEAT + SJ_NCOEF + lane] = fs;
            }
            const SjTables *G = a.tables;
            lg[lane] = (fm - __ldg(&G->sv_off[lane])) * __ldg(&G->sv_scale[lane]);
            lg[lane + SJ_NCOEF] = (fs - __ldg(&G->sv_off[lane + SJ_NCOEF])) * __ldg(&G->sv_scale[lane + SJ_NCOEF]);
        }
        __syncwarp();
        if (a.do_svm) {
            const SvmOut r = svm_eval(a.tables, lg, lane);
            if (lane == 0) {
                if (a.labels) a.labels[clip] = r.label;
                if (a.conf) a.conf[clip] = r.conf;
                if (a.score) a.score[clip] = r.score;
            }
        }
        __syncwarp();
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

template <int WARPS> size_t smem_bytes(int maxseg)
{
    size_t off = (kHotBytes + 127) & ~size_t(127);
    off += (size_t(WARPS) * 2 * 8 + 127) & ~size_t(127);
    return off + size_t(WARPS) * warp_floats(maxseg) * sizeof(float);
}

constexpr int kWarps = 12;  // 14 fit too (128 registers) but measured no faster: the kernel is issue/LSU bound
constexpr int kMelFast = 27; // bins per lane of the 16 kHz Slaney bank: fully unrolled mel loop

template <bool TMA, int MELCH> cudaError_t setup_one(int maxseg)
{
    return cudaFuncSetAttribute(cepstrum_kernel<kWarps, TMA, MELCH>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)smem_bytes<kWarps>(maxseg));
}

template <bool TMA, int MELCH>
cudaError_t launch_one(const SjCepstrumArgs &a, int maxseg, int num_sms, cudaStream_t stream)
{
    long ctas = (a.n_clips + kWarps - 1) / kWarps;
    if (ctas > num_sms) ctas = num_sms;
    cepstrum_kernel<kWarps, TMA, MELCH><<<(unsigned)ctas, kWarps * 32, smem_bytes<kWarps>(maxseg), stream>>>(a, maxseg);
    return cudaGetLastError();
}

} // namespace

cudaError_t sj_cepstrum_setup(int maxseg)
{
    cudaError_t e = setup_one<true, kMelFast>(maxseg);
    if (e == cudaSuccess) e = setup_one<false, kMelFast>(maxseg);
    if (e == cudaSuccess) e = setup_one<true, 0>(maxseg);
    if (e == cudaSuccess) e = setup_one<false, 0>(maxseg);
    return e;
}

cudaError_t sj_launch_cepstrum(const SjCepstrumArgs &a, int maxseg, int mel_ch, int num_sms, cudaStream_t stream)
{
    if (a.n_clips <= 0) return cudaSuccess;
    // the bulk-copy path needs 16-byte aligned hop addresses and sizes
    const bool tma = (a.n_samples % 4 == 0) && ((reinterpret_cast<uintptr_t>(a.clips) & 15) == 0);
    if (mel_ch == kMelFast)
        return tma ? launch_one<true, kMelFast>(a, maxseg, num_sms, stream) : launch_one<false, kMelFast>(a, maxseg, num_sms, stream);
    return tma ? launch_one<true, 0>(a, maxseg, num_sms, stream) : launch_one<false, 0>(a, maxseg, num_sms, stream);
}

cudaError_t sj_launch_svm(const float *feats, long n, const SjTables *tables, int32_t *labels, float *conf,
                          float *score, cudaStream_t stream)
{
    if (n <= 0) return cudaSuccess;
    long blocks = (n + 7) / 8;
    if (blocks > 148 * 8) blocks = 148 * 8;
    svm_kernel<<<(unsigned)blocks, 256, 0, stream>>>(feats, n, tables, labels, conf, score);
    return cudaGetLastError();
}
