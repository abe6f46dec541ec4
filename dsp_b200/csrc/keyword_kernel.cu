// keyword_kernel.cu -- fused keyword ("stop") detector for sm_100a: for a batch of clips, the whole body of
// classify_signal (2fa/audio/word/c/stop_detector.c:12-55): compute_mfcc (mfcc.c:108-232: frames of 400 samples
// every 160, Hann, 512-point real FFT, power spectrum, 40-band mel, power_to_db with ref = max and top_db = 80,
// 13 x 40 DCT), the coefficient-major re-layout padded to 500 frames, and audio_classifier_predict
// (audio_classifier_inference.c:38-90: scaler, dense 6500-4-2-2-1 with ReLU, sigmoid).
//
// Mapping: one CTA (8 warps) per clip, persistent over clips.  A HALF-WARP transforms one frame: the 512 real
// samples are packed as 256 complex points, split 16 x 16 (16-point register FFT per lane, twiddle, transpose
// through shared memory, second 16-point FFT), untangled with the partner lane's value fetched by a shuffle,
// and the power spectrum goes to shared memory.  The full warp then finishes its two frames one after the
// other: sparse mel sums (the dense 40 x 257 table of the reference has 494 non-zeros, <= 2 adjacent filters per
// bin), dB conversion with warp-wide max reductions, DCT split over 26 lanes.  MFCCs never leave the SM: they
// land in a 13 x 500 shared-memory feature plane that the CTA reduces against the first dense layer.
// The clip is read once from HBM (frames overlap 60 %: the re-reads hit L1/L2).
#include <cuda_runtime.h>
#include <math.h>
#include <stddef.h>
#include <stdint.h>

#define SJ_PACKED_F32 1
#include "sj_pk.h"
#include "fft_gen.inc"
#include "keyword_kernel.h"

namespace {

constexpr unsigned kFull = 0xffffffffu;
constexpr int kWarps = 8;
constexpr int kZStride = 17;                  // padded row stride (float2) of the 16 x 16 transpose buffer
constexpr int kPwFloats = 32 * KW_MEL_CH;     // 288: power spectrum of one frame, padded to the mel chunks
constexpr size_t kHotBytes = offsetof(KwTables, mean);
static_assert(kHotBytes % 16 == 0, "tables are staged with 16-byte copies");

__host__ __device__ constexpr int kw_p_floats(int maxseg) { return (32 * maxseg + 4) & ~3; }
__host__ __device__ constexpr int kw_warp_floats(int maxseg)
{
    return 2 * 2 * 16 * kZStride + 2 * kPwFloats + kw_p_floats(maxseg) + 64;
}

// 10 * log10(x) for x >= 1e-10 on one MUFU.LG2
__device__ __forceinline__ float db10(float x)
{
    float l;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l) : "f"(x));
    return l * 3.0102999566398120f;
}

__global__ void __launch_bounds__(kWarps * 32, 2) keyword_kernel(const KwArgs a, const int maxseg)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, h = lane & 15, half = lane >> 4;
    KwTables *T = reinterpret_cast<KwTables *>(smem_raw);
    float *feat = reinterpret_cast<float *>(smem_raw + ((kHotBytes + 15) & ~size_t(15))); // [13][500]
    float *red = feat + KW_INPUT;                                                         // [8][4] reduction scratch
    float *wbase = red + 32 + size_t(warp) * kw_warp_floats(maxseg);
    float2 *zb = reinterpret_cast<float2 *>(wbase) + half * 16 * kZStride; // this half-warp's transpose buffer
    float *pw = wbase + 2 * 2 * 16 * kZStride;                             // [2][288] power spectra
    float *P = pw + 2 * kPwFloats;                                         // mel partial sums
    float *lg = P + kw_p_floats(maxseg);                                   // 40 dB mel energies

    {
        const int4 *src = reinterpret_cast<const int4 *>(a.tables);
        int4 *dst = reinterpret_cast<int4 *>(T);
        for (int i = tid; i < int(kHotBytes / 16); i += kWarps * 32) dst[i] = __ldg(src + i);
    }
    for (int i = lane; i < kw_p_floats(maxseg); i += 32) P[i] = 0.f;
    for (int i = lane; i < 2 * kPwFloats; i += 32) pw[i] = 0.f; // the padding bins 257..287 stay zero
    __syncthreads();

    const int n = a.n_samples;
    int nfr = n < KW_FRAME ? 0 : (n - KW_FRAME) / KW_HOP + 1;
    if (nfr > KW_MAX_FRAMES) nfr = KW_MAX_FRAMES;
    const int ncomb = T->mel_ncomb;
    const int hw = warp * 2 + half; // half-warp number: frames hw, hw + 16, ...
    const int pl = (lane & 16) | ((16 - h) & 15); // lane holding the mirror bins of this lane's bins
    const KwTables *G = a.tables;

    for (long clip = blockIdx.x; clip < a.n_clips; clip += gridDim.x) {
        const float *x = a.signals + clip * (long)n;
        for (int i = tid; i < KW_INPUT; i += kWarps * 32) feat[i] = 0.f; // frames beyond nfr stay zero (stop_detector.c:40-47)
        __syncthreads();

        for (int t0 = 0; t0 < nfr; t0 += 2 * kWarps) {
            const int t = t0 + hw;
            const bool active = t < nfr;
            const float *xs = x + (long)(active ? t : 0) * KW_HOP;
            float2 v[16];
            // ---- frame * Hann, packed z[m] = (y[2m], y[2m+1]), m = 16*m1 + h; samples 400..511 are the zero padding
#pragma unroll
            for (int m1 = 0; m1 < 16; ++m1) {
                const int m = 16 * m1 + h;
                if (m1 < 13 && m < KW_FRAME / 2) {
                    const KwPair w = T->hann2[m];
                    v[m1] = make_float2(__ldg(xs + 2 * m) * w.x, __ldg(xs + 2 * m + 1) * w.y);
                } else v[m1] = make_float2(0.f, 0.f);
            }
            // ---- 256-point complex FFT, 16 x 16 split
            sj_fft16(v); // over m1: Y[k1] in v[SJ_BREV16(k1)]
#pragma unroll
            for (int k1 = 1; k1 < 16; ++k1) {
                const KwPair w = T->tw16[k1][h];
                float2 &y = v[SJ_BREV16(k1)];
                const float yr = y.x * w.x - y.y * w.y;
                y.y = fmaf(y.x, w.y, y.y * w.x);
                y.x = yr;
            }
#pragma unroll
            for (int k1 = 0; k1 < 16; ++k1) zb[k1 * kZStride + h] = v[SJ_BREV16(k1)];
            __syncwarp();
#pragma unroll
            for (int m2 = 0; m2 < 16; ++m2) v[m2] = zb[h * kZStride + m2];
            __syncwarp();
            sj_fft16(v); // over m2: Z[h + 16*k2] in v[SJ_BREV16(k2)]

            // ---- untangle the real FFT and take the power spectrum: bins k = h + 16 j and 256 - k, j < 8.
            // The values formed are 2 X[k]; the factor 4 on the power is folded into the mel weights.
            float *pwh = pw + half * kPwFloats;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float2 A = v[SJ_BREV16(j)];
                const float2 S = v[SJ_BREV16(15 - j)];
                float bx = __shfl_sync(kFull, S.x, pl), by = __shfl_sync(kFull, S.y, pl);
                if (h == 0) { bx = v[SJ_BREV16((16 - j) & 15)].x; by = v[SJ_BREV16((16 - j) & 15)].y; }
                const float ex = A.x + bx, ey = A.y - by, dx = A.x - bx, dy = A.y + by;
                const KwPair w = T->tp[h + 16 * j];
                const float tx = fmaf(w.x, dy, w.y * dx), ty = fmaf(w.y, dy, -w.x * dx);
                const float xa = ex + tx, ya = ey + ty, xb = ex - tx, yb = ey - ty;
                pwh[h + 16 * j] = fmaf(xa, xa, ya * ya);
                pwh[256 - h - 16 * j] = fmaf(xb, xb, yb * yb);
            }
            if (h == 0) {
                const float2 c = v[SJ_BREV16(8)];
                pwh[128] = 4.f * fmaf(c.x, c.x, c.y * c.y);
            }
            __syncwarp();

            // ---- the full warp finishes its two frames, one after the other
            for (int fr = 0; fr < 2; ++fr) {
                const int tf = t0 + warp * 2 + fr;
                if (tf >= nfr) break; // warp-uniform
                const float *pf = pw + fr * kPwFloats;
                {   // sparse mel sums over this lane's 9-bin chunk (see cepstrum_kernel's generic path)
                    float lo = 0.f, hi = 0.f;
                    float *Pw = P + lane * maxseg;
#pragma unroll
                    for (int i = 0; i < KW_MEL_CH; ++i) {
                        const float av = pf[lane * KW_MEL_CH + i];
                        const KwPair w = T->melw[i][lane];
                        const bool step = __float_as_int(w.x) < 0;
                        if (step) *Pw = lo;
                        Pw += step;
                        lo = fmaf(fabsf(w.x), av, step ? hi : lo);
                        hi = fmaf(w.y, av, step ? 0.f : hi);
                    }
                    Pw[0] = lo;
                    Pw[1] = hi;
                }
                __syncwarp();
                float e0 = 0.f, e1 = 0.f; // mel energy of filter `lane` and, on lanes 0..7, of filter lane + 32
                for (int e = 0; e < ncomb; ++e) {
                    const uint32_t c = T->comb2[e][lane];
                    e0 += P[c & 0xffffu];
                    e1 += P[c >> 16];
                }
                // power_to_db(ref = max, amin = 1e-10, top_db = 80): mfcc.c:171-210
                const float amin = 1e-10f;
                float ref = fmaxf(e0, lane < 8 ? e1 : 0.f);
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) ref = fmaxf(ref, __shfl_xor_sync(kFull, ref, o));
                ref = fmaxf(ref, amin);
                const float log_ref = db10(ref);
                float d0 = db10(fmaxf(e0, amin)) - log_ref, d1 = db10(fmaxf(e1, amin)) - log_ref;
                float mx = fmaxf(d0, lane < 8 ? d1 : -INFINITY);
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(kFull, mx, o));
                const float floor_db = mx - 80.0f;
                lg[lane] = fmaxf(d0, floor_db);
                if (lane < 8) lg[32 + lane] = fmaxf(d1, floor_db);
                __syncwarp();
                // DCT (mfcc.c:213-220): lane l < 26 sums mel rows 20*(l/13) .. +19 of coefficient l % 13
                float acc = 0.f;
                // (the 20 dB values of a half as five 128-bit words: two addresses per warp)
                const float4 *lgh = reinterpret_cast<const float4 *>(lg + (lane >= 13 ? 20 : 0));
#pragma unroll
                for (int q = 0; q < 5; ++q) {
                    const float4 e = lgh[q];
                    acc = fmaf(T->dct[4 * q][lane], e.x, acc);
                    acc = fmaf(T->dct[4 * q + 1][lane], e.y, acc);
                    acc = fmaf(T->dct[4 * q + 2][lane], e.z, acc);
                    acc = fmaf(T->dct[4 * q + 3][lane], e.w, acc);
                }
                acc += __shfl_down_sync(kFull, acc, 13);
                if (lane < KW_NMFCC) {
                    feat[lane * KW_MAX_FRAMES + tf] = acc;
                    if (a.mfcc) a.mfcc[(clip * KW_MAX_FRAMES + tf) * KW_NMFCC + lane] = acc;
                }
                __syncwarp();
            }
        }
        __syncthreads();

        // ---- scaler + first dense layer over the 6500 features, then the 4-2-2-1 tail
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
        for (int i = tid; i < KW_INPUT; i += kWarps * 32) {
            float sc = __ldg(&G->scale[i]);
            if (sc == 0.0f) sc = 1.0f;
            const float xi = __fdiv_rn(feat[i] - __ldg(&G->mean[i]), sc);
            const float4 w = __ldg(reinterpret_cast<const float4 *>(&G->w1[i][0]));
            s0 = fmaf(w.x, xi, s0);
            s1 = fmaf(w.y, xi, s1);
            s2 = fmaf(w.z, xi, s2);
            s3 = fmaf(w.w, xi, s3);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            s0 += __shfl_xor_sync(kFull, s0, o);
            s1 += __shfl_xor_sync(kFull, s1, o);
            s2 += __shfl_xor_sync(kFull, s2, o);
            s3 += __shfl_xor_sync(kFull, s3, o);
        }
        if (lane == 0) { red[warp * 4 + 0] = s0; red[warp * 4 + 1] = s1; red[warp * 4 + 2] = s2; red[warp * 4 + 3] = s3; }
        __syncthreads();
        if (tid == 0) {
            float h1[4], h2[2], h3[2];
            for (int j = 0; j < 4; ++j) {
                float s = __ldg(&G->b1[j]);
                for (int w = 0; w < kWarps; ++w) s += red[w * 4 + j];
                h1[j] = s > 0.f ? s : 0.f;
            }
            for (int j = 0; j < 2; ++j) {
                float s = __ldg(&G->b2[j]);
                for (int i = 0; i < 4; ++i) s += __ldg(&G->w2[i][j]) * h1[i];
                h2[j] = s > 0.f ? s : 0.f;
            }
            for (int j = 0; j < 2; ++j) {
                float s = __ldg(&G->b3[j]);
                for (int i = 0; i < 2; ++i) s += __ldg(&G->w3[i][j]) * h2[i];
                h3[j] = s > 0.f ? s : 0.f;
            }
            float o = __ldg(&G->b4[0]);
            for (int i = 0; i < 2; ++i) o += __ldg(&G->w4[i]) * h3[i];
            a.prob[clip] = 1.0f / (1.0f + expf(-o));
            if (a.n_frames) a.n_frames[clip] = nfr;
        }
        if (a.mfcc) // compute_mfcc leaves the frames it did not compute untouched; the batched form zero-fills them
            for (int i = nfr * KW_NMFCC + tid; i < KW_INPUT; i += kWarps * 32) a.mfcc[clip * KW_INPUT + i] = 0.f;
        __syncthreads();
    }
}

size_t kw_smem_bytes(int maxseg)
{
    return ((kHotBytes + 15) & ~size_t(15)) + (KW_INPUT + 32 + size_t(kWarps) * kw_warp_floats(maxseg)) * sizeof(float);
}

} // namespace

cudaError_t kw_setup()
{
    return cudaFuncSetAttribute(keyword_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kw_smem_bytes(8));
}

cudaError_t kw_launch(const KwArgs &a, int maxseg, int num_sms, cudaStream_t stream)
{
    if (a.n_clips <= 0) return cudaSuccess;
    if (maxseg > 8) return cudaErrorInvalidValue;
    long ctas = a.n_clips < 2L * num_sms ? a.n_clips : 2L * num_sms;
    keyword_kernel<<<(unsigned)ctas, kWarps * 32, kw_smem_bytes(maxseg), stream>>>(a, maxseg);
    return cudaGetLastError();
}
