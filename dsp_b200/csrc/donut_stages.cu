// donut_stages.cu -- stand-alone stage kernels of the donut classifier for sm_100a (see donut_stages.h).
//
//  * iir_exact_kernel: butter_bandpass_filter (classifier.c:420-446 / classifier.cpp:193-219) for a batch of
//    signals, one direct-form-II recurrence per lane, every multiply and add rounded separately in the
//    reference's order => outputs bit-identical to the reference for any coefficient set.  The state can be
//    carried from call to call (continuous streams).
//  * iir_scan_*: the same filter as a block-wise state-space scan over time for FEW LONG signals (one stream is
//    the only parallelism): pass 1 runs every block of SJ_SCAN_BLOCK samples from the zero state and keeps the
//    final state, pass 2 chains the blocks (s_{k+1} = A^L s_k + f_k, A^L from the host), pass 3 re-runs every
//    block from its true initial state and writes the outputs.  Twice the arithmetic, n_samples / L-fold
//    parallelism; the block-boundary states differ from the sequential ones by rounding, so outputs agree to
//    ~1e-15 relative in double but are not bit-identical (this is why the clip path does not use it).
//  * spectrogram_kernel: compute_spectrogram (classifier.c:448-592 / classifier.cpp:221-368 / classifier.ino:245-395)
//    into a dense [signal][bin][frame] array.  One lane per frame, the frame's FFT in a shared-memory column:
//    sequential frame mean, Tukey window, radix-2 DIT with PlainFFT::Compute's butterfly arithmetic and twiddle
//    sequence (float: bit-identical PSD cells; double: FFTW's order is unknowable, cells agree to ~1e-15).
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "donut_stages.h"
#include "sj_pk.h"

namespace {

__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub_rn(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float div_rn(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double sub_rn(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double div_rn(double a, double b) { return __ddiv_rn(a, b); }

// one step of the reference's loop body; w[k] = w0 of time n-1-k (the reference's shifted array)
template <typename T> __device__ __forceinline__ T iir_step(T x, T (&w)[8], const T (&a)[9], const T (&b)[9])
{
    T w0 = x;
#pragma unroll
    for (int j = 1; j < 9; ++j) w0 = sub_rn(w0, mul_rn(a[j], w[j - 1]));
    T y = mul_rn(b[0], w0);
#pragma unroll
    for (int j = 1; j < 9; ++j) y = add_rn(y, mul_rn(b[j], w[j - 1]));
#pragma unroll
    for (int j = 7; j > 0; --j) w[j] = w[j - 1];
    w[0] = w0;
    return y;
}

// eight steps with the state in a rotating register file (no shifts): before step s, w0 of time n-j is r[(s-j)&7]
template <typename T>
__device__ __forceinline__ void iir_block8(const T (&x)[8], T (&r)[8], const T (&a)[9], const T (&b)[9], T (&y)[8])
{
#pragma unroll
    for (int s = 0; s < 8; ++s) {
        T w0 = x[s];
#pragma unroll
        for (int j = 1; j < 9; ++j) w0 = sub_rn(w0, mul_rn(a[j], r[(s - j + 8) & 7]));
        T yy = mul_rn(b[0], w0);
#pragma unroll
        for (int j = 1; j < 9; ++j) yy = add_rn(yy, mul_rn(b[j], r[(s - j + 8) & 7]));
        r[s] = w0;
        y[s] = yy;
    }
}

// float: a[j] w and b[j] w of the same delayed value from ONE packed multiply (mul.rn.f32x2 rounds each half like the
// scalar instruction; the ordered subtractions / sums stay scalar, so nothing is contracted and the results keep the
// reference's bits) -- 9 instead of 17 multiply issue slots per sample
__device__ __forceinline__ void iir_block8(const float (&x)[8], float (&r)[8], const float (&a)[9], const float (&b)[9],
                                           float (&y)[8])
{
#pragma unroll
    for (int s = 0; s < 8; ++s) {
        float2 p[8];
#pragma unroll
        for (int j = 1; j < 9; ++j) p[j - 1] = pk_mul(pk_bc(r[(s - j + 8) & 7]), make_float2(a[j], b[j]));
        float w0 = x[s];
#pragma unroll
        for (int j = 1; j < 9; ++j) w0 = sub_rn(w0, p[j - 1].x);
        float yy = mul_rn(b[0], w0);
#pragma unroll
        for (int j = 1; j < 9; ++j) yy = add_rn(yy, p[j - 1].y);
        r[s] = w0;
        y[s] = yy;
    }
}

// eight consecutive elements of one row: 16-byte accesses when the row is 16-byte aligned (each lane walks its own row,
// so a warp-level access touches 32 different lines either way -- wide accesses cut the number of such instructions)
template <typename E> __device__ __forceinline__ void load8(const E *p, E (&d)[8], bool vec)
{
    if (vec) {
        constexpr int kChunks = int(sizeof(E)) * 8 / 16;
        uint4 u[kChunks];
#pragma unroll
        for (int c = 0; c < kChunks; ++c) u[c] = __ldg(reinterpret_cast<const uint4 *>(p) + c);
        const E *e = reinterpret_cast<const E *>(u);
#pragma unroll
        for (int k = 0; k < 8; ++k) d[k] = e[k];
    } else {
#pragma unroll
        for (int k = 0; k < 8; ++k) d[k] = __ldg(p + k);
    }
}
template <typename E> __device__ __forceinline__ void store8(E *p, const E (&v)[8], bool vec)
{
    if (vec) {
        constexpr int kChunks = int(sizeof(E)) * 8 / 16;
        uint4 u[kChunks];
        E *e = reinterpret_cast<E *>(u);
#pragma unroll
        for (int k = 0; k < 8; ++k) e[k] = v[k];
#pragma unroll
        for (int c = 0; c < kChunks; ++c) reinterpret_cast<uint4 *>(p)[c] = u[c];
    } else {
#pragma unroll
        for (int k = 0; k < 8; ++k) p[k] = v[k];
    }
}

// run `n` samples of one signal from state w (shifted form); out may be NULL (state only)
template <typename T, bool WRITE, typename IN = T>
__device__ __forceinline__ void iir_run(const IN *__restrict__ in, T *__restrict__ out, long n, T (&w)[8],
                                        const T (&a)[9], const T (&b)[9], bool vec_in, bool vec_out)
{
    T r[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) r[7 - k] = w[k]; // w0 of time n-1-k = r[(0-(k+1))&7]
    // The recurrence is one dependent chain per lane, so nothing but the lane itself can hide the latency of its
    // own input: a register ring of kPre blocks of 8 samples, each requested kPre blocks before it is filtered.
    constexpr int kPre = 4;
    IN ring[kPre][8];
    const long nblk = n / 8;
#pragma unroll
    for (int j = 0; j < kPre; ++j) {
        if ((long)j < nblk) load8<IN>(in + 8 * j, ring[j], vec_in);
        else {
#pragma unroll
            for (int k = 0; k < 8; ++k) ring[j][k] = (IN)0;
        }
    }
    long blk = 0;
    for (; blk + kPre <= nblk; blk += kPre) {
#pragma unroll
        for (int j = 0; j < kPre; ++j) {
            T x[8], y[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) x[k] = (T)ring[j][k];
            if (blk + j + kPre < nblk) load8<IN>(in + 8 * (blk + j + kPre), ring[j], vec_in);
            iir_block8(x, r, a, b, y);
            if (WRITE) store8<T>(out + 8 * (blk + j), y, vec_out);
        }
    }
    long i = 8 * blk;
    for (; i + 8 <= n; i += 8) { // the last (< kPre) whole blocks
        T x[8], y[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) x[k] = (T)__ldg(in + i + k);
        iir_block8(x, r, a, b, y);
        if (WRITE) {
#pragma unroll
            for (int k = 0; k < 8; ++k) out[i + k] = y[k];
        }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) w[k] = r[7 - k];
    for (; i < n; ++i) {
        const T y = iir_step<T>((T)__ldg(in + i), w, a, b);
        if (WRITE) out[i] = y;
    }
}

template <typename T> __device__ __forceinline__ void load_coef(const void *coef, T (&a)[9], T (&b)[9])
{
    const T *c = reinterpret_cast<const T *>(coef);
#pragma unroll
    for (int j = 0; j < 9; ++j) { a[j] = __ldg(c + j); b[j] = __ldg(c + 9 + j); }
}

template <typename T> __global__ void __launch_bounds__(128) iir_exact_kernel(const SjIirArgs p)
{
    const long s = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= p.n_signals) return;
    T a[9], b[9], w[8];
    load_coef<T>(p.coef, a, b);
    T *st = p.state ? reinterpret_cast<T *>(p.state) + s * 8 : nullptr;
#pragma unroll
    for (int k = 0; k < 8; ++k) w[k] = st ? st[k] : (T)0;
    if (sizeof(T) == 8 && p.in_f32)
        iir_run<T, true, float>(reinterpret_cast<const float *>(p.in) + s * p.in_stride,
                                reinterpret_cast<T *>(p.out) + s * p.out_stride, p.n_samples, w, a, b, p.vec_in, p.vec_out);
    else
        iir_run<T, true>(reinterpret_cast<const T *>(p.in) + s * p.in_stride, reinterpret_cast<T *>(p.out) + s * p.out_stride,
                         p.n_samples, w, a, b, p.vec_in, p.vec_out);
    if (st) {
#pragma unroll
        for (int k = 0; k < 8; ++k) st[k] = w[k];
    }
}

// scan pass 1: final state of every block started from the zero state
template <typename T> __global__ void __launch_bounds__(128) iir_scan_local_kernel(const SjIirArgs p)
{
    const long id = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= p.n_signals * p.n_blocks) return;
    const long s = id / p.n_blocks, k = id - s * p.n_blocks;
    if (k == p.n_blocks - 1) return; // nothing follows the last block
    T a[9], b[9], w[8];
    load_coef<T>(p.coef, a, b);
#pragma unroll
    for (int j = 0; j < 8; ++j) w[j] = 0;
    iir_run<T, false>(reinterpret_cast<const T *>(p.in) + s * p.in_stride + k * SJ_SCAN_BLOCK, nullptr, SJ_SCAN_BLOCK, w, a, b,
                      p.vec_in, false);
    T *f = reinterpret_cast<T *>(p.blk_final) + id * 8;
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] = w[j];
}

// scan pass 2: chain the blocks of one signal; blk_final[k] becomes the INITIAL state of block k
template <typename T> __global__ void __launch_bounds__(128) iir_scan_chain_kernel(const SjIirArgs p)
{
    const long s = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= p.n_signals) return;
    const T *M = reinterpret_cast<const T *>(p.M);
    T *f = reinterpret_cast<T *>(p.blk_final) + s * p.n_blocks * 8;
    const T *st = p.state ? reinterpret_cast<const T *>(p.state) + s * 8 : nullptr;
    T cur[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) cur[j] = st ? st[j] : (T)0;
    for (long k = 0; k < p.n_blocks; ++k) {
        T fin[8], nxt[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) { fin[j] = f[k * 8 + j]; f[k * 8 + j] = cur[j]; }
        if (k == p.n_blocks - 1) break;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            T acc = fin[i];
#pragma unroll
            for (int j = 0; j < 8; ++j) acc = fma(__ldg(M + i * 8 + j), cur[j], acc);
            nxt[i] = acc;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) cur[j] = nxt[j];
    }
}

// scan pass 3: every block again, from its true initial state, writing the outputs
template <typename T> __global__ void __launch_bounds__(128) iir_scan_apply_kernel(const SjIirArgs p)
{
    const long id = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= p.n_signals * p.n_blocks) return;
    const long s = id / p.n_blocks, k = id - s * p.n_blocks;
    T a[9], b[9], w[8];
    load_coef<T>(p.coef, a, b);
    const T *f = reinterpret_cast<const T *>(p.blk_final) + id * 8;
#pragma unroll
    for (int j = 0; j < 8; ++j) w[j] = f[j];
    const long i0 = k * SJ_SCAN_BLOCK;
    const long n = (k == p.n_blocks - 1) ? p.n_samples - i0 : SJ_SCAN_BLOCK;
    iir_run<T, true>(reinterpret_cast<const T *>(p.in) + s * p.in_stride + i0,
                     reinterpret_cast<T *>(p.out) + s * p.out_stride + i0, n, w, a, b, p.vec_in, p.vec_out);
    if (p.state && k == p.n_blocks - 1) {
        T *st = reinterpret_cast<T *>(p.state) + s * 8;
#pragma unroll
        for (int j = 0; j < 8; ++j) st[j] = w[j];
    }
}

template <typename T> struct V2;
template <> struct V2<float> { typedef float2 type; };
template <> struct V2<double> { typedef double2 type; };

// PlainFFT::Compute's butterfly (sync/lib/PlainFFT.cpp:64-71): t = u * hi; hi = lo - t; lo = lo + t, ten roundings
__device__ __forceinline__ void bfly(float2 &lo, float2 &hi, float u1, float u2)
{
    const float2 m1 = pk_mul(pk_bc(u1), hi);
    const float2 m2 = pk_mul(pk_bc(u2), make_float2(hi.y, hi.x));
    // (scalar: ptxas would contract mul.rn.f32x2 + add.rn.f32x2 into one FFMA2 and drop two roundings)
    const float2 t = make_float2(sub_rn(m1.x, m2.x), add_rn(m1.y, m2.y));
    hi = pk_sub(lo, t);
    lo = pk_add(lo, t);
}
__device__ __forceinline__ void bfly(double2 &lo, double2 &hi, double u1, double u2)
{
    const double t1 = sub_rn(mul_rn(u1, hi.x), mul_rn(u2, hi.y));
    const double t2 = add_rn(mul_rn(u1, hi.y), mul_rn(u2, hi.x));
    hi.x = sub_rn(lo.x, t1);
    hi.y = sub_rn(lo.y, t2);
    lo.x = add_rn(lo.x, t1);
    lo.y = add_rn(lo.y, t2);
}

// stages L0 .. L0+K-1 of the radix-2 DIT network on the column col[i * 32]
template <typename T, int N, int L0, int K>
__device__ __forceinline__ void fft_pass(typename V2<T>::type *col, const SjStageTab<T> *tab)
{
    typedef typename V2<T>::type T2;
    constexpr int G = 1 << K, l1 = 1 << L0;
#pragma unroll 2
    for (int grp = 0; grp < N / G; ++grp) {
        const int lo = grp & (l1 - 1), i = ((grp >> L0) << (L0 + K)) | lo; // bits [L0, L0+K) of i are zero
        T2 e[G];
#pragma unroll
        for (int j = 0; j < G; ++j) e[j] = col[(i + j * l1) * 32];
#pragma unroll
        for (int s = 0; s < K; ++s) {
#pragma unroll
            for (int j = 0; j < G; ++j)
                if (!(j & (1 << s))) {
                    // stage L0+s pairs elements 2^(L0+s) apart; twiddle offset = (element index) mod 2^(L0+s)
                    const int m = lo + (j & ((1 << s) - 1)) * l1, ls = l1 << s;
                    bfly(e[j], e[j | (1 << s)], tab->tw[ls - 1 + m][0], tab->tw[ls - 1 + m][1]);
                }
        }
#pragma unroll
        for (int j = 0; j < G; ++j) col[(i + j * l1) * 32] = e[j];
    }
}

template <int N> __host__ __device__ constexpr int log2c() { return N == 1 ? 0 : 1 + log2c<N / 2>(); }
template <> __host__ __device__ constexpr int log2c<1>() { return 0; }

// CTA = one warp = 32 consecutive frames (frame index = signal * frames + t), lane = frame
template <typename T, int N> __global__ void __launch_bounds__(32) spectrogram_kernel(const SjSpecArgs p)
{
    typedef typename V2<T>::type T2;
    extern __shared__ __align__(16) unsigned char smem[];
    SjStageTab<T> *tab = reinterpret_cast<SjStageTab<T> *>(smem);
    T2 *col = reinterpret_cast<T2 *>(smem + ((sizeof(SjStageTab<T>) + 15) & ~size_t(15))) + threadIdx.x; // col[i * 32]
    {
        const uint32_t *src = reinterpret_cast<const uint32_t *>(p.tab);
        uint32_t *dst = reinterpret_cast<uint32_t *>(tab);
        for (int i = threadIdx.x; i < int(sizeof(SjStageTab<T>) / 4); i += 32) dst[i] = __ldg(src + i);
    }
    __syncwarp();
    constexpr int LOG = log2c<N>(), BINS = N / 2 + 1;
    const int hop = tab->hop;
    const T U = tab->U;
    const long total = p.n_signals * (long)p.frames;
    for (long g = (long)blockIdx.x * 32 + threadIdx.x; g < total; g += (long)gridDim.x * 32) {
        const long s = g / p.frames;
        const int t = (int)(g - s * p.frames);
        const T *x = reinterpret_cast<const T *>(p.signals) + s * p.stride + (long)t * hop;
        // mean: "sum += vReal[i]" in sample order (classifier.cpp:320-324)
        // samples past the end of a short signal read as zero (classifier.cpp:310-315)
        const long left = p.n_samples - (long)t * hop;
        const int valid = left < N ? (int)left : N;
        T sum = 0;
#pragma unroll 8
        for (int i = 0; i < N; ++i) sum = add_rn(sum, i < valid ? __ldg(x + i) : (T)0);
        const T mean = div_rn(sum, (T)N);
        // detrend, window, bit-reversed placement (PlainFFT::Compute starts with the reversal)
#pragma unroll 8
        for (int i = 0; i < N; ++i) {
            const int r = (int)(__brev((unsigned)i) >> (32 - LOG));
            T2 v;
            v.x = mul_rn(sub_rn(i < valid ? __ldg(x + i) : (T)0, mean), tab->window[i]);
            v.y = 0;
            col[r * 32] = v;
        }
        // the butterfly network, K stages at a time on 2^K elements held in registers (3 + 3 + 2 stages for N = 256):
        // every butterfly sees the operands and the twiddle it has in PlainFFT::Compute, only the visiting order differs
        fft_pass<T, N, 0, 3>(col, tab);
        fft_pass<T, N, 3, 3>(col, tab);
        if constexpr (LOG > 6) fft_pass<T, N, 6, LOG - 6>(col, tab);
        // PSD (classifier.cpp:349-364): mag^2 / U, inner bins doubled
        T *out = reinterpret_cast<T *>(p.sxx) + s * (long)BINS * p.frames + t;
        for (int f = 0; f < BINS; ++f) {
            const T2 v = col[f * 32];
            T sxx = div_rn(add_rn(mul_rn(v.x, v.x), mul_rn(v.y, v.y)), U);
            if (f > 0 && f < BINS - 1) sxx = mul_rn(sxx, (T)2);
            out[(long)f * p.frames] = sxx;
        }
    }
}

template <typename T, int N> constexpr size_t spec_smem()
{
    return ((sizeof(SjStageTab<T>) + 15) & ~size_t(15)) + sizeof(T) * 2 * N * 32;
}

__global__ void sum_intense_kernel(const float *plane, int frames, int f0, int f1, int t0, int t1, float *out)
{
    float total = 0.0f;
    for (int i = f0; i <= f1; ++i)
        for (int j = t0; j <= t1; ++j) {
            const float v = plane[(long)i * frames + j];
            if (!isnan(v)) total = add_rn(total, v);
        }
    *out = total;
}

// dB value exactly as the reference forms it: (T)(10 * log10((double)s / 1e-12))
template <typename T> __device__ __forceinline__ T to_db(T s) { return (T)(10.0 * log10((double)s / 1e-12)); }

constexpr int kEpThreads = 256;

template <typename T> __device__ T block_min(T v, T *red)
{
    for (int o = 16; o > 0; o >>= 1) { const T u = __shfl_xor_sync(0xffffffffu, v, o); v = u < v ? u : v; }
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    T r = red[0];
    for (int i = 1; i < kEpThreads / 32; ++i) r = red[i] < r ? red[i] : r;
    return r;
}
template <typename T> __device__ T block_max(T v, T *red)
{
    for (int o = 16; o > 0; o >>= 1) { const T u = __shfl_xor_sync(0xffffffffu, v, o); v = u > v ? u : v; }
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    T r = red[0];
    for (int i = 1; i < kEpThreads / 32; ++i) r = red[i] > r ? red[i] : r;
    return r;
}
__device__ double block_sum(double v, double *red)
{
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    double r = 0;
    for (int i = 0; i < kEpThreads / 32; ++i) r += red[i];
    return r;
}

// one CTA per signal
template <typename T> __global__ void __launch_bounds__(kEpThreads) donut_epilogue_kernel(const SjEpilogueArgs a)
{
    __shared__ SjEpilogueParams<T> P;
    __shared__ double red[kEpThreads / 32];
    __shared__ int sh_count, sh_label;
    __shared__ T sh_sums[3];
    {
        const uint32_t *src = reinterpret_cast<const uint32_t *>(a.params);
        uint32_t *dst = reinterpret_cast<uint32_t *>(&P);
        for (int i = threadIdx.x; i < int(sizeof(P) / 4); i += kEpThreads) dst[i] = __ldg(src + i);
    }
    __syncthreads();
    const int frames = a.frames, bins = P.bins;
    const long cells = (long)bins * frames;
    for (long s = blockIdx.x; s < a.n_signals; s += gridDim.x) {
        const T *band = reinterpret_cast<const T *>(a.sxx_band) + s * cells;
        const T *mid = reinterpret_cast<const T *>(a.sxx_mid) + s * cells;
        unsigned char *flags = a.flags + s * (long)frames;
        T *mids = reinterpret_cast<T *>(a.mids_t) + s * (long)P.max_mid;
        float *fmids = a.mids ? a.mids + s * (long)P.max_mid : nullptr;
        // find_midpoints, part 1: frames with any cell above the dB threshold (threshold moved to the PSD domain)
        for (int t = threadIdx.x; t < frames; t += kEpThreads) {
            bool any = false;
            for (int f = 0; f < bins; ++f) any = any || (mid[(long)f * frames + t] >= P.s_thr);
            flags[t] = any;
        }
        // global minimum / maximum of the positive cells of the band spectrogram (dB is monotone)
        T mn = (T)INFINITY, mx = 0;
        for (long i = threadIdx.x; i < cells; i += kEpThreads) {
            const T v = band[i];
            if (v > 0) { mn = v < mn ? v : mn; mx = v > mx ? v : mx; }
        }
        mn = block_min<T>(mn, reinterpret_cast<T *>(red));
        mx = block_max<T>(mx, reinterpret_cast<T *>(red));
        __syncthreads();
        // find_midpoints, part 2: clusters of flagged frames (gap <= gap_tol, duration >= min_dur), sequentially
        if (threadIdx.x == 0) {
            int count = 0, size = 0;
            T first = 0, last = 0, tsum = 0;
            for (int j = 0; j <= frames; ++j) {
                const bool on = j < frames && flags[j];
                const T tj = div_rn((T)(j * P.hop + P.nperseg / 2), P.fs);
                const bool joins = on && size > 0 && sub_rn(tj, last) <= P.gap_tol;
                if (size > 0 && (!on ? j == frames : !joins)) {
                    if (sub_rn(last, first) >= P.min_dur) {
                        if (count < P.max_mid) mids[count] = div_rn(tsum, (T)size);
                        ++count;
                    }
                    size = 0;
                }
                if (on) {
                    if (size == 0) { first = tj; tsum = 0; }
                    tsum = add_rn(tsum, tj);
                    last = tj;
                    ++size;
                }
            }
            if (fmids)
                for (int k = 0; k < P.max_mid; ++k) fmids[k] = k < count ? (float)mids[k] : __int_as_float(0x7fc00000);
            sh_count = count;
            sh_label = 0;
            sh_sums[0] = sh_sums[1] = sh_sums[2] = 0;
        }
        __syncthreads();
        const int count = sh_count, cap = count < P.max_mid ? count : P.max_mid;
        const T dmn = to_db<T>(mn), dmx = to_db<T>(mx), range = sub_rn(dmx, dmn);
        for (int m = 0; m < cap; ++m) {
            const T mp = mids[m];
            T sums[3];
            for (int b = 0; b < 3; ++b) {
                const int f0 = P.box_f[b][0], f1 = P.box_f[b][1];
                const T tlo = sub_rn(mp, P.half[b]), thi = add_rn(mp, P.half[b]);
                // sum_intense's time searches: frames before the window / after it (times are increasing)
                int nb = 0, na = 0;
                for (int j = threadIdx.x; j < frames; j += kEpThreads) {
                    const T tj = div_rn((T)(j * P.hop + P.nperseg / 2), P.fs);
                    nb += tj < tlo;
                    na += tj > thi;
                }
                nb = (int)block_sum((double)nb, red);
                __syncthreads();
                na = (int)block_sum((double)na, red);
                __syncthreads();
                int t0 = nb, t1 = frames - 1 - na;
                if (t0 >= frames) t0 = frames - 1;
                if (t1 < 0) t1 = 0;
                if (t0 > t1) { const int x = t0; t0 = t1; t1 = x; }
                const int nt = t1 - t0 + 1;
                const long ncell = (long)(f1 - f0 + 1) * nt;
                double acc = 0.0;
                if (range > 0)
                    for (long c = threadIdx.x; c < ncell; c += kEpThreads) {
                        const int i = (int)(c / nt), j = (int)(c - (long)i * nt);
                        const T v = band[(long)(f0 + i) * frames + t0 + j];
                        if (v > 0) {
                            const T nv = div_rn(sub_rn(to_db<T>(v), dmn), range);
                            if (nv > P.lo_norm && nv < P.hi_norm) acc += (double)nv;
                        }
                    }
                sums[b] = (T)block_sum(acc, red);
                __syncthreads();
            }
            const bool hit = sums[1] < P.thr_mid && sums[0] > P.thr_above && sums[2] > P.thr_below;
            if (threadIdx.x == 0) { sh_sums[0] = sums[0]; sh_sums[1] = sums[1]; sh_sums[2] = sums[2]; if (hit) sh_label = 1; }
            if (hit) break;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            a.labels[s] = sh_label;
            if (a.detail) {
                float *d = a.detail + s * 4;
                d[0] = (float)count; d[1] = (float)sh_sums[0]; d[2] = (float)sh_sums[1]; d[3] = (float)sh_sums[2];
            }
        }
        __syncthreads();
    }
}

} // namespace

cudaError_t sj_launch_donut_epilogue(const SjEpilogueArgs &a, bool f64, cudaStream_t stream)
{
    if (a.n_signals <= 0) return cudaSuccess;
    const unsigned grid = (unsigned)(a.n_signals < 148L * 8 ? a.n_signals : 148L * 8);
    if (f64) donut_epilogue_kernel<double><<<grid, kEpThreads, 0, stream>>>(a);
    else donut_epilogue_kernel<float><<<grid, kEpThreads, 0, stream>>>(a);
    return cudaGetLastError();
}

static SjIirArgs with_alignment(SjIirArgs a, bool f64)
{
    const size_t es_out = f64 ? 8 : 4, es_in = (f64 && !a.in_f32) ? 8 : 4;
    a.vec_in = (reinterpret_cast<uintptr_t>(a.in) % 16 == 0) && ((size_t)a.in_stride * es_in % 16 == 0);
    a.vec_out = (reinterpret_cast<uintptr_t>(a.out) % 16 == 0) && ((size_t)a.out_stride * es_out % 16 == 0);
    return a;
}

cudaError_t sj_launch_iir_exact(const SjIirArgs &a0, bool f64, cudaStream_t stream)
{
    const SjIirArgs a = with_alignment(a0, f64);
    if (a.n_signals <= 0 || a.n_samples <= 0) return cudaSuccess;
    const unsigned grid = (unsigned)((a.n_signals + 127) / 128);
    if (f64) iir_exact_kernel<double><<<grid, 128, 0, stream>>>(a);
    else iir_exact_kernel<float><<<grid, 128, 0, stream>>>(a);
    return cudaGetLastError();
}

cudaError_t sj_launch_iir_scan(const SjIirArgs &a0, bool f64, cudaStream_t stream)
{
    const SjIirArgs a = with_alignment(a0, f64);
    if (a.n_signals <= 0 || a.n_samples <= 0) return cudaSuccess;
    const long items = a.n_signals * a.n_blocks;
    const unsigned grid = (unsigned)((items + 127) / 128), grid_s = (unsigned)((a.n_signals + 127) / 128);
    if (f64) {
        iir_scan_local_kernel<double><<<grid, 128, 0, stream>>>(a);
        iir_scan_chain_kernel<double><<<grid_s, 128, 0, stream>>>(a);
        iir_scan_apply_kernel<double><<<grid, 128, 0, stream>>>(a);
    } else {
        iir_scan_local_kernel<float><<<grid, 128, 0, stream>>>(a);
        iir_scan_chain_kernel<float><<<grid_s, 128, 0, stream>>>(a);
        iir_scan_apply_kernel<float><<<grid, 128, 0, stream>>>(a);
    }
    return cudaGetLastError();
}

cudaError_t sj_stage_setup()
{
    cudaError_t e;
    e = cudaFuncSetAttribute(spectrogram_kernel<float, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)spec_smem<float, 256>());
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(spectrogram_kernel<double, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)spec_smem<double, 256>());
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(spectrogram_kernel<float, 64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)spec_smem<float, 64>());
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(spectrogram_kernel<double, 64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)spec_smem<double, 64>());
}

cudaError_t sj_launch_spectrogram(const SjSpecArgs &a, bool f64, int nperseg, cudaStream_t stream)
{
    const long total = a.n_signals * (long)a.frames;
    if (total <= 0) return cudaSuccess;
    long want = (total + 31) / 32;
    const unsigned grid = (unsigned)(want < 148L * 16 ? want : 148L * 16);
    if (nperseg == 256) {
        if (f64) spectrogram_kernel<double, 256><<<grid, 32, spec_smem<double, 256>(), stream>>>(a);
        else spectrogram_kernel<float, 256><<<grid, 32, spec_smem<float, 256>(), stream>>>(a);
    } else if (nperseg == 64) {
        if (f64) spectrogram_kernel<double, 64><<<grid, 32, spec_smem<double, 64>(), stream>>>(a);
        else spectrogram_kernel<float, 64><<<grid, 32, spec_smem<float, 64>(), stream>>>(a);
    } else return cudaErrorInvalidValue;
    return cudaGetLastError();
}

cudaError_t sj_launch_sum_intense(const float *d_plane, int bins, int frames, int f0, int f1, int t0, int t1,
                                  float *d_out, cudaStream_t stream)
{
    (void)bins;
    sum_intense_kernel<<<1, 1, 0, stream>>>(d_plane, frames, f0, f1, t0, t1, d_out);
    return cudaGetLastError();
}
