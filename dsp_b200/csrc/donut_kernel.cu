// donut_kernel.cu -- fused donut classifier for sm_100a: the per-clip body of
// `int classify(float *data, int data_size)` (sync/lib/classifier.cpp:9-136, float) and of main in
// donut-classifier/classifier.c:84-198 (double), for a whole batch of clips in one launch.
//
// Mapping.  A CTA of 3 (float) or 4 (double) warps owns 16 clips at a time, i.e. 32 (clip, filter) rows, row = 2*clip + filter.
// Warp 0 runs phase 1 (one row per lane); all warps share phase 2 (the row passes are dealt round-robin) and the
// band-sum epilogue (clips dealt round-robin), with CTA barriers between the phases:
//   phase 1 (IIR, lane-parallel over clips): every lane advances ITS direct-form-II recurrence
//     (butter_bandpass_filter, classifier.c:420-446) by one spectrogram hop, sequentially in time and
//     with the reference's exact operation order (separate multiply and add, no FMA), so the filtered
//     samples are bit-identical to the reference.  With ~10^6 independent clips there is ample
//     parallelism across clips; a time-parallel scan would change the rounding.  The 256 most recent
//     outputs of each lane live in a circular shared-memory row.
//   phase 2 (spectrogram, 4 lanes per frame): rows are taken 8 at a time; 4 lanes share one frame:
//     mean removal, Tukey window, 256-point real FFT as a 128-point complex FFT split 32 (register
//     FFT per lane) x 4 (two shuffle butterflies across the lane quad), untangling with the partner
//     lane, PSD scaling (compute_spectrogram, classifier.c:448-592).
//     filter 1 (3000-7500 Hz): PSD cells go to a per-CTA scratch [clip][t][bin] (L2/HBM) and a running
//     min/max of the positive cells is kept; filter 0 (1000-3000 Hz): only "any cell above the dB
//     threshold" per frame is kept (find_midpoints, classifier.c:657-727), tested against the exact
//     PSD value whose dB conversion first exceeds the threshold (no logarithm in the loop).
//   epilogue: blob clustering -> midpoints (classifier.c:748-800), then the three band sums around each
//     midpoint (sum_intense, :594-655) over dB-normalised cells recomputed from the scratch with the
//     reference's double-precision log10, and the verdict (:184).
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

#include "donut_kernel.h"
#include "sj_pk.h"
#include "fft_gen.inc"

namespace {

constexpr unsigned kFull = 0xffffffffu;
// warps per CTA.  Measured (clips/s, float / double): 1 warp 3.17 M / 1.01 M, 2: 3.93 M / 1.59 M, 3: 4.58 M / 1.65 M,
// 4: 4.44 M / 2.57 M -- the float variant deals 16 row passes (3 warps leave room for 4 CTAs per SM), the double
// variant 4 passes (one per warp).
template <typename T> struct DnWarps { static constexpr int value = sizeof(T) == 8 ? 4 : 3; };
constexpr int kRow = 264; // row stride in elements: = 8 (mod 32) floats / = 8 (mod 16) doubles => conflict-free 64/128-bit reads

template <typename T> struct V2;
template <> struct V2<float> { typedef float2 type; };
template <> struct V2<double> { typedef double2 type; };

// bit-pattern views for the threshold searches (positive IEEE values are ordered like their bit patterns)
template <typename T> struct Bits;
template <> struct Bits<float> {
    typedef uint32_t type;
    static __device__ __forceinline__ uint32_t of(float x) { return __float_as_uint(x); }
    static __device__ __forceinline__ float from(uint32_t b) { return __uint_as_float(b); }
    static __device__ __forceinline__ uint32_t shfl(uint32_t b, int src) { return __shfl_sync(0xffffffffu, b, src); }
};
template <> struct Bits<double> {
    typedef unsigned long long type;
    static __device__ __forceinline__ unsigned long long of(double x) { return (unsigned long long)__double_as_longlong(x); }
    static __device__ __forceinline__ double from(unsigned long long b) { return __longlong_as_double((long long)b); }
    static __device__ __forceinline__ unsigned long long shfl(unsigned long long b, int src) { return __shfl_sync(0xffffffffu, b, src); }
};

__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub_rn(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float div_rn(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double sub_rn(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double div_rn(double a, double b) { return __ddiv_rn(a, b); }

__device__ __forceinline__ void fft32(double2 (&v)[32]) { sj_fft32d(v); }

// 4- / 8-byte asynchronous copy global -> shared (LDGSTS): the band-sum epilogue stages a whole box of PSD cells with
// one round trip to L2 / DRAM instead of one dependent load per 32 cells
__device__ __forceinline__ void cp_async_elem(float *dst, const float *src)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_elem(double *dst, const double *src)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// dB value exactly as the reference forms it: (T)(10 * log10((double)s / 1e-12))
template <typename T> __device__ __forceinline__ T to_db(T s) { return (T)(10.0 * log10((double)s / 1e-12)); }

template <typename T> __host__ __device__ constexpr size_t smem_bytes_t()
{
    return sizeof(SjDonutTables<T>) + sizeof(T) * 32 * kRow + sizeof(T) * DN_CLIPS_PER_WARP * DN_MAX_MID + 64;
}

} // namespace

namespace {

// eight steps of the direct-form-II recurrence with the state in a rotating register file:
// before step s the value w[n-j] sits in w[(s - j) & 7]
template <typename T>
__device__ __forceinline__ void iir8(const T (&x)[8], T (&w)[8], const T (&a)[9], const T (&b)[9], T (&y)[8])
{
#pragma unroll
    for (int s = 0; s < 8; ++s) {
        T w0 = x[s];
#pragma unroll
        for (int j = 1; j < 9; ++j) w0 = sub_rn(w0, mul_rn(a[j], w[(s - j + 8) & 7]));
        T yy = mul_rn(b[0], w0);
#pragma unroll
        for (int j = 2; j < 9; j += 2) yy = add_rn(yy, mul_rn(b[j], w[(s - j + 8) & 7])); // b[odd] == 0
        w[s] = w0;
        y[s] = yy;
    }
}

// float: for even j the products a[j] w and b[j] w of the same delayed value come from ONE packed multiply
// (mul.rn.f32x2 rounds each half like the scalar instruction; the sums stay scalar and ordered, see bfly) -- four
// issue slots fewer per sample on a warp whose only other limit is the dependent chain
__device__ __forceinline__ void iir8(const float (&x)[8], float (&w)[8], const float (&a)[9], const float (&b)[9], float (&y)[8])
{
#pragma unroll
    for (int s = 0; s < 8; ++s) {
        float2 pe[4];
#pragma unroll
        for (int j = 2; j < 9; j += 2) pe[j / 2 - 1] = pk_mul(pk_bc(w[(s - j + 8) & 7]), make_float2(a[j], b[j]));
        float w0 = x[s];
#pragma unroll
        for (int j = 1; j < 9; ++j) w0 = sub_rn(w0, (j & 1) ? mul_rn(a[j], w[(s - j + 8) & 7]) : pe[j / 2 - 1].x);
        float yy = mul_rn(b[0], w0);
#pragma unroll
        for (int j = 2; j < 9; j += 2) yy = add_rn(yy, pe[j / 2 - 1].y); // b[odd] == 0
        w[s] = w0;
        y[s] = yy;
    }
}

template <typename T> struct RowState {
    T min_in[32], max_in[32];   // min / max of a = re^2 + im^2 over the cells with Sxx > 0: inner bins 1..127 (doubled)
    T min_ed[32], max_ed[32];   // ... and the edge bins 0 and 128
    T queue[DnWarps<T>::value][64];      // band-sum epilogue: a values of the cells that passed the level test (per warp)
    uint32_t flags[32][3];
    int n_mid[DN_CLIPS_PER_WARP];
    T mids[DN_CLIPS_PER_WARP][DN_MAX_MID];
    T mean[32];                 // per-row frame mean, summed sequentially like the reference
    float2 xbuf[DnWarps<T>::value][2][16 * 17]; // 16x16 transpose buffers of the exact float FFT (two frames per pass, per warp)
};

__host__ __device__ constexpr int brev4(int x) { return ((x & 1) << 3) | ((x & 2) << 1) | ((x & 4) >> 1) | ((x & 8) >> 3); }

// one butterfly of the reference's float FFT, operation for operation (PlainFFT.cpp:64-71)
// (packed fp32 pairs: mul.rn.f32x2 / add.rn.f32x2 round each half exactly like the scalar operations and are never
// contracted, so the ten roundings of the reference's butterfly are reproduced in five issue slots)
__device__ __forceinline__ void bfly(float2 &lo, float2 &hi, float u1, float u2)
{
    const float2 m1 = pk_mul(pk_bc(u1), hi);                      // (u1 hi.x, u1 hi.y)
    const float2 m2 = pk_mul(pk_bc(u2), make_float2(hi.y, hi.x)); // (u2 hi.y, u2 hi.x)
    // (scalar here: ptxas contracts mul.rn.f32x2 + add.rn.f32x2 into FFMA2, which would drop two roundings)
    const float2 t = make_float2(sub_rn(m1.x, m2.x), add_rn(m1.y, m2.y));
    hi = pk_sub(lo, t);
    lo = pk_add(lo, t);
}

// Float variant: spectrogram column t of all 32 rows with the reference's exact arithmetic
// (compute_spectrogram, sync/lib/classifier.cpp:307-366 + PlainFFT::Compute).  The butterfly network
// and its twiddles are data-independent, so running the same butterflies in a different order and in
// parallel gives bit-identical results.  16 lanes per frame (2 frames per pass, 16 passes): index
// i = 16*h + r; stages 0-3 act on r inside a lane (lane = h), a 16x16 transpose through shared memory,
// stages 4-7 act on h inside a lane (lane = r).
__device__ __forceinline__ void spectrogram_column_exact_f32(const float *rows, int pos, int t, int tb,
                                                             const SjDonutTables<float> *tab, RowState<float> *st,
                                                             float *scratch, int lane, int warp, int nact, bool pipe)
{
    const int half = lane >> 4, h = lane & 15;
    float2 *xb = st->xbuf[warp][half];
    const float a_thr_e = tab->a_thr[0], a_thr_i = tab->a_thr[1], a_pos = tab->a_pos;
    // pipe (at most one pass per warp, none on warp 0): the CTA barrier that frees the rows for the next hop sits
    // right after this warp has its frame in registers, so that warp 0 filters hop t+1 while the frame is transformed
    bool freed = false;
    for (int p = pipe ? warp - 1 : warp; p >= 0 && p < nact; p += pipe ? DnWarps<float>::value - 1 : DnWarps<float>::value) { // pass p = the two rows (filters) of clip p
        const int r = 2 * p + half;
        const float *rowp = rows + r * kRow;
        const float mean = st->mean[r];
        float2 v[16];
        // bit-reversed load: position i = 16h + rr holds sample brev8(i) = 16*brev4(rr) + brev4(h)
        const int hb = brev4(h);
#pragma unroll
        for (int rr = 0; rr < 16; ++rr) {
            const int n = 16 * brev4(rr) + hb;
            const float x = rowp[(pos + n) & 255];
            v[rr] = make_float2(mul_rn(sub_rn(x, mean), tab->window[n]), 0.f);
        }
        if (pipe) { __syncthreads(); freed = true; }
        // stages 0 and 1 act on purely real data with the twiddles (1, 0) and (0, -1) (exact in the reference's
        // recurrence): products with 0 and 1 and sums with 0 are exact, so the ten operations of a butterfly
        // collapse to the two that change a value (the sign of a zero is the only thing lost; it squares away)
#pragma unroll
        for (int i = 0; i < 16; i += 2) {
            const float t1 = v[i + 1].x;
            v[i + 1].x = sub_rn(v[i].x, t1);
            v[i].x = add_rn(v[i].x, t1);
        }
#pragma unroll
        for (int i = 0; i < 16; i += 4) {
            const float t1 = v[i + 2].x; // twiddle (1, 0)
            v[i + 2].x = sub_rn(v[i].x, t1);
            v[i].x = add_rn(v[i].x, t1);
            const float t2 = v[i + 3].x; // twiddle (0, -1): t = (hi.y, -hi.x) = (0, -t2)
            v[i + 3] = make_float2(v[i + 1].x, t2);
            v[i + 1].y = -t2;
        }
#pragma unroll
        for (int l = 2; l < 4; ++l) {
            const int l1 = 1 << l;
#pragma unroll
            for (int i = 0; i < 16; ++i)
                if ((i & l1) == 0) bfly(v[i], v[i + l1], tab->bourke[l1 - 1 + (i & (l1 - 1))][0], tab->bourke[l1 - 1 + (i & (l1 - 1))][1]);
        }
#pragma unroll
        for (int rr = 0; rr < 16; ++rr) xb[h * 17 + rr] = v[rr];
        __syncwarp();
#pragma unroll
        for (int hh = 0; hh < 16; ++hh) v[hh] = xb[hh * 17 + h]; // now lane = r (low index bits), register = h
        __syncwarp();
#pragma unroll
        for (int l = 4; l < 8; ++l) {
            const int l1h = 1 << (l - 4), l1 = 16 << (l - 4);
#pragma unroll
            for (int i = 0; i < 16; ++i)
                if ((i & l1h) == 0) {
                    const int m = (i & (l1h - 1)) * 16 + h;
                    bfly(v[i], v[i + l1h], tab->bourke[l1 - 1 + m][0], tab->bourke[l1 - 1 + m][1]);
                }
        }
        // PSD of bins k = 16*hh + lane16 <= 128 (classifier.cpp:349-364) and the per-row bookkeeping
        const bool is_band = r & 1;
        float *cell = scratch + (size_t(r >> 1) * tb + t) * DN_BINS;
        float lmin = INFINITY, lmax = 0.f, emin = INFINITY, emax = 0.f;
        float imax = 0.f, xmax = 0.f; // flag rows: largest a over the inner / the edge cells (any(a >= thr) = max(a) >= thr)
        // a = re^2 + im^2 with the reference's operations; "/ U", "* 2" (classifier.cpp:349-364) are applied
        // later, only where a value is needed; the tests on Sxx use exact thresholds on a.
        // The two half-warps of a pass hold a flag row and a band row: both kinds of bookkeeping are done by every lane
        // (selects, no branches -- a branch on the row kind diverges in every one of the nine rounds) and each row
        // keeps what it needs.
#pragma unroll
        for (int hh = 0; hh < 9; ++hh) {
            const int k = 16 * hh + h;
            if (hh < 8 || h == 0) {
                const float av = add_rn(mul_rn(v[hh].x, v[hh].x), mul_rn(v[hh].y, v[hh].y));
                if (is_band) cell[k] = av;
                const bool pos = av >= a_pos;
                const float vmin = pos ? av : INFINITY, vmax = pos ? av : 0.f;
                if (hh == 0 || hh == 8) { // bins 0 and 128 (lane h == 0) are the edge cells
                    const bool edge = h == 0;
                    emin = fminf(emin, edge ? vmin : INFINITY);
                    emax = fmaxf(emax, edge ? vmax : 0.f);
                    xmax = fmaxf(xmax, edge ? av : 0.f);
                    lmin = fminf(lmin, edge ? INFINITY : vmin);
                    lmax = fmaxf(lmax, edge ? 0.f : vmax);
                    imax = fmaxf(imax, edge ? 0.f : av);
                } else {
                    lmin = fminf(lmin, vmin);
                    lmax = fmaxf(lmax, vmax);
                    imax = fmaxf(imax, av);
                }
            }
        }
        const bool any = !is_band && (imax >= a_thr_i || xmax >= a_thr_e);
#pragma unroll
        for (int o = 1; o < 16; o <<= 1) {
            lmin = fminf(lmin, __shfl_xor_sync(kFull, lmin, o));
            lmax = fmaxf(lmax, __shfl_xor_sync(kFull, lmax, o));
        }
        const unsigned ballot = __ballot_sync(kFull, any);
        if (h == 0) {
            if (is_band) {
                st->min_in[r] = fminf(st->min_in[r], lmin);
                st->max_in[r] = fmaxf(st->max_in[r], lmax);
                st->min_ed[r] = fminf(st->min_ed[r], emin); // edge bins live in lane h == 0 only
                st->max_ed[r] = fmaxf(st->max_ed[r], emax);
            } else if ((ballot >> (16 * half)) & 0xFFFFu) st->flags[r][t >> 5] |= 1u << (t & 31);
        }
    }
    if (pipe && !freed) __syncthreads(); // warp 0 and warps without a pass
}

template <typename T, bool ALIGNED, bool PIPE>
__global__ void __launch_bounds__(DnWarps<T>::value * 32) donut_kernel(const SjDonutArgs a)
{
    typedef typename V2<T>::type T2;
    extern __shared__ __align__(16) unsigned char smem[];
    SjDonutTables<T> *tab = reinterpret_cast<SjDonutTables<T> *>(smem);
    constexpr size_t kTabBytes = (sizeof(SjDonutTables<T>) + 15) & ~size_t(15);
    T *rows = reinterpret_cast<T *>(smem + kTabBytes);
    RowState<T> *st = reinterpret_cast<RowState<T> *>(rows + 32 * kRow);
    constexpr int kDnWarps = DnWarps<T>::value;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    {
        const uint32_t *src = reinterpret_cast<const uint32_t *>(a.tables);
        uint32_t *dst = reinterpret_cast<uint32_t *>(tab);
        for (int i = threadIdx.x; i < int(sizeof(SjDonutTables<T>) / 4); i += kDnWarps * 32) dst[i] = __ldg(src + i);
    }
    __syncthreads();

    const int f = lane & 1, ci = lane >> 1, tb = a.tb, n_samples = a.n_samples;
    T ca[9], cb[9];
#pragma unroll
    for (int j = 0; j < 9; ++j) { ca[j] = tab->a[f][j]; cb[j] = tab->b[f][j]; }
    T *myrow = rows + lane * kRow;
    T *scratch = reinterpret_cast<T *>(a.scratch) + size_t(blockIdx.x) * DN_CLIPS_PER_WARP * tb * DN_BINS;
    const int q = lane >> 2, g = lane & 3;
    const int k2 = ((g & 1) << 1) | (g >> 1); // which quarter of the 128-point spectrum this lane ends up holding
    const T U = tab->U, a_thr_e = tab->a_thr[0], a_thr_i = tab->a_thr[1], a_pos = tab->a_pos;

    // a CTA owns `cpc` clips at a time: 16 for large batches; small batches are spread over the resident CTAs (down to
    // one clip per CTA), because a hop costs the same recurrence time however many lanes carry a clip, and the
    // spectrogram passes of absent clips are skipped -- the latency of a small batch is that of one clip
    const int cpc = a.clips_per_cta;
    // small groups (at most one spectrogram pass per warp with warp 0 left out): the recurrence of hop t+1 runs on
    // warp 0 while the other warps transform frame t (see spectrogram_column_exact_f32)
    constexpr bool pipe = PIPE; // chosen by sj_launch_donut: float cpc <= warps - 1, double cpc <= 4 (warps - 1)
    for (long base = long(blockIdx.x) * cpc; base < a.n_clips; base += long(gridDim.x) * cpc) {
        const int nact = a.n_clips - base < cpc ? (int)(a.n_clips - base) : cpc; // clips of this group
        long clip = base + ci;
        if (ci >= nact) clip = base + nact - 1; // idle lanes shadow the group's last clip; nothing is written for them
        // clip -> samples: contiguous clips, or second k of channel r of a stream array (sync/sync.cpp:190-203
        // classifies every full second of the microphone stream from zero state)
        const float *src = a.clips_per_row > 0
                               ? a.clips + (clip / a.clips_per_row) * a.row_stride + (clip % a.clips_per_row) * (long)n_samples
                               : a.clips + clip * (long)n_samples;
        T w[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) w[j] = 0;
        if (warp == 0) {
            st->min_in[lane] = st->min_ed[lane] = (T)INFINITY;
            st->max_in[lane] = st->max_ed[lane] = 0;
            st->flags[lane][0] = st->flags[lane][1] = st->flags[lane][2] = 0;
        }
        int pos = 0, sample = 0;
        // software pipeline of the input: a register ring of kPre blocks of 8 samples, each block requested
        // kPre blocks (~1 us of recurrence) before it is filtered -- one warp per CTA and a serial
        // recurrence leave nothing else to hide the DRAM latency
        constexpr int kPre = 4;
        float4 ring[kPre][2];
#pragma unroll
        for (int j = 0; j < kPre; ++j) {
            ring[j][0] = ring[j][1] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (ALIGNED && warp == 0) { // tb >= 1 guarantees 256 samples
                ring[j][0] = __ldg(reinterpret_cast<const float4 *>(src + 8 * j));
                ring[j][1] = __ldg(reinterpret_cast<const float4 *>(src + 8 * j + 4));
            }
        }
        T acc_cur = 0, acc_next = 0; // running sums of the current frame and of the next one (32-sample overlap)
        __syncthreads();

        for (int t = 0; t < tb; ++t) {
            // ---- phase 1 (warp 0): advance each lane's IIR by one hop (256 samples for the first frame)
            const int nnew = t == 0 ? DN_WIN : DN_HOP;
            if (warp == 0)
            for (int i0 = 0; i0 < nnew; i0 += 8 * kPre) { // 224 and 256 are multiples of 32
#pragma unroll
                for (int j = 0; j < kPre; ++j) {
                    const int i = i0 + 8 * j;
                    T x[8], y[8];
                    if (ALIGNED) {
                        const float4 u0 = ring[j][0], u1 = ring[j][1];
                        x[0] = u0.x; x[1] = u0.y; x[2] = u0.z; x[3] = u0.w; x[4] = u1.x; x[5] = u1.y; x[6] = u1.z; x[7] = u1.w;
                        if (sample + i + 8 * kPre + 8 <= n_samples) {
                            ring[j][0] = __ldg(reinterpret_cast<const float4 *>(src + sample + i + 8 * kPre));
                            ring[j][1] = __ldg(reinterpret_cast<const float4 *>(src + sample + i + 8 * kPre + 4));
                        }
                    } else {
#pragma unroll
                        for (int k = 0; k < 8; ++k) x[k] = __ldg(src + sample + i + k);
                    }
                    iir8(x, w, ca, cb, y); // float: the packed overload
                    // frame mean: "sum += vReal[i]" in sample order (classifier.cpp:320-324), kept exact
                    const bool tail = i >= nnew - 32; // the last 32 samples of a frame open the next frame
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        acc_cur = add_rn(acc_cur, y[k]);
                        if (tail) acc_next = add_rn(acc_next, y[k]);
                    }
                    T *dst = myrow + ((pos + i) & 255);
#pragma unroll
                    for (int k = 0; k < 8; k += 2) *reinterpret_cast<T2 *>(dst + k) = T2{y[k], y[k + 1]};
                }
            }
            sample += nnew;
            pos = (pos + nnew) & 255; // = offset of the oldest of the 256 most recent samples = frame start
            if (warp == 0) st->mean[lane] = div_rn(acc_cur, (T)DN_WIN);
            acc_cur = acc_next;
            acc_next = 0;
            __syncthreads();

            // ---- phase 2 (all warps): spectrogram column t of all 32 rows, the row passes dealt round-robin
            if constexpr (sizeof(T) == 4) {
                spectrogram_column_exact_f32(rows, pos, t, tb, tab, st, scratch, lane, warp, nact, pipe);
            } else
            // double: 8 rows per pass, 4 lanes per row (no bit-exact FFT to follow: FFTW's order is unknown)
            {
            bool freed = false;
            // pass bq = the rows of clips 4 bq .. 4 bq + 3: four passes at most and four warps, so a warp has one pass or none
            static_assert(sizeof(T) == 4 || kDnWarps * 4 >= DN_CLIPS_PER_WARP, "one pass per warp");
            // (large groups run all four passes: the test on the group size would cost this 255-register kernel a spill)
            if (const int bq = pipe ? warp - 1 : warp; !pipe || (bq >= 0 && 4 * bq < nact)) {
                const int r = 8 * bq + q;
                const T *rowp = rows + r * kRow;
                T2 v[32];
                const T mean = st->mean[r];
#pragma unroll
                for (int n = 0; n < 32; ++n) {
                    v[n] = *reinterpret_cast<const T2 *>(rowp + ((pos + 8 * n + 2 * g) & 255));
                    const T2 wv = *reinterpret_cast<const T2 *>(&tab->window[8 * n + 2 * g]);
                    v[n].x = mul_rn(sub_rn(v[n].x, mean), wv.x);
                    v[n].y = mul_rn(sub_rn(v[n].y, mean), wv.y);
                }
                if (pipe) { __syncthreads(); freed = true; }
                fft32(v); // over n: Y_g[k1] in v[SJ_BREV32(k1)]
#pragma unroll
                for (int k1 = 1; k1 < 32; ++k1) {
                    const T wr = tab->tw128[k1][g][0], wi = tab->tw128[k1][g][1];
                    T2 &y = v[SJ_BREV32(k1)];
                    const T yr = y.x * wr - y.y * wi;
                    y.y = y.x * wi + y.y * wr;
                    y.x = yr;
                }
                // 4-point DFT across the lane quad (g = n mod 4): afterwards this lane holds Z[k1 + 32*k2]
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    T ox = __shfl_xor_sync(kFull, v[i].x, 2), oy = __shfl_xor_sync(kFull, v[i].y, 2);
                    if (g & 2) { v[i].x = ox - v[i].x; v[i].y = oy - v[i].y; } else { v[i].x += ox; v[i].y += oy; }
                    if (g == 3) { const T tx = v[i].y; v[i].y = -v[i].x; v[i].x = tx; } // times -i
                    ox = __shfl_xor_sync(kFull, v[i].x, 1); oy = __shfl_xor_sync(kFull, v[i].y, 1);
                    if (g & 1) { v[i].x = ox - v[i].x; v[i].y = oy - v[i].y; } else { v[i].x += ox; v[i].y += oy; }
                }
                // untangle the real FFT, PSD scaling, and per-row bookkeeping
                const bool is_band = r & 1; // filter 1 = 3000-7500 Hz: cells are kept; filter 0: midpoint flags
                T *cell = scratch + (size_t(r >> 1) * tb + t) * DN_BINS;
                T lmin = (T)INFINITY, lmax = 0, emin = (T)INFINITY, emax = 0;
                bool any = false;
                auto emit = [&](int bin, T x2, T y2) { // (x2, y2) = 2 X[bin]
                    const T re = (T)0.5 * x2, im = (T)0.5 * y2;
                    const T av = add_rn(mul_rn(re, re), mul_rn(im, im)); // Sxx = av / U (* 2 for inner bins), applied later
                    const bool edge = bin == 0 || bin == DN_BINS - 1;
                    if (is_band) {
                        cell[bin] = av;
                        if (av >= a_pos) {
                            if (edge) { emin = av < emin ? av : emin; emax = av > emax ? av : emax; }
                            else { lmin = av < lmin ? av : lmin; lmax = av > lmax ? av : lmax; }
                        }
                    } else any = any || (av >= (edge ? a_thr_e : a_thr_i));
                };
                auto pair = [&](T2 A, T bx, T by, int k, T &xa, T &ya, T &xb, T &yb) {
                    // A = Z[k], (bx, by) = Z[128 - k]; returns 2 X[k] and 2 conj-partner for bin 128 - k
                    const T ex = A.x + bx, ey = A.y - by, dx = A.x - bx, dy = A.y + by;
                    const T wr = tab->tw256[k][0], wi = tab->tw256[k][1];
                    const T tx = wr * dy + wi * dx, ty = wi * dy - wr * dx;
                    xa = ex + tx; ya = ey + ty; xb = ex - tx; yb = ey - ty;
                };
#pragma unroll
                for (int k1 = 1; k1 < 16; ++k1) {
                    const T2 S = v[SJ_BREV32(32 - k1)];
                    const T bx = __shfl_xor_sync(kFull, S.x, 3), by = __shfl_xor_sync(kFull, S.y, 3);
                    const int k = k1 + 32 * k2;
                    T xa, ya, xb, yb;
                    pair(v[SJ_BREV32(k1)], bx, by, k, xa, ya, xb, yb);
                    emit(k, xa, ya);
                    emit(DN_WIN / 2 - k, xb, yb);
                }
                {   // k1 = 16: partner Z[16 + 32 (3 - k2)] sits in the mirror lane, same register
                    const T2 S = v[SJ_BREV32(16)];
                    const T bx = __shfl_xor_sync(kFull, S.x, 3), by = __shfl_xor_sync(kFull, S.y, 3);
                    T xa, ya, xb, yb;
                    pair(S, bx, by, 16 + 32 * k2, xa, ya, xb, yb);
                    emit(16 + 32 * k2, xa, ya);
                }
                {   // k1 = 0: bins 0, 32, 64, 96 (and 128 from the k2 = 0 lane)
                    const T2 S = v[0];
                    T bx = __shfl_xor_sync(kFull, S.x, 1), by = __shfl_xor_sync(kFull, S.y, 1);
                    if (g < 2) { bx = S.x; by = S.y; } // k2 = 0 and k2 = 2 are their own partners
                    T xa, ya, xb, yb;
                    pair(S, bx, by, 32 * k2, xa, ya, xb, yb);
                    emit(32 * k2, xa, ya);
                    if (k2 == 0) emit(DN_WIN / 2, xb, yb);
                }
                // combine the quad's partial results; the quad leader updates the row state
                lmin = fmin(lmin, __shfl_xor_sync(kFull, lmin, 1)); lmin = fmin(lmin, __shfl_xor_sync(kFull, lmin, 2));
                lmax = fmax(lmax, __shfl_xor_sync(kFull, lmax, 1)); lmax = fmax(lmax, __shfl_xor_sync(kFull, lmax, 2));
                emin = fmin(emin, __shfl_xor_sync(kFull, emin, 1)); emin = fmin(emin, __shfl_xor_sync(kFull, emin, 2));
                emax = fmax(emax, __shfl_xor_sync(kFull, emax, 1)); emax = fmax(emax, __shfl_xor_sync(kFull, emax, 2));
                const unsigned ballot = __ballot_sync(kFull, any);
                if (g == 0) {
                    if (is_band) {
                        st->min_in[r] = fmin(st->min_in[r], lmin);
                        st->max_in[r] = fmax(st->max_in[r], lmax);
                        st->min_ed[r] = fmin(st->min_ed[r], emin);
                        st->max_ed[r] = fmax(st->max_ed[r], emax);
                    } else if ((ballot >> (4 * q)) & 0xFu) st->flags[r][t >> 5] |= 1u << (t & 31);
                }
            }
            if (pipe && !freed) __syncthreads(); // warp 0 and warps without a pass
            }
            if (!pipe) __syncthreads(); // the rows are overwritten by the next hop; cells and row state become visible to all warps
        }
        if (pipe) __syncthreads(); // the last frame's cells and row state

        // ---- epilogue A: blob clustering -> midpoints, one lane per clip (find_midpoints, classifier.c:729-800)
        const T fs = tab->fs;
        if (warp == 0 && lane < nact) {
            const uint32_t *fl = st->flags[2 * lane];
            const T tol = tab->gap_tol, mind = tab->min_dur;
            int count = 0, size = 0;
            T first = 0, last = 0, tsum = 0;
            for (int j = 0; j <= tb; ++j) {
                const bool on = j < tb && ((fl[j >> 5] >> (j & 31)) & 1u);
                const T tj = div_rn((T)(j * DN_HOP + DN_WIN / 2), fs);
                const bool joins = on && size > 0 && sub_rn(tj, last) <= tol;
                if (size > 0 && (!on ? j == tb : !joins)) {
                    // close the running cluster (also reached at j == tb)
                    if (sub_rn(last, first) >= mind) {
                        if (count < DN_MAX_MID) st->mids[lane][count] = div_rn(tsum, (T)size);
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
            st->n_mid[lane] = count;
        }
        __syncthreads();

        // ---- epilogue B: band sums around each midpoint and the verdict, the warp cooperating per clip.
        // A cell counts when its normalised dB level nv lies strictly inside (lo_norm, hi_norm)
        // (classifier.c:128-156).  nv is a monotone function of a, so the two level tests become
        // a_lo <= a < a_hi with per-clip thresholds found by a 32-way search over bit patterns that evaluates
        // the reference's own expression (double log10 included); the logarithm is then only taken for the
        // cells that pass, compacted through a small queue so that all 32 lanes work on passing cells.
        for (int c2 = warp; c2 < nact; c2 += kDnWarps) { // clips dealt round-robin to the warps
            const int nm = st->n_mid[c2];
            int label = 0;
            T s_above = 0, s_mid = 0, s_below = 0;
            if (nm > 0) {
                const int rb = 2 * c2 + 1;
                T smin = (T)INFINITY, smax = 0;
                if (st->max_in[rb] > 0) { smin = mul_rn(div_rn(st->min_in[rb], U), (T)2); smax = mul_rn(div_rn(st->max_in[rb], U), (T)2); }
                if (st->max_ed[rb] > 0) {
                    const T e0 = div_rn(st->min_ed[rb], U), e1 = div_rn(st->max_ed[rb], U);
                    smin = e0 < smin ? e0 : smin;
                    smax = e1 > smax ? e1 : smax;
                }
                const T mn = to_db<T>(smin), mx = to_db<T>(smax);
                const T range = sub_rn(mx, mn), lo = tab->lo_norm, hi = tab->hi_norm;
                auto level = [&](T av) -> T { // normalised dB level of an inner-bin cell, the reference's operations
                    const T sv = mul_rn(div_rn(av, U), (T)2);
                    return div_rn(sub_rn(to_db<T>(sv), mn), range);
                };
                // smallest a in [a_pos, max] with pred true (pred monotone); "none" = +inf
                auto first_true = [&](auto pred) -> T {
                    typedef typename Bits<T>::type UI;
                    UI blo = Bits<T>::of(a_pos), bhi = Bits<T>::of(st->max_in[rb]);
                    if (!(st->max_in[rb] >= a_pos) || !pred(st->max_in[rb])) return (T)INFINITY;
                    while (blo < bhi) {
                        const UI step = (bhi - blo) / 32 + 1;
                        UI cand = blo + step * (UI)(lane + 1) - 1;
                        if (cand > bhi || cand < blo) cand = bhi;
                        const unsigned bal = __ballot_sync(kFull, pred(Bits<T>::from(cand)));
                        const int first = __ffs(bal) - 1; // lane 31 holds bhi, where pred is true
                        const UI at = Bits<T>::shfl(cand, first), before = Bits<T>::shfl(cand, first > 0 ? first - 1 : 0);
                        bhi = at;
                        if (first > 0) blo = before + 1;
                    }
                    return Bits<T>::from(bhi);
                };
                T a_lo = (T)INFINITY, a_hi = (T)INFINITY;
                if (range > 0) { // range == 0 makes every level NaN/inf in the reference: nothing passes
                    a_lo = first_true([&](T av) { return level(av) > lo; });
                    a_hi = first_true([&](T av) { return !(level(av) < hi); });
                }
                const T *scr = scratch + size_t(c2) * tb * DN_BINS;
                T *qu = st->queue[warp];
                // staging area of the box cells: this warp's transpose buffer (idle during the epilogue)
                T *stage = reinterpret_cast<T *>(st->xbuf[warp]);
                constexpr int kStage = (int)(sizeof(st->xbuf[0]) / sizeof(T)) / 32 * 32;
                auto box = [&](int band, T half, T midpoint) -> T {
                    const int f0 = tab->box_f[band][0], f1 = tab->box_f[band][1];
                    // time index search exactly as sum_intense (classifier.c:619-640), evaluated lane-parallel:
                    // times are increasing, so the loops count the frames before / after the window
                    const T tlo = sub_rn(midpoint, half), thi = add_rn(midpoint, half);
                    int nb = 0, na = 0;
                    for (int j = lane; j < DN_MAX_TB; j += 32) {
                        const T tj = div_rn((T)(j * DN_HOP + DN_WIN / 2), fs);
                        nb += __popc(__ballot_sync(kFull, j < tb && tj < tlo));
                        na += __popc(__ballot_sync(kFull, j < tb && tj > thi));
                    }
                    int t0 = nb, t1 = tb - 1 - na;
                    if (t0 >= tb) t0 = tb - 1;
                    if (t1 < 0) t1 = 0;
                    if (t0 > t1) { const int x = t0; t0 = t1; t1 = x; }
                    double acc = 0.0;
                    const int nf = f1 - f0 + 1, ncell = nf * (t1 - t0 + 1);
                    int nq = 0; // queued cells (warp-uniform)
                    // cell c = c0 + lane sits at (frame j, bin i) = (c / nf, c % nf): one division per box, then the pair
                    // advances by 32 cells per round.  The whole box (<= 41 x 27 cells) is first staged in shared memory
                    // by asynchronous copies -- one round trip to L2 for the box instead of one per 32 cells (the dependent
                    // loads of this loop held 20 % of the transform warps' time, profiles/r2c) -- in chunks of kStage cells.
                    const int q32 = 32 / nf, r32 = 32 - q32 * nf;
                    int cj = lane / nf, ci = lane - cj * nf;
                    for (int c0 = 0; c0 < ncell; c0 += 32) {
                        if (c0 % kStage == 0) {
                            __syncwarp();
                            int sj = cj, si = ci;
                            for (int c = c0 + lane; c < ncell && c < c0 + kStage; c += 32) {
                                cp_async_elem(stage + (c - c0), scr + (t0 + sj) * DN_BINS + f0 + si);
                                si += r32;
                                sj += q32;
                                if (si >= nf) { si -= nf; ++sj; }
                            }
                            cp_async_wait_all();
                            __syncwarp();
                        }
                        const int c = c0 + lane;
                        const T av = c < ncell ? stage[c % kStage] : (T)0;
                        ci += r32;
                        cj += q32;
                        if (ci >= nf) { ci -= nf; ++cj; }
                        const bool pass = c < ncell && av >= a_lo && av < a_hi;
                        const unsigned bal = __ballot_sync(kFull, pass);
                        if (pass) qu[nq + __popc(bal & ((1u << lane) - 1u))] = av;
                        nq += __popc(bal);
                        __syncwarp();
                        if (nq >= 32) {
                            acc += (double)level(qu[lane]);
                            nq -= 32;
                            __syncwarp();
                            if (lane < nq) qu[lane] = qu[32 + lane];
                            __syncwarp();
                        }
                    }
                    if (lane < nq) acc += (double)level(qu[lane]);
                    __syncwarp();
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(kFull, acc, o);
                    return (T)acc;
                };
                const int cap = nm < DN_MAX_MID ? nm : DN_MAX_MID;
                for (int m = 0; m < cap; ++m) {
                    const T mp = st->mids[c2][m];
                    s_above = box(0, (T)0.18, mp);
                    s_mid = box(1, (T)0.05, mp);
                    s_below = box(2, (T)0.18, mp);
                    if (s_mid < tab->thr_mid && s_above > tab->thr_above && s_below > tab->thr_below) { label = 1; break; }
                }
            }
            if (lane == 0) {
                if (a.labels) a.labels[base + c2] = label;
                if (a.track) a.track[base + c2] = (int8_t)(label + 1); // imu_window[].classification, sync/sync.cpp:202-203
                if (a.detail) {
                    float *d = a.detail + (base + c2) * 4;
                    d[0] = (float)nm; d[1] = (float)s_above; d[2] = (float)s_mid; d[3] = (float)s_below;
                }
            }
            if (a.mids && lane < DN_MAX_MID) // find_midpoints' return array (classifier.cpp:433-598), NaN-padded
                a.mids[(base + c2) * DN_MAX_MID + lane] = lane < nm ? (float)st->mids[c2][lane] : __int_as_float(0x7fc00000);
        }
        __syncthreads();
    }
}


// ================================================================================================
// Warp-specialised kernels (round 2), float and double.  One persistent CTA of 16 warps per SM in three roles -- kIir
// "filter" warps (4 float / 3 double), 8 "transform" warps, 4 "epilogue" warps -- coupled per clip group by two named
// barriers ("full", "empty") and two shared-memory counters; no CTA-wide barrier in the steady state.  The roles live in
// separate warpgroups and trade registers with setmaxnreg (filter ~100, transform 128-144, epilogue what is left).
//   * A filter warp owns a group of 16 clips = 32 (clip, filter) rows, one direct-form-II recurrence per
//     lane in the reference's exact operation order (iir8: bit-identical samples, classifier.cpp:193-219).
//     One filter warp sits on each of the SM's four schedulers: the recurrence is a dependent chain (4 cycles per
//     FP32, 8 per FP64 instruction), so a step costs its scheduler ~50 (float) / ~75 (double) cycles whatever it does;
//     independent groups in flight are what fills the pipes, and the rest of each pipe goes to the other roles.
//     (A first version with both filters of a clip packed in one lane -- 13 cross-filter FMUL2 + two scalar chains --
//     needed 107 cycles per step on two warps and lost to the round-1 kernel: 4.85 M against 5.8 M clips/s.)
//   * The warp writes one spectrogram hop (224 new samples per row) into circular 256-sample rows
//     (row = filter * 16 + clip, stride 258 elements: the warp's column stores and the row reads of the passes are
//     bank-conflict-free), signals "full", and waits for "empty" before the next hop.  A transform warp arrives at
//     "empty" as soon as its last pass of the frame has its rows in registers, so the next hop is filtered while the
//     frame is transformed; the groups drift out of phase, so the transform warps always have a frame to work on.
//   * The transform warps compute the spectrogram column of the group's 32 rows: float = PlainFFT's butterfly network
//     replayed exactly, 2 rows per pass (16 lanes per row, the arithmetic of spectrogram_column_exact_f32), a pass
//     holding two rows of the SAME kind (flag rows or band rows) so that the per-cell bookkeeping is specialised per
//     kind; double = 4 rows per pass, 8 lanes per row (pass_f64).  Warp w takes the passes w, w + 8, ...
//   * After a group's last frame the epilogue warps cluster the flagged frames into midpoints, sum the three boxes around
//     each and write the verdict, from the group's copy of the clip state and PSD scratch (two copies per group slot),
//     while the filter and transform warps are already busy with the slot's next group.  With the epilogue on the
//     transform warps every group end stalled the whole pipeline: 26 % of the float kernel's time for 8 % of its
//     instructions.
//   * The three code streams run at the same time and must share the 32 KB instruction cache: the double pass is one
//     body for both row kinds, level() (double log10) is out of line, the filter loop handles 16 samples per trip behind
//     an L2 prefetch.  (Two copies of a 50 KB pass body cost 3.3 "no instruction" stall cycles per issued instruction.)
namespace ws {

#ifndef SJ_DONUT_IIR_WARPS
#define SJ_DONUT_IIR_WARPS 4
#endif
// double: a group's 32 rows of doubles take 66 KB of shared memory, so three groups are in flight
#ifndef SJ_DONUT_IIR_WARPS_F64
#define SJ_DONUT_IIR_WARPS_F64 3
#endif
// registers per thread of the transform / filter / epilogue warpgroups (multiples of 8; 8 T + 4 F + 4 E <= 2048)
#ifndef SJ_DONUT_REG_T_F32
#define SJ_DONUT_REG_T_F32 144
#define SJ_DONUT_REG_F_F32 104
#define SJ_DONUT_REG_E_F32 120
#endif
#ifndef SJ_DONUT_REG_T_F64
#define SJ_DONUT_REG_T_F64 144
#define SJ_DONUT_REG_F_F64 136
#define SJ_DONUT_REG_E_F64 80
#endif
constexpr int kGroup = 16;         // clips per group (32 rows: lane = filter * 16 + clip)
constexpr int kRS = 258;           // row stride in elements
// Warp roles, by warpgroup (setmaxnreg moves registers between whole warpgroups): warps 0-7 transform, 8-11 epilogue,
// 12-15 filter (the double kernel leaves warp 15 idle).  The kernel is compiled for 512 threads = 128 registers; the
// warpgroups then grow or shrink to the Cfg<T>::kReg* of their role (8 T + 4 F + 4 E <= 2048 = 64 K registers / 32 lanes).
constexpr int kFft = 8, kEpi = 4, kFilterWarp0 = kFft + kEpi, kThreads = 512;
template <typename T> struct Cfg;
template <> struct Cfg<float> {
    static constexpr int kIir = SJ_DONUT_IIR_WARPS; // filter warps = clip groups in flight per CTA
    static constexpr int kPasses = 2 * kGroup / 2;  // passes per group frame: two rows each (16 lanes per row)
    static constexpr int kStage = 512;              // epilogue staging area (cells) per epilogue warp
    static constexpr int kRegT = SJ_DONUT_REG_T_F32, kRegF = SJ_DONUT_REG_F_F32, kRegE = SJ_DONUT_REG_E_F32;
};
template <> struct Cfg<double> {
    static constexpr int kIir = SJ_DONUT_IIR_WARPS_F64;
    static constexpr int kPasses = 2 * kGroup / 4;  // four rows each (8 lanes per row)
    static constexpr int kStage = 96;
    static constexpr int kRegT = SJ_DONUT_REG_T_F64, kRegF = SJ_DONUT_REG_F_F64, kRegE = SJ_DONUT_REG_E_F64;
};
static_assert(8 * Cfg<float>::kRegT + 4 * Cfg<float>::kRegE + 4 * Cfg<float>::kRegF <= 2048, "register file");
static_assert(8 * Cfg<double>::kRegT + 4 * Cfg<double>::kRegE + 4 * Cfg<double>::kRegF <= 2048, "register file");
static_assert(Cfg<float>::kIir <= 4 && Cfg<double>::kIir <= 4, "one warpgroup of filter warps");
// named barriers (0 = __syncthreads): full[g] = 1 + g, the filter warp arrives and the transform warps wait; empty[g] = 5 + g,
// every transform warp arrives once per frame (when its last pass of the frame has its rows in registers) and the filter
// warp waits.  (An mbarrier polled with naps did the same at equal speed but 15 % more issued instructions.)
constexpr int kBarFull = 1, kBarEmpty = 5;
constexpr int kPair = 32 + 32 * kFft; // threads of a full / empty barrier: one filter warp + the transform warps

// per-clip results of a group's frames, read by the epilogue warps after the last frame: two copies per group slot, so
// that the next group of the slot is filtered and transformed while the epilogue of the previous one runs
template <typename T> struct ClipState {
    T min_in[kGroup], max_in[kGroup], min_ed[kGroup], max_ed[kGroup]; // band rows: min / max of a over the cells with Sxx > 0
    uint32_t flags[kGroup][3];                                // flag rows: frames with any cell above the dB threshold
    int n_mid[kGroup];
    T mids[kGroup][DN_MAX_MID];
};
template <typename T> struct GroupState {
    T mean[2 * kGroup];            // per row: frame mean, summed sequentially like the reference
    ClipState<T> cs[2];
    int frames_done;               // transform warps that have finished the last frame of a group, summed over the slot's groups
    int epilogues_done;            // epilogue warps that have finished a group, summed over the slot's groups
};
template <typename T> struct FftWarpState;
template <> struct FftWarpState<float> {
    float2 xbuf[2][16 * 17]; // 16x16 transpose buffers of the exact float FFT (two frames per pass)
};
template <> struct FftWarpState<double> { // nothing: the double pass transposes through the rows
    double pad[2];
};
template <typename T> struct EpiWarpState {
    T queue[64];                 // band-sum epilogue: a values of the cells that passed the level test
    T stage[Cfg<T>::kStage];     // ... and the staging area of a box's cells
};

__device__ __forceinline__ void bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void bar_arrive(int id, int n) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory"); }
template <typename T> constexpr size_t kTabBytesT = (sizeof(SjDonutTables<T>) + 15) & ~size_t(15);
template <typename T>
constexpr size_t kSmemT = kTabBytesT<T> + sizeof(T) * Cfg<T>::kIir * 2 * kGroup * kRS + sizeof(GroupState<T>) * Cfg<T>::kIir +
                          sizeof(FftWarpState<T>) * kFft + sizeof(EpiWarpState<T>) * kEpi + sizeof(T) * DN_MAX_TB;
static_assert(kSmemT<float> <= 232448 && kSmemT<double> <= 232448, "shared memory per CTA");

// normalised dB level of an inner-bin cell with the reference's operations.  Out of line on purpose: the double-precision
// log10 is ~150 instructions, and five inlined copies made the epilogue too large to share the 32 KB instruction cache
// with the filter loop and the passes that run at the same time.
template <typename T> __device__ __noinline__ T level_of(T av, T U, T mn, T range)
{
    const T sv = mul_rn(div_rn(av, U), (T)2);
    return div_rn(sub_rn(to_db<T>(sv), mn), range);
}

// volatile poll of a shared-memory counter (the hand-overs to and from the epilogue warps: rare, never on the frame path)
#ifndef SJ_DONUT_EPI_NAP_NS
#define SJ_DONUT_EPI_NAP_NS 1000
#endif
__device__ __forceinline__ void wait_count(const int *ctr, int at_least)
{
    while (*reinterpret_cast<const volatile int *>(ctr) < at_least) __nanosleep(SJ_DONUT_EPI_NAP_NS);
    __threadfence_block();
}
template <int N> __device__ __forceinline__ void reg_alloc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N> __device__ __forceinline__ void reg_dealloc() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }

// one pass = spectrogram column t of two rows of one kind (BAND: 3000-7500 Hz rows whose cells are kept; otherwise the
// 1000-3000 Hz rows that only flag frames), the clips c0 and c0 + 8 of the group; PlainFFT replayed as in
// spectrogram_column_exact_f32
// (BAND as a run-time flag -- one copy of the FFT code for both row kinds, 9 KB less in the instruction cache -- was measured:
// +2 % with the long filter loop, -1.5 % with the short one that is shipped.)
template <bool BAND>
__device__ __forceinline__ void pass_exact_f32(const float *rows, int pos, int t, int tb, const SjDonutTables<float> *tab,
                                               const float *mean_tab, ClipState<float> *st, FftWarpState<float> *fw, float *scratch,
                                               int lane, int c0, const float2 (&tw)[15], const float (&win)[16], int release)
{
    const int half = lane >> 4, h = lane & 15;
    const int clip = c0 + 8 * half, r = (BAND ? kGroup : 0) + clip;
    float2 *xb = fw->xbuf[half];
    const float *rowp = rows + r * kRS;
    const float mean = mean_tab[r];
    float2 v[16];
    const int hb = brev4(h);
#pragma unroll
    for (int rr = 0; rr < 16; ++rr) {
        const int n = 16 * brev4(rr) + hb;
        const float x = rowp[(pos + n) & 255];
        v[rr] = make_float2(mul_rn(sub_rn(x, mean), win[rr]), 0.f); // win[rr] = tab->window[n], held in registers
    }
    // the two rows are in registers: if this is the warp's last pass of the frame, the filter warp may overwrite the
    // group's rows with the next hop as far as this warp is concerned
    if (release) bar_arrive(release, kPair);
#pragma unroll
    for (int i = 0; i < 16; i += 2) {
        const float t1 = v[i + 1].x;
        v[i + 1].x = sub_rn(v[i].x, t1);
        v[i].x = add_rn(v[i].x, t1);
    }
#pragma unroll
    for (int i = 0; i < 16; i += 4) {
        const float t1 = v[i + 2].x; // twiddle (1, 0)
        v[i + 2].x = sub_rn(v[i].x, t1);
        v[i].x = add_rn(v[i].x, t1);
        const float t2 = v[i + 3].x; // twiddle (0, -1): t = (hi.y, -hi.x) = (0, -t2)
        v[i + 3] = make_float2(v[i + 1].x, t2);
        v[i + 1].y = -t2;
    }
#pragma unroll
    for (int l = 2; l < 4; ++l) {
        const int l1 = 1 << l;
#pragma unroll
        for (int i = 0; i < 16; ++i)
            if ((i & l1) == 0) bfly(v[i], v[i + l1], tab->bourke[l1 - 1 + (i & (l1 - 1))][0], tab->bourke[l1 - 1 + (i & (l1 - 1))][1]);
    }
#pragma unroll
    for (int rr = 0; rr < 16; ++rr) xb[h * 17 + rr] = v[rr];
    __syncwarp();
#pragma unroll
    for (int hh = 0; hh < 16; ++hh) v[hh] = xb[hh * 17 + h];
    __syncwarp();
#pragma unroll
    for (int l = 4; l < 8; ++l) {
        const int l1h = 1 << (l - 4);
#pragma unroll
        for (int i = 0; i < 16; ++i)
            if ((i & l1h) == 0) {
                // twiddle (l1 - 1) + (i & (l1h - 1)) * 16 + h of PlainFFT's table: the same 15 pairs for every pass of this
                // lane, held in registers for the whole kernel (they were 32 per-lane shared-memory loads per pass)
                const float2 u = tw[(l1h - 1) + (i & (l1h - 1))];
                bfly(v[i], v[i + l1h], u.x, u.y);
            }
    }
    // a = re^2 + im^2 of bins k = 16 hh + h <= 128 with the reference's operations (classifier.cpp:349-364); "/ U" and "* 2"
    // are applied later, only where a value is needed; the tests on Sxx use exact thresholds on a
    if (BAND) {
        float *cell = scratch + (size_t(clip) * tb + t) * DN_BINS;
        const float a_pos = tab->a_pos;
        float lmin = INFINITY, lmax = 0.f, emin = INFINITY, emax = 0.f;
#pragma unroll
        for (int hh = 0; hh < 9; ++hh) {
            if (hh < 8 || h == 0) {
                const float av = add_rn(mul_rn(v[hh].x, v[hh].x), mul_rn(v[hh].y, v[hh].y));
                cell[16 * hh + h] = av;
                const bool posv = av >= a_pos;
                const float vmin = posv ? av : INFINITY, vmax = posv ? av : 0.f;
                if (hh == 0 || hh == 8) { // bins 0 and 128 (lane h == 0) are the edge cells
                    const bool edge = h == 0;
                    emin = fminf(emin, edge ? vmin : INFINITY);
                    emax = fmaxf(emax, edge ? vmax : 0.f);
                    lmin = fminf(lmin, edge ? INFINITY : vmin);
                    lmax = fmaxf(lmax, edge ? 0.f : vmax);
                } else {
                    lmin = fminf(lmin, vmin);
                    lmax = fmaxf(lmax, vmax);
                }
            }
        }
#pragma unroll
        for (int o = 1; o < 16; o <<= 1) {
            lmin = fminf(lmin, __shfl_xor_sync(kFull, lmin, o));
            lmax = fmaxf(lmax, __shfl_xor_sync(kFull, lmax, o));
        }
        if (h == 0) {
            st->min_in[clip] = fminf(st->min_in[clip], lmin);
            st->max_in[clip] = fmaxf(st->max_in[clip], lmax);
            st->min_ed[clip] = fminf(st->min_ed[clip], emin); // the edge bins live in lane h == 0 only
            st->max_ed[clip] = fmaxf(st->max_ed[clip], emax);
        }
    } else {
        float imax = 0.f, xmax = 0.f; // largest a over the inner / the edge cells (any(a >= thr) = max(a) >= thr)
#pragma unroll
        for (int hh = 0; hh < 9; ++hh) {
            if (hh < 8 || h == 0) {
                const float av = add_rn(mul_rn(v[hh].x, v[hh].x), mul_rn(v[hh].y, v[hh].y));
                if (hh == 0 || hh == 8) {
                    const bool edge = h == 0;
                    xmax = fmaxf(xmax, edge ? av : 0.f);
                    imax = fmaxf(imax, edge ? 0.f : av);
                } else imax = fmaxf(imax, av);
            }
        }
        const bool any = imax >= tab->a_thr[1] || xmax >= tab->a_thr[0];
        const unsigned ballot = __ballot_sync(kFull, any);
        if (h == 0 && ((ballot >> (16 * half)) & 0xFFFFu)) st->flags[clip][t >> 5] |= 1u << (t & 31);
    }
}

// double: one pass = spectrogram column t of four rows of one kind, 8 lanes per row (no bit-exact FFT to follow: FFTW's
// order is unknown).  The 256-point real FFT is a 128-point complex FFT of z[n] = x[2n] + i x[2n+1], split 16 x 8:
//   lane l of a row holds z[l + 8m], m = 0..15: FFT16 over m in registers -> Y_l[k1]; times W128^(l k1);
//   an 8 x 16 transpose (in two halves of 8 x 8) -> lane j holds T_l[j] and T_l[j + 8] for l = 0..7; two FFT8 over l ->
//   Z[k], k = j + 8c, c = 0..15, in lane j; the real-FFT untangling pairs Z[k] with Z[128 - k] = element 15 - c of lane 8 - j.
// The transposes go through the part of the row's own circular buffer that the next hop overwrites anyway (224 of the
// 256 samples; the transform warps own the rows until they signal "empty"): no shared memory besides the rows.
// A pass is ~900 instructions (14 KB of code) against the ~3100 (50 KB) of the round-1 scheme (4 lanes per row, FFT32 per
// lane, two shuffle butterflies across the lane quad), which did not fit the 32 KB instruction cache: 3.3 "no instruction"
// stall cycles per issued instruction with warps at different places of the pass (profiles/r2l_*).
// (BAND is a run-time flag, not a template parameter as in the float pass: one copy of the code.)
__device__ __forceinline__ void pass_f64(const bool BAND, double *rows, int pos, int t, int tb, const SjDonutTables<double> *tab,
                                         const double *mean_tab, ClipState<double> *st, double *scratch, int lane, int c0, int release)
{
    typedef double T;
    typedef double2 T2;
    const int q = lane >> 3, l = lane & 7;
    const int clip = c0 + q, r = (BAND ? kGroup : 0) + clip;
    T *rowp = rows + r * kRS;
    T2 v[16];
    const T mean = mean_tab[r];
#pragma unroll
    for (int m = 0; m < 16; ++m) { // the 8 lanes of a row read 8 consecutive 16-byte units: conflict-free
        v[m] = *reinterpret_cast<const T2 *>(rowp + ((pos + 2 * l + 16 * m) & 255));
        const T2 wv = *reinterpret_cast<const T2 *>(&tab->window[2 * l + 16 * m]);
        v[m].x = mul_rn(sub_rn(v[m].x, mean), wv.x);
        v[m].y = mul_rn(sub_rn(v[m].y, mean), wv.y);
    }
    __syncwarp(); // every lane has its samples: the row's first 224 positions are free from here on
    sj_fft16d(v); // over m: Y_l[k1] in v[SJ_BREV16(k1)]
#pragma unroll
    for (int k1 = 1; k1 < 16; ++k1) { // W128^(l k1) = W256^(2 e) for e = l k1 < 64, minus that of e - 64 otherwise
        const int e = l * k1;
        const T2 w0 = *reinterpret_cast<const T2 *>(tab->tw256[2 * (e & 63)]);
        const int flip = (e & 64) << 25; // sign bit of the high word
        const T wr = __hiloint2double(__double2hiint(w0.x) ^ flip, __double2loint(w0.x));
        const T wi = __hiloint2double(__double2hiint(w0.y) ^ flip, __double2loint(w0.y));
        T2 &y = v[SJ_BREV16(k1)];
        const T yr = y.x * wr - y.y * wi;
        y.y = y.x * wi + y.y * wr;
        y.x = yr;
    }
    // transpose, half by half: element (l, k1) at unit 9 l + (k1 & 7) of the row's free part (16-byte units; the odd
    // stride keeps the column writes of the 8 lanes and their row reads on distinct bank groups)
    T2 w[16]; // w[8 h + l'] = T_l'[j + 8 h]
    const int sbase = pos + 18 * l, lbase = pos + 2 * l;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
#pragma unroll
        for (int k = 0; k < 8; ++k) *reinterpret_cast<T2 *>(rowp + ((sbase + 2 * k) & 255)) = v[SJ_BREV16(8 * h + k)];
        __syncwarp();
#pragma unroll
        for (int k = 0; k < 8; ++k) w[8 * h + k] = *reinterpret_cast<const T2 *>(rowp + ((lbase + 18 * k) & 255));
        __syncwarp();
    }
    if (release) bar_arrive(release, kPair); // done with the rows (the warp's last pass of the frame): the next hop may be written
    {
        T2(&u0)[8] = *reinterpret_cast<T2(*)[8]>(&w[0]);
        T2(&u1)[8] = *reinterpret_cast<T2(*)[8]>(&w[8]);
        sj_fft8d(u0); // over l: Z[j + 16 k2] in w[SJ_BREV8(k2)]
        sj_fft8d(u1); // Z[j + 8 + 16 k2] in w[8 + SJ_BREV8(k2)]
    }
    // Z[j + 8 c] = w[ZI(c)]
#define SJ_ZI(c) ((((c) & 1) << 3) + SJ_BREV8((c) >> 1))
    const T a_thr_e = tab->a_thr[0], a_thr_i = tab->a_thr[1], a_pos = tab->a_pos;
    T *cell = scratch + (size_t(clip) * tb + t) * DN_BINS;
    T lmin = (T)INFINITY, lmax = 0, emin = (T)INFINITY, emax = 0;
    bool any = false;
    auto emit = [&](int bin, bool edge, T x2, T y2) { // (x2, y2) = 2 X[bin]; a = re^2 + im^2 ("/ U", "* 2" applied later)
        const T re = (T)0.5 * x2, im = (T)0.5 * y2;
        const T av = add_rn(mul_rn(re, re), mul_rn(im, im));
        if (BAND) {
            cell[bin] = av;
            if (av >= a_pos) {
                if (edge) { emin = av < emin ? av : emin; emax = av > emax ? av : emax; }
                else { lmin = av < lmin ? av : lmin; lmax = av > lmax ? av : lmax; }
            }
        } else any = any || (av >= (edge ? a_thr_e : a_thr_i));
    };
    auto pair = [&](T2 A, T bx, T by, int k, T &xa, T &ya, T &xb, T &yb) {
        // A = Z[k], (bx, by) = Z[128 - k]; returns 2 X[k] and twice the conjugate partner for bin 128 - k
        const T ex = A.x + bx, ey = A.y - by, dx = A.x - bx, dy = A.y + by;
        const T2 tw = *reinterpret_cast<const T2 *>(tab->tw256[k]);
        const T tx = tw.x * dy + tw.y * dx, ty = tw.y * dy - tw.x * dx;
        xa = ex + tx; ya = ey + ty; xb = ex - tx; yb = ey - ty;
    };
    const int src = (lane & 24) | ((8 - l) & 7); // the lane that holds Z[128 - k] for this lane's k (lane 0 and 4: itself)
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        // partner element: 15 - c of lane 8 - j; lane 0 pairs with itself, element (16 - c) mod 16
        const T2 S = w[SJ_ZI(15 - c)], O = w[SJ_ZI((16 - c) & 15)];
        T bx = __shfl_sync(kFull, S.x, src), by = __shfl_sync(kFull, S.y, src);
        if (l == 0) { bx = O.x; by = O.y; }
        const int k = l + 8 * c;
        T xa, ya, xb, yb;
        pair(w[SJ_ZI(c)], bx, by, k, xa, ya, xb, yb);
        emit(k, c == 0 && l == 0, xa, ya);           // bin 0 is an edge cell
        emit(DN_WIN / 2 - k, c == 0 && l == 0, xb, yb); // ... and so is bin 128 = 128 - 0
    }
    if (l == 0) { // bin 64 pairs with itself
        const T2 A = w[SJ_ZI(8)];
        T xa, ya, xb, yb;
        pair(A, A.x, A.y, 64, xa, ya, xb, yb);
        emit(64, false, xa, ya);
    }
#undef SJ_ZI
    if (BAND) { // combine the partial results of the row's 8 lanes; lane 0 of the row (which alone saw the edge cells) updates the clip's state
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) {
            lmin = fmin(lmin, __shfl_xor_sync(kFull, lmin, o));
            lmax = fmax(lmax, __shfl_xor_sync(kFull, lmax, o));
        }
        if (l == 0) {
            st->min_in[clip] = fmin(st->min_in[clip], lmin);
            st->max_in[clip] = fmax(st->max_in[clip], lmax);
            st->min_ed[clip] = fmin(st->min_ed[clip], emin);
            st->max_ed[clip] = fmax(st->max_ed[clip], emax);
        }
    } else {
        const unsigned ballot = __ballot_sync(kFull, any);
        if (l == 0 && ((ballot >> (8 * q)) & 0xFFu)) st->flags[clip][t >> 5] |= 1u << (t & 31);
    }
}

template <typename T, bool ALIGNED>
__global__ void __launch_bounds__(kThreads, 1) donut_ws_kernel(const SjDonutArgs a)
{
    typedef typename V2<T>::type T2;
    constexpr int kIir = Cfg<T>::kIir, kPasses = Cfg<T>::kPasses;
    extern __shared__ __align__(16) unsigned char smem[];
    SjDonutTables<T> *tab = reinterpret_cast<SjDonutTables<T> *>(smem);
    T *rows_all = reinterpret_cast<T *>(smem + kTabBytesT<T>);
    GroupState<T> *gs_all = reinterpret_cast<GroupState<T> *>(rows_all + kIir * 2 * kGroup * kRS);
    FftWarpState<T> *fw_all = reinterpret_cast<FftWarpState<T> *>(gs_all + kIir);
    EpiWarpState<T> *ew_all = reinterpret_cast<EpiWarpState<T> *>(fw_all + kFft);
    T *frame_time = reinterpret_cast<T *>(ew_all + kEpi); // centre time of frame j with the reference's operations (classifier.c:586-590)
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    {
        const uint32_t *src = reinterpret_cast<const uint32_t *>(a.tables);
        uint32_t *dst = reinterpret_cast<uint32_t *>(tab);
        for (int i = threadIdx.x; i < int(sizeof(SjDonutTables<T>) / 4); i += kThreads) dst[i] = __ldg(src + i);
        if (threadIdx.x < kIir) gs_all[threadIdx.x].frames_done = gs_all[threadIdx.x].epilogues_done = 0;
    }
    __syncthreads();
    if (threadIdx.x < DN_MAX_TB) frame_time[threadIdx.x] = div_rn((T)(threadIdx.x * DN_HOP + DN_WIN / 2), tab->fs);
    __syncthreads();

    const int tb = a.tb, n_samples = a.n_samples, cpg = a.clips_per_cta; // clips per group: 16, fewer for small batches
    const long n_groups = (a.n_clips + cpg - 1) / cpg;
    static_assert(kPasses >= kFft && kPasses % kFft == 0 && kIir <= 4, "every transform warp has the same number of passes in every frame; barrier ids");
    // scratch of group slot g, buffer b: [16][tb][129] PSD cells of the band rows
    auto scratch_of = [&](int g, int b) {
        return reinterpret_cast<T *>(a.scratch) + ((size_t(blockIdx.x) * kIir + g) * 2 + b) * kGroup * tb * DN_BINS;
    };

    if (warp >= kFilterWarp0) {
        // ------------------------------------------------------------------ filter warp: groups q = g, g + kIir, ... of this CTA
        // (the highest warp ids: the SM's warp arbiter favours them, and the recurrence is the latency-critical chain)
        if constexpr (Cfg<T>::kRegF < 128) reg_dealloc<Cfg<T>::kRegF>(); else if constexpr (Cfg<T>::kRegF > 128) reg_alloc<Cfg<T>::kRegF>();
        const int g = warp - kFilterWarp0;
        if (g >= kIir) return;
        T *rows = rows_all + g * 2 * kGroup * kRS;
        GroupState<T> *st = gs_all + g;
        const int f = lane >> 4, ci = lane & 15; // row = lane = filter * 16 + clip
        T ca[9], cb[9];
#pragma unroll
        for (int j = 0; j < 9; ++j) { ca[j] = tab->a[f][j]; cb[j] = tab->b[f][j]; }
        T *myrow = rows + lane * kRS;
        for (long r = 0;; ++r) {
            const long gid = blockIdx.x + (r * kIir + g) * gridDim.x;
            if (gid >= n_groups) break;
            const long base = gid * cpg;
            const int nact = a.n_clips - base < cpg ? (int)(a.n_clips - base) : cpg;
            const long clip = base + (ci < nact ? ci : nact - 1); // idle lanes shadow the group's last clip
            const float *src = a.clips_per_row > 0
                                   ? a.clips + (clip / a.clips_per_row) * a.row_stride + (clip % a.clips_per_row) * (long)n_samples
                                   : a.clips + clip * (long)n_samples;
            T w[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) w[j] = 0;
            // the slot's state copy r & 1 was last used by group r - 2: its epilogue must be over
            if (r >= 2) wait_count(&st->epilogues_done, kEpi * (int)(r - 1));
            if (lane < kGroup) {
                ClipState<T> *cs = &st->cs[r & 1];
                cs->min_in[lane] = cs->min_ed[lane] = (T)INFINITY;
                cs->max_in[lane] = cs->max_ed[lane] = 0;
                cs->flags[lane][0] = cs->flags[lane][1] = cs->flags[lane][2] = 0;
            }
// input samples: a register ring of kPre blocks of 8 samples, requested kPre blocks ahead, behind an L2 prefetch of one
// 128-byte line per lane SJ_DONUT_L2_AHEAD samples ahead.  (Measured, float / double: ring of 4 alone 7.05 / 3.81 M clips/s;
// ring of 2 + L2 prefetch 256 ahead 7.82 / 4.50 M -- the loop body is half as long, and the filter loop, the passes and the
// epilogue must share the 32 KB instruction cache; ring of 1: 7.53 / 4.22 M; 128 / 512 ahead: 7.79 / 7.74 M.)
#ifndef SJ_DONUT_PRE
#define SJ_DONUT_PRE 2
#endif
#ifndef SJ_DONUT_L2_AHEAD
#define SJ_DONUT_L2_AHEAD 256
#endif
            constexpr int kPre = SJ_DONUT_PRE; // register ring of kPre blocks of 8 samples, requested kPre blocks ahead
            float4 ring[kPre][2];
#pragma unroll
            for (int j = 0; j < kPre; ++j) {
                ring[j][0] = ring[j][1] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (ALIGNED) { // tb >= 1 guarantees 256 samples
                    ring[j][0] = __ldg(reinterpret_cast<const float4 *>(src + 8 * j));
                    ring[j][1] = __ldg(reinterpret_cast<const float4 *>(src + 8 * j + 4));
                }
            }
            T acc_cur = 0, acc_next = 0;
            int pos = 0, sample = 0;
            for (int t = 0; t < tb; ++t) {
                if (t > 0) bar_sync(kBarEmpty + g, kPair); // every pass of frame t - 1 has read its rows
                const int nnew = t == 0 ? DN_WIN : DN_HOP;
#ifdef SJ_DONUT_SKIP_FILTER /* timing experiment only (results are wrong): the filter warps run the barrier protocol alone */
                if (t == 0) for (int i = 0; i < 256; ++i) myrow[i] = (T)0.001 * (T)((i * 37 + lane * 11) & 63);
                for (int i0 = 0; i0 < 0; i0 += 8 * kPre) {
#else
#pragma unroll 1
                for (int i0 = 0; i0 < nnew; i0 += 8 * kPre) {
#endif
#if SJ_DONUT_L2_AHEAD > 0
                    if (ALIGNED && (i0 & 31) == 0 && sample + i0 + SJ_DONUT_L2_AHEAD < n_samples)
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(src + sample + i0 + SJ_DONUT_L2_AHEAD));
#endif
#pragma unroll
                    for (int j = 0; j < kPre; ++j) {
                        const int i = i0 + 8 * j;
                        T x[8], y[8];
                        if (ALIGNED) {
                            const float4 u0 = ring[j][0], u1 = ring[j][1];
                            x[0] = u0.x; x[1] = u0.y; x[2] = u0.z; x[3] = u0.w; x[4] = u1.x; x[5] = u1.y; x[6] = u1.z; x[7] = u1.w;
                            if (sample + i + 8 * kPre + 8 <= n_samples) {
                                ring[j][0] = __ldg(reinterpret_cast<const float4 *>(src + sample + i + 8 * kPre));
                                ring[j][1] = __ldg(reinterpret_cast<const float4 *>(src + sample + i + 8 * kPre + 4));
                            }
                        } else {
#pragma unroll
                            for (int k = 0; k < 8; ++k) x[k] = __ldg(src + sample + i + k);
                        }
#ifdef SJ_DONUT_SKIP_IIR /* timing experiment only: results are wrong */
#pragma unroll
                        for (int k = 0; k < 8; ++k) y[k] = x[k];
#else
                        iir8(x, w, ca, cb, y); // float: the packed overload, (a[j] w, b[j] w) from one FMUL2 for even j
#endif
                        // frame mean: "sum += vReal[i]" in sample order (classifier.cpp:320-324), kept exact
                        const bool tail = i >= nnew - 32; // the last 32 samples of a frame open the next frame
#pragma unroll
                        for (int k = 0; k < 8; ++k) {
                            acc_cur = add_rn(acc_cur, y[k]);
                            if (tail) acc_next = add_rn(acc_next, y[k]);
                        }
                        T *dst = myrow + ((pos + i) & 255);
#pragma unroll
                        for (int k = 0; k < 8; k += 2) *reinterpret_cast<T2 *>(dst + k) = T2{y[k], y[k + 1]};
                    }
                }
                sample += nnew;
                pos = (pos + nnew) & 255;
                st->mean[lane] = div_rn(acc_cur, (T)DN_WIN);
                acc_cur = acc_next;
                acc_next = 0;
                __threadfence_block();
                bar_arrive(kBarFull + g, kPair);
            }
            bar_sync(kBarEmpty + g, kPair); // every pass of the last frame has read its rows: they are free for the next group
        }
    } else if (warp < kFft) {
        // ------------------------------------------------------------------ transform warps
        if constexpr (Cfg<T>::kRegT > 128) reg_alloc<Cfg<T>::kRegT>();
        const int w = warp;
        FftWarpState<T> *fw = fw_all + w;
        float2 tw[15]; // float: this lane's twiddles of FFT stages 4..7 (lane & 15 = the low index digit after the transpose)
        if constexpr (sizeof(T) == 4) {
            const int hh = lane & 15;
            auto twid = [&](int l1, int q) { const int m = q * 16 + hh; return make_float2(tab->bourke[l1 - 1 + m][0], tab->bourke[l1 - 1 + m][1]); };
            tw[0] = twid(16, 0);
            tw[1] = twid(32, 0); tw[2] = twid(32, 1);
            tw[3] = twid(64, 0); tw[4] = twid(64, 1); tw[5] = twid(64, 2); tw[6] = twid(64, 3);
            tw[7] = twid(128, 0); tw[8] = twid(128, 1); tw[9] = twid(128, 2); tw[10] = twid(128, 3);
            tw[11] = twid(128, 4); tw[12] = twid(128, 5); tw[13] = twid(128, 6); tw[14] = twid(128, 7);
        }
        // float: the window values of this lane's 16 samples of a row, the same for every pass (16 shared-memory loads
        // less per pass: 7.99 M against 7.85 M clips/s; the 12 twiddle pairs of stages 2 and 3 in registers as well: 7.7-7.8 M)
        float win[16];
        if constexpr (sizeof(T) == 4) {
#pragma unroll
            for (int rr = 0; rr < 16; ++rr) win[rr] = tab->window[16 * brev4(rr) + brev4(lane & 15)];
        }
        for (long r = 0;; ++r) {
            if (blockIdx.x + r * kIir * gridDim.x >= n_groups) break;
            for (int t = 0; t < tb; ++t) {
                const int pos = (DN_HOP * t) & 255; // offset of the frame's first sample in the circular rows
#pragma unroll 1
                for (int g = 0; g < kIir; ++g) {
                    const long gid = blockIdx.x + (r * kIir + g) * gridDim.x;
                    if (gid >= n_groups) break;
                    const long base = gid * cpg;
                    const int nact = a.n_clips - base < cpg ? (int)(a.n_clips - base) : cpg;
                    T *rows = rows_all + g * 2 * kGroup * kRS;
                    GroupState<T> *st = gs_all + g;
                    ClipState<T> *cs = &st->cs[r & 1];
                    T *scratch = scratch_of(g, (int)(r & 1));
                    bar_sync(kBarFull + g, kPair);
                    // float: 16 passes, p < 8 the flag rows, p >= 8 the band rows; pass j of a kind holds the clips j, j + 8.
                    // double: 8 passes, p < 4 the flag rows; pass j of a kind holds the clips 4 j .. 4 j + 3.
                    // Warp w takes one pass per round k of kFft passes: (w + 1 + 4 k) mod kFft of round k.  (A shared-memory
                    // ticket counter -- atomicAdd, later a compare-and-swap that never claims beyond the frame -- cost 6 % of
                    // the float kernel's throughput.)  The offsets matter for small batches only, where a group holds few clips:
                    // + 1 keeps the passes of a group's first clip off the scheduler of the group's filter warp (warp 12 + g),
                    // + 4 k gives the flag pass and the band pass of a clip to different warps, so that they run side by side
                    // and each warp reports "rows read" right after its loads (one clip: 645 -> 583 us; round-1 kernel 536 us).
                    auto pass_of = [&](int k) { return (w + 1 + 4 * k) % kFft + kFft * k; };
                    auto clip0_of = [&](int p) { return sizeof(T) == 4 ? (p & 7) : 4 * (p & 3); };
                    // the warp's last pass with clips in it: the filter warp is told "rows read" after that pass's loads (at
                    // once if the warp has nothing to do in this frame)
                    int last_k = -1;
#pragma unroll
                    for (int k = 0; k < kPasses / kFft; ++k)
                        if (clip0_of(pass_of(k)) < nact) last_k = k;
#ifdef SJ_DONUT_SKIP_FFT /* timing experiment only: results are wrong */
                    last_k = -1;
#endif
                    if (last_k < 0) bar_arrive(kBarEmpty + g, kPair);
#pragma unroll
                    for (int k = 0; k < kPasses / kFft; ++k) {
                        const int p = pass_of(k), c0 = clip0_of(p);
                        if (c0 >= nact || last_k < 0) continue;
                        const int release = k == last_k ? kBarEmpty + g : 0;
                        if constexpr (sizeof(T) == 4) {
                            if (p < 8) pass_exact_f32<false>(rows, pos, t, tb, tab, st->mean, cs, fw, scratch, lane, c0, tw, win, release);
                            else pass_exact_f32<true>(rows, pos, t, tb, tab, st->mean, cs, fw, scratch, lane, c0, tw, win, release);
                        } else {
                            pass_f64(p >= 4, rows, pos, t, tb, tab, st->mean, cs, scratch, lane, c0, release);
                        }
                    }
                    if (t == tb - 1) { // the group's last frame: hand its clips to the epilogue warps
                        __syncwarp();
                        if (lane == 0) { __threadfence_block(); atomicAdd(&st->frames_done, 1); }
                    }
                }
            }
        }
    } else {
        // ------------------------------------------------------------------ epilogue warps: clustering, band sums, verdict
        if constexpr (Cfg<T>::kRegE < 128) reg_dealloc<Cfg<T>::kRegE>(); else if constexpr (Cfg<T>::kRegE > 128) reg_alloc<Cfg<T>::kRegE>();
        const int e = warp - kFft;
        EpiWarpState<T> *ew = ew_all + e;
        const T U = tab->U, a_pos = tab->a_pos, fs = tab->fs;
        for (long r = 0;; ++r) {
            if (blockIdx.x + r * kIir * gridDim.x >= n_groups) break;
#pragma unroll 1
            for (int g = 0; g < kIir; ++g) {
                const long gid = blockIdx.x + (r * kIir + g) * gridDim.x;
                if (gid >= n_groups) break;
                const long base = gid * cpg;
                const int nact = a.n_clips - base < cpg ? (int)(a.n_clips - base) : cpg;
                GroupState<T> *gs = gs_all + g;
                ClipState<T> *st = &gs->cs[r & 1];
                const T *scratch = scratch_of(g, (int)(r & 1));
                wait_count(&gs->frames_done, kFft * (int)(r + 1)); // every pass of the group's last frame is done
                // epilogue A: blob clustering -> midpoints, one lane per clip (find_midpoints, classifier.c:729-800)
                {
                    const int c = e + kEpi * lane;
                    if (c < nact) {
                        const uint32_t *fl = st->flags[c];
                        const T tol = tab->gap_tol, mind = tab->min_dur;
                        int count = 0, size = 0;
                        T first = 0, last = 0, tsum = 0;
                        for (int j = 0; j <= tb; ++j) {
                            const bool on = j < tb && ((fl[j >> 5] >> (j & 31)) & 1u);
                            const T tj = frame_time[j < DN_MAX_TB ? j : DN_MAX_TB - 1]; // (j == tb only closes the last cluster)
                            const bool joins = on && size > 0 && sub_rn(tj, last) <= tol;
                            if (size > 0 && (!on ? j == tb : !joins)) {
                                if (sub_rn(last, first) >= mind) {
                                    if (count < DN_MAX_MID) st->mids[c][count] = div_rn(tsum, (T)size);
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
                        st->n_mid[c] = count;
                    }
                }
                __syncwarp();
                // epilogue B: band sums around each midpoint and the verdict, the warp cooperating per clip (see donut_kernel)
                for (int c2 = e; c2 < nact; c2 += kEpi) {
                    const int nm = st->n_mid[c2];
                    int label = 0;
                    T s_above = 0, s_mid = 0, s_below = 0;
#ifdef SJ_DONUT_SKIP_EPILOGUE /* timing experiment only: results are wrong */
                    if (false) {
#else
                    if (nm > 0) {
#endif
                    const int cap = nm < DN_MAX_MID ? nm : DN_MAX_MID;
                    const T *scr = scratch + size_t(c2) * tb * DN_BINS;
                    // The boxes' cells went to L2 up to 71 frames ago and most have been evicted to DRAM since (64 clips
                    // in flight per SM x 148 SMs do not fit L2): ask for the frames around every midpoint now, one 128-byte
                    // line per lane and instruction, so that they arrive while the thresholds are searched.
                    for (int m = 0; m < cap; ++m) {
                        const T mp = st->mids[c2][m];
                        int f_lo = (int)((mp - (T)0.18) * fs - (T)(DN_WIN / 2)) / DN_HOP - 1, f_hi = (int)((mp + (T)0.18) * fs - (T)(DN_WIN / 2)) / DN_HOP + 2;
                        f_lo = f_lo < 0 ? 0 : f_lo;
                        f_hi = f_hi > tb ? tb : f_hi;
                        const char *p0 = reinterpret_cast<const char *>(scr + f_lo * DN_BINS), *p1 = reinterpret_cast<const char *>(scr + f_hi * DN_BINS);
                        for (const char *q = p0 + 128 * lane; q < p1; q += 128 * 32) asm volatile("prefetch.global.L2 [%0];" ::"l"(q));
                    }
                    T smin = (T)INFINITY, smax = 0;
                    if (st->max_in[c2] > 0) { smin = mul_rn(div_rn(st->min_in[c2], U), (T)2); smax = mul_rn(div_rn(st->max_in[c2], U), (T)2); }
                    if (st->max_ed[c2] > 0) {
                        const T e0 = div_rn(st->min_ed[c2], U), e1 = div_rn(st->max_ed[c2], U);
                        smin = e0 < smin ? e0 : smin;
                        smax = e1 > smax ? e1 : smax;
                    }
                    const T mn = to_db<T>(smin), mx = to_db<T>(smax);
                    const T range = sub_rn(mx, mn), lo = tab->lo_norm, hi = tab->hi_norm;
                    auto level = [&](T av) -> T { return level_of<T>(av, U, mn, range); };
                    const T amax = st->max_in[c2];
                    // smallest a in [a_pos, amax] with pred true (pred monotone); "none" = +inf.  `seed` is where the
                    // threshold should lie if level() were evaluated without rounding: the search starts in a window of
                    // +-2^10 bit patterns around it when the predicate brackets there (two rounds instead of seven / thirteen),
                    // and in the whole range otherwise -- the result is the exact threshold either way.
                    auto first_true = [&](const bool upper, T seed) -> T {
                        auto pred = [&](T av) { const T lv = level(av); return upper ? !(lv < hi) : lv > lo; };
                        typedef typename Bits<T>::type UI;
                        UI blo = Bits<T>::of(a_pos), bhi = Bits<T>::of(amax);
                        if (!(amax >= a_pos) || !pred(amax)) return (T)INFINITY;
                        if (seed > a_pos && seed < amax) {
                            const UI sb = Bits<T>::of(seed), wlo = sb - 1024 > blo ? sb - 1024 : blo, whi = sb + 1024 < bhi ? sb + 1024 : bhi;
                            // lane 0 tests the pattern below the window (must be false), lane 1 its top (must be true)
                            const unsigned bal = __ballot_sync(kFull, pred(Bits<T>::from(lane == 0 ? wlo - 1 : whi)));
                            if (wlo > blo && !(bal & 1u)) blo = wlo;
                            if (bal & 2u) bhi = whi;
                        }
                        while (blo < bhi) {
                            const UI step = (bhi - blo) / 32 + 1;
                            UI cand = blo + step * (UI)(lane + 1) - 1;
                            if (cand > bhi || cand < blo) cand = bhi;
                            const unsigned bal = __ballot_sync(kFull, pred(Bits<T>::from(cand)));
                            const int first = __ffs(bal) - 1;
                            const UI at = Bits<T>::shfl(cand, first), before = Bits<T>::shfl(cand, first > 0 ? first - 1 : 0);
                            bhi = at;
                            if (first > 0) blo = before + 1;
                        }
                        return Bits<T>::from(bhi);
                    };
                    T a_lo = (T)INFINITY, a_hi = (T)INFINITY;
                    if (range > 0) {
                        // level(a) = (10 log10(2 a / U / 1e-12) - mn) / range = x  <=>  a = U / 2 * 1e-12 * 10^((x range + mn) / 10)
                        const double k = 0.5e-12 * (double)U;
                        a_lo = first_true(false, (T)(k * exp10(((double)lo * (double)range + (double)mn) * 0.1)));
                        a_hi = first_true(true, (T)(k * exp10(((double)hi * (double)range + (double)mn) * 0.1)));
                    }
                    T *qu = ew->queue, *stage = ew->stage;
                    constexpr int kStage = Cfg<T>::kStage;
                    // time index search exactly as sum_intense (classifier.c:619-640), lane-parallel over the frame-time
                    // table: times are increasing, so the loops count the frames before / after the window
                    auto window = [&](T halfw, T midpoint, int &t0, int &t1) {
                        const T tlo = sub_rn(midpoint, halfw), thi = add_rn(midpoint, halfw);
                        int nb = 0, na = 0;
                        for (int j = lane; j < DN_MAX_TB; j += 32) {
                            const T tj = frame_time[j];
                            nb += __popc(__ballot_sync(kFull, j < tb && tj < tlo));
                            na += __popc(__ballot_sync(kFull, j < tb && tj > thi));
                        }
                        t0 = nb; t1 = tb - 1 - na;
                        if (t0 >= tb) t0 = tb - 1;
                        if (t1 < 0) t1 = 0;
                        if (t0 > t1) { const int x = t0; t0 = t1; t1 = x; }
                    };
                    auto box = [&](int band, int t0, int t1) -> T {
                        const int f0 = tab->box_f[band][0], f1 = tab->box_f[band][1];
                        double acc = 0.0;
                        const int nf = f1 - f0 + 1, ncell = nf * (t1 - t0 + 1);
                        int nq = 0;
                        // the box is staged in shared memory by asynchronous copies first (see donut_kernel)
                        const int q32 = 32 / nf, r32 = 32 - q32 * nf;
                        int cj = lane / nf, ci = lane - cj * nf;
                        for (int cc = 0; cc < ncell; cc += 32) {
                            if (cc % kStage == 0) {
                                __syncwarp();
                                int sj = cj, si = ci;
                                for (int c = cc + lane; c < ncell && c < cc + kStage; c += 32) {
                                    cp_async_elem(stage + (c - cc), scr + (t0 + sj) * DN_BINS + f0 + si);
                                    si += r32;
                                    sj += q32;
                                    if (si >= nf) { si -= nf; ++sj; }
                                }
                                cp_async_wait_all();
                                __syncwarp();
                            }
                            const int c = cc + lane;
                            const T av = c < ncell ? stage[c % kStage] : (T)0;
                            ci += r32;
                            cj += q32;
                            if (ci >= nf) { ci -= nf; ++cj; }
                            const bool pass = c < ncell && av >= a_lo && av < a_hi;
                            const unsigned bal = __ballot_sync(kFull, pass);
                            if (pass) qu[nq + __popc(bal & ((1u << lane) - 1u))] = av;
                            nq += __popc(bal);
                            __syncwarp();
                            if (nq >= 32) {
                                acc += (double)level(qu[lane]);
                                nq -= 32;
                                __syncwarp();
                                if (lane < nq) qu[lane] = qu[32 + lane];
                                __syncwarp();
                            }
                        }
                        if (lane < nq) acc += (double)level(qu[lane]);
                        __syncwarp();
#pragma unroll
                        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(kFull, acc, o);
                        return (T)acc;
                    };
                    for (int m = 0; m < cap; ++m) {
                        const T mp = st->mids[c2][m];
                        int t0, t1, u0, u1;
                        window((T)0.18, mp, t0, t1); // the above and below boxes share their time window
                        window((T)0.05, mp, u0, u1);
                        s_above = box(0, t0, t1);
                        s_mid = box(1, u0, u1);
                        s_below = box(2, t0, t1);
                        if (s_mid < tab->thr_mid && s_above > tab->thr_above && s_below > tab->thr_below) { label = 1; break; }
                    }
                }
                    if (lane == 0) {
                        if (a.labels) a.labels[base + c2] = label;
                        if (a.track) a.track[base + c2] = (int8_t)(label + 1);
                        if (a.detail) {
                            float *d = a.detail + (base + c2) * 4;
                            d[0] = (float)nm; d[1] = (float)s_above; d[2] = (float)s_mid; d[3] = (float)s_below;
                        }
                    }
                    if (a.mids && lane < DN_MAX_MID)
                        a.mids[(base + c2) * DN_MAX_MID + lane] = lane < nm ? (float)st->mids[c2][lane] : __int_as_float(0x7fc00000);
                }
                __syncwarp();
                if (lane == 0) { __threadfence_block(); atomicAdd(&gs->epilogues_done, 1); } // the slot's state copy and scratch buffer are free
            }
        }
    }
}

} // namespace ws

template <typename T> size_t smem_total()
{
    return ((sizeof(SjDonutTables<T>) + 15) & ~size_t(15)) + sizeof(T) * 32 * kRow + sizeof(RowState<T>);
}

} // namespace

// Which kernel runs.  The warp-specialised kernels win on throughput, the round-1 kernels (CTA-wide barriers, warp 0 filters
// alone) on the latency of small batches: up to about one round of groups per CTA (60 clips per SM) a float batch takes the
// round-1 kernel, whose results are the same bit for bit (tests/test_donut_gpu.py: ..._agree_with_the_warp_specialised_ones).
// The double kernels organise the FFT differently (equal to rounding only), and a clip's result must not depend on the size
// of the batch it came in: double always takes the warp-specialised kernel.
// SCRUBJAY_DONUT_LEGACY=1 forces the round-1 kernels for both precisions, =f32 / =f64 for one of them, =0 never uses them.
static int donut_legacy_mask() // bit 0 / 1: always round-1 for float / double; bit 2: never
{
    static const int mask = [] {
        const char *e = getenv("SCRUBJAY_DONUT_LEGACY");
        if (!e || !e[0]) return 0;
        if (e[0] == '0') return 4;
        if (e[0] == 'f') return e[1] == '6' ? 2 : 1;
        return 3;
    }();
    return mask;
}
static bool use_round1(bool f64, long n_clips, int num_sms)
{
    const int m = donut_legacy_mask();
    if (m & (f64 ? 2 : 1)) return true;
    if (m & 4) return false;
    return !f64 && n_clips <= 60L * num_sms;
}

cudaError_t sj_donut_setup()
{
    cudaError_t e = cudaSuccess;
#define SJ_DN_ATTR(T, AL, PI)                                                                                          \
    if (e == cudaSuccess)                                                                                              \
    e = cudaFuncSetAttribute(donut_kernel<T, AL, PI>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_total<T>())
    SJ_DN_ATTR(float, true, false); SJ_DN_ATTR(float, false, false); SJ_DN_ATTR(float, true, true); SJ_DN_ATTR(float, false, true);
    SJ_DN_ATTR(double, true, false); SJ_DN_ATTR(double, false, false); SJ_DN_ATTR(double, true, true); SJ_DN_ATTR(double, false, true);
#undef SJ_DN_ATTR
#define SJ_WS_ATTR(T, AL)                                                                                              \
    if (e == cudaSuccess)                                                                                              \
    e = cudaFuncSetAttribute(ws::donut_ws_kernel<T, AL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ws::kSmemT<T>)
    SJ_WS_ATTR(float, true); SJ_WS_ATTR(float, false); SJ_WS_ATTR(double, true); SJ_WS_ATTR(double, false);
#undef SJ_WS_ATTR
    return e;
}

static int round1_grid(int num_sms, bool f64)
{
    static int per_sm_cache[2] = {0, 0};
    int &per_sm = per_sm_cache[f64 ? 1 : 0];
    if (per_sm < 1) {
        if (f64) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, donut_kernel<double, true, false>, DnWarps<double>::value * 32, smem_total<double>());
        else cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, donut_kernel<float, true, false>, DnWarps<float>::value * 32, smem_total<float>());
        if (per_sm < 1) per_sm = 1;
    }
    return num_sms * per_sm;
}

// PSD scratch for the clips in flight, sized for whichever kernel a launch may take
size_t sj_donut_scratch_bytes(int num_sms, int tb, bool f64)
{
    // warp-specialised: two buffers per group slot (a group's epilogue runs while the slot's next group is transformed)
    const size_t ws_bytes = size_t(num_sms) * (f64 ? ws::Cfg<double>::kIir : ws::Cfg<float>::kIir) * 2 * ws::kGroup * tb * DN_BINS * (f64 ? 8 : 4);
    const size_t r1_bytes = size_t(round1_grid(num_sms, f64)) * DN_CLIPS_PER_WARP * tb * DN_BINS * (f64 ? 8 : 4);
    return ws_bytes > r1_bytes ? ws_bytes : r1_bytes;
}

template <typename T> static cudaError_t launch_ws(SjDonutArgs a, int grid, cudaStream_t stream)
{
    // groups of up to 16 clips, kIir in flight per CTA; small batches are spread over the CTAs in smaller groups
    constexpr int kIir = ws::Cfg<T>::kIir;
    long cpg = (a.n_clips + (long)grid * kIir - 1) / ((long)grid * kIir);
    if (cpg > ws::kGroup) cpg = ws::kGroup;
    if (cpg < 1) cpg = 1;
    a.clips_per_cta = (int)cpg;
    const long groups = (a.n_clips + cpg - 1) / cpg;
    if (groups < grid) grid = (int)groups;
    const bool al = (a.n_samples % 4 == 0) && ((reinterpret_cast<uintptr_t>(a.clips) & 15) == 0) &&
                    (a.clips_per_row <= 0 || a.row_stride % 4 == 0);
    if (al) ws::donut_ws_kernel<T, true><<<grid, ws::kThreads, ws::kSmemT<T>, stream>>>(a);
    else ws::donut_ws_kernel<T, false><<<grid, ws::kThreads, ws::kSmemT<T>, stream>>>(a);
    return cudaGetLastError();
}

cudaError_t sj_launch_donut(const SjDonutArgs &a_in, bool f64, int num_sms, cudaStream_t stream)
{
    if (a_in.n_clips <= 0) return cudaSuccess;
    SjDonutArgs a = a_in;
    if (!use_round1(f64, a.n_clips, num_sms)) return f64 ? launch_ws<double>(a, num_sms, stream) : launch_ws<float>(a, num_sms, stream);
    int grid = round1_grid(num_sms, f64);
    long cpc = (a.n_clips + grid - 1) / grid; // spread small batches over the resident CTAs
    if (cpc > DN_CLIPS_PER_WARP) cpc = DN_CLIPS_PER_WARP;
    if (cpc < 1) cpc = 1;
    a.clips_per_cta = (int)cpc;
    long groups = (a.n_clips + cpc - 1) / cpc;
    if (groups < grid) grid = (int)groups;
    const bool aligned = (a.n_samples % 4 == 0) && ((reinterpret_cast<uintptr_t>(a.clips) & 15) == 0) &&
                         (a.clips_per_row <= 0 || a.row_stride % 4 == 0);
    const bool pipe = f64 ? cpc <= 4 * (DnWarps<double>::value - 1) : cpc <= DnWarps<float>::value - 1;
#define SJ_DN_LAUNCH(T, AL, PI) donut_kernel<T, AL, PI><<<grid, DnWarps<T>::value * 32, smem_total<T>(), stream>>>(a)
    if (f64) {
        if (aligned) { if (pipe) SJ_DN_LAUNCH(double, true, true); else SJ_DN_LAUNCH(double, true, false); }
        else { if (pipe) SJ_DN_LAUNCH(double, false, true); else SJ_DN_LAUNCH(double, false, false); }
    } else {
        if (aligned) { if (pipe) SJ_DN_LAUNCH(float, true, true); else SJ_DN_LAUNCH(float, true, false); }
        else { if (pipe) SJ_DN_LAUNCH(float, false, true); else SJ_DN_LAUNCH(float, false, false); }
    }
#undef SJ_DN_LAUNCH
    return cudaGetLastError();
}
