// fp64_pipe.cu -- micro-benchmark of the FP64 pipe on sm_100a, the resource the double donut kernel lives on:
//  (1) dependent latency of DADD / DMUL / DFMA (one warp, a serial chain);
//  (2) issue interval of independent DADDs from one warp (8 chains) and from W warps per SM (W = 4, 8, 16: one, two,
//      four warps per scheduler) -- is the pipe per scheduler, and how many cycles does a warp instruction hold it;
//  (3) the recurrence's own pattern: DMUL of the newest state, then eight dependent DSUBs, with 1 / 2 / 4 independent
//      recurrences per lane.
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o fp64_pipe fp64_pipe.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int KIND> __global__ void lat(double *out, int iters, long long *cyc, double b, double c)
{
    double a = threadIdx.x * 0.001 + 1.0;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 64; ++u) {
            if (KIND == 0) a = __dadd_rn(a, b);
            if (KIND == 1) a = __dmul_rn(a, b);
            if (KIND == 2) a = __fma_rn(a, b, c);
        }
    }
    const long long t1 = clock64();
    out[threadIdx.x] = a;
    if (threadIdx.x == 0) *cyc = t1 - t0;
}

// CH independent chains per lane, W warps in the CTA; cycles per warp instruction per SCHEDULER (warp id mod 4)
template <int CH> __global__ void thr(double *out, int iters, long long *cyc, double b)
{
    double a[CH];
#pragma unroll
    for (int j = 0; j < CH; ++j) a[j] = threadIdx.x * 0.001 + j;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 16; ++u)
#pragma unroll
            for (int j = 0; j < CH; ++j) a[j] = __dadd_rn(a[j], b);
    }
    __syncthreads();
    const long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int j = 0; j < CH; ++j) s += a[j];
    out[threadIdx.x] = s;
    if (threadIdx.x == 0) *cyc = t1 - t0;
}

// NR independent recurrences per lane: w0 = x - a1 w1 - ... - a8 w8 (DMUL of each delayed value, 8 dependent DSUBs)
template <int NR> __global__ void rec(double *out, int iters, long long *cyc, double b, double c)
{
    double w[NR][8];
#pragma unroll
    for (int r = 0; r < NR; ++r)
#pragma unroll
        for (int j = 0; j < 8; ++j) w[r][j] = threadIdx.x * 1e-3 + r + j * 0.01;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int s = 0; s < 8; ++s)
#pragma unroll
            for (int r = 0; r < NR; ++r) {
                double w0 = c;
#pragma unroll
                for (int j = 1; j < 9; ++j) w0 = __dsub_rn(w0, __dmul_rn(b, w[r][(s - j + 8) & 7]));
                w[r][s] = w0;
            }
    }
    const long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int r = 0; r < NR; ++r)
#pragma unroll
        for (int j = 0; j < 8; ++j) s += w[r][j];
    out[threadIdx.x] = s;
    if (threadIdx.x == 0) *cyc = t1 - t0;
}

// the kernel's whole step: the recurrence plus y = b0 w0 + b2 w2 + b4 w4 + b6 w6 + b8 w8 (b[odd] = 0) and the frame-mean sum.
// LAG = 1 forms y of step n - 1 during step n (its inputs are all old values: nothing of it sits behind the new w0)
template <int LAG> __global__ void full(double *out, int iters, long long *cyc, double b, double c)
{
    double w[8], acc = 0, wprev = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) w[j] = threadIdx.x * 1e-3 + j * 0.01;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int s = 0; s < 8; ++s) {
            double w0 = c;
#pragma unroll
            for (int j = 1; j < 9; ++j) w0 = __dsub_rn(w0, __dmul_rn(b, w[(s - j + 8) & 7]));
            if (LAG == 0) {
                double yy = __dmul_rn(c, w0);
#pragma unroll
                for (int j = 2; j < 9; j += 2) yy = __dadd_rn(yy, __dmul_rn(b, w[(s - j + 8) & 7]));
                acc = __dadd_rn(acc, yy);
            } else { // y of the previous step: w0(n-1) = wprev, delayed values shifted by one
                double yy = __dmul_rn(c, wprev);
#pragma unroll
                for (int j = 2; j < 9; j += 2) yy = __dadd_rn(yy, __dmul_rn(b, w[(s - 1 - j + 16) & 7]));
                acc = __dadd_rn(acc, yy);
                wprev = w0;
            }
            w[s] = w0;
        }
    }
    const long long t1 = clock64();
    double sum = acc;
#pragma unroll
    for (int j = 0; j < 8; ++j) sum += w[j];
    out[threadIdx.x] = sum;
    if (threadIdx.x == 0) *cyc = t1 - t0;
}

static double *out;
static long long *cyc;
static long long get() { long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); return h; }

int main()
{
    cudaMalloc(&out, 1 << 16); cudaMalloc(&cyc, 8);
    const int iters = 1000;
    printf("{\n");
    lat<0><<<1, 32>>>(out, iters, cyc, 1.0001, 1e-3); printf(" \"dadd_latency_cycles\": %.2f,\n", (double)get() / (iters * 64.0));
    lat<1><<<1, 32>>>(out, iters, cyc, 1.0001, 1e-3); printf(" \"dmul_latency_cycles\": %.2f,\n", (double)get() / (iters * 64.0));
    lat<2><<<1, 32>>>(out, iters, cyc, 1.0001, 1e-3); printf(" \"dfma_latency_cycles\": %.2f,\n", (double)get() / (iters * 64.0));
    const int ws[] = {1, 4, 8, 16};
    for (int w : ws) {
        thr<8><<<1, 32 * w>>>(out, iters, cyc, 1.0001);
        const double per_sched = (double)get() / (iters * 16.0 * 8.0 * ((w + 3) / 4));
        printf(" \"dadd_cycles_per_warp_instr_per_scheduler_%dwarps_8chains\": %.2f,\n", w, per_sched);
    }
    thr<2><<<1, 32>>>(out, iters, cyc, 1.0001); printf(" \"dadd_cycles_per_instr_1warp_2chains\": %.2f,\n", (double)get() / (iters * 32.0));
    thr<4><<<1, 32>>>(out, iters, cyc, 1.0001); printf(" \"dadd_cycles_per_instr_1warp_4chains\": %.2f,\n", (double)get() / (iters * 64.0));
    rec<1><<<1, 32>>>(out, iters, cyc, 0.1, 1e-3); printf(" \"recurrence_step_cycles_1_per_lane\": %.1f,\n", (double)get() / (iters * 8.0));
    rec<2><<<1, 32>>>(out, iters, cyc, 0.1, 1e-3); printf(" \"recurrence_step_cycles_2_per_lane\": %.1f,\n", (double)get() / (iters * 8.0));
    rec<4><<<1, 32>>>(out, iters, cyc, 0.1, 1e-3); printf(" \"recurrence_step_cycles_4_per_lane\": %.1f,\n", (double)get() / (iters * 8.0));
    full<0><<<1, 32>>>(out, iters, cyc, 0.1, 1e-3); printf(" \"full_step_cycles\": %.1f,\n", (double)get() / (iters * 8.0));
    full<1><<<1, 32>>>(out, iters, cyc, 0.1, 1e-3); printf(" \"full_step_cycles_output_one_step_behind\": %.1f,\n", (double)get() / (iters * 8.0));
    rec<1><<<1, 128>>>(out, iters, cyc, 0.1, 1e-3); printf(" \"recurrence_step_cycles_1_per_lane_4warps\": %.1f,\n", (double)get() / (iters * 8.0));
    rec<1><<<1, 256>>>(out, iters, cyc, 0.1, 1e-3); printf(" \"recurrence_step_cycles_1_per_lane_8warps\": %.1f\n", (double)get() / (iters * 8.0));
    printf("}\n");
    return cudaDeviceSynchronize() != cudaSuccess;
}
