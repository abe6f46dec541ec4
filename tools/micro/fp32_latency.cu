// fp32_latency.cu -- micro-benchmark: dependent-issue latency of the FP32 instructions the donut recurrence is made
// of, one warp alone on an SM: a serial chain of FADD, of FMUL, FMUL2 -> FADD (packed product consumed by a scalar
// add), and the recurrence's own pattern (FMUL of the newest state, then eight dependent FADDs).
#include <cstdio>
#include <cuda_runtime.h>
#include "../../dsp_b200/csrc/sj_pk.h"

template <int KIND> __global__ void k(float *out, int iters, long long *cyc, float b, float c)
{
    float a = threadIdx.x * 0.001f + 1.f;
    float2 p = make_float2(a, a);
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 64; ++u) {
            if (KIND == 0) a = __fadd_rn(a, b);
            if (KIND == 1) a = __fmul_rn(a, b);
            if (KIND == 2) { p = pk_mul(make_float2(a, a), make_float2(b, c)); a = __fadd_rn(p.x, c); } // FMUL2 -> FADD
            if (KIND == 3) { // one recurrence step: FMUL of the newest value, then 8 dependent subtractions
                float w = __fmul_rn(a, b);
                float acc = c;
#pragma unroll
                for (int j = 0; j < 8; ++j) acc = __fsub_rn(acc, j == 0 ? w : b);
                a = acc;
            }
            if (KIND == 4) p = pk_add(p, make_float2(b, c));
        }
    }
    const long long t1 = clock64();
    out[threadIdx.x] = a + p.x + p.y;
    if (threadIdx.x == 0) *cyc = t1 - t0;
}

template <int KIND> void run(const char *name, int ops)
{
    float *out; long long *cyc;
    cudaMalloc(&out, 4096); cudaMalloc(&cyc, 8);
    const int iters = 1000;
    k<KIND><<<1, 32>>>(out, iters, cyc, 1.0001f, 1e-3f);
    long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-44s %.2f cycles per dependent instruction (%.1f per unit)\n", name, (double)h / ((double)iters * 64 * ops), (double)h / ((double)iters * 64));
    cudaFree(out); cudaFree(cyc);
}

int main()
{
    run<0>("FADD -> FADD", 1); run<1>("FMUL -> FMUL", 1); run<2>("FMUL2 -> FADD -> FMUL2 ...", 2); run<3>("FMUL + 8 x FADD (recurrence step)", 9);
    run<4>("FADD2 -> FADD2", 1);
    return 0;
}
