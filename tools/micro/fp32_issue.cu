// fp32_issue.cu -- micro-benchmark: how many FP32 instructions per cycle can ONE scheduler (SMSP) of a B200 issue, as a
// function of the number of resident warps on it and of the instruction kind (scalar FADD / FMUL / FFMA, packed
// FADD2 / FMUL2 / FFMA2 = add/mul/fma.rn.f32x2)?  Each warp runs 8 independent dependency chains, fully unrolled, so
// a single warp is never latency-bound (8 chains x 1 instr >= 4-cycle latency).  Prints cycles per warp-instruction
// per SMSP.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp32_issue fp32_issue.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "../../dsp_b200/csrc/sj_pk.h"

template <int KIND> __device__ __forceinline__ void op(float2 &a, float2 b, float2 c)
{
    if (KIND == 0) a.x = __fadd_rn(a.x, b.x);
    if (KIND == 1) a.x = __fmul_rn(a.x, b.x);
    if (KIND == 2) a.x = __fmaf_rn(a.x, b.x, c.x);
    if (KIND == 3) a = pk_add(a, b);
    if (KIND == 4) a = pk_mul(a, b);
    if (KIND == 5) a = pk_fma(a, b, c);
    if (KIND == 6) { a.x = __fadd_rn(a.x, b.x); a.y = __int_as_float(__float_as_int(a.y) + __float_as_int(b.y)); } // FADD + IADD
    if (KIND == 7) { a = pk_mul(a, b); a.x = __fadd_rn(a.x, c.x); } // FMUL2 then a dependent scalar FADD (the donut recurrence's pattern)
}

template <int KIND> __global__ void k(float2 *out, int iters, long long *cyc, float2 b, float2 c)
{
    float2 v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = make_float2(threadIdx.x * 0.001f + i, i * 0.5f);
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u)
#pragma unroll
            for (int i = 0; i < 8; ++i) op<KIND>(v[i], b, c);
    }
    const long long t1 = clock64();
    float2 s = make_float2(0.f, 0.f);
#pragma unroll
    for (int i = 0; i < 8; ++i) { s.x += v[i].x; s.y += v[i].y; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int KIND> void run(const char *name, int instr_per_op)
{
    float2 *out; long long *cyc;
    cudaMalloc(&out, 1 << 20); cudaMalloc(&cyc, 8);
    const int iters = 2000;
    printf("%-34s", name);
    for (int wps : {1, 2, 3, 4, 8}) { // warps per scheduler: one CTA of 4 * wps warps on one SM
        k<KIND><<<1, 128 * wps>>>(out, iters, cyc, make_float2(1.0001f, 0.9999f), make_float2(1e-3f, 1e-3f));
        long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
        // per SMSP: wps warps x iters x 64 ops x instr_per_op instructions
        printf("  %dw: %.2f", wps, (double)h / ((double)wps * iters * 64 * instr_per_op));
    }
    printf("   cycles per warp-instruction per scheduler\n");
    cudaFree(out); cudaFree(cyc);
}

int main()
{
    run<0>("FADD", 1); run<1>("FMUL", 1); run<2>("FFMA", 1); run<3>("FADD2 (add.rn.f32x2)", 1); run<4>("FMUL2 (mul.rn.f32x2)", 1);
    run<5>("FFMA2 (fma.rn.f32x2)", 1); run<6>("FADD + IADD pair", 2); run<7>("FMUL2 + dependent FADD", 2);
    return 0;
}
