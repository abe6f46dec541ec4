// iir_step.cu -- micro-benchmark: cycles per sample step of the donut recurrence (the float iir8 of donut_kernel.cu:
// direct-form II, 8 ordered subtractions + 4 ordered additions, products of even taps from packed multiplies), one warp
// alone on a scheduler, (a) on register inputs only, (b) with the circular-row stores, (c) variant with all products
// scalar, (d) two independent recurrences per lane interleaved.
#include <cstdio>
#include <cuda_runtime.h>
#include "../../dsp_b200/csrc/sj_pk.h"

__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub_rn(float a, float b) { return __fsub_rn(a, b); }

template <bool PACKED>
__device__ __forceinline__ void iir8(const float (&x)[8], float (&w)[8], const float (&a)[9], const float (&b)[9], float (&y)[8])
{
#pragma unroll
    for (int s = 0; s < 8; ++s) {
        float2 pe[4];
#pragma unroll
        for (int j = 2; j < 9; j += 2)
            pe[j / 2 - 1] = PACKED ? pk_mul(pk_bc(w[(s - j + 8) & 7]), make_float2(a[j], b[j]))
                                   : make_float2(mul_rn(w[(s - j + 8) & 7], a[j]), mul_rn(w[(s - j + 8) & 7], b[j]));
        float w0 = x[s];
#pragma unroll
        for (int j = 1; j < 9; ++j) w0 = sub_rn(w0, (j & 1) ? mul_rn(a[j], w[(s - j + 8) & 7]) : pe[j / 2 - 1].x);
        float yy = mul_rn(b[0], w0);
#pragma unroll
        for (int j = 2; j < 9; j += 2) yy = add_rn(yy, pe[j / 2 - 1].y);
        w[s] = w0;
        y[s] = yy;
    }
}

template <int MODE> __global__ void k(float *out, int iters, long long *cyc, const float *coef)
{
    __shared__ float rows[32 * 258];
    float a[9], b[9];
    for (int j = 0; j < 9; ++j) { a[j] = coef[j]; b[j] = coef[9 + j]; }
    float w[8], w2[8], x[8], y[8], y2[8];
    for (int j = 0; j < 8; ++j) { w[j] = 0.f; w2[j] = 0.f; x[j] = 0.001f * (threadIdx.x + j); }
    float acc = 0.f;
    float *row = rows + threadIdx.x * 258;
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (MODE == 2) iir8<false>(x, w, a, b, y); else iir8<true>(x, w, a, b, y);
            if (MODE == 3) iir8<true>(x, w2, a, b, y2);
#pragma unroll
            for (int j = 0; j < 8; ++j) { acc = add_rn(acc, y[j]); if (MODE == 3) acc = add_rn(acc, y2[j]); x[j] = y[j] * 0.5f + 0.001f; }
            if (MODE == 1) {
                float *dst = row + ((it * 32 + u * 8) & 255);
#pragma unroll
                for (int j = 0; j < 8; j += 2) *reinterpret_cast<float2 *>(dst + j) = make_float2(y[j], y[j + 1]);
            }
        }
    }
    const long long t1 = clock64();
    out[threadIdx.x] = MODE == 1 ? acc + rows[threadIdx.x * 258] : acc;
    if (threadIdx.x == 0) *cyc = t1 - t0;
}

template <int MODE> void run(const char *name, const float *coef, int per)
{
    float *out; long long *cyc;
    cudaMalloc(&out, 4096); cudaMalloc(&cyc, 8);
    const int iters = 500;
    k<MODE><<<1, 32>>>(out, iters, cyc, coef);
    cudaError_t e1 = cudaGetLastError();
    long long h = -1; cudaError_t e2 = cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-52s %.1f cycles per sample step  (%s / %s)\n", name, (double)h / ((double)iters * 32 * per), cudaGetErrorString(e1), cudaGetErrorString(e2));
    cudaFree(out); cudaFree(cyc);
}

int main()
{
    const float h[18] = {1.f, -4.56803686f, 9.95922498f, -13.49912589f, 12.43979269f, -7.94997696f, 3.43760562f, -0.92305481f, 0.1203896f,
                         0.01020948f, 0.f, -0.04083792f, 0.f, 0.06125688f, 0.f, -0.04083792f, 0.f, 0.01020948f};
    float *coef; cudaMalloc(&coef, sizeof h); cudaMemcpy(coef, h, sizeof h, cudaMemcpyHostToDevice);
    run<0>("iir8 packed even taps, registers only", coef, 1);
    run<1>("iir8 packed + circular-row stores", coef, 1);
    run<2>("iir8 all products scalar", coef, 1);
    run<3>("two independent recurrences per lane (per step of each)", coef, 2);
    return 0;
}
