// Packed fp32 pairs on sm_100a: add/sub/mul/fma.rn.f32x2 (SASS FADD2 / FMUL2 / FFMA2) do two IEEE
// round-to-nearest fp32 operations in one issue slot; ptxas folds half swaps, broadcasts of one 32-bit
// register and per-half negation into the operand modifiers.
#pragma once
#include <cuda_runtime.h>

#define SJ_PK_OP2(NAME, OP)                                                                                   \
    __device__ __forceinline__ float2 NAME(float2 a, float2 b)                                                \
    {                                                                                                         \
        float2 r;                                                                                             \
        asm("{.reg .b64 pa, pb, pr; mov.b64 pa, {%2,%3}; mov.b64 pb, {%4,%5}; " OP                            \
            ".rn.f32x2 pr, pa, pb; mov.b64 {%0,%1}, pr;}"                                                     \
            : "=f"(r.x), "=f"(r.y)                                                                            \
            : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));                                                        \
        return r;                                                                                             \
    }
SJ_PK_OP2(pk_add, "add")
SJ_PK_OP2(pk_sub, "sub")
SJ_PK_OP2(pk_mul, "mul")
#undef SJ_PK_OP2

__device__ __forceinline__ float2 pk_fma(float2 a, float2 b, float2 c)
{
    float2 r;
    asm("{.reg .b64 pa, pb, pc, pr; mov.b64 pa, {%2,%3}; mov.b64 pb, {%4,%5}; mov.b64 pc, {%6,%7}; "
        "fma.rn.f32x2 pr, pa, pb, pc; mov.b64 {%0,%1}, pr;}"
        : "=f"(r.x), "=f"(r.y)
        : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(c.x), "f"(c.y));
    return r;
}
__device__ __forceinline__ float2 pk_bc(float x) { return make_float2(x, x); }
__device__ __forceinline__ float2 pk_neg(float2 a) { return make_float2(-a.x, -a.y); }
