#!/usr/bin/env python3
"""Generate straight-line register FFTs for the CUDA kernels.

Emits dsp_b200/csrc/fft_gen.inc with, for each N in SIZES, a macro-free device function

    __device__ __forceinline__ void sj_fft<N>(float2 (&v)[N]);   (also a double2 flavour)

Forward transform (e^{-i..}), unnormalised, radix-2 decimation in time on a register array whose
indices are all compile-time constants.  Input  v[n] = x[n] in natural order; the DIT network is
run on the bit-reversed alias u(i) = v[brev(i)], so no data moves; output X[k] is left in
v[brev(k)] (callers index with the generated constexpr SJ_BREV<N> tables).
Butterfly a' = a + w b, b' = 2a - a' costs 6 FMA-class instructions for a general twiddle,
4 add/sub for w in {1, -i}.
"""
import math
import os

SIZES = (8, 16, 32)


def brev(i, bits):
    r = 0
    for _ in range(bits):
        r = (r << 1) | (i & 1)
        i >>= 1
    return r


def lit(x, dbl):
    s = repr(float(x))
    return s if dbl else s + "f"


def gen_packed(n, tw=False):
    """float flavour on Blackwell's packed fp32 pipe: every butterfly works on (re, im) register pairs with
    add/sub/fma.rn.f32x2 (SASS FADD2 / FFMA2): 2 issue slots for w in {1, -i}, 3 for a general twiddle
    (p = cs*b + a; a' = (sn,-sn)*(b.y,b.x) + p; b' = 2a - a'), same roundings as the scalar network."""
    bits = n.bit_length() - 1
    out = []
    name = "sj_fft%d%s" % (n, "tw" if tw else "")
    if tw:
        out.append("__device__ __forceinline__ void %s(float2 (&v)[%d], const float4 *tw) {" % (name, n))
        out.append("    float2 t, y;")
        for j in range(n // 2):
            a = "v[%d]" % j
            c = "v[%d]" % (j + n // 2)
            out.append("    { const float4 w = tw[%d];" % (32 * j))
            if j == 0:
                out.append("      y = %s;" % a)
            else:
                out.append("      y = pk_fma(pk_bc(%s.y), make_float2(-w.y, w.x), pk_mul(pk_bc(%s.x), make_float2(w.x, w.y)));"
                           % (a, a))
            out.append("      t = pk_fma(pk_bc(%s.y), make_float2(-w.w, w.z), pk_fma(pk_bc(%s.x), make_float2(w.z, w.w), y));"
                       % (c, c))
            out.append("      %s = pk_fma(pk_bc(2.0f), y, pk_neg(t)); %s = t; }" % (c, a))
    else:
        out.append("__device__ __forceinline__ void %s(float2 (&v)[%d]) {" % (name, n))
        out.append("    float2 t;")
    half = 2 if tw else 1
    while half < n:
        span = 2 * half
        for b in range(0, n, span):
            for j in range(half):
                a = "v[%d]" % brev(b + j, bits)
                c = "v[%d]" % brev(b + j + half, bits)
                if j == 0:
                    out.append("    t = %s; %s = pk_sub(%s, t); %s = pk_add(%s, t);" % (c, c, a, a, a))
                elif 4 * j == span:  # w = -i : w*b = (b.y, -b.x)
                    out.append("    t = make_float2(%s.y, -%s.x); %s = pk_sub(%s, t); %s = pk_add(%s, t);"
                               % (c, c, c, a, a, a))
                else:
                    ang = 2.0 * math.pi * j / span
                    cs, sn = math.cos(ang), math.sin(ang)
                    out.append("    t = pk_fma(make_float2(%s, %s), make_float2(%s.y, %s.x), pk_fma(pk_bc(%s), %s, %s));"
                               % (lit(sn, False), lit(-sn, False), c, c, lit(cs, False), c, a))
                    out.append("    %s = pk_fma(pk_bc(2.0f), %s, pk_neg(t)); %s = t;" % (c, a, a))
        half = span
    out.append("}")
    return "\n".join(out)


def gen(n, dbl, tw=False):
    """tw=True: the inputs carry a pending per-element twiddle (step 2 of a two-pass split): element n
    still has to be multiplied by w_n, read as float4 (w_n, w_{n+n/2}) from tw[n * 32].  The multiply is
    merged into the first butterfly stage: y = w_n a (4 ops), a' = y + w_{n+n/2} b (4), b' = 2y - a' (2)
    instead of 2 x 4 + 4."""
    bits = n.bit_length() - 1
    T = "double" if dbl else "float"
    V = "double2" if dbl else "float2"
    fma = "fma" if dbl else "fmaf"
    out = []
    name = "sj_fft%d%s%s" % (n, "d" if dbl else "", "tw" if tw else "")
    if tw:
        out.append("__device__ __forceinline__ void %s(%s (&v)[%d], const float4 *tw) {" % (name, V, n))
        out.append("    %s tx, ty, yx, yy;" % T)
        for j in range(n // 2):
            a = "v[%d]" % j
            c = "v[%d]" % (j + n // 2)
            out.append("    { const float4 w = tw[%d];" % (32 * j))
            if j == 0:
                out.append("      yx = %s.x; yy = %s.y;" % (a, a))
            else:
                out.append("      yx = %s(-%s.y, w.y, %s.x * w.x); yy = %s(%s.y, w.x, %s.x * w.y);" % (fma, a, a, fma, a, a))
            out.append("      tx = %s(-w.w, %s.y, %s(w.z, %s.x, yx)); ty = %s(w.w, %s.x, %s(w.z, %s.y, yy));"
                       % (fma, c, fma, c, fma, c, fma, c))
            out.append("      %s.x = %s(%s, yx, -tx); %s.y = %s(%s, yy, -ty); %s.x = tx; %s.y = ty; }"
                       % (c, fma, lit(2.0, dbl), c, fma, lit(2.0, dbl), a, a))
    else:
        out.append("__device__ __forceinline__ void %s(%s (&v)[%d]) {" % (name, V, n))
        out.append("    %s tx, ty;" % T)
    half = 2 if tw else 1
    while half < n:
        span = 2 * half
        for b in range(0, n, span):
            for j in range(half):
                i0 = brev(b + j, bits)
                i1 = brev(b + j + half, bits)
                a = "v[%d]" % i0
                c = "v[%d]" % i1
                # w = exp(-2*pi*i*j/span) = (cs, -sn)
                if j == 0:
                    out.append("    tx = %s.x; ty = %s.y; %s.x = %s.x - tx; %s.y = %s.y - ty; %s.x += tx; %s.y += ty;"
                               % (c, c, c, a, c, a, a, a))
                elif 4 * j == span:  # w = -i : w*b = (b.y, -b.x)
                    out.append("    tx = %s.y; ty = %s.x; %s.x = %s.x - tx; %s.y = %s.y + ty; %s.x += tx; %s.y -= ty;"
                               % (c, c, c, a, c, a, a, a))
                else:
                    ang = 2.0 * math.pi * j / span
                    cs, sn = math.cos(ang), math.sin(ang)
                    out.append("    tx = %s(%s, %s.y, %s(%s, %s.x, %s.x)); ty = %s(%s, %s.x, %s(%s, %s.y, %s.y));"
                               % (fma, lit(sn, dbl), c, fma, lit(cs, dbl), c, a,
                                  fma, lit(-sn, dbl), c, fma, lit(cs, dbl), c, a))
                    out.append("    %s.x = %s(%s, %s.x, -tx); %s.y = %s(%s, %s.y, -ty); %s.x = tx; %s.y = ty;"
                               % (c, fma, lit(2.0, dbl), a, c, fma, lit(2.0, dbl), a, a, a))
        half = span
    out.append("}")
    return "\n".join(out)


def main():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    parts = ["// generated by tools/gen_fft.py -- do not edit", "#pragma once"]
    for n in SIZES:
        bits = n.bit_length() - 1
        parts.append("// X[k] of sj_fft%d is found in v[SJ_BREV%d(k)]" % (n, n))
        parts.append("__host__ __device__ constexpr int SJ_BREV%d(int k) { return %s; }" % (
            n, " | ".join("(((k >> %d) & 1) << %d)" % (b, bits - 1 - b) for b in range(bits))))
        parts.append("#ifdef SJ_PACKED_F32")
        parts.append(gen_packed(n))
        parts.append("#else")
        parts.append(gen(n, False))
        parts.append("#endif")
        parts.append(gen(n, True))
        if n == 32:
            parts.append("#ifdef SJ_PACKED_F32")
            parts.append(gen_packed(n, tw=True))
            parts.append("#else")
            parts.append(gen(n, False, tw=True))
            parts.append("#endif")
    with open(os.path.join(root, "dsp_b200", "csrc", "fft_gen.inc"), "w") as f:
        f.write("\n".join(parts) + "\n")


if __name__ == "__main__":
    main()
