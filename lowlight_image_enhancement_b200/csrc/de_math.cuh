// Number types and elementary functions of the CIEDE2000 kernels (csrc/deltae.cu), round 2.
//
// The dE00 kernels are bound by instruction issue and by the quarter-rate special-function unit, not by HBM
// (DESIGN.md, dE00 section).  Everything here serves two goals: fewer issue slots per pixel and fewer MUFU operations,
// inside the error budget of the path (1e-4 absolute on dE00 values of order 1..100):
//   * one generic formula source (`template <class V>`) instantiated for V = float (one pixel per thread), V = F2 (two
//     pixels per thread in the halves of a 64-bit register pair: add / mul / fma are the packed sm_100 instructions
//     FADD2 / FMUL2 / FFMA2, one issue slot for two pixels) and V = Dual3<float> (value + three tangents, the loss
//     gradient);
//   * square roots, reciprocals and divisions for operands that are known to be positive and normal: one MUFU
//     (sqrt / rsqrt / rcp .approx, <= 1-2 ulp) plus at most two FMAs of Newton / Markstein refinement, no guarded
//     slow path, no branch.  The IEEE forms cost 8-14 instructions and a convergence barrier each;
//   * no inverse trigonometry on the hue path (see de_core.cuh): hue differences and mean hues are formed from unit
//     vectors; the single angle the formula needs as a number (the exponent of the rotation term) comes from a
//     degree-7 polynomial arctangent (1.4e-7 absolute in fp32).
#pragma once
#include <cuda_runtime.h>

namespace llie {
namespace dem {

constexpr float kPiF = 3.14159265358979323846f;

// ------------------------------------------------------------------------------------------------ scalar primitives
__device__ __forceinline__ float mufu_rcp(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float mufu_rsqrt(float x) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float mufu_sqrt(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float mufu_ex2(float x) { float r; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float mufu_lg2(float x) { float r; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

// ------------------------------------------------------------------------------------------------ packed pair
struct F2 {
    unsigned long long v;
};
__device__ __forceinline__ F2 f2(float lo, float hi) {
    F2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r.v) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ float lo(F2 a) { return __uint_as_float((unsigned)(a.v & 0xffffffffull)); }
__device__ __forceinline__ float hi(F2 a) { return __uint_as_float((unsigned)(a.v >> 32)); }
__device__ __forceinline__ F2 operator+(F2 a, F2 b) { F2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
__device__ __forceinline__ F2 operator*(F2 a, F2 b) { F2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
__device__ __forceinline__ F2 fma_(F2 a, F2 b, F2 c) { F2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.v) : "l"(a.v), "l"(b.v), "l"(c.v)); return r; }
__device__ __forceinline__ F2 operator-(F2 a) { return f2(-lo(a), -hi(a)); }
__device__ __forceinline__ F2 operator-(F2 a, F2 b) { return fma_(b, f2(-1.f, -1.f), a); }
__device__ __forceinline__ F2 operator*(float s, F2 a) { return f2(s, s) * a; }
__device__ __forceinline__ F2 operator+(F2 a, float s) { return a + f2(s, s); }
__device__ __forceinline__ F2 operator-(F2 a, float s) { return a + f2(-s, -s); }
__device__ __forceinline__ F2 operator-(float s, F2 a) { return fma_(a, f2(-1.f, -1.f), f2(s, s)); }

__device__ __forceinline__ float fma_(float a, float b, float c) { return fmaf(a, b, c); }

// a b - c d with BOTH products rounded (no FMA contraction): exactly 0 when (a, b) == (d, c) -- the sine of the
// angle between two identical unit vectors must vanish so that dE00(x, x) is exactly 0
__device__ __forceinline__ float cross2(float a, float b, float c, float d) {
    return __fsub_rn(__fmul_rn(a, b), __fmul_rn(c, d));
}
__device__ __forceinline__ F2 cross2(F2 a, F2 b, F2 c, F2 d) { return (a * b) - (c * d); }

// constant broadcast
template <class V> __device__ __forceinline__ V cst(float c);
template <> __device__ __forceinline__ float cst<float>(float c) { return c; }
template <> __device__ __forceinline__ F2 cst<F2>(float c) { return f2(c, c); }

// fma with scalar coefficients (Horner steps, affine maps)
__device__ __forceinline__ float fmac(float a, float s, float c) { return fmaf(a, s, c); }       // a * s + c
__device__ __forceinline__ F2 fmac(F2 a, float s, float c) { return fma_(a, f2(s, s), f2(c, c)); }
__device__ __forceinline__ float fmav(float a, float s, float c) { return fmaf(a, s, c); }       // a * s + c, c a value
__device__ __forceinline__ F2 fmav(F2 a, float s, F2 c) { return fma_(a, f2(s, s), c); }

// per-half application of a scalar function / select helpers
#define DEM_MAP1(NAME, EXPR)                                                                   \
    __device__ __forceinline__ float NAME(float x) { return EXPR; }                              \
    __device__ __forceinline__ F2 NAME(F2 a) {                                                   \
        float x = lo(a);                                                                         \
        const float r0 = EXPR;                                                                   \
        x = hi(a);                                                                               \
        const float r1 = EXPR;                                                                   \
        return f2(r0, r1);                                                                       \
    }
DEM_MAP1(vrcp0, mufu_rcp(x))       // raw MUFU reciprocal (~1 ulp)
DEM_MAP1(vrsqrt0, mufu_rsqrt(x))   // raw MUFU reciprocal square root
DEM_MAP1(vsqrt0, mufu_sqrt(x))     // raw MUFU square root (max relative error 2^-23)
DEM_MAP1(vex2, mufu_ex2(x))
DEM_MAP1(vlg2, mufu_lg2(x))
DEM_MAP1(vabs, fabsf(x))
DEM_MAP1(vsat, __saturatef(x))     // clamp to [0, 1] (NaN -> 0: callers check finiteness before)
#undef DEM_MAP1

__device__ __forceinline__ float vmax(float a, float b) { return fmaxf(a, b); }
__device__ __forceinline__ float vmin(float a, float b) { return fminf(a, b); }
__device__ __forceinline__ F2 vmax(F2 a, F2 b) { return f2(fmaxf(lo(a), lo(b)), fmaxf(hi(a), hi(b))); }
__device__ __forceinline__ F2 vmin(F2 a, F2 b) { return f2(fminf(lo(a), lo(b)), fminf(hi(a), hi(b))); }

// masks: per-lane booleans
struct M2 {
    bool a, b;
};
__device__ __forceinline__ bool m_gt(float x, float y) { return x > y; }
__device__ __forceinline__ M2 m_gt(F2 x, F2 y) { return {lo(x) > lo(y), hi(x) > hi(y)}; }
__device__ __forceinline__ bool m_lt(float x, float y) { return x < y; }
__device__ __forceinline__ M2 m_lt(F2 x, F2 y) { return {lo(x) < lo(y), hi(x) < hi(y)}; }
__device__ __forceinline__ bool m_ge(float x, float y) { return x >= y; }   // false for NaN
__device__ __forceinline__ M2 m_ge(F2 x, F2 y) { return {lo(x) >= lo(y), hi(x) >= hi(y)}; }
__device__ __forceinline__ bool m_le(float x, float y) { return x <= y; }
__device__ __forceinline__ M2 m_le(F2 x, F2 y) { return {lo(x) <= lo(y), hi(x) <= hi(y)}; }
__device__ __forceinline__ bool m_neg(float x) { return (__float_as_uint(x) >> 31) != 0u; }  // sign bit (-0 counts)
__device__ __forceinline__ M2 m_neg(F2 x) { return {m_neg(lo(x)), m_neg(hi(x))}; }
__device__ __forceinline__ bool m_and(bool p, bool q) { return p && q; }
__device__ __forceinline__ M2 m_and(M2 p, M2 q) { return {p.a && q.a, p.b && q.b}; }
__device__ __forceinline__ bool m_or(bool p, bool q) { return p || q; }
__device__ __forceinline__ M2 m_or(M2 p, M2 q) { return {p.a || q.a, p.b || q.b}; }
__device__ __forceinline__ bool m_xor(bool p, bool q) { return p != q; }
__device__ __forceinline__ M2 m_xor(M2 p, M2 q) { return {p.a != q.a, p.b != q.b}; }
__device__ __forceinline__ bool m_not(bool p) { return !p; }
__device__ __forceinline__ M2 m_not(M2 p) { return {!p.a, !p.b}; }
__device__ __forceinline__ float vsel(bool m, float x, float y) { return m ? x : y; }
__device__ __forceinline__ F2 vsel(M2 m, F2 x, F2 y) { return f2(m.a ? lo(x) : lo(y), m.b ? hi(x) : hi(y)); }

// ------------------------------------------------------------------------------------------------ refined primitives
// 1 / d, d > 0 normal: MUFU + one Newton step (<= 1 ulp)
template <class V>
__device__ __forceinline__ V rcp_n(V d) {
    const V r = vrcp0(d);
    return fma_(fma_(-d, r, cst<V>(1.f)), r, r);
}
// n / d with a reciprocal r of d accurate to ~1 ulp: Markstein correction, the correctly rounded quotient but for rare
// 1-ulp cases
template <class V>
__device__ __forceinline__ V div_r(V n, V d, V r) {
    const V q = n * r;
    return fma_(fma_(-d, q, n), r, q);
}
template <class V>
__device__ __forceinline__ V div_n(V n, V d) { return div_r(n, d, rcp_n(d)); }
// sqrt(x) and 1 / sqrt(x) for x > 0 normal from ONE MUFU: r = rsqrt(x), s = x r refined by one Newton step
template <class V>
__device__ __forceinline__ void sqrt_rsqrt(V x, V &s, V &r) {
    const V r0 = vrsqrt0(x);
    const V s0 = x * r0;
    const V e = fma_(-s0, s0, x);                 // x - s0^2 (exact residual up to one rounding)
    s = fma_(e, cst<V>(0.5f) * r0, s0);           // s0 + e / (2 s0)
    r = r0;                                        // ~2 ulp, used only as a scale factor
}

}  // namespace dem
}  // namespace llie
