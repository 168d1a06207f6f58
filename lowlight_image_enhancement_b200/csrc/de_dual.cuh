// Forward-mode dual numbers with N tangents over S = float or dem::F2 (two pixels per thread), so that the generic dE00
// formula of de_core.cuh yields value AND gradient from one evaluation (DeltaE00Loss, NewBP_model/losses.py:92-143, whose
// autograd graph saves ~40 planes and replays ~90 kernels).  The loss-gradient kernel differentiates the chroma / hue
// part with N = 2 (d/da1, d/db1) and the lightness part with N = 1 (d/dL1): the two parts share no input, and a tangent
// that is identically zero still costs its multiply-adds (x * 0 cannot be folded without fast-math).
// Rules are the ones ATen's autograd applies to the same elementary operations; comparisons, selects, abs / min / max
// pick a branch and differentiate it (torch.where semantics: the unselected branch contributes nothing).
#pragma once
#include "de_math.cuh"

namespace llie {
namespace dem {

template <class S, int N>
struct Dual {
    S v;
    S t[N];
};
#define DEM_FOR for (int k_ = 0; k_ < N; ++k_)
#define DEM_T template <class S, int N> __device__ __forceinline__

DEM_T Dual<S, N> dual_const(S v) {
    Dual<S, N> r;
    r.v = v;
#pragma unroll
    DEM_FOR r.t[k_] = cst<S>(0.f);
    return r;
}
// independent variable number `which`
DEM_T Dual<S, N> dual_var(S v, int which) {
    Dual<S, N> r = dual_const<S, N>(v);
#pragma unroll
    DEM_FOR if (k_ == which) r.t[k_] = cst<S>(1.f);
    return r;
}
template <> __device__ __forceinline__ Dual<float, 1> cst<Dual<float, 1>>(float c) { return dual_const<float, 1>(c); }
template <> __device__ __forceinline__ Dual<float, 2> cst<Dual<float, 2>>(float c) { return dual_const<float, 2>(c); }
template <> __device__ __forceinline__ Dual<F2, 1> cst<Dual<F2, 1>>(float c) { return dual_const<F2, 1>(f2(c, c)); }
template <> __device__ __forceinline__ Dual<F2, 2> cst<Dual<F2, 2>>(float c) { return dual_const<F2, 2>(f2(c, c)); }

// chain rule helper: value v, derivative dv of the outer function at a.v
DEM_T Dual<S, N> chain(S v, S dv, Dual<S, N> a) {
    Dual<S, N> r;
    r.v = v;
#pragma unroll
    DEM_FOR r.t[k_] = dv * a.t[k_];
    return r;
}
#define DEM_BIN(OP)                                        \
    DEM_T Dual<S, N> operator OP(Dual<S, N> a, Dual<S, N> b) { \
        Dual<S, N> r;                                      \
        r.v = a.v OP b.v;                                  \
        _Pragma("unroll") DEM_FOR r.t[k_] = a.t[k_] OP b.t[k_]; \
        return r;                                          \
    }
DEM_BIN(+)
DEM_BIN(-)
#undef DEM_BIN
DEM_T Dual<S, N> operator-(Dual<S, N> a) {
    Dual<S, N> r;
    r.v = -a.v;
#pragma unroll
    DEM_FOR r.t[k_] = -a.t[k_];
    return r;
}
DEM_T Dual<S, N> operator+(Dual<S, N> a, float s) { a.v = a.v + s; return a; }
DEM_T Dual<S, N> operator-(Dual<S, N> a, float s) { a.v = a.v - s; return a; }
DEM_T Dual<S, N> operator-(float s, Dual<S, N> a) { Dual<S, N> r = -a; r.v = s - a.v; return r; }
DEM_T Dual<S, N> operator*(float s, Dual<S, N> a) {
    a.v = s * a.v;
#pragma unroll
    DEM_FOR a.t[k_] = s * a.t[k_];
    return a;
}
DEM_T Dual<S, N> operator*(Dual<S, N> a, Dual<S, N> b) {
    Dual<S, N> r;
    r.v = a.v * b.v;
#pragma unroll
    DEM_FOR r.t[k_] = fma_(a.t[k_], b.v, a.v * b.t[k_]);
    return r;
}
DEM_T Dual<S, N> fma_(Dual<S, N> a, Dual<S, N> b, Dual<S, N> c) {
    Dual<S, N> r;
    r.v = fma_(a.v, b.v, c.v);
#pragma unroll
    DEM_FOR r.t[k_] = fma_(a.t[k_], b.v, fma_(a.v, b.t[k_], c.t[k_]));
    return r;
}
DEM_T Dual<S, N> fmac(Dual<S, N> a, float s, float c) {
    a.v = fmac(a.v, s, c);
#pragma unroll
    DEM_FOR a.t[k_] = s * a.t[k_];
    return a;
}
DEM_T Dual<S, N> fmav(Dual<S, N> a, float s, Dual<S, N> c) {
    c.v = fmav(a.v, s, c.v);
#pragma unroll
    DEM_FOR c.t[k_] = fmav(a.t[k_], s, c.t[k_]);
    return c;
}
DEM_T Dual<S, N> cross2(Dual<S, N> a, Dual<S, N> b, Dual<S, N> c, Dual<S, N> d) {
    Dual<S, N> r;
    r.v = cross2(a.v, b.v, c.v, d.v);
#pragma unroll
    DEM_FOR r.t[k_] = fma_(a.t[k_], b.v, a.v * b.t[k_]) - fma_(c.t[k_], d.v, c.v * d.t[k_]);
    return r;
}

// elementary functions (one MUFU for the value, the derivative from the value where possible)
DEM_T Dual<S, N> vsqrt0(Dual<S, N> a) {   // a.v > 0
    const S r = vrsqrt0(a.v);
    const S s0 = a.v * r;
    const S s = fma_(fma_(-s0, s0, a.v), cst<S>(0.5f) * r, s0);
    return chain(s, cst<S>(0.5f) * r, a);
}
DEM_T Dual<S, N> vrsqrt0(Dual<S, N> a) {
    const S r = vrsqrt0(a.v);
    return chain(r, cst<S>(-0.5f) * r * (r * r), a);
}
DEM_T Dual<S, N> vrcp0(Dual<S, N> a) {
    const S r0 = vrcp0(a.v);
    const S r = fma_(fma_(-a.v, r0, cst<S>(1.f)), r0, r0);
    return chain(r, -(r * r), a);
}
DEM_T Dual<S, N> rcp_n(Dual<S, N> a) { return vrcp0(a); }
DEM_T Dual<S, N> vex2(Dual<S, N> a) {
    const S e = vex2(a.v);
    return chain(e, cst<S>(0.6931471805599453f) * e, a);
}
DEM_T Dual<S, N> vabs(Dual<S, N> a) {
    const auto neg = m_neg(a.v);
    a.v = vabs(a.v);
#pragma unroll
    DEM_FOR a.t[k_] = vsel(neg, -a.t[k_], a.t[k_]);
    return a;
}
DEM_T auto m_gt(Dual<S, N> a, Dual<S, N> b) { return m_gt(a.v, b.v); }
DEM_T auto m_lt(Dual<S, N> a, Dual<S, N> b) { return m_lt(a.v, b.v); }
DEM_T auto m_neg(Dual<S, N> a) { return m_neg(a.v); }
template <class S, int N, class M>
__device__ __forceinline__ Dual<S, N> vsel(M m, Dual<S, N> a, Dual<S, N> b) {
    a.v = vsel(m, a.v, b.v);
#pragma unroll
    DEM_FOR a.t[k_] = vsel(m, a.t[k_], b.t[k_]);
    return a;
}
DEM_T Dual<S, N> vmax(Dual<S, N> a, Dual<S, N> b) { return vsel(m_gt(a.v, b.v), a, b); }
DEM_T Dual<S, N> vmin(Dual<S, N> a, Dual<S, N> b) { return vsel(m_lt(a.v, b.v), a, b); }
#undef DEM_FOR
#undef DEM_T

}  // namespace dem
}  // namespace llie
