// Forward-mode dual numbers with three tangents over S = float or dem::F2 (two pixels per thread), so that the generic
// dE00 formula of de_core.cuh yields value AND gradient w.r.t. (L1, a1, b1) from one evaluation (DeltaE00Loss,
// NewBP_model/losses.py:92-143, whose autograd graph saves ~40 planes and replays ~90 kernels).
// Rules are the ones ATen's autograd applies to the same elementary operations; comparisons, selects, abs / min / max
// pick a branch and differentiate it (torch.where semantics: the unselected branch contributes nothing).
#pragma once
#include "de_math.cuh"

namespace llie {
namespace dem {

template <class S>
struct Dual3 {
    S v, x, y, z;
};
template <class S> struct ScalarOf { using type = S; };
template <class S> struct ScalarOf<Dual3<S>> { using type = S; };

template <class S> __device__ __forceinline__ Dual3<S> dual_const(S v) { return {v, cst<S>(0.f), cst<S>(0.f), cst<S>(0.f)}; }
template <> __device__ __forceinline__ Dual3<float> cst<Dual3<float>>(float c) { return dual_const<float>(c); }
template <> __device__ __forceinline__ Dual3<F2> cst<Dual3<F2>>(float c) { return dual_const<F2>(f2(c, c)); }

// chain rule helper: value v, derivative dv of the outer function at a.v
template <class S>
__device__ __forceinline__ Dual3<S> chain(S v, S dv, Dual3<S> a) { return {v, dv * a.x, dv * a.y, dv * a.z}; }

template <class S> __device__ __forceinline__ Dual3<S> operator+(Dual3<S> a, Dual3<S> b) { return {a.v + b.v, a.x + b.x, a.y + b.y, a.z + b.z}; }
template <class S> __device__ __forceinline__ Dual3<S> operator-(Dual3<S> a, Dual3<S> b) { return {a.v - b.v, a.x - b.x, a.y - b.y, a.z - b.z}; }
template <class S> __device__ __forceinline__ Dual3<S> operator-(Dual3<S> a) { return {-a.v, -a.x, -a.y, -a.z}; }
template <class S> __device__ __forceinline__ Dual3<S> operator+(Dual3<S> a, float s) { return {a.v + s, a.x, a.y, a.z}; }
template <class S> __device__ __forceinline__ Dual3<S> operator-(Dual3<S> a, float s) { return {a.v - s, a.x, a.y, a.z}; }
template <class S> __device__ __forceinline__ Dual3<S> operator-(float s, Dual3<S> a) { return {s - a.v, -a.x, -a.y, -a.z}; }
template <class S> __device__ __forceinline__ Dual3<S> operator*(float s, Dual3<S> a) { return {s * a.v, s * a.x, s * a.y, s * a.z}; }
template <class S>
__device__ __forceinline__ Dual3<S> operator*(Dual3<S> a, Dual3<S> b) {
    return {a.v * b.v, fma_(a.x, b.v, a.v * b.x), fma_(a.y, b.v, a.v * b.y), fma_(a.z, b.v, a.v * b.z)};
}
template <class S>
__device__ __forceinline__ Dual3<S> fma_(Dual3<S> a, Dual3<S> b, Dual3<S> c) {
    return {fma_(a.v, b.v, c.v), fma_(a.x, b.v, fma_(a.v, b.x, c.x)), fma_(a.y, b.v, fma_(a.v, b.y, c.y)),
            fma_(a.z, b.v, fma_(a.v, b.z, c.z))};
}
template <class S>
__device__ __forceinline__ Dual3<S> fmac(Dual3<S> a, float s, float c) { return {fmac(a.v, s, c), s * a.x, s * a.y, s * a.z}; }
template <class S>
__device__ __forceinline__ Dual3<S> fmav(Dual3<S> a, float s, Dual3<S> c) {
    return {fmav(a.v, s, c.v), fmav(a.x, s, c.x), fmav(a.y, s, c.y), fmav(a.z, s, c.z)};
}
// mixed: dual with a plain (constant) value of the same scalar type
template <class S> __device__ __forceinline__ Dual3<S> operator*(Dual3<S> a, S b) { return {a.v * b, a.x * b, a.y * b, a.z * b}; }
template <class S> __device__ __forceinline__ Dual3<S> operator*(S b, Dual3<S> a) { return a * b; }

template <class S>
__device__ __forceinline__ Dual3<S> cross2(Dual3<S> a, Dual3<S> b, Dual3<S> c, Dual3<S> d) {
    return {cross2(a.v, b.v, c.v, d.v),
            fma_(a.x, b.v, a.v * b.x) - fma_(c.x, d.v, c.v * d.x),
            fma_(a.y, b.v, a.v * b.y) - fma_(c.y, d.v, c.v * d.y),
            fma_(a.z, b.v, a.v * b.z) - fma_(c.z, d.v, c.v * d.z)};
}

// elementary functions (one MUFU for the value, the derivative from the value where possible)
template <class S>
__device__ __forceinline__ Dual3<S> vsqrt0(Dual3<S> a) {   // a.v > 0
    const S r = vrsqrt0(a.v);
    const S s0 = a.v * r;
    const S s = fma_(fma_(-s0, s0, a.v), cst<S>(0.5f) * r, s0);
    return chain(s, cst<S>(0.5f) * r, a);
}
template <class S>
__device__ __forceinline__ Dual3<S> vrsqrt0(Dual3<S> a) {
    const S r = vrsqrt0(a.v);
    return chain(r, cst<S>(-0.5f) * r * (r * r), a);
}
template <class S>
__device__ __forceinline__ Dual3<S> vrcp0(Dual3<S> a) {
    const S r0 = vrcp0(a.v);
    const S r = fma_(fma_(-a.v, r0, cst<S>(1.f)), r0, r0);
    return chain(r, -(r * r), a);
}
template <class S>
__device__ __forceinline__ Dual3<S> rcp_n(Dual3<S> a) { return vrcp0(a); }
template <class S>
__device__ __forceinline__ Dual3<S> vex2(Dual3<S> a) {
    const S e = vex2(a.v);
    return chain(e, cst<S>(0.6931471805599453f) * e, a);
}
template <class S> __device__ __forceinline__ Dual3<S> vabs(Dual3<S> a) {
    const auto neg = m_neg(a.v);
    return {vabs(a.v), vsel(neg, -a.x, a.x), vsel(neg, -a.y, a.y), vsel(neg, -a.z, a.z)};
}
template <class S> __device__ __forceinline__ auto m_gt(Dual3<S> a, Dual3<S> b) { return m_gt(a.v, b.v); }
template <class S> __device__ __forceinline__ auto m_lt(Dual3<S> a, Dual3<S> b) { return m_lt(a.v, b.v); }
template <class S> __device__ __forceinline__ auto m_neg(Dual3<S> a) { return m_neg(a.v); }
template <class S, class M>
__device__ __forceinline__ Dual3<S> vsel(M m, Dual3<S> a, Dual3<S> b) {
    return {vsel(m, a.v, b.v), vsel(m, a.x, b.x), vsel(m, a.y, b.y), vsel(m, a.z, b.z)};
}
template <class S> __device__ __forceinline__ Dual3<S> vmax(Dual3<S> a, Dual3<S> b) { return vsel(m_gt(a.v, b.v), a, b); }
template <class S> __device__ __forceinline__ Dual3<S> vmin(Dual3<S> a, Dual3<S> b) { return vsel(m_lt(a.v, b.v), a, b); }

}  // namespace dem
}  // namespace llie
