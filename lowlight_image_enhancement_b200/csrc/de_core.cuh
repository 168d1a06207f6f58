// sRGB -> Lab (kornia 0.6.12 formula) and the reference's two CIEDE2000 variants as ONE generic formula source
// (template <class V>, see de_math.cuh for the number types).
//
// Hue without inverse trigonometry.  The reference forms h_i = atan2(b_i, a_i'), a wrapped difference dh, the mean hue
// hbar by case rules, then sin(dh / 2), four cosines of multiples of hbar and exp(-((deg hbar - 275) / 25)^2).  With the
// unit vectors u_i = (a_i', b_i) / |(a_i', b_i)|:
//   * sin(dh / 2) = sign(u1 x u2) |u2 - u1| / 2          (dh is the signed short-arc angle from h1 to h2);
//   * w = u1 + u2 points along the bisector of the short arc, which IS the reference's hbar modulo 2 pi in every case
//     of both variants (|h1 - h2| <= pi: the plain mean; otherwise the mean shifted by pi), so cos hbar = w_x / |w|,
//     sin hbar = w_y / |w| and T follows from multiple-angle identities;
//   * only the Gaussian rotation term needs hbar as a NUMBER, with the reference's (non-periodic) branch choice:
//     theta = atan2(w_y, w_x) in (-pi, pi] from a degree-7 polynomial, then
//       metric variant (hues in (-pi, pi], color_error.py:169-181):  hbar = theta, + 2 pi when the short arc crosses the
//                                                                    cut at +-pi and theta < 0;
//       loss variant   (hues in [0, 2 pi), losses.py:112-121):       hbar = theta, + 2 pi when the short arc does NOT
//                                                                    cross the cut at 0 and theta < 0.
//     "Crosses the cut" is decided from the signs of b1, b2 and of s = b1 a2' - a1' b2 = |.||.| sin(h1 - h2).
// A colour with a' = b = 0 has hue atan2(0, 0) = 0 in the reference: its unit vector is (1, 0).  Opposite hues
// (|w| -> 0) are a discontinuity of the reference formula itself (the mean hue jumps by pi): there, as in the
// reference's own float32 evaluation, the branch is decided by rounding.
#pragma once
#include "de_math.cuh"

namespace llie {
namespace dem {

// x / C for a compile-time constant C, correctly rounded (Markstein; see deltae.cu div_const)
template <class V>
__device__ __forceinline__ V divc(V x, float c, float rc) {
    const V q = x * cst<V>(rc);
    const V r = fma_(cst<V>(-c), q, x);
    return fma_(r, cst<V>(rc), q);
}
#define DEM_DIVC(x, C) divc((x), (C), 1.0f / (C))

// b^2.4 for b in [0.05, 1.06]: (b * b^(1/5))^2 with a Newton-refined fifth root carried as an unevaluated sum
// (< 1 ulp against double pow; same algorithm as round 1, generic in V)
template <class V>
__device__ __forceinline__ V pow24(V b) {
    const V y = vex2(cst<V>(0.2f) * vlg2(b));
    const V y2 = y * y, e2 = fma_(y, y, -y2);
    const V y4 = y2 * y2;
    const V e4 = fma_(y2, y2, -y4) + cst<V>(2.0f) * y2 * e2;
    const V r = fma_(-y4, y, b) - e4 * y;
    const V dy = r * vrcp0(cst<V>(5.0f) * y4);
    const V p = b * y, ep = fma_(b, y, -p);
    const V lo_ = fma_(b, dy, ep);
    return fma_(p, p, cst<V>(2.0f) * p * lo_);
}
template <class V>
__device__ __forceinline__ V srgb_to_linear(V c) {
    const V hi_ = pow24(DEM_DIVC(c + 0.055f, 1.055f));
    const V lo_ = DEM_DIVC(c, 12.92f);
    return vsel(m_gt(c, cst<V>(0.04045f)), hi_, lo_);
}
// t^(1/3) for t in (0.0088, 1.1]: y0 = t * 2^(-2/3 log2 t), one Newton step on y^3 = t with 1 / (3 y^2) ~ z / 3
// (2 MUFU, <= 1 ulp; lanes with t <= 0.008856 produce garbage / NaN that the caller's select drops)
template <class V>
__device__ __forceinline__ V cbrt_fast(V t) {
    const V z = vex2(cst<V>(-2.0f / 3.0f) * vlg2(t));   // t^(-2/3)
    const V y0 = t * z;
    const V e = fma_(y0 * y0, y0, -t);                   // y0^3 - t
    return fma_(-e, cst<V>(1.0f / 3.0f) * z, y0);
}
// value and derivative t^(-2/3) / 3 (the derivative is the Newton scale factor, no division)
template <class V>
__device__ __forceinline__ V cbrt_fast_d(V t, V &dt) {
    const V z = vex2(cst<V>(-2.0f / 3.0f) * vlg2(t));
    const V y0 = t * z;
    const V e = fma_(y0 * y0, y0, -t);
    dt = cst<V>(1.0f / 3.0f) * z;
    return fma_(-e, dt, y0);
}
template <class V>
__device__ __forceinline__ V lab_f_d(V t, V &dt) {
    V dc;
    const V c = cbrt_fast_d(t, dc);
    const auto hi_ = m_gt(t, cst<V>(0.008856f));
    dt = vsel(hi_, dc, cst<V>(7.787f));
    return vsel(hi_, c, fmac(t, 7.787f, 4.0f / 29.0f));
}
// sRGB channel -> linear with derivative (autograd of where(c > .04045, pow(.., 2.4), c / 12.92))
template <class V>
__device__ __forceinline__ V srgb_to_linear_d(V c, V &dc) {
    const V base = DEM_DIVC(c + 0.055f, 1.055f);
    const V v = pow24(base);
    const auto hi_ = m_gt(c, cst<V>(0.04045f));
    dc = vsel(hi_, cst<V>(2.4f / 1.055f) * (v * rcp_n(base)), cst<V>(1.0f / 12.92f));   // 2.4 base^1.4 / 1.055
    return vsel(hi_, v, DEM_DIVC(c, 12.92f));
}
__device__ __forceinline__ float cbrt_exact(float t) { return cbrtf(t); }
__device__ __forceinline__ F2 cbrt_exact(F2 t) { return f2(cbrtf(lo(t)), cbrtf(hi(t))); }

template <class V, bool EXACT = false>
__device__ __forceinline__ V lab_f(V t) {
    const V c = EXACT ? cbrt_exact(t) : cbrt_fast(t);
    return vsel(m_gt(t, cst<V>(0.008856f)), c, fmac(t, 7.787f, 4.0f / 29.0f));
}

// kornia 0.6.12 rgb_to_lab; EXACT_L: f(Y) through cbrtf (1 ulp, the value round 1 shipped) so that the L plane the
// edge mask is built from keeps its bit pattern
template <class V, bool EXACT_L>
__device__ __forceinline__ void rgb_to_lab(V r, V g, V b, V &L, V &A, V &B) {
    r = srgb_to_linear(r);
    g = srgb_to_linear(g);
    b = srgb_to_linear(b);
    const V x = fmav(b, 0.180423f, fmav(g, 0.357580f, cst<V>(0.412453f) * r));
    const V y = fmav(b, 0.072169f, fmav(g, 0.715160f, cst<V>(0.212671f) * r));
    const V z = fmav(b, 0.950227f, fmav(g, 0.119193f, cst<V>(0.019334f) * r));
    const V fx = lab_f<V, false>(DEM_DIVC(x, 0.95047f));
    const V fy = lab_f<V, EXACT_L>(y);
    const V fz = lab_f<V, false>(DEM_DIVC(z, 1.08883f));
    L = fmac(fy, 116.0f, -16.0f);
    A = cst<V>(500.0f) * (fx - fy);
    B = cst<V>(200.0f) * (fy - fz);
}

template <class V>
__device__ __forceinline__ V pow7(V v) {
    const V v2 = v * v, v4 = v2 * v2;
    return v4 * v2 * v;
}

// atan2(y, x) in (-pi, pi], 1.4e-7 absolute (degree-7 minimax of atan(q) / q in q^2 on [0, 1]); atan2(0, 0) = 0
template <class V>
__device__ __forceinline__ V atan2_poly(V y, V x) {
    const V ax = vabs(x), ay = vabs(y);
    const V mx = vmax(ax, ay), mn = vmin(ax, ay);
    const V q = mn * vrcp0(vmax(mx, cst<V>(1e-30f)));
    const V t = q * q;
    V p = cst<V>(-3.962627612e-03f);
    p = fma_(p, t, cst<V>(2.151858062e-02f));
    p = fma_(p, t, cst<V>(-5.539653823e-02f));
    p = fma_(p, t, cst<V>(9.602877498e-02f));
    p = fma_(p, t, cst<V>(-1.389259100e-01f));
    p = fma_(p, t, cst<V>(1.994321346e-01f));
    p = fma_(p, t, cst<V>(-3.332955241e-01f));
    p = fma_(p, t, cst<V>(9.999992251e-01f));
    V r = q * p;
    r = vsel(m_gt(ay, ax), cst<V>(0.5f * kPiF) - r, r);
    r = vsel(m_neg(x), cst<V>(kPiF) - r, r);
    return vsel(m_neg(y), -r, r);
}

// sin(x) for x in [0, 1.1]: Taylor to x^9 (4.2e-8 at pi / 3)
template <class V>
__device__ __forceinline__ V sin_small(V x) {
    const V x2 = x * x;
    V p = cst<V>(1.0f / 362880.0f);
    p = fma_(p, x2, cst<V>(-1.0f / 5040.0f));
    p = fma_(p, x2, cst<V>(1.0f / 120.0f));
    p = fma_(p, x2, cst<V>(-1.0f / 6.0f));
    return fma_(p * x2, x, x);
}

// T = 1 - 0.17 cos(h - 30deg) + 0.24 cos 2h + 0.32 cos(3h + 6deg) - 0.20 cos(4h - 63deg) from (cos h, sin h)
template <class V>
__device__ __forceinline__ V t_weight(V c1, V s1) {
    const V two = cst<V>(2.0f);
    const V c2 = fma_(two * c1, c1, cst<V>(-1.0f)), s2 = two * s1 * c1;
    const V f = cst<V>(4.0f) * c1 * c1;
    const V c3 = c1 * (f - 3.0f), s3 = s1 * (f - 1.0f);
    const V c4 = fma_(two * c2, c2, cst<V>(-1.0f)), s4 = two * s2 * c2;
    constexpr float C30 = 0.86602540378443865f, S30 = 0.5f;
    constexpr float C6 = 0.99452189536827334f, S6 = 0.10452846326765347f;
    constexpr float C63 = 0.45399049973954675f, S63 = 0.89100652418836786f;
    // 1 - 0.17 (c1 C30 + s1 S30) + 0.24 c2 + 0.32 (c3 C6 - s3 S6) - 0.20 (c4 C63 + s4 S63)
    V t = fmac(c1, -0.17f * C30, 1.0f);
    t = fmav(s1, -0.17f * S30, t);
    t = fmav(c2, 0.24f, t);
    t = fmav(c3, 0.32f * C6, t);
    t = fmav(s3, -0.32f * S6, t);
    t = fmav(c4, -0.20f * C63, t);
    return fmav(s4, -0.20f * S63, t);
}

struct DeConsts {
    float kL, kC, kH, eps;
};

// VARIANT 0: metrics/color_error.py:105-210 (eps inside sqrt / denominators, hue in (-pi, pi], sin(2 dtheta),
//            clamp >= 0 under the final sqrt);  VARIANT 1: NewBP_model/losses.py:99-136 (eps = 1e-6 inside every
//            sqrt and under the final sqrt, hue in [0, 2 pi), sin(dtheta), plain divisions by S_L, S_C, S_H)
// Chroma / hue part: Q = (dC' / (kC S_C))^2 + (dH' / (kH S_H))^2 + R_T (dC' / ..)(dH' / ..) -- independent of L1, L2
template <int VARIANT, class V>
__device__ __forceinline__ V de00_chroma_hue(V a1, V b1, V a2, V b2, const DeConsts k) {
    const V eps = cst<V>(k.eps);
    const float k25 = 6103515625.0f;
    const V c1 = vsqrt0(fma_(a1, a1, fma_(b1, b1, eps))), c2 = vsqrt0(fma_(a2, a2, fma_(b2, b2, eps)));
    const V cb = cst<V>(0.5f) * (c1 + c2);
    const V cb7 = pow7(cb);
    const V g1 = fmac(vsqrt0(cb7 * vrcp0(cb7 + k25)), -0.5f, 1.5f);   // 1 + G = 1.5 - 0.5 sqrt(.)
    const V a1p = g1 * a1, a2p = g1 * a2;
    // chroma' (eps inside the root as the reference) and EXACT unit hue vectors (the loss variant's eps = 1e-6 would
    // shorten them visibly below chroma 0.03)
    const V n1 = fma_(a1p, a1p, b1 * b1), n2 = fma_(a2p, a2p, b2 * b2);
    const V c1p = vsqrt0(n1 + eps), c2p = vsqrt0(n2 + eps);
    const V r1 = vrsqrt0(vmax(n1, cst<V>(1e-30f))), r2 = vrsqrt0(vmax(n2, cst<V>(1e-30f)));
    const V zero = cst<V>(0.0f);
    // (both components through the select: a grey colour is the CONSTANT (1, 0), zero gradient like ATen's atan2(0, 0))
    const auto chro1 = m_gt(n1, zero), chro2 = m_gt(n2, zero);
    const V u1x = vsel(chro1, a1p * r1, cst<V>(1.0f)), u1y = vsel(chro1, b1 * r1, zero);
    const V u2x = vsel(chro2, a2p * r2, cst<V>(1.0f)), u2y = vsel(chro2, b2 * r2, zero);
    const V dC = c2p - c1p;
    // signed half-angle sine and mean-hue direction from w = u1 + u2 (|w| = 2 |cos(dh / 2)|) and d = u2 - u1
    // (|d| = 2 |sin(dh / 2)|, |w|^2 + |d|^2 = 4).  The larger of the two carries the precision:
    //   hues less than a right angle apart: sin(dh / 2) = sin(dh) / (2 cos(dh / 2)) = -s / |w|,  direction = w
    //   else:                               sin(dh / 2) = sgn |d| / 2,  direction = sgn (d_y, -d_x)  (d turned by 90deg
    //                                       to the side the signed half-angle says)
    // One reciprocal square root of max(|w|^2, |d|^2) >= 2 serves both; every branch is smooth in the inputs (the dual
    // numbers of the loss gradient differentiate the selected one), no square root of a possibly zero argument.
    const V s = cross2(u1y, u2x, u1x, u2y);                 // sin(h1 - h2) (unit vectors: a grey colour counts as hue 0)
    const V ddx = u2x - u1x, ddy = u2y - u1y;
    const V wx = u1x + u2x, wy = u1y + u2y;
    const V wn2 = fma_(wx, wx, wy * wy), dn2 = fma_(ddx, ddx, ddy * ddy);
    const auto use_d = m_gt(dn2, wn2);
    const V sg = vsel(m_gt(s, zero), cst<V>(-1.0f), cst<V>(1.0f));
    const V vx = vsel(use_d, sg * ddy, wx), vy = vsel(use_d, -(sg * ddx), wy);
    const V rw = vrsqrt0(vmax(dn2, wn2));
    const V sin_half = vsel(use_d, cst<V>(0.5f) * sg * dn2, -s) * rw;
    const V dH = cst<V>(2.0f) * vsqrt0(fma_(c1p, c2p, eps)) * sin_half;
    const V ch = vx * rw, sh = vy * rw;
    const V T = t_weight(ch, sh);
    const V th = atan2_poly(vy, vx);
    const auto neg1 = m_neg(b1), neg2 = m_neg(b2);
    const auto mixed = m_xor(neg1, neg2);
    // the short arc between the hues crosses the cut of the variant's hue range
    //   metric (cut at +-pi): b1 >= 0 > b2 and s < 0, or b2 >= 0 > b1 and s > 0
    //   loss   (cut at 0):    the opposite sign of s
    const auto spos = m_gt(s, zero), sneg = m_lt(s, zero);
    const auto crosses = VARIANT == 0 ? m_and(mixed, m_or(m_and(neg2, sneg), m_and(neg1, spos)))
                                      : m_and(mixed, m_or(m_and(neg2, spos), m_and(neg1, sneg)));
    const auto lift = VARIANT == 0 ? m_and(crosses, m_neg(th)) : m_and(m_not(crosses), m_neg(th));
    const V hb = vsel(lift, th + 2.0f * kPiF, th);
    const V Cb = cst<V>(0.5f) * (c1p + c2p);
    const V q = fmac(hb, 180.0f / (kPiF * 25.0f), -11.0f);   // (deg(hbar) - 275) / 25
    const V ex = vex2(q * q * cst<V>(-1.4426950408889634f));
    const V Cb7 = pow7(Cb);
    const V rc = cst<V>(2.0f) * vsqrt0(Cb7 * vrcp0(Cb7 + k25));
    const V rt = VARIANT == 0 ? -(sin_small(cst<V>(kPiF / 3.0f) * ex) * rc)      // -sin(2 dtheta) RC
                              : -(sin_small(cst<V>(kPiF / 6.0f) * ex) * rc);     // -sin(dtheta) RC
    const V sC = fmac(Cb, 0.045f, 1.0f);
    const V sH = fma_(cst<V>(0.015f) * Cb, T, cst<V>(1.0f));
    V ct, ht;
    if (VARIANT == 0) {
        ct = dC * rcp_n(fmav(sC, k.kC, eps));
        ht = dH * rcp_n(fmav(sH, k.kH, eps));
    } else {
        ct = dC * rcp_n(sC);
        ht = dH * rcp_n(sH);
    }
    return fma_(ct, ct, fma_(ht, ht, rt * ct * ht));
}

// Lightness part: dL' / (kL S_L)
template <int VARIANT, class V>
__device__ __forceinline__ V de00_lightness(V L1, V L2, const DeConsts k) {
    const V eps = cst<V>(k.eps);
    const V dL = L2 - L1, Lb = cst<V>(0.5f) * (L1 + L2);
    const V l50 = (Lb - 50.0f) * (Lb - 50.0f);
    const V sL = fma_(cst<V>(0.015f) * l50, vrsqrt0(l50 + 20.0f + eps), cst<V>(1.0f));
    return VARIANT == 0 ? dL * rcp_n(fmav(sL, k.kL, eps)) : dL * rcp_n(sL);
}

template <int VARIANT, class V>
__device__ __forceinline__ V de00(V L1, V a1, V b1, V L2, V a2, V b2, const DeConsts k) {
    const V lt = de00_lightness<VARIANT>(L1, L2, k);
    const V sum = fma_(lt, lt, de00_chroma_hue<VARIANT>(a1, b1, a2, b2, k));
    return VARIANT == 0 ? vsqrt0(vmax(sum, cst<V>(0.0f))) : vsqrt0(sum + cst<V>(k.eps));
}

}  // namespace dem
}  // namespace llie
