// sRGB -> Lab(D65/2deg) -> CIEDE2000 maps and sums (sm_100a), both variants of the reference, fused into
// one pass over the two images (24 B/px read) instead of ~50 eager pointwise kernels:
//   * metrics/color_error.py:105-210  _deltaE00_lab_map  METRIC variant (eps 1e-12, unwrapped hue,
//                                      piecewise dh / hbar under `valid`, sin(2 dtheta), clamp >= 0)
//   * NewBP_model/losses.py:99-136    DeltaE00Loss._ciede2000  LOSS variant (eps 1e-6 everywhere, hue
//                                      mod 2pi, boolean-mask wraps, sin(dtheta), + eps under the sqrt)
//   * kornia 0.6.12 rgb_to_lab (color_error.py:79-85, losses.py:140-141) restated, see oracle/colour.py
// plus the Sobel magnitude of the L plane (color_error.py:296-301) for the edge mask.
// Precise transcendentals (no fast-math): the tolerance is 1e-4 absolute on dE00.  The kernel is bound by
// instruction issue (FP32 + SFU), not by HBM, so the three expensive groups of libm calls are replaced by
// cheaper forms of at least the same accuracy: x^2.4 by a Newton-refined fifth root (pow24, < 1 ulp),
// t^(1/3) by cbrtf (1 ulp; powf is 4), and the four cosines of T by one sincosf plus multiple-angle
// identities.
#include <math.h>

#include "common.cuh"
#include "de_core.cuh"
#include "de_dual.cuh"

// grid = LLIE_DE_WAVES waves of resident blocks: one wave leaves the slowest SM's blocks as the tail (402 us per 12 Mpx
// pair), 1.6 waves as before 383, four waves let the block scheduler even it out (366); 8 is no better
#ifndef LLIE_DE_WAVES
#define LLIE_DE_WAVES 4
#endif
#ifndef LLIE_DEGRAD_WAVES
#define LLIE_DEGRAD_WAVES 2
#endif
namespace llie {

constexpr float kPi = 3.14159265358979323846f;

// x / C for a compile-time constant C, correctly rounded, in three instructions (Markstein): with y = RN(1 / C) folded
// at compile time, q = RN(x y) is within an ulp of the quotient, r = x - C q is exact in one FMA and RN(q + r y) is the
// IEEE quotient -- for every finite x whose quotient is a normal number and every C whose significand is not all ones
// (llie_selftest_div_const + tests/test_gpu_ssim_colour.py compare it with the division instruction over hundreds of
// millions of operands).  The division instruction costs ~12 issue slots with its FCHK / slow-path branch; this
// kernel divides by constants 18 times per pixel.
__device__ __forceinline__ float div_const(float x, float c, float rc) {
    const float q = x * rc;
    const float r = fmaf(-c, q, x);
    return fmaf(r, rc, q);
}
#define DIVC(x, C) div_const((x), (C), 1.0f / (C))

struct DeArgs {
    llie_deltae_desc d;
    const float *x, *y;
    double *out_sum;
    float *out_map, *out_l;
    int *out_flags;
    double *part;  // [n][G]
    int *pflag;    // [n][G]
    unsigned *ticket;
    double *out_sq;   // NULL or [n]: sum over the 3 channels of (x - y)^2 in float64 (PSNR rides on this pass)
    double *out_sqc;  // NULL or [n][3]: the same sums per channel (CPSNR / RGB-PSNR table, channelwise.py:113-222)
    double *part_sq;  // [3][n][G]
};

// One pixel group of the dE00 pass: V = float (1 pixel) or dem::F2 (2 pixels).  Returns dE00 and L of x.
template <int VARIANT, bool EXACT_L, class V>
__device__ __forceinline__ V de_pixel(const llie_deltae_desc &d, V p0, V p1, V p2, V t0, V t1, V t2, V &Lx) {
    V L1, A1, B1, L2, A2, B2;
    if (d.input_lab) {
        L1 = p0; A1 = p1; B1 = p2;
        L2 = t0; A2 = t1; B2 = t2;
    } else {
        if (d.clamp_in) {
            p0 = dem::vsat(p0); p1 = dem::vsat(p1); p2 = dem::vsat(p2);
            t0 = dem::vsat(t0); t1 = dem::vsat(t1); t2 = dem::vsat(t2);
        }
        // the SAME conversion for both images (dE00(x, x) must be exactly 0) whatever outputs are requested
        dem::rgb_to_lab<V, false>(p0, p1, p2, L1, A1, B1);
        dem::rgb_to_lab<V, false>(t0, t1, t2, L2, A2, B2);
    }
    Lx = L1;
    return dem::de00<VARIANT>(L1, A1, B1, L2, A2, B2, dem::DeConsts{d.kL, d.kC, d.kH, d.eps});
}

// VEC: 4 consecutive pixels per thread and step (128-bit loads / stores, two packed-pair evaluations: the FP32
// add / mul / fma of two pixels share one FADD2 / FMUL2 / FFMA2 issue slot); needs plane % 4 == 0 and 16-byte aligned
// tensors.  EXACT_L (only when the L plane is written): f(Y) of x through cbrtf, see dem::rgb_to_lab.
template <int VARIANT, bool VEC, bool EXACT_L>
__global__ void __launch_bounds__(256) deltae_kernel(const DeArgs a) {
    __shared__ double red[32];
    const llie_deltae_desc &d = a.d;
    const int n = blockIdx.y, G = gridDim.x;
    const long long plane = (long long)d.h * d.w;
    const float *x0 = a.x + (long long)n * 3 * plane, *y0 = a.y + (long long)n * 3 * plane;
    float acc = 0.f, chk_x = 0.f, chk_y = 0.f;
    double sq[3] = {0.0, 0.0, 0.0};
    const bool want_sq = a.out_sq != nullptr || a.out_sqc != nullptr;
    if (VEC) {
        using dem::F2;
        const long long groups = plane >> 2;
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < groups; i += (long long)G * blockDim.x) {
            float4 p[3], t[3];
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                p[c] = __ldcs(reinterpret_cast<const float4 *>(x0 + c * plane) + i);
                t[c] = __ldcs(reinterpret_cast<const float4 *>(y0 + c * plane) + i);
            }
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                chk_x = fmaf(p[c].x, 0.f, fmaf(p[c].y, 0.f, fmaf(p[c].z, 0.f, fmaf(p[c].w, 0.f, chk_x))));
                chk_y = fmaf(t[c].x, 0.f, fmaf(t[c].y, 0.f, fmaf(t[c].z, 0.f, fmaf(t[c].w, 0.f, chk_y))));
                if (want_sq) {  // metrics/psnr.py:39-88: both images promoted to float64 before the difference
                    const double e0 = (double)p[c].x - (double)t[c].x, e1 = (double)p[c].y - (double)t[c].y;
                    const double e2 = (double)p[c].z - (double)t[c].z, e3 = (double)p[c].w - (double)t[c].w;
                    sq[c] = fma(e0, e0, fma(e1, e1, fma(e2, e2, fma(e3, e3, sq[c]))));
                }
            }
            F2 La, Lb;
            const F2 da = de_pixel<VARIANT, EXACT_L, F2>(d, dem::f2(p[0].x, p[0].y), dem::f2(p[1].x, p[1].y),
                                                         dem::f2(p[2].x, p[2].y), dem::f2(t[0].x, t[0].y),
                                                         dem::f2(t[1].x, t[1].y), dem::f2(t[2].x, t[2].y), La);
            const F2 db = de_pixel<VARIANT, EXACT_L, F2>(d, dem::f2(p[0].z, p[0].w), dem::f2(p[1].z, p[1].w),
                                                         dem::f2(p[2].z, p[2].w), dem::f2(t[0].z, t[0].w),
                                                         dem::f2(t[1].z, t[1].w), dem::f2(t[2].z, t[2].w), Lb);
            const float4 de = make_float4(dem::lo(da), dem::hi(da), dem::lo(db), dem::hi(db));
            acc += (de.x + de.y) + (de.z + de.w);
            if (a.out_map) __stcs(reinterpret_cast<float4 *>(a.out_map + (long long)n * plane) + i, de);
            if (a.out_l)
                __stcs(reinterpret_cast<float4 *>(a.out_l + (long long)n * plane) + i,
                       make_float4(dem::lo(La), dem::hi(La), dem::lo(Lb), dem::hi(Lb)));
        }
    } else {
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < plane;
             i += (long long)G * blockDim.x) {
            const float p0 = __ldcs(x0 + i), p1 = __ldcs(x0 + plane + i), p2 = __ldcs(x0 + 2 * plane + i);
            const float t0 = __ldcs(y0 + i), t1 = __ldcs(y0 + plane + i), t2 = __ldcs(y0 + 2 * plane + i);
            chk_x = fmaf(p0, 0.f, fmaf(p1, 0.f, fmaf(p2, 0.f, chk_x)));
            chk_y = fmaf(t0, 0.f, fmaf(t1, 0.f, fmaf(t2, 0.f, chk_y)));
            if (want_sq) {
                const double d0 = (double)p0 - (double)t0, d1 = (double)p1 - (double)t1, d2 = (double)p2 - (double)t2;
                sq[0] = fma(d0, d0, sq[0]);
                sq[1] = fma(d1, d1, sq[1]);
                sq[2] = fma(d2, d2, sq[2]);
            }
            float L1;
            const float de = de_pixel<VARIANT, EXACT_L, float>(d, p0, p1, p2, t0, t1, t2, L1);
            acc += de;
            if (a.out_map) a.out_map[(long long)n * plane + i] = de;
            if (a.out_l) a.out_l[(long long)n * plane + i] = L1;
        }
    }
    const double bs = block_sum<double>((double)acc, red);
    double bq[3] = {0.0, 0.0, 0.0};
    if (want_sq) {
#pragma unroll
        for (int c = 0; c < 3; ++c) bq[c] = block_sum<double>(sq[c], red);
    }
    int flag = 0;
    if (!(chk_x == 0.f)) flag |= 1;
    if (!(chk_y == 0.f)) flag |= 2;
    const int f = block_or(flag, reinterpret_cast<int *>(red));
    if (threadIdx.x == 0) {
        a.part[(size_t)n * G + blockIdx.x] = bs;
        if (want_sq) {
#pragma unroll
            for (int c = 0; c < 3; ++c) a.part_sq[((size_t)c * d.n + n) * G + blockIdx.x] = bq[c];
        }
        a.pflag[(size_t)n * G + blockIdx.x] = f;
    }
    if (!last_block_done(a.ticket, gridDim.x * gridDim.y)) return;
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    // fixed-order sums, one warp per (quantity, image): quantity 0 = dE00, 1..3 = squared error of channel 0..2
    __shared__ double s_sq[3][8];
    for (int img0 = 0; img0 < d.n; img0 += 8) {
        const int imgs = min(8, d.n - img0);
        for (int s = wrp; s < imgs * (want_sq ? 4 : 1); s += 8) {
            const int q = s / imgs, img = img0 + s % imgs;
            const double *p = (q == 0 ? a.part : a.part_sq + (size_t)(q - 1) * d.n * G) + (size_t)img * G;
            double s0 = 0.0, s1 = 0.0;
            int b = lane;
            for (; b + 32 < G; b += 64) {
                s0 += __ldcg(p + b);
                s1 += __ldcg(p + b + 32);
            }
            for (; b < G; b += 32) s0 += __ldcg(p + b);
            const double t = warp_sum(s0 + s1);
            if (lane == 0) {
                if (q == 0) a.out_sum[img] = t;
                else s_sq[q - 1][img - img0] = t;
            }
        }
        __syncthreads();
        if (want_sq && threadIdx.x < imgs) {
            const int img = img0 + threadIdx.x;
            const double c0 = s_sq[0][threadIdx.x], c1 = s_sq[1][threadIdx.x], c2 = s_sq[2][threadIdx.x];
            if (a.out_sq) a.out_sq[img] = (c0 + c1) + c2;
            if (a.out_sqc) {
                a.out_sqc[(size_t)img * 3] = c0;
                a.out_sqc[(size_t)img * 3 + 1] = c1;
                a.out_sqc[(size_t)img * 3 + 2] = c2;
            }
        }
        __syncthreads();
    }
    int fl = 0;
    for (int b = threadIdx.x; b < G * d.n; b += blockDim.x) fl |= __ldcg(a.pflag + b);
    const int ff = block_or(fl, reinterpret_cast<int *>(red));
    if (threadIdx.x == 0) *a.out_flags = ff;
}

// ---------------------------------------------------------------------------------------------------
// LOSS variant with its gradient w.r.t. x: the generic formula of de_core.cuh evaluated on forward-mode duals with three
// tangents (d/dL1, d/da1, d/db1; de_dual.cuh), then chained through the closed-form Lab Jacobian and the clamp mask, so
// value and gradient come out of ONE pass (24 B/px read, 12 B/px written) and backward is a scale by the upstream
// scalar.  ATen's rules where they matter for parity: zero hue gradient for a grey colour (atan2(0, 0)), where()-style
// clamp masks.
struct DeGradArgs {
    int n, h, w, clamp_in;
    float eps, scale;
    const float *x, *y;
    double *out_sum;
    float *out_grad;
    int *out_flags;
    double *part;
    int *pflag;
    unsigned *ticket;
};

// one pixel group (S = float: 1 pixel, S = dem::F2: 2 pixels): dE00 value and scale * d dE00 / d (r, g, b) of x
template <class S>
__device__ __forceinline__ S de_grad_pixel(const DeGradArgs &a, S r0, S r1, S r2, S t0, S t1, S t2, S &g0, S &g1, S &g2) {
    using namespace dem;
    using D2 = Dual<S, 2>;
    using D1 = Dual<S, 1>;
    S p0 = r0, p1 = r1, p2 = r2;
    const S zero = cst<S>(0.f), one = cst<S>(1.f);
    auto m0 = m_ge(one, zero), m1 = m0, m2 = m0;  // clamp pass-through masks (ATen: min <= x <= max), all true
    if (a.clamp_in) {
        m0 = m_and(m_ge(r0, zero), m_le(r0, one));
        m1 = m_and(m_ge(r1, zero), m_le(r1, one));
        m2 = m_and(m_ge(r2, zero), m_le(r2, one));
        p0 = vsat(p0); p1 = vsat(p1); p2 = vsat(p2);
        t0 = vsat(t0); t1 = vsat(t1); t2 = vsat(t2);
    }
    S d0, d1, d2;
    const S l0 = srgb_to_linear_d(p0, d0), l1 = srgb_to_linear_d(p1, d1), l2 = srgb_to_linear_d(p2, d2);
    const S X = fmav(l2, 0.180423f, fmav(l1, 0.357580f, cst<S>(0.412453f) * l0));
    const S Y = fmav(l2, 0.072169f, fmav(l1, 0.715160f, cst<S>(0.212671f) * l0));
    const S Z = fmav(l2, 0.950227f, fmav(l1, 0.119193f, cst<S>(0.019334f) * l0));
    S dfx, dfy, dfz;
    const S fx = lab_f_d(DEM_DIVC(X, 0.95047f), dfx), fy = lab_f_d(Y, dfy), fz = lab_f_d(DEM_DIVC(Z, 1.08883f), dfz);
    const S L1 = fmac(fy, 116.0f, -16.0f), A1 = cst<S>(500.0f) * (fx - fy), B1 = cst<S>(200.0f) * (fy - fz);
    // the target goes through the SAME value arithmetic (identical images -> identical Lab)
    S u0, u1, u2, e0, e1, e2;
    const S k0 = srgb_to_linear_d(t0, u0), k1 = srgb_to_linear_d(t1, u1), k2 = srgb_to_linear_d(t2, u2);
    const S X2 = fmav(k2, 0.180423f, fmav(k1, 0.357580f, cst<S>(0.412453f) * k0));
    const S Y2 = fmav(k2, 0.072169f, fmav(k1, 0.715160f, cst<S>(0.212671f) * k0));
    const S Z2 = fmav(k2, 0.950227f, fmav(k1, 0.119193f, cst<S>(0.019334f) * k0));
    const S fx2 = lab_f_d(DEM_DIVC(X2, 0.95047f), e0), fy2 = lab_f_d(Y2, e1), fz2 = lab_f_d(DEM_DIVC(Z2, 1.08883f), e2);
    const S L2 = fmac(fy2, 116.0f, -16.0f), A2 = cst<S>(500.0f) * (fx2 - fy2), B2 = cst<S>(200.0f) * (fy2 - fz2);
    // dE = sqrt(lt^2 + Q + eps): Q (chroma / hue) with tangents (d/da1, d/db1), lt (lightness) with d/dL1
    const DeConsts kc{1.f, 1.f, 1.f, a.eps};
    const D2 Q = de00_chroma_hue<1>(dual_var<S, 2>(A1, 0), dual_var<S, 2>(B1, 1), dual_const<S, 2>(A2),
                                    dual_const<S, 2>(B2), kc);
    const D1 lt = de00_lightness<1>(dual_var<S, 1>(L1, 0), dual_const<S, 1>(L2), kc);
    const S sum = fma_(lt.v, lt.v, Q.v) + a.eps;
    const S rE = vrsqrt0(sum);
    const S E0 = sum * rE;
    const S dE = fma_(fma_(-E0, E0, sum), cst<S>(0.5f) * rE, E0);     // sqrt(sum), Newton-refined
    const S hE = cst<S>(0.5f) * rE;                                    // 1 / (2 sqrt(sum))
    const S deL = cst<S>(2.0f) * lt.v * lt.t[0] * hE, deA = Q.t[0] * hE, deB = Q.t[1] * hE;
    // reverse through Lab -> f -> XYZ -> linear -> sRGB
    const S gfx = cst<S>(500.0f) * deA;
    const S gfy = fmav(deB, 200.0f, fmav(deA, -500.0f, cst<S>(116.0f) * deL));
    const S gfz = cst<S>(-200.0f) * deB;
    const S gX = DEM_DIVC(gfx * dfx, 0.95047f), gY = gfy * dfy, gZ = DEM_DIVC(gfz * dfz, 1.08883f);
    const S gl0 = fmav(gZ, 0.019334f, fmav(gY, 0.212671f, cst<S>(0.412453f) * gX));
    const S gl1 = fmav(gZ, 0.119193f, fmav(gY, 0.715160f, cst<S>(0.357580f) * gX));
    const S gl2 = fmav(gZ, 0.950227f, fmav(gY, 0.072169f, cst<S>(0.180423f) * gX));
    // where()-style masking: a clamped-away channel gets exactly 0 even when the chain holds NaN
    const S sc = cst<S>(a.scale);
    g0 = vsel(m0, sc * gl0 * d0, zero);
    g1 = vsel(m1, sc * gl1 * d1, zero);
    g2 = vsel(m2, sc * gl2 * d2, zero);
    return dE;
}

// losses.py:138-143 value (per-image sums) + `scale` * d dE / d x, one pass, one pixel per thread and step in scalar
// FP32.  The dual arithmetic is bound by the FP32 pipe itself (68 % pipe-active, 50 % issue-active in ncu): FFMA2
// issues at half rate, so packing two pixels per thread saves issue slots this kernel is not short of and costs
// occupancy (174 vs 75 registers: 184 vs 175 us at 16x3x512x512; four scalar pixels per thread behind 128-bit loads:
// 104 registers, 180 us).
__global__ void __launch_bounds__(256) deltae_loss_grad_kernel(const DeGradArgs a) {
    __shared__ double red[32];
    const int n = blockIdx.y, G = gridDim.x;
    const long long plane = (long long)a.h * a.w;
    const float *x0 = a.x + (long long)n * 3 * plane, *y0 = a.y + (long long)n * 3 * plane;
    float *g0 = a.out_grad + (long long)n * 3 * plane;
    float acc = 0.f, chk_x = 0.f, chk_y = 0.f;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < plane; i += (long long)G * blockDim.x) {
        const float r0 = __ldcs(x0 + i), r1 = __ldcs(x0 + plane + i), r2 = __ldcs(x0 + 2 * plane + i);
        const float t0 = __ldcs(y0 + i), t1 = __ldcs(y0 + plane + i), t2 = __ldcs(y0 + 2 * plane + i);
        chk_x = fmaf(r0, 0.f, fmaf(r1, 0.f, fmaf(r2, 0.f, chk_x)));
        chk_y = fmaf(t0, 0.f, fmaf(t1, 0.f, fmaf(t2, 0.f, chk_y)));
        float q0, q1, q2;
        acc += de_grad_pixel<float>(a, r0, r1, r2, t0, t1, t2, q0, q1, q2);
        __stcs(g0 + i, q0);
        __stcs(g0 + plane + i, q1);
        __stcs(g0 + 2 * plane + i, q2);
    }
    const double bs = block_sum<double>((double)acc, red);
    int flag = 0;
    if (!(chk_x == 0.f)) flag |= 1;
    if (!(chk_y == 0.f)) flag |= 2;
    const int f = block_or(flag, reinterpret_cast<int *>(red));
    if (threadIdx.x == 0) {
        a.part[(size_t)n * G + blockIdx.x] = bs;
        a.pflag[(size_t)n * G + blockIdx.x] = f;
    }
    if (!last_block_done(a.ticket, gridDim.x * gridDim.y)) return;
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    for (int s = wrp; s < a.n; s += 8) {
        const double *p = a.part + (size_t)s * G;
        double s0 = 0.0;
        for (int b = lane; b < G; b += 32) s0 += __ldcg(p + b);
        const double t = warp_sum(s0);
        if (lane == 0) a.out_sum[s] = t;
    }
    int fl = 0;
    for (int b = threadIdx.x; b < G * a.n; b += blockDim.x) fl |= __ldcg(a.pflag + b);
    const int ff = block_or(fl, reinterpret_cast<int *>(red));
    if (threadIdx.x == 0) *a.out_flags = ff;
}

// Sobel magnitude of a [planes, H, W] stack, zero padding (color_error.py:23-30, 296-301).  Products and
// sums are kept un-contracted (no FMA) so that the result is the IEEE value torch computes.
// One thread per column of a 32-row band: the three rows of the window roll through registers and the left /
// right neighbours come from warp shuffles (the two edge lanes of a warp load theirs), so each value is read
// from memory ~1.1 times instead of 9 and the kernel is bound by HBM, not by address arithmetic.
constexpr int kSobelRows = 32;
#ifndef LLIE_SOBEL4_ROWS
#define LLIE_SOBEL4_ROWS 12
#endif
constexpr int kSobel4Rows = LLIE_SOBEL4_ROWS;  // rows per block of the four-column kernel (even): 32 rows left
// 20 warps per SM in flight on a 12 Mpx plane (37 us); 16 / 12 / 8 rows: 35 / 31 / 31 us (tools/time_sobel.py)

__device__ __forceinline__ float sobel_mag(float v00, float v01, float v02, float v10, float v12, float v20, float v21,
                                           float v22) {
    // gx = sum_k Kx[k] * v[k] in row-major tap order, gy likewise (zero taps contribute exactly 0)
    float gx = __fmul_rn(-1.f, v00);
    gx = __fadd_rn(gx, v02);
    gx = __fadd_rn(gx, __fmul_rn(-2.f, v10));
    gx = __fadd_rn(gx, __fmul_rn(2.f, v12));
    gx = __fadd_rn(gx, __fmul_rn(-1.f, v20));
    gx = __fadd_rn(gx, v22);
    float gy = __fmul_rn(-1.f, v00);
    gy = __fadd_rn(gy, __fmul_rn(-2.f, v01));
    gy = __fadd_rn(gy, __fmul_rn(-1.f, v02));
    gy = __fadd_rn(gy, v20);
    gy = __fadd_rn(gy, __fmul_rn(2.f, v21));
    gy = __fadd_rn(gy, v22);
    const float s = __fadd_rn(__fadd_rn(__fmul_rn(gx, gx), __fmul_rn(gy, gy)), 1e-12f);
    return __fsqrt_rn(s);
}

// grid (ceil(W / 256), ceil(H / kSobelRows), planes), 256 threads
__global__ void __launch_bounds__(256) sobel_kernel(const float *L, float *g, int H, int W, int rows) {
    const int x = blockIdx.x * 256 + threadIdx.x;
    const int y0 = blockIdx.y * rows, y1 = min(y0 + rows, H);
    const long long plane = (long long)H * W;
    const float *p = L + (long long)blockIdx.z * plane;
    float *o = g + (long long)blockIdx.z * plane;
    const int lane = threadIdx.x & 31;
    const bool in = x < W;
    auto row3 = [&](int yy, float &l, float &c, float &r) {
        // (l, c, r) = L[yy][x-1 .. x+1] with zeros outside the image
        c = 0.f;
        float el = 0.f, er = 0.f;
        if (yy >= 0 && yy < H) {
            const float *rp = p + (long long)yy * W;
            if (in) c = __ldg(rp + x);
            if (lane == 0 && x >= 1 && x - 1 < W) el = __ldg(rp + x - 1);
            if (lane == 31 && x + 1 < W) er = __ldg(rp + x + 1);
        }
        l = __shfl_up_sync(0xffffffffu, c, 1);
        r = __shfl_down_sync(0xffffffffu, c, 1);
        if (lane == 0) l = el;
        if (lane == 31) r = er;
    };
    float a0, a1, a2, b0, b1, b2;
    row3(y0 - 1, a0, a1, a2);
    row3(y0, b0, b1, b2);
    // four rows per step: their loads are independent and issue back to back (one 4-byte load in flight per thread
    // left the kernel latency-bound at 2.2 TB/s); the arithmetic per output is unchanged
    for (int y = y0; y < y1; y += 4) {
        float n0[4], n1[4], n2[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) row3(y + 1 + u, n0[u], n1[u], n2[u]);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (in && y + u < y1) o[(long long)(y + u) * W + x] = sobel_mag(a0, a1, a2, b0, b2, n0[u], n1[u], n2[u]);
            a0 = b0; a1 = b1; a2 = b2;
            b0 = n0[u]; b1 = n1[u]; b2 = n2[u];
        }
    }
}

// Four columns per thread (W % 4 == 0, 16-byte aligned planes): one 128-bit load and store per row instead of four
// 32-bit ones, the outside neighbours by shuffle (warp-edge lanes load them); same sobel_mag per pixel.
// grid (ceil(W / 256), ceil(H / kSobel4Rows), planes), 64 threads
__global__ void __launch_bounds__(64) sobel4_kernel(const float *L, float *g, int H, int W) {
    const int x = (blockIdx.x * 64 + threadIdx.x) * 4;
    const int y0 = blockIdx.y * kSobel4Rows, y1 = min(y0 + kSobel4Rows, H);
    const long long plane = (long long)H * W;
    const float *p = L + (long long)blockIdx.z * plane;
    float *o = g + (long long)blockIdx.z * plane;
    const int lane = threadIdx.x & 31;
    const bool in = x < W;
    auto row6 = [&](int yy, float (&v)[6]) {  // v = L[yy][x-1 .. x+4], zeros outside the image
        float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
        float el = 0.f, er = 0.f;
        if (yy >= 0 && yy < H) {
            const float *rp = p + (long long)yy * W;
            if (in) c = __ldg(reinterpret_cast<const float4 *>(rp + x));
            if (lane == 0 && x >= 1 && x - 1 < W) el = __ldg(rp + x - 1);
            if (lane == 31 && x + 4 < W) er = __ldg(rp + x + 4);
        }
        const float l = __shfl_up_sync(0xffffffffu, c.w, 1), r = __shfl_down_sync(0xffffffffu, c.x, 1);
        v[0] = lane == 0 ? el : l;
        v[1] = c.x; v[2] = c.y; v[3] = c.z; v[4] = c.w;
        v[5] = lane == 31 ? er : r;
    };
    float a[6], b[6];
    row6(y0 - 1, a);
    row6(y0, b);
    for (int y = y0; y < y1; y += 2) {
        float n0[6], n1[6];
        row6(y + 1, n0);
        row6(y + 2, n1);
        if (in) {
            float4 r;
            r.x = sobel_mag(a[0], a[1], a[2], b[0], b[2], n0[0], n0[1], n0[2]);
            r.y = sobel_mag(a[1], a[2], a[3], b[1], b[3], n0[1], n0[2], n0[3]);
            r.z = sobel_mag(a[2], a[3], a[4], b[2], b[4], n0[2], n0[3], n0[4]);
            r.w = sobel_mag(a[3], a[4], a[5], b[3], b[5], n0[3], n0[4], n0[5]);
            *reinterpret_cast<float4 *>(o + (long long)y * W + x) = r;
            if (y + 1 < y1) {
                r.x = sobel_mag(b[0], b[1], b[2], n0[0], n0[2], n1[0], n1[1], n1[2]);
                r.y = sobel_mag(b[1], b[2], b[3], n0[1], n0[3], n1[1], n1[2], n1[3]);
                r.z = sobel_mag(b[2], b[3], b[4], n0[2], n0[4], n1[2], n1[3], n1[4]);
                r.w = sobel_mag(b[3], b[4], b[5], n0[3], n0[5], n1[3], n1[4], n1[5]);
                *reinterpret_cast<float4 *>(o + (long long)(y + 1) * W + x) = r;
            }
        }
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            a[k] = n0[k];
            b[k] = n1[k];
        }
    }
}

}  // namespace llie

using namespace llie;

static int deltae_launch(const llie_deltae_desc *desc, const float *x, const float *y, double *out_sum,
                         double *out_sq, double *out_sqc, float *out_map, float *out_l, int32_t *out_flags, void *workspace,
                         void *stream) {
    LLIE_REQUIRE(desc && x && y && out_sum && out_flags && workspace, "deltae: null pointer");
    LLIE_REQUIRE(desc->n > 0 && desc->h > 0 && desc->w > 0, "deltae: empty tensor");
    LLIE_REQUIRE(desc->n <= 65535, "deltae: batch too large for grid.y");
    LLIE_REQUIRE(desc->variant == 0 || desc->variant == 1, "deltae: variant must be 0 (metric) or 1 (loss)");
    DeArgs a{};
    a.d = *desc;
    a.x = x; a.y = y; a.out_sum = out_sum; a.out_map = out_map; a.out_l = out_l; a.out_flags = out_flags;
    const long long plane = (long long)desc->h * desc->w;
    const bool vec = (plane & 3) == 0 && aligned16(x) && aligned16(y) && (out_map == nullptr || aligned16(out_map)) &&
                     (out_l == nullptr || aligned16(out_l));
    const bool exact_l = out_l != nullptr;
    void (*kern)(const DeArgs) = nullptr;
    const int key = (desc->variant ? 4 : 0) | (vec ? 2 : 0) | (exact_l ? 1 : 0);
    switch (key) {
        case 0: kern = deltae_kernel<0, false, false>; break;
        case 1: kern = deltae_kernel<0, false, true>; break;
        case 2: kern = deltae_kernel<0, true, false>; break;
        case 3: kern = deltae_kernel<0, true, true>; break;
        case 4: kern = deltae_kernel<1, false, false>; break;
        case 5: kern = deltae_kernel<1, false, true>; break;
        case 6: kern = deltae_kernel<1, true, false>; break;
        default: kern = deltae_kernel<1, true, true>; break;
    }
    static int occ_tab[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (occ_tab[key] == 0) occ_tab[key] = resident_blocks(kern, 256);
    const int occ = occ_tab[key];
    int G = kNumSM * occ * LLIE_DE_WAVES / desc->n;
    const long long maxg = ((vec ? plane / 4 : plane) + 255) / 256;
    if (G > maxg) G = (int)maxg;
    if (G < 1) G = 1;
    Workspace ws(workspace);
    const size_t slots = (size_t)G * desc->n;
    LLIE_REQUIRE(slots * (4 * sizeof(double) + sizeof(int)) <= ws.scratch_bytes, "deltae: workspace too small");
    a.part = reinterpret_cast<double *>(ws.scratch);
    a.part_sq = a.part + slots;
    a.pflag = reinterpret_cast<int *>(ws.scratch + 4 * slots * sizeof(double));
    a.out_sq = out_sq;
    a.out_sqc = out_sqc;
    a.ticket = ws.tickets;
    dim3 grid(G, desc->n);
    kern<<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(a);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

extern "C" int llie_deltae(const llie_deltae_desc *desc, const float *x, const float *y, double *out_sum,
                           float *out_map, float *out_l, int32_t *out_flags, void *workspace, void *stream) {
    return deltae_launch(desc, x, y, out_sum, nullptr, nullptr, out_map, out_l, out_flags, workspace, stream);
}

extern "C" int llie_deltae_sqerr(const llie_deltae_desc *desc, const float *x, const float *y, double *out_sum,
                                 double *out_sqerr, float *out_map, float *out_l, int32_t *out_flags,
                                 void *workspace, void *stream) {
    LLIE_REQUIRE(out_sqerr, "deltae_sqerr: null pointer");
    return deltae_launch(desc, x, y, out_sum, out_sqerr, nullptr, out_map, out_l, out_flags, workspace, stream);
}

extern "C" int llie_deltae_sqerr_rgb(const llie_deltae_desc *desc, const float *x, const float *y, double *out_sum,
                                     double *out_sqerr, double *out_sqerr_rgb, float *out_map, float *out_l,
                                     int32_t *out_flags, void *workspace, void *stream) {
    LLIE_REQUIRE(out_sqerr_rgb, "deltae_sqerr_rgb: null pointer");
    return deltae_launch(desc, x, y, out_sum, out_sqerr, out_sqerr_rgb, out_map, out_l, out_flags, workspace, stream);
}

extern "C" int llie_deltae_loss_grad(const llie_deltae_desc *desc, const float *x, const float *y, float grad_scale,
                                     double *out_sum, float *out_grad, int32_t *out_flags, void *workspace,
                                     void *stream) {
    LLIE_REQUIRE(desc && x && y && out_sum && out_grad && out_flags && workspace, "deltae_loss_grad: null pointer");
    LLIE_REQUIRE(desc->n > 0 && desc->h > 0 && desc->w > 0, "deltae_loss_grad: empty tensor");
    LLIE_REQUIRE(desc->n <= 65535, "deltae_loss_grad: batch too large for grid.y");
    LLIE_REQUIRE(desc->variant == 1 && !desc->input_lab, "deltae_loss_grad: loss variant on sRGB inputs only");
    DeGradArgs a{};
    a.n = desc->n; a.h = desc->h; a.w = desc->w; a.clamp_in = desc->clamp_in;
    a.eps = desc->eps; a.scale = grad_scale;
    a.x = x; a.y = y; a.out_sum = out_sum; a.out_grad = out_grad; a.out_flags = out_flags;
    const long long plane = (long long)desc->h * desc->w;
    static const int occ = resident_blocks(deltae_loss_grad_kernel, 256);
    int G = kNumSM * occ * LLIE_DEGRAD_WAVES / desc->n;
    const long long maxg = (plane + 255) / 256;
    if (G > maxg) G = (int)maxg;
    if (G < 1) G = 1;
    Workspace ws(workspace);
    const size_t slots = (size_t)G * desc->n;
    LLIE_REQUIRE(slots * (sizeof(double) + sizeof(int)) <= ws.scratch_bytes, "deltae_loss_grad: workspace too small");
    a.part = reinterpret_cast<double *>(ws.scratch);
    a.pflag = reinterpret_cast<int *>(ws.scratch + slots * sizeof(double));
    a.ticket = ws.tickets;
    dim3 grid(G, desc->n);
    deltae_loss_grad_kernel<<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(a);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

namespace llie {
__global__ void srgb_to_linear_kernel(const float *in, float *out, long long n) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        out[i] = dem::srgb_to_linear<float>(in[i]);
}
}  // namespace llie

// self-test of div_const: operands are ALL float bit patterns first_bits .. first_bits + count - 1; out_mismatch counts
// the operands whose three-instruction quotient differs from the division instruction (NaN results compare equal)
namespace llie {
template <int WHICH>
__device__ __forceinline__ bool divc_differs(float x) {
    float a, b;
    if (WHICH == 0) { a = DIVC(x, 1.055f); b = __fdiv_rn(x, 1.055f); }
    else if (WHICH == 1) { a = DIVC(x, 12.92f); b = __fdiv_rn(x, 12.92f); }
    else if (WHICH == 2) { a = DIVC(x, 0.95047f); b = __fdiv_rn(x, 0.95047f); }
    else if (WHICH == 3) { a = DIVC(x, 1.08883f); b = __fdiv_rn(x, 1.08883f); }
    else if (WHICH == 4) { a = DIVC(x, 25.0f); b = __fdiv_rn(x, 25.0f); }
    else if (WHICH == 5) { a = DIVC(x, kPi); b = __fdiv_rn(x, kPi); }
    else { a = DIVC(x, 65535.0f); b = __fdiv_rn(x, 65535.0f); }  // the uint16 decode of ingest.cu
    return __float_as_uint(a) != __float_as_uint(b) && !(a != a && b != b);
}
__global__ void div_const_selftest_kernel(int which, unsigned first_bits, unsigned long long count,
                                          unsigned long long *out_mismatch) {
    unsigned long long bad = 0;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < count;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        const float x = __uint_as_float(first_bits + (unsigned)i);
        bool d;
        switch (which) {
            case 0: d = divc_differs<0>(x); break;
            case 1: d = divc_differs<1>(x); break;
            case 2: d = divc_differs<2>(x); break;
            case 3: d = divc_differs<3>(x); break;
            case 4: d = divc_differs<4>(x); break;
            case 5: d = divc_differs<5>(x); break;
            default: d = divc_differs<6>(x); break;
        }
        bad += d ? 1ull : 0ull;
    }
    bad = warp_sum(bad);
    if ((threadIdx.x & 31) == 0 && bad) atomicAdd(out_mismatch, bad);
}
}  // namespace llie

extern "C" int llie_selftest_div_const(int32_t which, uint32_t first_bits, uint64_t count, uint64_t *out_mismatch,
                                       void *stream) {
    LLIE_REQUIRE(out_mismatch && which >= 0 && which <= 6 && count > 0, "selftest_div_const: bad arguments");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    LLIE_CUDA_OK(cudaMemsetAsync(out_mismatch, 0, sizeof(uint64_t), st));
    div_const_selftest_kernel<<<kNumSM * 16, 256, 0, st>>>(which, first_bits, count,
                                                           reinterpret_cast<unsigned long long *>(out_mismatch));
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

extern "C" int llie_selftest_srgb_to_linear(const float *in, float *out, int64_t n, void *stream) {
    LLIE_REQUIRE(in && out && n > 0, "selftest_srgb_to_linear: bad arguments");
    long long blocks = (n + 255) / 256;
    if (blocks > kNumSM * 16) blocks = kNumSM * 16;
    srgb_to_linear_kernel<<<(int)blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(in, out, n);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

extern "C" int llie_sobel_magnitude(const float *l_planes, float *out_g, int64_t planes, int32_t h, int32_t w,
                                    void *stream) {
    LLIE_REQUIRE(l_planes && out_g && planes > 0 && h > 0 && w > 0, "sobel: bad arguments");
    const bool four = (w & 3) == 0 && aligned16(l_planes) && aligned16(out_g);
    const int rows = four ? kSobel4Rows : kSobelRows;  // (a one-wave grid of 44-row blocks was slower: 59 vs 45 us)
    LLIE_REQUIRE(planes <= 65535 && (h + rows - 1) / rows <= 65535, "sobel: grid too large");
    dim3 grid((w + 255) / 256, (h + rows - 1) / rows, (unsigned)planes);
    if (four)
        sobel4_kernel<<<grid, 64, 0, reinterpret_cast<cudaStream_t>(stream)>>>(l_planes, out_g, h, w);
    else
        sobel_kernel<<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(l_planes, out_g, h, w, rows);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}
