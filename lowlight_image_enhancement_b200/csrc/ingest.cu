// Input decode / normalise step in front of the loss and metric path.
//
//   decode_u16   16-bit PNG samples (interleaved [n, h, w, c], as the image decoder leaves them) ->
//                planar float32 raw = u / 65535 and aligned = clip(raw * rho, 0, 1), optional crop window and
//                B<->R swap.  HBM-bound: 2 B read + 4 (or 8) B written per element.
//   linear_to_srgb   IEC 61966-2-1 OETF of the evaluation helpers, with the LPIPS range map 2 s - 1 as an optional
//                second output of the same pass.
//
// Both follow the host arithmetic of the reference operation by operation (one IEEE rounding per numpy / ATen
// operation, no FMA contraction), so decode_u16 is bit-exact and linear_to_srgb differs from ATen only by the
// rounding of x^(1/2.4) (here cbrt(x) * sqrt(sqrt(cbrt(x))), < 3 ulp).
#include "common.cuh"

#ifndef LLIE_INGEST_WAVES
#define LLIE_INGEST_WAVES 12
#endif
namespace llie {

constexpr int kDecThreads = 256;

struct DecArgs {
    llie_decode_desc d;
    const uint16_t *src;
    const float *rho;
    float rho_scalar;
    float *raw, *aligned;
};

// u / 65535 as q = u * rc, r = fma(-65535, q, u), q' = fma(r, rc, q): the correctly rounded quotient (same bits as
// the division instruction; llie_selftest_div_const(6) checks every float in [2^-100, 2^100]) in 3 instructions.
__device__ __forceinline__ float dec_raw(unsigned u) {
    const float x = (float)u, rc = 1.0f / 65535.0f;
    const float q = __fmul_rn(x, rc);
    return __fmaf_rn(__fmaf_rn(-65535.0f, q, x), rc, q);
}
__device__ __forceinline__ float dec_aligned(float raw, float rho) {
    return fminf(fmaxf(__fmul_rn(raw, rho), 0.f), 1.f);
}

// C == 3, crop width % 4 == 0: one thread = 4 pixels of a row = 12 samples (24 B in, 3 x 16 B out per plane set).
// VEC: the 24 B are three aligned 8-byte loads; otherwise twelve 2-byte loads (same sectors, L1-coalesced).
// The flat item index is split into (row, quad) once per thread and then advanced by the grid stride with adds and
// compares only (step_rows / step_quads come from the host); two items are in flight per thread.
template <bool VEC>
__device__ __forceinline__ void dec_load12(const uint16_t *sp, unsigned (&s)[12]) {
    if (VEC) {
        const uint2 *sv = reinterpret_cast<const uint2 *>(sp);
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const uint2 v = __ldg(sv + k);
            s[4 * k] = v.x & 0xffffu;
            s[4 * k + 1] = v.x >> 16;
            s[4 * k + 2] = v.y & 0xffffu;
            s[4 * k + 3] = v.y >> 16;
        }
    } else {
#pragma unroll
        for (int k = 0; k < 12; ++k) s[k] = __ldg(sp + k);
    }
}

__device__ __forceinline__ void dec_store12(const DecArgs &a, const unsigned (&s)[12], float rho, long long o,
                                            long long plane) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {  // c = source channel; it lands in plane c, or 2 - c under the B<->R swap
        const long long oc = o + (a.d.swap_rb ? 2 - c : c) * plane;
        float v[4];
#pragma unroll
        for (int p = 0; p < 4; ++p) v[p] = dec_raw(s[3 * p + c]);
        if (a.raw != nullptr) __stcs(reinterpret_cast<float4 *>(a.raw + oc), make_float4(v[0], v[1], v[2], v[3]));
        if (a.aligned != nullptr)
            __stcs(reinterpret_cast<float4 *>(a.aligned + oc),
                   make_float4(dec_aligned(v[0], rho), dec_aligned(v[1], rho), dec_aligned(v[2], rho),
                               dec_aligned(v[3], rho)));
    }
}

struct DecPos {
    int row, qx;  // row = n * crop_h + y
    __device__ __forceinline__ void advance(int step_rows, int step_quads, int quads) {
        row += step_rows;
        qx += step_quads;
        if (qx >= quads) {
            qx -= quads;
            ++row;
        }
    }
};

template <bool VEC>
__global__ void __launch_bounds__(kDecThreads) decode_u16_c3_kernel(const DecArgs a, int step_rows, int step_quads) {
    const llie_decode_desc &d = a.d;
    const int quads = d.crop_w >> 2;
    const int rows = d.n * d.crop_h;
    const long long plane = (long long)d.crop_h * d.crop_w;
    const unsigned t0 = blockIdx.x * kDecThreads + threadIdx.x;
    DecPos p{(int)(t0 / (unsigned)quads), (int)(t0 % (unsigned)quads)};
    auto src_of = [&](const DecPos &q, int &n, int &y) {
        n = d.n == 1 ? 0 : q.row / d.crop_h;
        y = q.row - n * d.crop_h;
        return a.src + (((long long)n * d.h + d.top + y) * d.w + d.left + 4 * q.qx) * 3;
    };
    while (p.row < rows) {
        DecPos p2 = p;
        p2.advance(step_rows, step_quads, quads);
        const bool two = p2.row < rows;
        int n0, y0, n1 = 0, y1 = 0;
        unsigned s0[12], s1[12];
        dec_load12<VEC>(src_of(p, n0, y0), s0);
        if (two) dec_load12<VEC>(src_of(p2, n1, y1), s1);
        const float rho0 = a.rho != nullptr ? __ldg(a.rho + n0) : a.rho_scalar;
        dec_store12(a, s0, rho0, ((long long)n0 * 3) * plane + (long long)y0 * d.crop_w + 4 * p.qx, plane);
        if (two) {
            const float rho1 = a.rho != nullptr ? __ldg(a.rho + n1) : a.rho_scalar;
            dec_store12(a, s1, rho1, ((long long)n1 * 3) * plane + (long long)y1 * d.crop_w + 4 * p2.qx, plane);
        }
        p = p2;
        p.advance(step_rows, step_quads, quads);
    }
}

// any channel count / width: one thread per output element (planar order, coalesced stores)
__global__ void __launch_bounds__(kDecThreads) decode_u16_generic_kernel(const DecArgs a) {
    const llie_decode_desc &d = a.d;
    const long long plane = (long long)d.crop_h * d.crop_w;
    const long long total = (long long)d.n * d.c * plane;
    for (long long t = (long long)blockIdx.x * kDecThreads + threadIdx.x; t < total;
         t += (long long)gridDim.x * kDecThreads) {
        const int x = (int)(t % d.crop_w);
        long long r = t / d.crop_w;
        const int y = (int)(r % d.crop_h);
        r /= d.crop_h;
        const int c = (int)(r % d.c), n = (int)(r / d.c);
        const int sc = (d.swap_rb && d.c >= 3 && c < 3) ? 2 - c : c;
        const unsigned u = __ldg(a.src + (((long long)n * d.h + d.top + y) * d.w + d.left + x) * d.c + sc);
        const float v = dec_raw(u);
        if (a.raw != nullptr) a.raw[t] = v;
        if (a.aligned != nullptr) a.aligned[t] = dec_aligned(v, a.rho != nullptr ? __ldg(a.rho + n) : a.rho_scalar);
    }
}

// ---------------------------------------------------------------------------------------------------
// x -> where(x <= 0.0031308, 12.92 x, 1.055 * max(x, 0.0031308)^(1/2.4) - 0.055) after clamp(x, 0, 1);
// NaN propagates as in torch.clamp / torch.where.
__device__ __forceinline__ float srgb_oetf(float x) {
    if (x != x) return x;
    x = fminf(fmaxf(x, 0.f), 1.f);
    const float xt = fmaxf(x, 0.0031308f);
    const float t = cbrtf(xt);                     // x^(1/3)
    const float p = __fmul_rn(t, sqrtf(sqrtf(t)));  // x^(1/3) * x^(1/12) = x^(5/12)
    const float high = __fsub_rn(__fmul_rn(1.055f, p), 0.055f);
    return x <= 0.0031308f ? __fmul_rn(12.92f, x) : high;
}

__global__ void __launch_bounds__(256) linear_to_srgb_kernel(const float *__restrict__ x, long long n,
                                                             float *__restrict__ out_srgb,
                                                             float *__restrict__ out_lpips, bool vec) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long t0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (vec) {
        const long long n4 = n >> 2;
        for (long long i = t0; i < n4; i += stride) {
            const float4 v = ldg_stream4(x + 4 * i);
            const float4 s = make_float4(srgb_oetf(v.x), srgb_oetf(v.y), srgb_oetf(v.z), srgb_oetf(v.w));
            if (out_srgb != nullptr) __stcs(reinterpret_cast<float4 *>(out_srgb) + i, s);
            if (out_lpips != nullptr)
                __stcs(reinterpret_cast<float4 *>(out_lpips) + i,
                       make_float4(__fsub_rn(__fmul_rn(s.x, 2.f), 1.f), __fsub_rn(__fmul_rn(s.y, 2.f), 1.f),
                                   __fsub_rn(__fmul_rn(s.z, 2.f), 1.f), __fsub_rn(__fmul_rn(s.w, 2.f), 1.f)));
        }
        for (long long i = (n4 << 2) + t0; i < n; i += stride) {
            const float s = srgb_oetf(__ldg(x + i));
            if (out_srgb != nullptr) out_srgb[i] = s;
            if (out_lpips != nullptr) out_lpips[i] = __fsub_rn(__fmul_rn(s, 2.f), 1.f);
        }
    } else {
        for (long long i = t0; i < n; i += stride) {
            const float s = srgb_oetf(__ldg(x + i));
            if (out_srgb != nullptr) out_srgb[i] = s;
            if (out_lpips != nullptr) out_lpips[i] = __fsub_rn(__fmul_rn(s, 2.f), 1.f);
        }
    }
}

static int grid_for(long long work_items, int threads, int per_sm) {
    const long long blocks = (work_items + threads - 1) / threads;
    const long long cap = (long long)kNumSM * per_sm;
    return (int)(blocks < cap ? (blocks < 1 ? 1 : blocks) : cap);
}

}  // namespace llie

using namespace llie;

extern "C" int llie_decode_u16(const llie_decode_desc *desc, const uint16_t *src, const float *rho, float rho_scalar,
                               float *out_raw, float *out_aligned, void *stream) {
    LLIE_REQUIRE(desc != nullptr && src != nullptr, "llie_decode_u16: null descriptor / source");
    const llie_decode_desc &d = *desc;
    LLIE_REQUIRE(d.n >= 1 && d.h >= 1 && d.w >= 1 && d.c >= 1, "llie_decode_u16: empty source [%d,%d,%d,%d]", d.n, d.h,
                 d.w, d.c);
    LLIE_REQUIRE(d.crop_h >= 1 && d.crop_w >= 1 && d.top >= 0 && d.left >= 0 && d.top + d.crop_h <= d.h &&
                     d.left + d.crop_w <= d.w,
                 "llie_decode_u16: crop window (%d,%d)+(%d,%d) outside the %dx%d source", d.top, d.left, d.crop_h,
                 d.crop_w, d.h, d.w);
    LLIE_REQUIRE(out_raw != nullptr || out_aligned != nullptr, "llie_decode_u16: no output requested");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    DecArgs a{d, src, rho, rho_scalar, out_raw, out_aligned};
    const bool out_ok = (out_raw == nullptr || aligned16(out_raw)) && (out_aligned == nullptr || aligned16(out_aligned));
    if (d.c == 3 && (d.crop_w & 3) == 0 && out_ok) {
        const long long items = (long long)d.n * d.crop_h * (d.crop_w >> 2);
        LLIE_REQUIRE(items < (1ll << 31) && (long long)d.n * d.crop_h < (1ll << 30),
                     "llie_decode_u16: %lld four-pixel items exceed the 32-bit item index", items);
        const bool vec = (reinterpret_cast<uintptr_t>(src) & 7u) == 0 && (d.w & 3) == 0 && (d.left & 3) == 0;
        // LLIE_INGEST_WAVES (12) waves of the blocks that fit: these write-dominated passes gain from finer items
        // (1 / 2 / 4 / 8 / 12 / 16 waves: 68.1 / 67.9 / 64.6 / 61.5 / 60.3 / 60.0 us per 12 Mpx frame for raw + aligned,
        // raw only best at 8-12; tools/time_ingest.py)
        static const int occ_v = resident_blocks(decode_u16_c3_kernel<true>, kDecThreads);
        static const int occ_s = resident_blocks(decode_u16_c3_kernel<false>, kDecThreads);
        const int grid = grid_for(items, kDecThreads, (vec ? occ_v : occ_s) * LLIE_INGEST_WAVES);
        const int quads = d.crop_w >> 2;
        const long long stride = (long long)grid * kDecThreads;
        const int step_rows = (int)(stride / quads), step_quads = (int)(stride % quads);
        if (vec)
            decode_u16_c3_kernel<true><<<grid, kDecThreads, 0, st>>>(a, step_rows, step_quads);
        else
            decode_u16_c3_kernel<false><<<grid, kDecThreads, 0, st>>>(a, step_rows, step_quads);
    } else {
        const long long items = (long long)d.n * d.c * d.crop_h * d.crop_w;
        decode_u16_generic_kernel<<<grid_for(items, kDecThreads, 8), kDecThreads, 0, st>>>(a);
    }
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

extern "C" int llie_linear_to_srgb(const float *x, int64_t n, float *out_srgb, float *out_lpips, void *stream) {
    LLIE_REQUIRE(x != nullptr && n >= 0, "llie_linear_to_srgb: null input");
    LLIE_REQUIRE(out_srgb != nullptr || out_lpips != nullptr, "llie_linear_to_srgb: no output requested");
    if (n == 0) return LLIE_OK;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const bool vec = aligned16(x) && (out_srgb == nullptr || aligned16(out_srgb)) &&
                     (out_lpips == nullptr || aligned16(out_lpips));
    const long long items = vec ? (n + 3) / 4 : n;
    static const int occ = resident_blocks(linear_to_srgb_kernel, 256);
    linear_to_srgb_kernel<<<grid_for(items, 256, occ * LLIE_INGEST_WAVES), 256, 0, st>>>(x, n, out_srgb, out_lpips, vec);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

// BT.601 luma of [n, 3, plane] float32 images -> [n, 1, plane] (metrics/ssim.py:116-127 `_to_luma_bt601`):
//   y = 0.2989 r + 0.5870 g + 0.1140 b   with every product and sum rounded separately (no FMA contraction), i.e. the
// value ATen's eager `0.2989 * r + 0.5870 * g + 0.1140 * b` produces.  One streaming pass, 16 B per pixel.
namespace llie {
__global__ void __launch_bounds__(256) luma_bt601_kernel(const float *__restrict__ x, long long plane, float *__restrict__ y,
                                                         bool vec) {
    const int n = blockIdx.y;
    const float *r = x + (long long)n * 3 * plane, *g = r + plane, *b = g + plane;
    float *o = y + (long long)n * plane;
    auto f = [](float rv, float gv, float bv) {
        return __fadd_rn(__fadd_rn(__fmul_rn(0.2989f, rv), __fmul_rn(0.5870f, gv)), __fmul_rn(0.1140f, bv));
    };
    const long long stride = (long long)gridDim.x * blockDim.x;
    if (vec) {
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < plane / 4; i += stride) {
            const float4 rv = ldg_stream4(r + 4 * i), gv = ldg_stream4(g + 4 * i), bv = ldg_stream4(b + 4 * i);
            *reinterpret_cast<float4 *>(o + 4 * i) =
                make_float4(f(rv.x, gv.x, bv.x), f(rv.y, gv.y, bv.y), f(rv.z, gv.z, bv.z), f(rv.w, gv.w, bv.w));
        }
    } else {
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < plane; i += stride)
            o[i] = f(__ldg(r + i), __ldg(g + i), __ldg(b + i));
    }
}
}  // namespace llie

extern "C" int llie_luma_bt601(const float *x, int32_t n, int64_t plane, float *y, void *stream) {
    LLIE_REQUIRE(x != nullptr && y != nullptr && n > 0 && n <= 65535 && plane > 0, "llie_luma_bt601: bad arguments");
    const bool vec = llie::aligned16(x) && llie::aligned16(y) && (plane % 4) == 0;
    const long long items = vec ? plane / 4 : plane;
    long long gx = (items + 255) / 256;
    const long long cap = (long long)llie::kNumSM * 8 * 4 / n + 1;
    if (gx > cap) gx = cap;
    dim3 grid((unsigned)gx, (unsigned)n);
    llie::luma_bt601_kernel<<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(x, plane, y, vec);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}
