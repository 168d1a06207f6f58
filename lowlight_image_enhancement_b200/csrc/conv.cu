// NewBP operator: y = conv2d(x, K, zero padding, groups in {1, C}) and its explicit adjoint
// gx = conv_transpose2d(gy, K, padding, groups).  Replaces NewBP_model/newbp_layer.py:9-21
// (NewBPFunction) and :117-126 (CrosstalkPSF.forward).  float32 and float64 (the reference's
// gradient tests run CrosstalkPSF in double, core_tests/test_physics_loss_grad.py:65-135).
//
// Generic kernel: one thread per output element, weights staged in shared memory (any kh x kw, any zero padding, dense
// or depthwise, float32 / float64).  The configuration the reference runs -- depthwise 3x3, padding 1, float32 -- takes
// the four-column fast path below.
#include "common.cuh"

namespace llie {

struct ConvArgs {
    llie_conv_desc d;
    const void *src, *ker;
    void *dst;
    int ho, wo;  // forward output size
};

template <typename T, bool ADJOINT>
__global__ void __launch_bounds__(256) newbp_conv_kernel(const ConvArgs a) {
    extern __shared__ unsigned char smraw[];
    T *sw = reinterpret_cast<T *>(smraw);
    const llie_conv_desc &d = a.d;
    const int kk = d.kh * d.kw;
    const int nw = (d.depthwise ? d.c_out : d.c_out * d.c_in) * kk;
    for (int i = threadIdx.x; i < nw; i += blockDim.x) sw[i] = static_cast<const T *>(a.ker)[i];
    __syncthreads();
    const T *src = static_cast<const T *>(a.src);
    T *dst = static_cast<T *>(a.dst);
    const int n = blockIdx.y;
    if (!ADJOINT) {
        const long long oplane = (long long)a.ho * a.wo, iplane = (long long)d.h * d.w;
        const long long total = oplane * d.c_out;
        for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
             idx += (long long)gridDim.x * blockDim.x) {
            const int co = (int)(idx / oplane);
            const int rem = (int)(idx - co * oplane);
            const int oy = rem / a.wo, ox = rem - oy * a.wo;
            T acc = T(0);
            const int ci0 = d.depthwise ? co : 0, ci1 = d.depthwise ? co + 1 : d.c_in;
            for (int ci = ci0; ci < ci1; ++ci) {
                const T *pl = src + ((long long)n * d.c_in + ci) * iplane;
                const T *wk = sw + (d.depthwise ? co : co * d.c_in + ci) * kk;
                for (int ky = 0; ky < d.kh; ++ky) {
                    const int sy = oy + ky - d.pad_h;
                    if (sy < 0 || sy >= d.h) continue;
                    for (int kx = 0; kx < d.kw; ++kx) {
                        const int sx = ox + kx - d.pad_w;
                        if (sx < 0 || sx >= d.w) continue;
                        acc = fma(wk[ky * d.kw + kx], pl[(long long)sy * d.w + sx], acc);
                    }
                }
            }
            dst[((long long)n * d.c_out + co) * oplane + rem] = acc;
        }
    } else {
        const long long oplane = (long long)a.ho * a.wo, iplane = (long long)d.h * d.w;
        const long long total = iplane * d.c_in;
        for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
             idx += (long long)gridDim.x * blockDim.x) {
            const int ci = (int)(idx / iplane);
            const int rem = (int)(idx - ci * iplane);
            const int y = rem / d.w, x = rem - y * d.w;
            T acc = T(0);
            const int co0 = d.depthwise ? ci : 0, co1 = d.depthwise ? ci + 1 : d.c_out;
            for (int co = co0; co < co1; ++co) {
                const T *pl = src + ((long long)n * d.c_out + co) * oplane;
                const T *wk = sw + (d.depthwise ? co : co * d.c_in + ci) * kk;
                for (int ky = 0; ky < d.kh; ++ky) {
                    const int oy = y - ky + d.pad_h;
                    if (oy < 0 || oy >= a.ho) continue;
                    for (int kx = 0; kx < d.kw; ++kx) {
                        const int ox = x - kx + d.pad_w;
                        if (ox < 0 || ox >= a.wo) continue;
                        acc = fma(wk[ky * d.kw + kx], pl[(long long)oy * a.wo + ox], acc);
                    }
                }
            }
            dst[((long long)n * d.c_in + ci) * iplane + rem] = acc;
        }
    }
}

// Fast path for what the reference actually runs (CrosstalkPSF: depthwise 3x3, padding 1, float32; newbp_layer.py:117-126):
// four adjacent outputs per thread, three 128-bit row loads rolling down a 12-row block, the outside neighbours by warp
// shuffle (warp-edge lanes load them), one 128-bit store per row -- 8 B per element, HBM-bound.  Taps are accumulated in
// the generic kernel's order (ky major, kx minor, from 0; an out-of-image tap multiplies a zero instead of being
// skipped), so the results are bit-identical to it.  The adjoint of a same-size zero-padded depthwise correlation is the
// correlation with the flipped kernel: the same code with the taps walked backwards.
constexpr int kConvRows = 12;

template <bool ADJOINT>
__global__ void __launch_bounds__(64) newbp_dw3_kernel(const float *__restrict__ src, const float *__restrict__ ker,
                                                       float *__restrict__ dst, int C, int H, int W) {
    const int x = (blockIdx.x * 64 + threadIdx.x) * 4;
    const int y0 = blockIdx.y * kConvRows, y1 = min(y0 + kConvRows, H);
    const long long plane = (long long)H * W;
    const float *p = src + (long long)blockIdx.z * plane;
    float *o = dst + (long long)blockIdx.z * plane;
    const int lane = threadIdx.x & 31;
    const bool in = x < W;
    float w[9];
#pragma unroll
    for (int q = 0; q < 9; ++q) w[q] = __ldg(ker + (blockIdx.z % C) * 9 + q);
    auto row6 = [&](int yy, float (&v)[6]) {  // v = src[yy][x-1 .. x+4], zeros outside the image
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
    // one output from the rows above / at / below it; i = column offset of the output inside the six-wide window
    auto tap9 = [&](const float (&up)[6], const float (&mid)[6], const float (&dn)[6], int i) {
        float acc = 0.f;
        if (!ADJOINT) {  // ky = 0, 1, 2 reads rows y-1, y, y+1; kx = 0, 1, 2 reads columns x-1, x, x+1
            acc = fmaf(w[0], up[i - 1], acc); acc = fmaf(w[1], up[i], acc); acc = fmaf(w[2], up[i + 1], acc);
            acc = fmaf(w[3], mid[i - 1], acc); acc = fmaf(w[4], mid[i], acc); acc = fmaf(w[5], mid[i + 1], acc);
            acc = fmaf(w[6], dn[i - 1], acc); acc = fmaf(w[7], dn[i], acc); acc = fmaf(w[8], dn[i + 1], acc);
        } else {         // ky = 0, 1, 2 reads rows y+1, y, y-1; kx = 0, 1, 2 reads columns x+1, x, x-1
            acc = fmaf(w[0], dn[i + 1], acc); acc = fmaf(w[1], dn[i], acc); acc = fmaf(w[2], dn[i - 1], acc);
            acc = fmaf(w[3], mid[i + 1], acc); acc = fmaf(w[4], mid[i], acc); acc = fmaf(w[5], mid[i - 1], acc);
            acc = fmaf(w[6], up[i + 1], acc); acc = fmaf(w[7], up[i], acc); acc = fmaf(w[8], up[i - 1], acc);
        }
        return acc;
    };
    float a[6], b[6];
    row6(y0 - 1, a);
    row6(y0, b);
    for (int y = y0; y < y1; y += 2) {
        float n0[6], n1[6];
        row6(y + 1, n0);
        row6(y + 2, n1);
        if (in) {
            *reinterpret_cast<float4 *>(o + (long long)y * W + x) =
                make_float4(tap9(a, b, n0, 1), tap9(a, b, n0, 2), tap9(a, b, n0, 3), tap9(a, b, n0, 4));
            if (y + 1 < y1)
                *reinterpret_cast<float4 *>(o + (long long)(y + 1) * W + x) =
                    make_float4(tap9(b, n0, n1, 1), tap9(b, n0, n1, 2), tap9(b, n0, n1, 3), tap9(b, n0, n1, 4));
        }
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            a[k] = n0[k];
            b[k] = n1[k];
        }
    }
}

static int launch_conv(const llie_conv_desc *d, const void *src, const void *ker, void *dst, void *stream,
                       bool adjoint) {
    LLIE_REQUIRE(d && src && ker && dst, "newbp: null pointer");
    LLIE_REQUIRE(d->n > 0 && d->n <= 65535 && d->c_in > 0 && d->c_out > 0 && d->h > 0 && d->w > 0, "newbp: bad shape");
    LLIE_REQUIRE(d->kh >= 1 && d->kw >= 1 && d->pad_h >= 0 && d->pad_w >= 0, "newbp: bad kernel/padding");
    LLIE_REQUIRE(!d->depthwise || d->c_in == d->c_out, "newbp: depthwise needs c_in == c_out");
    LLIE_REQUIRE(d->dtype == 0 || d->dtype == 1, "newbp: dtype must be 0 (f32) or 1 (f64)");
    ConvArgs a{};
    a.d = *d;
    a.src = src; a.ker = ker; a.dst = dst;
    a.ho = d->h + 2 * d->pad_h - d->kh + 1;
    a.wo = d->w + 2 * d->pad_w - d->kw + 1;
    LLIE_REQUIRE(a.ho > 0 && a.wo > 0, "newbp: empty output");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (d->dtype == 0 && d->depthwise && d->kh == 3 && d->kw == 3 && d->pad_h == 1 && d->pad_w == 1 && (d->w & 3) == 0 &&
        aligned16(src) && aligned16(dst) && (long long)d->n * d->c_out <= 65535 && div_up(d->h, kConvRows) <= 65535) {
        dim3 grid(div_up(d->w, 256), div_up(d->h, kConvRows), (unsigned)(d->n * d->c_out));
        if (adjoint)
            newbp_dw3_kernel<true><<<grid, 64, 0, st>>>(static_cast<const float *>(src), static_cast<const float *>(ker),
                                                        static_cast<float *>(dst), d->c_out, d->h, d->w);
        else
            newbp_dw3_kernel<false><<<grid, 64, 0, st>>>(static_cast<const float *>(src), static_cast<const float *>(ker),
                                                         static_cast<float *>(dst), d->c_out, d->h, d->w);
        LLIE_CUDA_OK(cudaGetLastError());
        return LLIE_OK;
    }
    const size_t esz = d->dtype ? 8 : 4;
    const size_t smem = (size_t)(d->depthwise ? d->c_out : d->c_out * d->c_in) * d->kh * d->kw * esz;
    LLIE_REQUIRE(smem <= 48 * 1024, "newbp: kernel bank larger than 48 KiB");
    const long long items = adjoint ? (long long)d->c_in * d->h * d->w : (long long)d->c_out * a.ho * a.wo;
    int gx = (int)((items + 255) / 256);
    const int cap = div_up(kNumSM * 8, d->n);
    if (gx > cap) gx = cap;
    dim3 grid(gx, d->n);
    if (d->dtype == 0) {
        if (adjoint) newbp_conv_kernel<float, true><<<grid, 256, smem, st>>>(a);
        else newbp_conv_kernel<float, false><<<grid, 256, smem, st>>>(a);
    } else {
        if (adjoint) newbp_conv_kernel<double, true><<<grid, 256, smem, st>>>(a);
        else newbp_conv_kernel<double, false><<<grid, 256, smem, st>>>(a);
    }
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

}  // namespace llie

extern "C" int llie_newbp_forward(const llie_conv_desc *desc, const void *x, const void *kernel, void *y,
                                  void *stream) {
    return llie::launch_conv(desc, x, kernel, y, stream, false);
}
extern "C" int llie_newbp_adjoint(const llie_conv_desc *desc, const void *gy, const void *kernel, void *gx,
                                  void *stream) {
    return llie::launch_conv(desc, gy, kernel, gx, stream, true);
}
