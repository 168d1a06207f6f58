// NewBP operator: y = conv2d(x, K, zero padding, groups in {1, C}) and its explicit adjoint
// gx = conv_transpose2d(gy, K, padding, groups).  Replaces NewBP_model/newbp_layer.py:9-21
// (NewBPFunction) and :117-126 (CrosstalkPSF.forward).  float32 and float64 (the reference's
// gradient tests run CrosstalkPSF in double, core_tests/test_physics_loss_grad.py:65-135).
//
// One thread per output element, weights staged in shared memory; the fused loss kernels in
// phys.cu are the bandwidth-critical users of the PSF, this entry point serves the stand-alone
// operator (any kh x kw, any zero padding, dense or depthwise).
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
    const size_t esz = d->dtype ? 8 : 4;
    const size_t smem = (size_t)(d->depthwise ? d->c_out : d->c_out * d->c_in) * d->kh * d->kw * esz;
    LLIE_REQUIRE(smem <= 48 * 1024, "newbp: kernel bank larger than 48 KiB");
    const long long items = adjoint ? (long long)d->c_in * d->h * d->w : (long long)d->c_out * a.ho * a.wo;
    int gx = (int)((items + 255) / 256);
    const int cap = div_up(kNumSM * 8, d->n);
    if (gx > cap) gx = cap;
    dim3 grid(gx, d->n);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
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
