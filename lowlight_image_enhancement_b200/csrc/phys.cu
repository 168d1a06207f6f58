// Physics-consistency projector, forward and adjoint backward (sm_100a).
//
//   r = clipY(sY * (K (*) padY(pred))) - clipA(sA * obs)      loss map = |r| or sqrt(r^2+eps^2)
//
// Replaces metrics/phys_consistency.py:193-255 (_phys_cons_core) and the physics loss terms
// NewBP_model/losses.py:173-192, :217-220 of the reference, fusing pad + conv + rho + clip + crop +
// residual + robust penalty + per-sample reduction (+ finiteness flags) into ONE pass over pred/obs,
// and the residual derivative + conv_transpose2d adjoint (newbp_layer.py:16-21) into one backward pass.
//
// Two code paths, both parity-tested:
//   * row ring     (phys_ring.cu) depthwise 3x3, W % 4 == 0, 16-byte aligned planes: persistent CTAs stream rows
//                  through a TMA ring in shared memory, HBM-bound.
//   * generic path (this file) any kh x kw, dense cross-channel PSFs, arbitrary W / alignment / rho strides.
#include "phys_common.cuh"
#include <atomic>

namespace llie {

// conv value at output (n, co, oy, ox), any layout / size / pad mode (generic path).  KH, KW > 0: compile-time PSF size
// with the tap loops unrolled and the padded source coordinates formed once per pixel; KH == 0: run-time size.
template <int KH, int KW, bool DENSE>
__device__ __forceinline__ float conv_at(const PhysArgs &a, int n, int co, int oy, int ox) {
    const llie_phys_desc &d = a.d;
    const int kh = KH ? KH : d.kh, kw = KH ? KW : d.kw;
    const int ph = kh >> 1, pw = kw >> 1;
    float acc = 0.f;
    const int ci0 = DENSE ? 0 : co;
    const int ci1 = DENSE ? d.c_in : co + 1;
    if constexpr (KH != 0) {
        int sy[KH], sx[KW];
#pragma unroll
        for (int ky = 0; ky < KH; ++ky) sy[ky] = pad_index(oy + ky - ph, d.h, d.pad_mode);
#pragma unroll
        for (int kx = 0; kx < KW; ++kx) sx[kx] = pad_index(ox + kx - pw, d.w, d.pad_mode);
        for (int ci = ci0; ci < ci1; ++ci) {
            const float *plane = a.pred + ((long long)n * d.c_in + ci) * d.h * d.w;
            const float *wk = a.psf + (DENSE ? ((long long)co * d.c_in + ci) : co) * (KH * KW);
#pragma unroll
            for (int ky = 0; ky < KH; ++ky) {
                if (sy[ky] < 0) continue;
                const float *row = plane + (long long)sy[ky] * d.w;
#pragma unroll
                for (int kx = 0; kx < KW; ++kx)
                    if (sx[kx] >= 0) acc = fmaf(__ldg(wk + ky * KW + kx), __ldg(row + sx[kx]), acc);
            }
        }
        return acc;
    }
    for (int ci = ci0; ci < ci1; ++ci) {
        const float *plane = a.pred + ((long long)n * d.c_in + ci) * d.h * d.w;
        const float *wk = a.psf + (DENSE ? ((long long)co * d.c_in + ci) : co) * kh * kw;
        for (int ky = 0; ky < kh; ++ky) {
            int sy = pad_index(oy + ky - ph, d.h, d.pad_mode);
            if (sy < 0) continue;
            for (int kx = 0; kx < kw; ++kx) {
                int sx = pad_index(ox + kx - pw, d.w, d.pad_mode);
                if (sx < 0) continue;
                acc = fmaf(__ldg(wk + ky * kw + kx), __ldg(plane + (long long)sy * d.w + sx), acc);
            }
        }
    }
    return acc;
}

// Interior pixels (every tap inside the image): no padding logic, taps from registers for depthwise PSFs.  Same tap order
// as conv_at, so a pixel gets the same bits on either path.  `centre` = this pixel in the first input plane the output
// channel reads (plane co for a depthwise PSF, plane 0 for a dense one).
template <int KH, int KW, bool DENSE>
__device__ __forceinline__ float conv_interior(const float *centre, long long plane, int w, int c_in,
                                               const float *wk_dense, const float *wreg) {
    float acc = 0.f;
    const float *p = centre - (KH / 2) * w - (KW / 2);
    const int nci = DENSE ? c_in : 1;
    for (int ci = 0; ci < nci; ++ci, p += plane, wk_dense += KH * KW) {
#pragma unroll
        for (int ky = 0; ky < KH; ++ky)
#pragma unroll
            for (int kx = 0; kx < KW; ++kx)
                acc = fmaf(DENSE ? __ldg(wk_dense + ky * KW + kx) : wreg[ky * KW + kx], __ldg(p + ky * w + kx), acc);
    }
    return acc;
}

// RB consecutive interior rows of one column, depthwise PSF: the (RB + KH - 1) x KW window is loaded first, with no
// branch between the loads (one row at a time leaves ~8 KB per SM in flight and the kernel latency-bound at 1 TB/s),
// then RB tap sums in conv_at's order.  TRANSPOSED: the adjoint stencil (flipped taps), for the gather pass.
template <int KH, int KW, int RB, bool TRANSPOSED>
__device__ __forceinline__ void conv_rows(const float *centre, int w, const float *wreg, float (&out)[RB]) {
    float win[RB + KH - 1][KW];
    const float *p = centre - (KH / 2) * w - (KW / 2);
#pragma unroll
    for (int r = 0; r < RB + KH - 1; ++r)
#pragma unroll
        for (int kx = 0; kx < KW; ++kx) win[r][kx] = __ldg(p + r * w + kx);
#pragma unroll
    for (int i = 0; i < RB; ++i) {
        float acc = 0.f;
#pragma unroll
        for (int ky = 0; ky < KH; ++ky)
#pragma unroll
            for (int kx = 0; kx < KW; ++kx)
                acc = fmaf(wreg[ky * KW + kx], TRANSPOSED ? win[i + KH - 1 - ky][KW - 1 - kx] : win[i + ky][kx], acc);
        out[i] = acc;
    }
}
template <int KH>
struct RowBatch { static constexpr int value = KH == 3 ? 4 : 2; };

// Work decomposition of the generic kernels: one unit = kRows rows x 256 consecutive columns of one plane, one thread
// per column marching down the rows (coalesced, the index divisions once per unit); units are dealt round-robin to the
// blocks of the sample.
constexpr int kStrip = 256, kRows = 8;
struct Units {
    int strips, rblocks, count;
    __host__ __device__ Units(int planes, int h, int w)
        : strips((w + kStrip - 1) / kStrip), rblocks((h + kRows - 1) / kRows) { count = planes * strips * rblocks; }
    __device__ __forceinline__ void at(int u, int &c, int &y0, int &x) const {
        const int strip = u % strips, t = u / strips;
        y0 = (t % rblocks) * kRows;
        c = t / rblocks;
        x = strip * kStrip + threadIdx.x;
    }
};

// rho of one (sample, channel) when it does not vary over the plane; `spatial` tells the caller to use rho_at instead
__device__ __forceinline__ float rho_plane(const PhysArgs &a, int n, int c, bool &spatial) {
    spatial = a.rho != nullptr && (a.d.rho_sh != 0 || a.d.rho_sw != 0);
    return spatial ? 0.f : rho_at(a, n, c, 0, 0);
}

// The pixels whose taps (margin = pad) or whose padding fold (margin = pad + 1, gather pass) can touch the image border,
// enumerated compactly so that they can be given one thread each after the interior sweep: in the sweep they would cost
// every edge warp the full padding logic on every row.  T top rows, B bottom rows, L / R columns of the rows between.
struct Border {
    int h, w, T, B, L, R, count;  // count per plane
    __device__ Border(int h_, int w_, int my, int mx) : h(h_), w(w_) {
        T = min(my, h);
        B = min(my, h - T);
        L = min(mx, w);
        R = min(mx, w - L);
        count = (T + B) * w + (h - T - B) * (L + R);
    }
    __device__ __forceinline__ bool row_inside(int y) const { return y >= T && y < h - B; }
    __device__ __forceinline__ bool col_inside(int x) const { return x >= L && x < w - R; }
    __device__ __forceinline__ void at(int k, int &y, int &x) const {
        if (k < (T + B) * w) {
            y = k / w;
            x = k - y * w;
            if (y >= T) y += h - T - B;
        } else {
            k -= (T + B) * w;
            const int c = k % (L + R);
            y = T + k / (L + R);
            x = c < L ? c : w - R + (c - L);
        }
    }
};

// ------------------------------------------------------------------------------------------------
// Finalize: last block sums the per-block partials in a fixed order and writes the outputs.
// ------------------------------------------------------------------------------------------------
__device__ void phys_finalize(const PhysArgs &a, double *red) {
    const int G = gridDim.x, N = a.d.n;
    double batch = 0.0, total = 0.0;
    int flags = 0;
    for (int n = 0; n < N; ++n) {
        double s = 0.0;
        for (int b = threadIdx.x; b < G; b += blockDim.x) {
            s += __ldcg(a.part_sum + (size_t)n * G + b);
            flags |= __ldcg(a.part_flag + (size_t)n * G + b);
        }
        s = block_sum<double, true>(s, red);
        if (threadIdx.x == 0) {
            float m = (float)(s * a.inv_count);
            a.out_mean[n] = m;
            if (a.out_sum) a.out_sum[n] = s;
            batch += (double)m;
            total += s;
        }
    }
    int f = block_or(flags, reinterpret_cast<int *>(red));
    if (threadIdx.x == 0) {
        a.out_mean[N] = (float)(batch / N);
        a.out_mean[N + 1] = (float)batch;
        if (a.out_sum) a.out_sum[N] = total;
        *a.out_flags = f;
    }
}

// ------------------------------------------------------------------------------------------------
// Generic forward: blockIdx.y = sample; an interior sweep in units of kRows x 256 pixels, then the border pixels.
// ------------------------------------------------------------------------------------------------
// one compared pixel of the generic forward (padding-aware taps)
template <int KH, int KW, bool DENSE>
__device__ __forceinline__ void fwd_pixel(const PhysArgs &a, int n, int co, int y, int x, double &acc, int &flag) {
    const llie_phys_desc &d = a.d;
    const long long plane = (long long)d.h * d.w, rem = (long long)y * d.w + x;
    const float av = __ldg(a.obs + ((long long)n * d.c_out + co) * plane + rem);
    if (!finite_f(av)) flag |= 2;
    if (!(y >= a.y_lo && y < a.y_lo + a.ho && x >= a.x_lo && x < a.x_lo + a.wo)) {
        // cropped border: only the finiteness contract of the reference applies here
        if (co < d.c_in && !finite_f(__ldg(a.pred + ((long long)n * d.c_in + co) * plane + rem))) flag |= 1;
        return;
    }
    const float cv = conv_at<KH, KW, DENSE>(a, n, co, y, x);
    if (!finite_f(cv)) flag |= 1;
    float r;
    acc += (double)residual(d, cv, av, rho_at(a, n, co, y, x), r);
    if (a.out_map) a.out_map[(((long long)n * d.c_out + co) * a.ho + (y - a.y_lo)) * a.wo + (x - a.x_lo)] = fabsf(r);
}

template <int KH, int KW, bool DENSE>
__global__ void __launch_bounds__(kStrip, 4) phys_fwd_generic(const PhysArgs a) {
    __shared__ double red[32];
    const llie_phys_desc &d = a.d;
    const int n = blockIdx.y;
    const long long plane = (long long)d.h * d.w;
    // run-time PSF size: every pixel takes the per-pixel path
    const Border bd(d.h, d.w, KH ? KH / 2 : d.h, KH ? KW / 2 : d.w);
    double acc = 0.0;
    int flag = 0;
    if constexpr (KH != 0) {
        // interior sweep: every tap inside the image, hence (for either crop) inside the compared domain
        const Units un(d.c_out, d.h, d.w);
        const int w = d.w;
        for (int u = blockIdx.x; u < un.count; u += gridDim.x) {
            int co, y0, x;
            un.at(u, co, y0, x);
            if (!bd.col_inside(x)) continue;
            const int y1 = min(y0 + kRows, d.h);
            float wreg[DENSE ? 1 : KH * KW];
            if constexpr (!DENSE)
#pragma unroll
                for (int t = 0; t < KH * KW; ++t) wreg[t] = __ldg(a.psf + co * (KH * KW) + t);
            bool rho_spatial;
            const float rho_c = rho_plane(a, n, co, rho_spatial);
            float part = 0.f;
            // running pointers: this column at row y in obs, in the first input plane of the channel, in the map
            const float *op = a.obs + ((long long)n * d.c_out + co) * plane + (long long)y0 * w + x;
            const float *pp = a.pred + ((long long)n * d.c_in + (DENSE ? 0 : co)) * plane + (long long)y0 * w + x;
            const float *wk_dense = a.psf + (long long)co * d.c_in * (KH * KW);
            long long moff = (((long long)n * d.c_out + co) * a.ho + (y0 - a.y_lo)) * a.wo + (x - a.x_lo);
            constexpr int RB = DENSE ? 1 : RowBatch<KH>::value;
            if (!DENSE && bd.row_inside(y0) && y0 + kRows <= d.h && bd.row_inside(y0 + kRows - 1)) {
                for (int yb = 0; yb < kRows; yb += RB, op += RB * w, pp += RB * w, moff += (long long)RB * a.wo) {
                    float av[RB], cv[RB];
#pragma unroll
                    for (int i = 0; i < RB; ++i) av[i] = __ldg(op + i * w);
                    conv_rows<KH, KW, RB, false>(pp, w, wreg, cv);
#pragma unroll
                    for (int i = 0; i < RB; ++i) {
                        if (!finite_f(av[i])) flag |= 2;
                        if (!finite_f(cv[i])) flag |= 1;
                        float r;
                        part += residual(d, cv[i], av[i], rho_spatial ? rho_at(a, n, co, y0 + yb + i, x) : rho_c, r);
                        if (a.out_map) a.out_map[moff + (long long)i * a.wo] = fabsf(r);
                    }
                }
            } else {
                for (int y = y0; y < y1; ++y, op += w, pp += w, moff += a.wo) {
                    if (!bd.row_inside(y)) continue;
                    const float av = __ldg(op);
                    const float cv = conv_interior<KH, KW, DENSE>(pp, plane, w, d.c_in, wk_dense, wreg);
                    if (!finite_f(av)) flag |= 2;
                    if (!finite_f(cv)) flag |= 1;
                    float r;
                    part += residual(d, cv, av, rho_spatial ? rho_at(a, n, co, y, x) : rho_c, r);
                    if (a.out_map) a.out_map[moff] = fabsf(r);
                }
            }
            acc += (double)part;  // at most kRows addends in fp32, everything above in fp64
        }
    }
    // border pixels, one thread each
    for (int k = blockIdx.x * kStrip + threadIdx.x; k < d.c_out * bd.count; k += gridDim.x * kStrip) {
        const int co = k / bd.count;
        int y, x;
        bd.at(k - co * bd.count, y, x);
        fwd_pixel<KH, KW, DENSE>(a, n, co, y, x, acc, flag);
    }
    acc = block_sum<double>(acc, red);
    const int f = block_or(flag, reinterpret_cast<int *>(red));
    if (threadIdx.x == 0) {
        a.part_sum[(size_t)n * gridDim.x + blockIdx.x] = acc;
        a.part_flag[(size_t)n * gridDim.x + blockIdx.x] = f;
    }
    if (last_block_done(a.ticket, gridDim.x * gridDim.y)) phys_finalize(a, red);
}

// ------------------------------------------------------------------------------------------------
// Backward, generic path: (1) q map into scratch, (2) transposed gather with padding fold.
// ------------------------------------------------------------------------------------------------
template <int KH, int KW, bool DENSE>
__global__ void __launch_bounds__(kStrip, 4) phys_bwd_q_generic(const PhysArgs a) {
    const llie_phys_desc &d = a.d;
    const int n = blockIdx.y;
    const long long plane = (long long)d.h * d.w;
    const Border bd(d.h, d.w, KH ? KH / 2 : d.h, KH ? KW / 2 : d.w);
    const float gn = __ldg(a.gout + a.gout_stride * n) * a.gscale;
    if constexpr (KH != 0) {
        const Units un(d.c_out, d.h, d.w);
        const int w = d.w;
        for (int u = blockIdx.x; u < un.count; u += gridDim.x) {
            int co, y0, x;
            un.at(u, co, y0, x);
            if (!bd.col_inside(x)) continue;
            const int y1 = min(y0 + kRows, d.h);
            float wreg[DENSE ? 1 : KH * KW];
            if constexpr (!DENSE)
#pragma unroll
                for (int t = 0; t < KH * KW; ++t) wreg[t] = __ldg(a.psf + co * (KH * KW) + t);
            bool rho_spatial;
            const float rho_c = rho_plane(a, n, co, rho_spatial);
            const long long base = ((long long)n * d.c_out + co) * plane + (long long)y0 * w + x;
            const float *op = a.obs + base;
            float *qp = a.scratch + base;
            const float *pp = a.pred + ((long long)n * d.c_in + (DENSE ? 0 : co)) * plane + (long long)y0 * w + x;
            const float *wk_dense = a.psf + (long long)co * d.c_in * (KH * KW);
            constexpr int RB = DENSE ? 1 : RowBatch<KH>::value;
            if (!DENSE && bd.row_inside(y0) && y0 + kRows <= d.h && bd.row_inside(y0 + kRows - 1)) {
                for (int yb = 0; yb < kRows; yb += RB, op += RB * w, pp += RB * w, qp += RB * w) {
                    float av[RB], cv[RB];
#pragma unroll
                    for (int i = 0; i < RB; ++i) av[i] = __ldg(op + i * w);
                    conv_rows<KH, KW, RB, false>(pp, w, wreg, cv);
#pragma unroll
                    for (int i = 0; i < RB; ++i)
                        qp[i * w] = gn * dresidual(d, cv[i], av[i],
                                                   rho_spatial ? rho_at(a, n, co, y0 + yb + i, x) : rho_c);
                }
            } else {
                for (int y = y0; y < y1; ++y, op += w, pp += w, qp += w) {
                    if (!bd.row_inside(y)) continue;
                    const float cv = conv_interior<KH, KW, DENSE>(pp, plane, w, d.c_in, wk_dense, wreg);
                    *qp = gn * dresidual(d, cv, __ldg(op), rho_spatial ? rho_at(a, n, co, y, x) : rho_c);
                }
            }
        }
    }
    for (int k = blockIdx.x * kStrip + threadIdx.x; k < d.c_out * bd.count; k += gridDim.x * kStrip) {
        const int co = k / bd.count;
        int y, x;
        bd.at(k - co * bd.count, y, x);
        const long long rem = (long long)y * d.w + x;
        float q = 0.f;
        if (y >= a.y_lo && y < a.y_lo + a.ho && x >= a.x_lo && x < a.x_lo + a.wo) {
            const float cv = conv_at<KH, KW, DENSE>(a, n, co, y, x);
            const float av = __ldg(a.obs + ((long long)n * d.c_out + co) * plane + rem);
            q = gn * dresidual(d, cv, av, rho_at(a, n, co, y, x));
        }
        a.scratch[((long long)n * d.c_out + co) * plane + rem] = q;
    }
}

template <int KH, int KW, bool DENSE>
__global__ void __launch_bounds__(kStrip, 4) phys_bwd_gather_generic(const PhysArgs a) {
    const llie_phys_desc &d = a.d;
    const int n = blockIdx.y;
    const int kh = KH ? KH : d.kh, kw = KH ? KW : d.kw;
    const int ph = kh >> 1, pw = kw >> 1;
    const long long plane = (long long)d.h * d.w;
    // strictly inside the band that padding can alias onto or that the transposed taps can leave
    const Border bd(d.h, d.w, KH ? KH / 2 + 1 : d.h, KH ? KW / 2 + 1 : d.w);
    if constexpr (KH != 0) {
        const Units un(d.c_in, d.h, d.w);
        const int w = d.w;
        for (int u = blockIdx.x; u < un.count; u += gridDim.x) {
            int ci, y0, x;
            un.at(u, ci, y0, x);
            if (!bd.col_inside(x)) continue;
            const int y1 = min(y0 + kRows, d.h);
            float wreg[DENSE ? 1 : KH * KW];
            if constexpr (!DENSE)
#pragma unroll
                for (int t = 0; t < KH * KW; ++t) wreg[t] = __ldg(a.psf + ci * (KH * KW) + t);
            // running pointers: this column at row y in the gradient plane and in the first q plane the channel receives from
            float *gp = a.grad + ((long long)n * d.c_in + ci) * plane + (long long)y0 * w + x;
            const float *qc = a.scratch + ((long long)n * d.c_out + (DENSE ? 0 : ci)) * plane + (long long)y0 * w + x;
            constexpr int RB = DENSE ? 1 : RowBatch<KH>::value;
            if (!DENSE && bd.row_inside(y0) && y0 + kRows <= d.h && bd.row_inside(y0 + kRows - 1)) {
                for (int yb = 0; yb < kRows; yb += RB, gp += RB * w, qc += RB * w) {
                    float g[RB];
                    conv_rows<KH, KW, RB, true>(qc, w, wreg, g);
#pragma unroll
                    for (int i = 0; i < RB; ++i) gp[i * w] = g[i];
                }
            } else {
                for (int y = y0; y < y1; ++y, gp += w, qc += w) {
                    if (!bd.row_inside(y)) continue;
                    float acc = 0.f;
                    const float *qp = qc + (KH / 2) * w + (KW / 2);
                    const float *wk = a.psf + (long long)ci * (KH * KW);  // DENSE only: + co * c_in * KH * KW
                    const int nco = DENSE ? d.c_out : 1;
                    for (int co = 0; co < nco; ++co, qp += plane, wk += (long long)d.c_in * (KH * KW)) {
#pragma unroll
                        for (int ky = 0; ky < KH; ++ky)
#pragma unroll
                            for (int kx = 0; kx < KW; ++kx)
                                acc = fmaf(DENSE ? __ldg(wk + ky * KW + kx) : wreg[ky * KW + kx],
                                           __ldg(qp - ky * w - kx), acc);
                    }
                    *gp = acc;
                }
            }
        }
    }
    for (int k = blockIdx.x * kStrip + threadIdx.x; k < d.c_in * bd.count; k += gridDim.x * kStrip) {
        const int ci = k / bd.count;
        int y, x;
        bd.at(k - ci * bd.count, y, x);
        float acc = 0.f;
        const int co0 = DENSE ? 0 : ci;
        const int co1 = DENSE ? d.c_out : ci + 1;
        // candidate padded coordinates that may alias onto (y, x): the pixel itself plus the
        // out-of-image bands [-p, 0) and [size, size+p)
        for (int ty = 0; ty <= 2 * ph; ++ty) {
            const int jy = (ty == 0) ? y : (ty <= ph ? -ty : d.h - 1 + (ty - ph));
            if (ty != 0 && pad_index(jy, d.h, d.pad_mode) != y) continue;
            for (int tx = 0; tx <= 2 * pw; ++tx) {
                const int jx = (tx == 0) ? x : (tx <= pw ? -tx : d.w - 1 + (tx - pw));
                if (tx != 0 && pad_index(jx, d.w, d.pad_mode) != x) continue;
                for (int co = co0; co < co1; ++co) {
                    const float *wk = a.psf + (DENSE ? ((long long)co * d.c_in + ci) : co) * kh * kw;
                    const float *qp = a.scratch + ((long long)n * d.c_out + co) * plane;
                    for (int ky = 0; ky < kh; ++ky) {
                        const int oy = jy - ky + ph;
                        if (oy < 0 || oy >= d.h) continue;
                        for (int kx = 0; kx < kw; ++kx) {
                            const int ox = jx - kx + pw;
                            if (ox < 0 || ox >= d.w) continue;
                            acc = fmaf(__ldg(wk + ky * kw + kx), __ldg(qp + (long long)oy * d.w + ox), acc);
                        }
                    }
                }
            }
        }
        a.grad[((long long)n * d.c_in + ci) * plane + (long long)y * d.w + x] = acc;
    }
}

// PSF sizes with their own instantiation; everything else runs the run-time-size body
#define PHYS_GENERIC_DISPATCH(KERNEL, GRID, STREAM, ARGS)                                              \
    do {                                                                                               \
        const bool dense_ = (ARGS).d.psf_layout == LLIE_PSF_DENSE;                                     \
        const int kh_ = (ARGS).d.kh, kw_ = (ARGS).d.kw;                                                \
        if (kh_ == 3 && kw_ == 3 && !dense_) KERNEL<3, 3, false><<<GRID, kStrip, 0, STREAM>>>(ARGS);   \
        else if (kh_ == 3 && kw_ == 3) KERNEL<3, 3, true><<<GRID, kStrip, 0, STREAM>>>(ARGS);          \
        else if (kh_ == 5 && kw_ == 5 && !dense_) KERNEL<5, 5, false><<<GRID, kStrip, 0, STREAM>>>(ARGS); \
        else if (dense_) KERNEL<0, 0, true><<<GRID, kStrip, 0, STREAM>>>(ARGS);                        \
        else KERNEL<0, 0, false><<<GRID, kStrip, 0, STREAM>>>(ARGS);                                   \
    } while (0)

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static int validate(const llie_phys_desc *d) {
    LLIE_REQUIRE(d != nullptr, "phys: null descriptor");
    LLIE_REQUIRE(d->n > 0 && d->c_in > 0 && d->c_out > 0 && d->h > 0 && d->w > 0, "phys: empty tensor");
    LLIE_REQUIRE(d->kh >= 1 && d->kw >= 1 && (d->kh & 1) && (d->kw & 1), "phys: PSF must be odd-sized");
    LLIE_REQUIRE(d->psf_layout == LLIE_PSF_DENSE || d->c_in == d->c_out, "phys: depthwise PSF needs c_in == c_out");
    LLIE_REQUIRE(d->pad_mode >= LLIE_PAD_ZEROS && d->pad_mode <= LLIE_PAD_REPLICATE, "phys: bad pad mode");
    LLIE_REQUIRE(d->pad_mode != LLIE_PAD_REFLECT || (d->kh / 2 < d->h && d->kw / 2 < d->w),
                 "phys: reflect padding needs pad < size");
    LLIE_REQUIRE(d->n <= 65535, "phys: batch too large for grid.y");
    if (d->crop_valid) LLIE_REQUIRE(d->h > 2 * (d->kh / 2) && d->w > 2 * (d->kw / 2), "phys: valid crop leaves nothing");
    LLIE_REQUIRE((long long)(d->c_in > d->c_out ? d->c_in : d->c_out) * d->h * d->w < (1ll << 31),
                 "phys: more than 2^31 elements per sample");
    return LLIE_OK;
}

static void fill_domain(PhysArgs &a) {
    const llie_phys_desc &d = a.d;
    a.y_lo = d.crop_valid ? d.kh / 2 : 0;
    a.x_lo = d.crop_valid ? d.kw / 2 : 0;
    a.ho = d.h - 2 * a.y_lo;
    a.wo = d.w - 2 * a.x_lo;
    a.inv_count = 1.0 / ((double)d.c_out * a.ho * a.wo);
}

static int grid_x_for(int n, int work_items) {
    // persistent-style grid: one wave of the 4 resident CTAs of 256 threads per SM, split evenly over samples
    int per_sample = div_up(kNumSM * 4, n);
    if (per_sample > work_items) per_sample = work_items;
    if (per_sample < 1) per_sample = 1;
    return per_sample;
}

void ssim_tune(int key, int value);  // ssim.cu

static std::atomic<int> g_phys_path{0};  // 0 TMA row ring (default) | 2 generic kernels only (llie_tune, diagnostic)
static inline bool ring_enabled() { return g_phys_path.load(std::memory_order_relaxed) == 0; }

}  // namespace llie

using namespace llie;

extern "C" int llie_tune(int key, int value) {
    switch (key) {
        case LLIE_TUNE_PHYS_PATH:
            if (value != 0 && value != 2) {
                llie::set_error("llie_tune: LLIE_TUNE_PHYS_PATH takes 0 (row ring) or 2 (generic kernels), got %d", value);
                return LLIE_EINVAL;
            }
            g_phys_path.store(value, std::memory_order_relaxed);
            return LLIE_OK;
        case LLIE_TUNE_PHYS_ROWS: phys_ring_set_rows(value); return LLIE_OK;
        case LLIE_TUNE_PHYS_PDL: phys_ring_set_pdl(value); return LLIE_OK;
        case LLIE_TUNE_PHYS_PREFETCH: phys_ring_set_prefetch(value); return LLIE_OK;
        case 6:
            llie::set_error("llie_tune: key 6 (process-wide stream overlap) was removed in ABI 3; set "
                            "llie_phys_desc.stream_overlap on the calls that meet the contract");
            return LLIE_EINVAL;
        case LLIE_TUNE_SSIM_PATH:
        case LLIE_TUNE_SSIM_SEG_ROWS: ssim_tune(key, value); return LLIE_OK;
        default: break;
    }
    llie::set_error("llie_tune: unknown key %d", key);
    return LLIE_EINVAL;
}

extern "C" int llie_phys_forward(const llie_phys_desc *desc, const float *pred, const float *obs, const float *psf,
                                 const float *rho, float *out_mean, double *out_sum, float *out_map,
                                 int32_t *out_flags, void *workspace, void *stream) {
    if (int rc = validate(desc)) return rc;
    LLIE_REQUIRE(pred && obs && psf && out_mean && out_flags && workspace, "phys_forward: null pointer");
    PhysArgs a{};
    a.d = *desc;
    a.pred = pred; a.obs = obs; a.psf = psf; a.rho = rho;
    a.out_mean = out_mean; a.out_sum = out_sum; a.out_map = out_map; a.out_flags = out_flags;
    fill_domain(a);
    Workspace ws(workspace);
    a.ticket = ws.tickets;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (ring_enabled() && phys_ring_eligible(a)) return launch_phys_fwd_ring(a, workspace, st);
    const int gx = grid_x_for(desc->n, Units(desc->c_out, desc->h, desc->w).count);
    const size_t need = (size_t)desc->n * gx * (sizeof(double) + sizeof(int));
    LLIE_REQUIRE(need <= ws.scratch_bytes, "phys_forward: workspace too small");
    a.part_sum = reinterpret_cast<double *>(ws.scratch);
    a.part_flag = reinterpret_cast<int *>(ws.scratch + (size_t)desc->n * gx * sizeof(double));
    dim3 grid(gx, desc->n);
    PHYS_GENERIC_DISPATCH(phys_fwd_generic, grid, st, a);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

extern "C" int llie_phys_backward_needs_scratch(const llie_phys_desc *desc, const float *pred, const float *obs,
                                               const float *rho, const float *grad_pred) {
    if (validate(desc)) return 1;
    PhysArgs a{};
    a.d = *desc;
    a.pred = pred; a.obs = obs; a.rho = rho;
    return (ring_enabled() && phys_ring_eligible(a) && aligned16(grad_pred)) ? 0 : 1;
}

extern "C" int llie_phys_backward(const llie_phys_desc *desc, const float *pred, const float *obs, const float *psf,
                                  const float *rho, const float *gout, int64_t gout_stride, float gscale,
                                  float *grad_pred, float *scratch, void *stream) {
    if (int rc = validate(desc)) return rc;
    LLIE_REQUIRE(pred && obs && psf && gout && grad_pred, "phys_backward: null pointer");
    PhysArgs a{};
    a.d = *desc;
    a.pred = pred; a.obs = obs; a.psf = psf; a.rho = rho;
    a.gout = gout; a.gout_stride = gout_stride; a.gscale = gscale;
    a.grad = grad_pred; a.scratch = scratch;
    fill_domain(a);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (ring_enabled() && phys_ring_eligible(a) && aligned16(grad_pred)) {
        return launch_phys_bwd_ring(a, st);
    } else {
        LLIE_REQUIRE(scratch != nullptr, "phys_backward: this PSF/shape needs the scratch buffer");
        LLIE_REQUIRE(desc->kh / 2 <= 16 && desc->kw / 2 <= 16, "phys_backward: PSF larger than 33x33 unsupported");
        dim3 grid(grid_x_for(desc->n, Units(desc->c_out, desc->h, desc->w).count), desc->n);
        PHYS_GENERIC_DISPATCH(phys_bwd_q_generic, grid, st, a);
        dim3 grid2(grid_x_for(desc->n, Units(desc->c_in, desc->h, desc->w).count), desc->n);
        PHYS_GENERIC_DISPATCH(phys_bwd_gather_generic, grid2, st, a);
    }
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

extern "C" int llie_phys_value_and_grad(const llie_phys_desc *desc, const float *pred, const float *obs,
                                        const float *psf, const float *rho, const float *gout, int64_t gout_stride,
                                        float gscale, float *out_mean, double *out_sum, int32_t *out_flags,
                                        float *grad_pred, float *scratch, void *workspace, void *stream) {
    if (int rc = validate(desc)) return rc;
    LLIE_REQUIRE(pred && obs && psf && gout && grad_pred && out_mean && out_flags && workspace,
                 "phys_value_and_grad: null pointer");
    PhysArgs a{};
    a.d = *desc;
    a.pred = pred; a.obs = obs; a.psf = psf; a.rho = rho;
    a.gout = gout; a.gout_stride = gout_stride; a.gscale = gscale;
    a.grad = grad_pred; a.scratch = scratch;
    a.out_mean = out_mean; a.out_sum = out_sum; a.out_map = nullptr; a.out_flags = out_flags;
    fill_domain(a);
    if (ring_enabled() && phys_ring_eligible(a) && aligned16(grad_pred))
        return launch_phys_fused_ring(a, workspace, reinterpret_cast<cudaStream_t>(stream));
    // shapes the fused kernel does not take: the two single-purpose passes, same results
    if (int rc = llie_phys_forward(desc, pred, obs, psf, rho, out_mean, out_sum, nullptr, out_flags, workspace, stream))
        return rc;
    return llie_phys_backward(desc, pred, obs, psf, rho, gout, gout_stride, gscale, grad_pred, scratch, stream);
}
