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
//   * tile path    depthwise 3x3, W % 4 == 0, 16-byte aligned planes: shared-memory halo tiles
//                  (128-float rows, 128-bit LDS, rolling 3-row register window), HBM-bound.
//   * generic path any kh x kw, dense cross-channel PSFs, arbitrary W / alignment / rho strides.
#include "common.cuh"

namespace llie {

struct PhysArgs {
    llie_phys_desc d;
    const float *pred, *obs, *psf, *rho;
    float *out_mean;
    double *out_sum;
    float *out_map;
    int *out_flags;
    double *part_sum;  // [n][gridDim.x]
    int *part_flag;    // [n][gridDim.x]
    unsigned *ticket;
    int y_lo, x_lo, ho, wo;  // compared (output) domain
    int tiles_x, tiles_y;
    double inv_count;        // 1 / (c_out*ho*wo)
    // backward only
    const float *gout;
    long long gout_stride;
    float gscale;
    float *grad;
    float *scratch;
};

// residual and loss for one output element; returns loss value, sets r
__device__ __forceinline__ float residual(const llie_phys_desc &d, float conv, float a, float rho, float &r) {
    float y = d.rho_side == 0 ? conv * rho : conv;
    float t = d.rho_side == 1 ? a * rho : a;
    if (d.clip_y) y = fminf(fmaxf(y, 0.f), 1.f);
    if (d.clip_a) t = fminf(fmaxf(t, 0.f), 1.f);
    r = y - t;
    return d.robust ? sqrtf(r * r + d.eps * d.eps) : fabsf(r);
}

// d loss / d conv for one output element (without the upstream weight)
__device__ __forceinline__ float dresidual(const llie_phys_desc &d, float conv, float a, float rho) {
    float y = d.rho_side == 0 ? conv * rho : conv;
    float t = d.rho_side == 1 ? a * rho : a;
    float pass = 1.f;
    if (d.clip_y) {
        // torch.clamp backward passes the gradient where lo <= y <= hi
        pass = (y >= 0.f && y <= 1.f) ? 1.f : 0.f;
        y = fminf(fmaxf(y, 0.f), 1.f);
    }
    if (d.clip_a) t = fminf(fmaxf(t, 0.f), 1.f);
    float r = y - t;
    float psi;
    if (d.robust)
        psi = r / sqrtf(r * r + d.eps * d.eps);
    else
        psi = (r > 0.f) ? 1.f : ((r < 0.f) ? -1.f : 0.f);
    return psi * pass * (d.rho_side == 0 ? rho : 1.f);
}

__device__ __forceinline__ float rho_at(const PhysArgs &a, int n, int c, int y, int x) {
    if (a.rho == nullptr) return a.d.rho_scalar;
    return __ldg(a.rho + (long long)n * a.d.rho_sn + (long long)c * a.d.rho_sc + (long long)y * a.d.rho_sh +
                 (long long)x * a.d.rho_sw);
}

// conv value at output (n, co, oy, ox), any layout / size / pad mode (generic path)
__device__ __forceinline__ float conv_at(const PhysArgs &a, int n, int co, int oy, int ox) {
    const llie_phys_desc &d = a.d;
    const int ph = d.kh >> 1, pw = d.kw >> 1;
    float acc = 0.f;
    const int ci0 = d.psf_layout == LLIE_PSF_DENSE ? 0 : co;
    const int ci1 = d.psf_layout == LLIE_PSF_DENSE ? d.c_in : co + 1;
    for (int ci = ci0; ci < ci1; ++ci) {
        const float *plane = a.pred + ((long long)n * d.c_in + ci) * d.h * d.w;
        const float *wk = a.psf + (d.psf_layout == LLIE_PSF_DENSE ? ((long long)co * d.c_in + ci) : co) * d.kh * d.kw;
        for (int ky = 0; ky < d.kh; ++ky) {
            int sy = pad_index(oy + ky - ph, d.h, d.pad_mode);
            if (sy < 0) continue;
            for (int kx = 0; kx < d.kw; ++kx) {
                int sx = pad_index(ox + kx - pw, d.w, d.pad_mode);
                if (sx < 0) continue;
                acc = fmaf(__ldg(wk + ky * d.kw + kx), __ldg(plane + (long long)sy * d.w + sx), acc);
            }
        }
    }
    return acc;
}

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
// Generic forward: one thread per element of the full [c_out, H, W] grid of sample blockIdx.y.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) phys_fwd_generic(const PhysArgs a) {
    __shared__ double red[32];
    const llie_phys_desc &d = a.d;
    const int n = blockIdx.y;
    const long long plane = (long long)d.h * d.w;
    const long long total = plane * d.c_out;
    double acc = 0.0;
    int flag = 0;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        const int co = (int)(idx / plane);
        const int rem = (int)(idx - (long long)co * plane);
        const int y = rem / d.w, x = rem - y * d.w;
        const float av = __ldg(a.obs + ((long long)n * d.c_out + co) * plane + rem);
        if (!finite_f(av)) flag |= 2;
        const bool in_dom = (y >= a.y_lo) && (y < a.y_lo + a.ho) && (x >= a.x_lo) && (x < a.x_lo + a.wo);
        if (!in_dom) {
            // cropped border: only the finiteness contract of the reference applies here
            if (co < d.c_in && !finite_f(__ldg(a.pred + ((long long)n * d.c_in + co) * plane + rem))) flag |= 1;
            continue;
        }
        const float cv = conv_at(a, n, co, y, x);
        if (!finite_f(cv)) flag |= 1;
        float r;
        const float l = residual(d, cv, av, rho_at(a, n, co, y, x), r);
        acc += (double)l;
        if (a.out_map)
            a.out_map[(((long long)n * d.c_out + co) * a.ho + (y - a.y_lo)) * a.wo + (x - a.x_lo)] = fabsf(r);
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
// Tile path (depthwise 3x3).  Tile = 120 x TH outputs; shared rows are 128 floats holding image
// columns x0-4 .. x0+123, so every row is 32 aligned float4 and lane L (1..30) owns outputs
// x0+4(L-1) .. x0+4(L-1)+3 reading float4 L-1, L, L+1 — conflict-free 128-bit LDS.
// ------------------------------------------------------------------------------------------------
constexpr int TWO = 120;  // outputs per tile row
constexpr int TWS = 128;  // shared row width (floats)
constexpr int THF = 32;   // forward tile height
constexpr int NTH = 256;

// Cooperative load of image rows [ya, ya+rows) x cols [x0-4, x0+124) of one plane into shared memory,
// applying the padding rule outside the image.  `vec_ok`: W % 4 == 0 and 16-byte aligned plane.
__device__ __forceinline__ void load_tile(float *sm, const float *plane, int H, int W, int ya, int rows, int x0,
                                          int pad_mode, bool vec_ok) {
    const int nvec = rows * (TWS / 4);
    for (int i = threadIdx.x; i < nvec; i += NTH) {
        const int r = i >> 5, v = i & 31;
        const int gx = x0 - 4 + 4 * v;
        const int sy = pad_index(ya + r, H, pad_mode);
        float4 val = make_float4(0.f, 0.f, 0.f, 0.f);
        if (sy >= 0) {
            const float *row = plane + (long long)sy * W;
            if (vec_ok && gx >= 0 && gx + 3 < W) {
                val = ldg_stream4(row + gx);
            } else {
                int s;
                s = pad_index(gx + 0, W, pad_mode); if (s >= 0) val.x = __ldg(row + s);
                s = pad_index(gx + 1, W, pad_mode); if (s >= 0) val.y = __ldg(row + s);
                s = pad_index(gx + 2, W, pad_mode); if (s >= 0) val.z = __ldg(row + s);
                s = pad_index(gx + 3, W, pad_mode); if (s >= 0) val.w = __ldg(row + s);
            }
        }
        reinterpret_cast<float4 *>(sm)[i] = val;
    }
}

struct Row6 {
    float v[6];  // columns 4(L-1)+3 .. 4(L+1)   (left neighbour, own 4, right neighbour)
};
__device__ __forceinline__ Row6 lds_row6(const float *srow, int lane) {
    const float4 *p = reinterpret_cast<const float4 *>(srow);
    const float4 l = p[lane - 1], m = p[lane], r = p[lane + 1];
    Row6 o;
    o.v[0] = l.w; o.v[1] = m.x; o.v[2] = m.y; o.v[3] = m.z; o.v[4] = m.w; o.v[5] = r.x;
    return o;
}
__device__ __forceinline__ void conv3_rows(const Row6 &r0, const Row6 &r1, const Row6 &r2, const float *w,
                                           float out[4]) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float s = w[0] * r0.v[i];
        s = fmaf(w[1], r0.v[i + 1], s);
        s = fmaf(w[2], r0.v[i + 2], s);
        s = fmaf(w[3], r1.v[i], s);
        s = fmaf(w[4], r1.v[i + 1], s);
        s = fmaf(w[5], r1.v[i + 2], s);
        s = fmaf(w[6], r2.v[i], s);
        s = fmaf(w[7], r2.v[i + 1], s);
        s = fmaf(w[8], r2.v[i + 2], s);
        out[i] = s;
    }
}

__global__ void __launch_bounds__(NTH) phys_fwd_tile3(const PhysArgs a) {
    __shared__ __align__(16) float sm[(THF + 2) * TWS];
    __shared__ double red[32];
    const llie_phys_desc &d = a.d;
    const int n = blockIdx.y;
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    const int H = d.h, W = d.w;
    const long long plane_sz = (long long)H * W;
    const int tiles_per_plane = a.tiles_x * a.tiles_y;
    const int tiles = d.c_out * tiles_per_plane;
    const bool map_vec = a.out_map != nullptr && (a.wo & 3) == 0 && a.x_lo == 0;
    float acc = 0.f;
    int flag = 0;

    for (int t = blockIdx.x; t < tiles; t += gridDim.x) {
        const int c = t / tiles_per_plane;
        const int tr = t - c * tiles_per_plane;
        const int ty = tr / a.tiles_x, tx = tr - ty * a.tiles_x;
        const int x0 = tx * TWO, y0 = ty * THF;
        const float *pplane = a.pred + ((long long)n * d.c_in + c) * plane_sz;
        const float *oplane = a.obs + ((long long)n * d.c_out + c) * plane_sz;
        float w[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) w[k] = __ldg(a.psf + c * 9 + k);
        float rho_nc = 1.f;
        if (a.rho == nullptr) rho_nc = d.rho_scalar;
        else rho_nc = __ldg(a.rho + (long long)n * d.rho_sn + (long long)c * d.rho_sc);

        __syncthreads();  // previous tile fully consumed
        load_tile(sm, pplane, H, W, y0 - 1, THF + 2, x0, d.pad_mode, true);
        __syncthreads();

        const int xx = x0 + 4 * (lane - 1);
        if (lane >= 1 && lane <= 30 && xx < W) {
            const int rbase = 4 * wrp;  // tile-local first output row of this warp
            Row6 r0 = lds_row6(sm + (rbase + 0) * TWS, lane);
            Row6 r1 = lds_row6(sm + (rbase + 1) * TWS, lane);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const Row6 r2 = lds_row6(sm + (rbase + j + 2) * TWS, lane);
                const int yy = y0 + rbase + j;
                if (yy < H) {
                    float cv[4];
                    conv3_rows(r0, r1, r2, w, cv);
                    const float4 ob = ldg_stream4(oplane + (long long)yy * W + xx);
                    const float av[4] = {ob.x, ob.y, ob.z, ob.w};
                    // finiteness of pred: the centre tap row carries pred itself
                    const bool row_in = (yy >= a.y_lo) && (yy < a.y_lo + a.ho);
                    float mv[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        if (!finite_f(av[i])) flag |= 2;
                        if (!finite_f(r1.v[i + 1])) flag |= 1;
                        const int x = xx + i;
                        const bool in_dom = row_in && (x >= a.x_lo) && (x < a.x_lo + a.wo);
                        float r;
                        const float l = residual(d, cv[i], av[i], rho_nc, r);
                        mv[i] = fabsf(r);
                        if (in_dom) acc += l;
                    }
                    if (a.out_map && row_in) {
                        float *mrow = a.out_map + (((long long)n * d.c_out + c) * a.ho + (yy - a.y_lo)) * a.wo;
                        if (map_vec) {
                            *reinterpret_cast<float4 *>(mrow + xx) = make_float4(mv[0], mv[1], mv[2], mv[3]);
                        } else {
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                const int xo = xx + i - a.x_lo;
                                if (xo >= 0 && xo < a.wo) mrow[xo] = mv[i];
                            }
                        }
                    }
                }
                r0 = r1;
                r1 = r2;
            }
        }
    }
    const double s = block_sum<double>((double)acc, red);
    const int f = block_or(flag, reinterpret_cast<int *>(red));
    if (threadIdx.x == 0) {
        a.part_sum[(size_t)n * gridDim.x + blockIdx.x] = s;
        a.part_flag[(size_t)n * gridDim.x + blockIdx.x] = f;
    }
    if (last_block_done(a.ticket, gridDim.x * gridDim.y)) phys_finalize(a, red);
}

// ------------------------------------------------------------------------------------------------
// Backward, tile path (depthwise 3x3).  pred tile with a 2-pixel halo -> q = dLoss/dconv on the tile
// plus a 1-pixel halo (zero outside the compared domain) -> grad = flipped-K stencil over q, plus the
// mirror/edge fold of reflect / replicate padding on image-border tiles.
// ------------------------------------------------------------------------------------------------
constexpr int THB = 32;

// 1-D set of padded coordinates that read image index i (the adjoint of the padding operator)
__device__ __forceinline__ int preimages(int i, int n, int p, int mode, int out[10]) {
    int k = 0;
    out[k++] = i;
    if (mode == LLIE_PAD_REFLECT) {
        if (i >= 1 && i <= p) out[k++] = -i;
        const int m = n - 1 - i;
        if (m >= 1 && m <= p) out[k++] = n - 1 + m;
    } else if (mode == LLIE_PAD_REPLICATE) {
        if (i == 0)
            for (int m = 1; m <= p; ++m) out[k++] = -m;
        if (i == n - 1)
            for (int m = 1; m <= p; ++m) out[k++] = n - 1 + m;
    }
    return k;
}

__global__ void __launch_bounds__(NTH) phys_bwd_tile3(const PhysArgs a) {
    __shared__ __align__(16) float sp[(THB + 4) * TWS];  // pred rows y0-2 .. y0+THB+1
    __shared__ __align__(16) float sq[(THB + 2) * TWS];  // q    rows y0-1 .. y0+THB
    const llie_phys_desc &d = a.d;
    const int n = blockIdx.y;
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    const int H = d.h, W = d.w;
    const long long plane_sz = (long long)H * W;
    const int tiles_per_plane = a.tiles_x * a.tiles_y;
    const int tiles = d.c_out * tiles_per_plane;
    const float gn = __ldg(a.gout + a.gout_stride * n) * a.gscale;

    for (int t = blockIdx.x; t < tiles; t += gridDim.x) {
        const int c = t / tiles_per_plane;
        const int tr = t - c * tiles_per_plane;
        const int ty = tr / a.tiles_x, tx = tr - ty * a.tiles_x;
        const int x0 = tx * TWO, y0 = ty * THB;
        const float *pplane = a.pred + ((long long)n * d.c_in + c) * plane_sz;
        const float *oplane = a.obs + ((long long)n * d.c_out + c) * plane_sz;
        float *gplane = a.grad + ((long long)n * d.c_in + c) * plane_sz;
        float w[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) w[k] = __ldg(a.psf + c * 9 + k);
        float rho_nc = (a.rho == nullptr) ? d.rho_scalar
                                          : __ldg(a.rho + (long long)n * d.rho_sn + (long long)c * d.rho_sc);

        __syncthreads();
        load_tile(sp, pplane, H, W, y0 - 2, THB + 4, x0, d.pad_mode, true);
        __syncthreads();

        // ---- phase 2: q on rows y0-1 .. y0+THB (34 rows), columns x0-4+4(L) .. for lanes 1..30
        // (columns x0 .. x0+119 plus one halo column each side are what phase 3 needs: lanes 1..30
        //  cover x0 .. x0+119; the halo columns x0-1 and x0+120 are produced by lanes 0 and 31 below)
        {
            // warp handles q rows [qa, qb) (tile-local, 0 .. THB+1)
            const int per = (THB + 2 + 7) / 8;  // 5
            const int qa = wrp * per;
            const int qb = min(qa + per, THB + 2);
            if (qa < qb) {
                const bool edge_lane = (lane == 0) || (lane == 31);
                // lanes 0 / 31 cannot read float4 lane-1 / lane+1: they compute only their inner column
                for (int qr = qa; qr < qb; ++qr) {
                    const int yy = y0 - 1 + qr;  // image row of this q row
                    float qv[4] = {0.f, 0.f, 0.f, 0.f};
                    const bool row_in = (yy >= a.y_lo) && (yy < a.y_lo + a.ho);
                    if (row_in) {
                        const float *s0 = sp + (qr + 0) * TWS;  // pred row yy-1
                        const float *s1 = sp + (qr + 1) * TWS;
                        const float *s2 = sp + (qr + 2) * TWS;
                        const int xx = x0 - 4 + 4 * lane;
                        if (!edge_lane) {
                            const Row6 r0 = lds_row6(s0, lane), r1 = lds_row6(s1, lane), r2 = lds_row6(s2, lane);
                            float cv[4];
                            conv3_rows(r0, r1, r2, w, cv);
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                const int x = xx + i;
                                if (x >= a.x_lo && x < a.x_lo + a.wo) {
                                    const float av = __ldg(oplane + (long long)yy * W + x);
                                    qv[i] = gn * dresidual(d, cv[i], av, rho_nc);
                                }
                            }
                        } else {
                            // single needed column: x0-1 (lane 0, i=3) or x0+120 (lane 31, i=0)
                            const int i = (lane == 0) ? 3 : 0;
                            const int sc = 4 * lane + i;
                            const int x = xx + i;
                            if (x >= a.x_lo && x < a.x_lo + a.wo) {
                                float s = 0.f;
#pragma unroll
                                for (int ky = 0; ky < 3; ++ky) {
                                    const float *sr = sp + (qr + ky) * TWS;
#pragma unroll
                                    for (int kx = 0; kx < 3; ++kx) s = fmaf(w[ky * 3 + kx], sr[sc + kx - 1], s);
                                }
                                const float av = __ldg(oplane + (long long)yy * W + x);
                                qv[i] = gn * dresidual(d, s, av, rho_nc);
                            }
                        }
                    }
                    reinterpret_cast<float4 *>(sq + qr * TWS)[lane] = make_float4(qv[0], qv[1], qv[2], qv[3]);
                }
            }
        }
        __syncthreads();

        // ---- phase 3: grad(y,x) = sum_k K[k] q(y - ky + 1, x - kx + 1)   (flipped stencil)
        const int xx = x0 + 4 * (lane - 1);
        const bool border_tile = d.pad_mode != LLIE_PAD_ZEROS &&
                                 (y0 <= 1 || y0 + THB + 1 >= H || x0 <= 1 || x0 + TWO + 1 >= W);
        if (lane >= 1 && lane <= 30 && xx < W) {
            const int rbase = 4 * wrp;
            float wf[9];
#pragma unroll
            for (int k = 0; k < 9; ++k) wf[k] = w[8 - k];
            Row6 r0 = lds_row6(sq + (rbase + 0) * TWS, lane);
            Row6 r1 = lds_row6(sq + (rbase + 1) * TWS, lane);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const Row6 r2 = lds_row6(sq + (rbase + j + 2) * TWS, lane);
                const int yy = y0 + rbase + j;
                if (yy < H) {
                    float g[4];
                    conv3_rows(r0, r1, r2, wf, g);
                    if (border_tile) {
                        // add the contributions of padded coordinates that alias onto (yy, x)
                        int py[10], px[10];
                        const int ny = preimages(yy, H, 1, d.pad_mode, py);
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const int x = xx + i;
                            if (x >= W) continue;
                            const int nx = preimages(x, W, 1, d.pad_mode, px);
                            if (ny == 1 && nx == 1) continue;
                            float extra = 0.f;
                            for (int iy = 0; iy < ny; ++iy)
                                for (int ix = 0; ix < nx; ++ix) {
                                    if (iy == 0 && ix == 0) continue;  // the pixel itself: already in g
                                    const int jy = py[iy], jx = px[ix];
                                    for (int ky = 0; ky < 3; ++ky) {
                                        const int oy = jy - ky + 1;
                                        if (oy < 0 || oy >= H) continue;
                                        for (int kx = 0; kx < 3; ++kx) {
                                            const int ox = jx - kx + 1;
                                            if (ox < 0 || ox >= W) continue;
                                            // q lives in shared: row oy-(y0-1), col ox-(x0-4)
                                            extra = fmaf(w[ky * 3 + kx], sq[(oy - y0 + 1) * TWS + (ox - x0 + 4)], extra);
                                        }
                                    }
                                }
                            g[i] += extra;
                        }
                    }
                    *reinterpret_cast<float4 *>(gplane + (long long)yy * W + xx) = make_float4(g[0], g[1], g[2], g[3]);
                }
                r0 = r1;
                r1 = r2;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Backward, generic path: (1) q map into scratch, (2) transposed gather with padding fold.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) phys_bwd_q_generic(const PhysArgs a) {
    const llie_phys_desc &d = a.d;
    const int n = blockIdx.y;
    const long long plane = (long long)d.h * d.w;
    const long long total = plane * d.c_out;
    const float gn = __ldg(a.gout + a.gout_stride * n) * a.gscale;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        const int co = (int)(idx / plane);
        const int rem = (int)(idx - (long long)co * plane);
        const int y = rem / d.w, x = rem - y * d.w;
        float q = 0.f;
        if (y >= a.y_lo && y < a.y_lo + a.ho && x >= a.x_lo && x < a.x_lo + a.wo) {
            const float cv = conv_at(a, n, co, y, x);
            const float av = __ldg(a.obs + ((long long)n * d.c_out + co) * plane + rem);
            q = gn * dresidual(d, cv, av, rho_at(a, n, co, y, x));
        }
        a.scratch[((long long)n * d.c_out + co) * plane + rem] = q;
    }
}

__global__ void __launch_bounds__(256) phys_bwd_gather_generic(const PhysArgs a) {
    const llie_phys_desc &d = a.d;
    const int n = blockIdx.y;
    const int ph = d.kh >> 1, pw = d.kw >> 1;
    const long long plane = (long long)d.h * d.w;
    const long long total = plane * d.c_in;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        const int ci = (int)(idx / plane);
        const int rem = (int)(idx - (long long)ci * plane);
        const int y = rem / d.w, x = rem - y * d.w;
        float acc = 0.f;
        const int co0 = d.psf_layout == LLIE_PSF_DENSE ? 0 : ci;
        const int co1 = d.psf_layout == LLIE_PSF_DENSE ? d.c_out : ci + 1;
        // candidate padded coordinates that may alias onto (y, x): the pixel itself plus the
        // out-of-image bands [-p, 0) and [size, size+p)
        for (int ty = 0; ty <= 2 * ph; ++ty) {
            const int jy = (ty == 0) ? y : (ty <= ph ? -ty : d.h - 1 + (ty - ph));
            if (ty != 0 && pad_index(jy, d.h, d.pad_mode) != y) continue;
            for (int tx = 0; tx <= 2 * pw; ++tx) {
                const int jx = (tx == 0) ? x : (tx <= pw ? -tx : d.w - 1 + (tx - pw));
                if (tx != 0 && pad_index(jx, d.w, d.pad_mode) != x) continue;
                for (int co = co0; co < co1; ++co) {
                    const float *wk =
                        a.psf + (d.psf_layout == LLIE_PSF_DENSE ? ((long long)co * d.c_in + ci) : co) * d.kh * d.kw;
                    const float *qp = a.scratch + ((long long)n * d.c_out + co) * plane;
                    for (int ky = 0; ky < d.kh; ++ky) {
                        const int oy = jy - ky + ph;
                        if (oy < 0 || oy >= d.h) continue;
                        for (int kx = 0; kx < d.kw; ++kx) {
                            const int ox = jx - kx + pw;
                            if (ox < 0 || ox >= d.w) continue;
                            acc = fmaf(__ldg(wk + ky * d.kw + kx), qp[(long long)oy * d.w + ox], acc);
                        }
                    }
                }
            }
        }
        a.grad[((long long)n * d.c_in + ci) * plane + rem] = acc;
    }
}

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

static bool tile_path_ok(const PhysArgs &a) {
    const llie_phys_desc &d = a.d;
    if (d.psf_layout != LLIE_PSF_DEPTHWISE || d.kh != 3 || d.kw != 3) return false;
    if ((d.w & 3) != 0 || d.w < 4) return false;
    if (!aligned16(a.pred) || !aligned16(a.obs)) return false;
    if (a.rho != nullptr && (d.rho_sh != 0 || d.rho_sw != 0)) return false;
    if (d.pad_mode == LLIE_PAD_REFLECT && (d.h < 3 || d.w < 3)) return false;
    return true;
}

static int grid_x_for(int n, int work_items) {
    // persistent-style grid: ~8 resident CTAs of 256 threads per SM, split evenly over samples
    int per_sample = div_up(kNumSM * 8, n);
    if (per_sample > work_items) per_sample = work_items;
    if (per_sample < 1) per_sample = 1;
    return per_sample;
}

}  // namespace llie

using namespace llie;

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
    const bool tile = tile_path_ok(a);
    int gx;
    if (tile) {
        a.tiles_x = div_up(desc->w, TWO);
        a.tiles_y = div_up(desc->h, THF);
        gx = grid_x_for(desc->n, desc->c_out * a.tiles_x * a.tiles_y);
    } else {
        long long items = (long long)desc->c_out * desc->h * desc->w;
        gx = grid_x_for(desc->n, (int)((items + 255) / 256));
    }
    const size_t need = (size_t)desc->n * gx * (sizeof(double) + sizeof(int));
    LLIE_REQUIRE(need <= ws.scratch_bytes, "phys_forward: workspace too small");
    a.part_sum = reinterpret_cast<double *>(ws.scratch);
    a.part_flag = reinterpret_cast<int *>(ws.scratch + (size_t)desc->n * gx * sizeof(double));
    dim3 grid(gx, desc->n);
    if (tile)
        phys_fwd_tile3<<<grid, NTH, 0, st>>>(a);
    else
        phys_fwd_generic<<<grid, 256, 0, st>>>(a);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
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
    if (tile_path_ok(a) && aligned16(grad_pred)) {
        a.tiles_x = div_up(desc->w, TWO);
        a.tiles_y = div_up(desc->h, THB);
        dim3 grid(grid_x_for(desc->n, desc->c_out * a.tiles_x * a.tiles_y), desc->n);
        phys_bwd_tile3<<<grid, NTH, 0, st>>>(a);
    } else {
        LLIE_REQUIRE(scratch != nullptr, "phys_backward: this PSF/shape needs the scratch buffer");
        LLIE_REQUIRE(desc->kh / 2 <= 16 && desc->kw / 2 <= 16, "phys_backward: PSF larger than 33x33 unsupported");
        long long items = (long long)desc->c_out * desc->h * desc->w;
        dim3 grid(grid_x_for(desc->n, (int)((items + 255) / 256)), desc->n);
        phys_bwd_q_generic<<<grid, 256, 0, st>>>(a);
        long long items2 = (long long)desc->c_in * desc->h * desc->w;
        dim3 grid2(grid_x_for(desc->n, (int)((items2 + 255) / 256)), desc->n);
        phys_bwd_gather_generic<<<grid2, 256, 0, st>>>(a);
    }
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}
