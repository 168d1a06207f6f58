// Gaussian-window SSIM statistics (sm_100a): per-(n,c) sums of the SSIM map (or of the DSSIM loss
// map) for the three SSIM conventions of the reference, one fused kernel instead of 2 pads + 5
// depthwise K x K convolutions + ~15 pointwise passes:
//   * metrics/linear.py:213-324   ssim_linear      selectable padding, "same", var clamp >= 0, eps
//   * metrics/ssim.py:246-311     torchmetrics 1.2 reflect pad, interior-only mean, no clamp, no eps
//   * NewBP_model/losses.py:146-155 kornia 0.6.12  reflect pad, "same", eps, inputs clamped to [0,1],
//                                                   loss map clamp((1-S)/2, 0, 1)
// The K x K window of all three is an outer product, so the filter runs separably on a shared-memory
// tile: (A) padded input tile -> smem, (B) horizontal pass of the 5 moment maps (x, y, x^2, y^2, xy)
// with 4 outputs per thread in registers, (C) vertical pass with 8 outputs per thread in registers,
// SSIM formula and the masked fp64 reduction.  Bank-conflict-free odd pitches.
// Roofline: this kernel is bound by the FP32 FMA pipe, not by HBM: 2 x 5 x K FMAs per element
// (110 for K = 11) against 8 bytes of traffic (SSIM section).
#include "common.cuh"
#include <atomic>
#include "ssim_common.cuh"

namespace llie {

constexpr int STH = 32, STW = 64, SNT = 256;

__device__ __forceinline__ float ssim_load(const SsimArgs &a, const float *base, long long plane_stride, int sy,
                                           int sx, float &chk) {
    const llie_ssim_desc &d = a.d;
    float v;
    if (d.luma) {
        const float *p = base + (long long)sy * d.w + sx;
        const float r = __ldg(p), g = __ldg(p + plane_stride), b = __ldg(p + 2 * plane_stride);
        chk = fmaf(r, 0.f, chk);
        chk = fmaf(g, 0.f, chk);
        chk = fmaf(b, 0.f, chk);
        v = 0.2989f * r + 0.5870f * g + 0.1140f * b;  // metrics/ssim.py:116-127 (BT.601)
    } else {
        v = __ldg(base + (long long)sy * d.w + sx);
        chk = fmaf(v, 0.f, chk);
    }
    if (d.clamp_in) v = fminf(fmaxf(v, 0.f), 1.f);
    return v;
}

// K > 0: window size known at compile time (taps in registers, fully unrolled passes).  K == 0: any odd window up to
// LLIE_SSIM_MAX_K read from the descriptor (taps in shared memory, rolled loops) -- the windows above 15 taps that the
// reference accepts but nobody trains with; same arithmetic, same summation order.
template <int K>
__global__ void __launch_bounds__(SNT, K ? 2 : 1) ssim_stats_kernel(const SsimArgs a) {
    const int KK = K ? K : a.d.ksize;
    const int P = KK / 2;
    const int XR = STH + 2 * P;           // rows of the input tile / of the horizontal-pass output
    const int XP = (STW + 2 * P) | 1;     // odd pitches: lanes walk rows without bank conflicts
    constexpr int HP = STW + 1;
    extern __shared__ __align__(16) float sm[];
    float *sX = sm, *sY = sX + XR * XP;
    float *sH = sY + XR * XP;  // [5][XR][HP]
    __shared__ double red[32];
    __shared__ float sg[K ? 1 : LLIE_SSIM_MAX_K];
    const llie_ssim_desc &d = a.d;
    const int H = d.h, W = d.w;
    const int tid = threadIdx.x;

    // tile coordinates
    const int tpp = a.tiles_x * a.tiles_y;
    const int plane = blockIdx.x / tpp;  // n * co + c
    const int trem = blockIdx.x - plane * tpp;
    const int ty = trem / a.tiles_x, tx = trem - ty * a.tiles_x;
    const int y0 = ty * STH, x0 = tx * STW;
    const int n = plane / a.co, c = plane - n * a.co;
    const long long plane_elems = (long long)H * W;
    const float *xb = a.x + ((long long)n * d.c + (d.luma ? 0 : c)) * plane_elems;
    const float *yb = a.y + ((long long)n * d.c + (d.luma ? 0 : c)) * plane_elems;

    float g[K ? K : 1];
#pragma unroll
    for (int j = 0; j < K; ++j) g[j] = __ldg(a.win + j);
    if (K == 0 && tid < KK) sg[tid] = __ldg(a.win + tid);
#define SSIM_TAP(j) (K ? g[K ? (j) : 0] : sg[j])

    // ---- (A) padded input tile
    float chk_x = 0.f, chk_y = 0.f;
    for (int idx = tid; idx < XR * (STW + 2 * P); idx += SNT) {
        const int r = idx / (STW + 2 * P), cc = idx - r * (STW + 2 * P);
        const int sy = pad_index(y0 - P + r, H, d.pad_mode), sx = pad_index(x0 - P + cc, W, d.pad_mode);
        float xv = 0.f, yv = 0.f;
        if (sy >= 0 && sx >= 0) {
            xv = ssim_load(a, xb, plane_elems, sy, sx, chk_x);
            yv = ssim_load(a, yb, plane_elems, sy, sx, chk_y);
        }
        sX[r * XP + cc] = xv;
        sY[r * XP + cc] = yv;
    }
    __syncthreads();

    // ---- (B) horizontal pass: task = (row, group of 4 columns); lanes walk rows
    for (int t = tid; t < XR * (STW / 4); t += SNT) {
        const int cg = t / XR, r = t - cg * XR;
        const float *px = sX + r * XP + cg * 4, *py = sY + r * XP + cg * 4;
        float ax[4], ay[4], axx[4], ayy[4], axy[4];
#pragma unroll
        for (int o = 0; o < 4; ++o) ax[o] = ay[o] = axx[o] = ayy[o] = axy[o] = 0.f;
#pragma unroll(K ? K + 3 : 1)
        for (int u = 0; u < KK + 3; ++u) {
            const float xv = px[u], yv = py[u];
            const float xx = xv * xv, yy = yv * yv, xy = xv * yv;
#pragma unroll
            for (int o = 0; o < 4; ++o) {
                const int j = u - o;
                if (j >= 0 && j < KK) {
                    const float gj = SSIM_TAP(j);
                    ax[o] = fmaf(gj, xv, ax[o]);
                    ay[o] = fmaf(gj, yv, ay[o]);
                    axx[o] = fmaf(gj, xx, axx[o]);
                    ayy[o] = fmaf(gj, yy, ayy[o]);
                    axy[o] = fmaf(gj, xy, axy[o]);
                }
            }
        }
        float *h = sH + r * HP + cg * 4;
#pragma unroll
        for (int o = 0; o < 4; ++o) {
            h[o] = ax[o];
            h[XR * HP + o] = ay[o];
            h[2 * XR * HP + o] = axx[o];
            h[3 * XR * HP + o] = ayy[o];
            h[4 * XR * HP + o] = axy[o];
        }
    }
    __syncthreads();

    // ---- (C) vertical pass: thread = (column, group of 8 rows), then the SSIM formula
    const int col = tid & (STW - 1), rg = tid / STW;  // SNT == STW * (STH / 8)
    float m[5][8];
#pragma unroll
    for (int q = 0; q < 5; ++q) {
        const float *h = sH + q * XR * HP + (rg * 8) * HP + col;
#pragma unroll
        for (int o = 0; o < 8; ++o) m[q][o] = 0.f;
#pragma unroll(K ? K + 7 : 1)
        for (int u = 0; u < KK + 7; ++u) {
            const float v = h[u * HP];
#pragma unroll
            for (int o = 0; o < 8; ++o) {
                const int j = u - o;
                if (j >= 0 && j < KK) m[q][o] = fmaf(SSIM_TAP(j), v, m[q][o]);
            }
        }
    }
    const int gx = x0 + col;
    const int lo_y = d.interior ? P : 0, hi_y = d.interior ? H - P : H;
    const int lo_x = d.interior ? P : 0, hi_x = d.interior ? W - P : W;
    double acc = 0.0;
    float facc = 0.f;
#pragma unroll
    for (int o = 0; o < 8; ++o) {
        const int gy = y0 + rg * 8 + o;
        const float mx = m[0][o], my = m[1][o];
        const float mx2 = mx * mx, my2 = my * my, mxy = mx * my;
        float sx = m[2][o] - mx2, sy = m[3][o] - my2;
        const float sxy = m[4][o] - mxy;
        if (d.clamp_var) {
            sx = fmaxf(sx, 0.f);
            sy = fmaxf(sy, 0.f);
        }
        const float num = (2.f * mxy + d.c1) * (2.f * sxy + d.c2);
        const float den = (mx2 + my2 + d.c1) * (sx + sy + d.c2) + d.eps;
        float s = num / den;
        if (a.out_map != nullptr && gy < H && gx < W) a.out_map[((long long)plane * H + gy) * W + gx] = s;
        if (a.coef != nullptr && gy < H && gx < W) {
            // S as a function of (mu_x, E[x^2], E[xy]) at this window:
            //   A3 = dS/dsigma_xy, A2 = dS/dsigma_x^2, A1 = dS/dmu_x - 2 mu_x A2 - mu_y A3
            const float l = (1.f - s) * 0.5f;
            const float up = (!d.loss_map) ? a.coef_scale : ((l >= 0.f && l <= 1.f) ? -0.5f * a.coef_scale : 0.f);
            const float inv = 1.f / den;
            const float n1 = 2.f * mxy + d.c1, n2 = 2.f * sxy + d.c2, d1 = mx2 + my2 + d.c1, d2 = sx + sy + d.c2;
            const float a3 = 2.f * n1 * inv;
            const float a2 = -s * d1 * inv;
            const float dmu = 2.f * (my * n2 - s * mx * d2) * inv;
            const float a1 = dmu - 2.f * mx * a2 - my * a3;
            const long long stride = (long long)d.n * a.co * plane_elems;
            float *cp = a.coef + ((long long)plane * H + gy) * W + gx;
            cp[0] = up * a1;
            cp[stride] = up * a2;
            cp[2 * stride] = up * a3;
        }
        if (d.loss_map) s = fminf(fmaxf((1.f - s) * 0.5f, 0.f), 1.f);
        if (gy >= lo_y && gy < hi_y && gx >= lo_x && gx < hi_x) facc += s;
    }
    acc = (double)facc;
    const double bs = block_sum<double>(acc, red);
    int flag = 0;
    if (!(chk_x == 0.f)) flag |= 1;
    if (!(chk_y == 0.f)) flag |= 2;
    const int f = block_or(flag, reinterpret_cast<int *>(red));
    if (tid == 0) {
        a.part[blockIdx.x] = bs;
        a.pflag[blockIdx.x] = f;
    }
    if (!last_block_done(a.ticket, gridDim.x)) return;

    // ---- finalize: per-plane sums in fixed tile order (one warp per plane, 4 loads in flight per lane)
    const int lane = tid & 31, wrp = tid >> 5;
    const int planes = d.n * a.co;
    for (int pl = wrp; pl < planes; pl += SNT / 32) {
        const double *p = a.part + (size_t)pl * tpp;
        double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
        int b = lane;
        for (; b + 96 < tpp; b += 128) {
            s0 += __ldcg(p + b);
            s1 += __ldcg(p + b + 32);
            s2 += __ldcg(p + b + 64);
            s3 += __ldcg(p + b + 96);
        }
        for (; b < tpp; b += 32) s0 += __ldcg(p + b);
        const double s = warp_sum((s0 + s1) + (s2 + s3));
        if (lane == 0) a.out_sum[pl] = s;
    }
    int fl = 0;
    for (int b = tid; b < (int)gridDim.x; b += SNT) fl |= __ldcg(a.pflag + b);
    const int ff = block_or(fl, reinterpret_cast<int *>(red));
    if (tid == 0) *a.out_flags = ff;
}

template <int K>
static int launch_ssim(const SsimArgs &a, int grid, cudaStream_t st) {
    const int P = (K ? K : a.d.ksize) / 2, XR = STH + 2 * P, XP = (STW + 2 * P) | 1, HP = STW + 1;
    const size_t smem = (size_t)(2 * XR * XP + 5 * XR * HP) * sizeof(float);
    LLIE_CUDA_OK(cudaFuncSetAttribute(ssim_stats_kernel<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ssim_stats_kernel<K><<<grid, SNT, smem, st>>>(a);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

// ---------------------------------------------------------------------------------------------------
// Streaming variant (default for non-luma inputs): one thread per output column of a 128-column strip, rows
// streamed top to bottom through a 4-row shared-memory ring filled by cp.async three rows ahead (no
// registers held by loads in flight, one __syncthreads per row).  Per input row a thread runs the K-tap
// horizontal pass of the five moment maps from shared memory and scatters the five values into K rolling
// vertical accumulators per map held in registers (slot = output row mod K, static after unrolling by K).
// Shared memory per CTA is 4.4 KB instead of the tile kernel's 80 KB, so occupancy is bound by registers only.
constexpr int SW = 128;
constexpr int SD = 4;  // ring depth (rows)

__device__ __forceinline__ void cp_async4(unsigned dst, const float *src, unsigned src_bytes) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

template <int K>
__global__ void __launch_bounds__(SW) ssim_stream_kernel(const SsimArgs a) {
    constexpr int P = K / 2;
    constexpr int RW = SW + 2 * P;
    __shared__ float sX[SD][RW], sY[SD][RW];
    __shared__ double red[32];
    const llie_ssim_desc &d = a.d;
    const int H = d.h, W = d.w, tid = threadIdx.x;
    const int tpp = a.tiles_x * a.tiles_y;  // strips x row segments per plane
    const int plane = blockIdx.x / tpp;
    const int trem = blockIdx.x - plane * tpp;
    const int seg = trem / a.tiles_x, strip = trem - seg * a.tiles_x;
    const int x0 = strip * SW, y0 = seg * a.seg_rows;
    const int rows_out = min(a.seg_rows, H - y0);
    const int nrows_in = rows_out + 2 * P;
    const long long plane_elems = (long long)H * W;
    const float *xb = a.x + (long long)plane * plane_elems;  // no luma here: plane == n * c + ch
    const float *yb = a.y + (long long)plane * plane_elems;

    float g[K];
#pragma unroll
    for (int j = 0; j < K; ++j) g[j] = __ldg(a.win + j);

    // padded source columns of this thread's ring slots (tid, and SW + tid for the right halo); columns no
    // in-image output needs read as zero
    const int gin0 = x0 - P + tid, gin1 = x0 - P + SW + tid;
    const int cmain = gin0 <= W - 1 + P ? pad_index(gin0, W, d.pad_mode) : -1;
    const bool has_ext = tid < 2 * P;
    const int cext = (has_ext && gin1 <= W - 1 + P) ? pad_index(gin1, W, d.pad_mode) : -1;
    const float *pxm = xb + max(cmain, 0), *pym = yb + max(cmain, 0);
    const float *pxe = xb + max(cext, 0), *pye = yb + max(cext, 0);
    const unsigned bm = cmain >= 0 ? 4u : 0u, be = cext >= 0 ? 4u : 0u;
    const unsigned sxa = (unsigned)__cvta_generic_to_shared(&sX[0][tid]);
    const unsigned sya = (unsigned)__cvta_generic_to_shared(&sY[0][tid]);

    auto issue_row = [&](int i) {  // i is uniform over the CTA; one commit group per call
        if (i < nrows_in) {
            const int sy = pad_index(y0 - P + i, H, d.pad_mode);
            const long long ro = (long long)max(sy, 0) * W;
            const unsigned so = (unsigned)((i & (SD - 1)) * RW * 4);
            const unsigned ok = sy >= 0 ? 4u : 0u;
            cp_async4(sxa + so, pxm + ro, min(bm, ok));
            cp_async4(sya + so, pym + ro, min(bm, ok));
            if (has_ext) {
                cp_async4(sxa + so + SW * 4, pxe + ro, min(be, ok));
                cp_async4(sya + so + SW * 4, pye + ro, min(be, ok));
            }
        }
        cp_async_commit();
    };

    const int gx = x0 + tid;
    const int lo_y = d.interior ? P : 0, hi_y = d.interior ? H - P : H;
    const bool col_in = gx < W;
    const bool col_ok = col_in && (!d.interior || (gx >= P && gx < W - P));
    const long long cstride = (long long)d.n * a.co * plane_elems;
    long long off = ((long long)plane * H + y0) * W + gx;  // output element of the next completed row
    float acc[5][K];
#pragma unroll
    for (int q = 0; q < 5; ++q)
#pragma unroll
        for (int j = 0; j < K; ++j) acc[q][j] = 0.f;
    float facc = 0.f, chk_x = 0.f, chk_y = 0.f;

#pragma unroll
    for (int i = 0; i < SD - 1; ++i) issue_row(i);
    for (int base = 0; base < nrows_in; base += K) {
#pragma unroll
        for (int ii = 0; ii < K; ++ii) {
            const int i = base + ii;
            if (i >= nrows_in) break;  // uniform
            float *rx = sX[i & (SD - 1)], *ry = sY[i & (SD - 1)];
            cp_async_wait<SD - 2>();  // this thread's copies of row i have landed
            {
                // own elements: finite check and the optional clamp, before they become visible to the CTA
                float vx = rx[tid], vy = ry[tid];
                chk_x = fmaf(vx, 0.f, chk_x);
                chk_y = fmaf(vy, 0.f, chk_y);
                if (d.clamp_in) {
                    rx[tid] = fminf(fmaxf(vx, 0.f), 1.f);
                    ry[tid] = fminf(fmaxf(vy, 0.f), 1.f);
                }
                if (has_ext) {
                    vx = rx[SW + tid];
                    vy = ry[SW + tid];
                    chk_x = fmaf(vx, 0.f, chk_x);
                    chk_y = fmaf(vy, 0.f, chk_y);
                    if (d.clamp_in) {
                        rx[SW + tid] = fminf(fmaxf(vx, 0.f), 1.f);
                        ry[SW + tid] = fminf(fmaxf(vy, 0.f), 1.f);
                    }
                }
            }
            __syncthreads();          // row i visible to all; everyone is done with row i - 1
            issue_row(i + SD - 1);    // refills the slot row i - 1 occupied
            float h[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int j = 0; j < K; ++j) {
                const float xv = rx[tid + j], yv = ry[tid + j];
                h[0] = fmaf(g[j], xv, h[0]);
                h[1] = fmaf(g[j], yv, h[1]);
                h[2] = fmaf(g[j], xv * xv, h[2]);
                h[3] = fmaf(g[j], yv * yv, h[3]);
                h[4] = fmaf(g[j], xv * yv, h[4]);
            }
            // input row i feeds output rows o = i - j with tap j; slot(o) = o mod K = (ii - j) mod K
#pragma unroll
            for (int j = 0; j < K; ++j) {
                const int sl = (ii - j + K) % K;
#pragma unroll
                for (int q = 0; q < 5; ++q) acc[q][sl] = (j == 0) ? g[0] * h[q] : fmaf(g[j], h[q], acc[q][sl]);
            }
            if (i >= K - 1) {  // output row o = i - (K - 1) is complete
                const int sl = (ii + 1) % K;
                const int gy = y0 + i - (K - 1);
                const float mx = acc[0][sl], my = acc[1][sl];
                const float mx2 = mx * mx, my2 = my * my, mxy = mx * my;
                float sx = acc[2][sl] - mx2, sy = acc[3][sl] - my2;
                const float sxy = acc[4][sl] - mxy;
                if (d.clamp_var) {
                    sx = fmaxf(sx, 0.f);
                    sy = fmaxf(sy, 0.f);
                }
                const float num = (2.f * mxy + d.c1) * (2.f * sxy + d.c2);
                const float den = (mx2 + my2 + d.c1) * (sx + sy + d.c2) + d.eps;
                float sv = num / den;
                if (a.out_map != nullptr && col_in) a.out_map[off] = sv;
                if (a.coef != nullptr && col_in) {
                    const float l = (1.f - sv) * 0.5f;
                    const float up =
                        (!d.loss_map) ? a.coef_scale : ((l >= 0.f && l <= 1.f) ? -0.5f * a.coef_scale : 0.f);
                    const float inv = 1.f / den;
                    const float n1 = 2.f * mxy + d.c1, n2 = 2.f * sxy + d.c2, d1 = mx2 + my2 + d.c1,
                                d2 = sx + sy + d.c2;
                    const float a3 = 2.f * n1 * inv;
                    const float a2 = -sv * d1 * inv;
                    const float dmu = 2.f * (my * n2 - sv * mx * d2) * inv;
                    const float a1 = dmu - 2.f * mx * a2 - my * a3;
                    float *cp = a.coef + off;
                    cp[0] = up * a1;
                    cp[cstride] = up * a2;
                    cp[2 * cstride] = up * a3;
                }
                off += W;
                if (d.loss_map) sv = fminf(fmaxf((1.f - sv) * 0.5f, 0.f), 1.f);
                if (col_ok && gy >= lo_y && gy < hi_y) facc += sv;
            }
        }
    }
    cp_async_wait<0>();
    const double bs = block_sum<double>((double)facc, red);
    int flag = 0;
    if (!(chk_x == 0.f)) flag |= 1;
    if (!(chk_y == 0.f)) flag |= 2;
    const int f = block_or(flag, reinterpret_cast<int *>(red));
    if (tid == 0) {
        a.part[blockIdx.x] = bs;
        a.pflag[blockIdx.x] = f;
    }
    if (!last_block_done(a.ticket, gridDim.x)) return;
    const int lane = tid & 31, wrp = tid >> 5;
    const int planes = d.n * a.co;
    for (int pl = wrp; pl < planes; pl += SW / 32) {
        const double *p = a.part + (size_t)pl * tpp;
        double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
        int b = lane;
        for (; b + 96 < tpp; b += 128) {
            s0 += __ldcg(p + b);
            s1 += __ldcg(p + b + 32);
            s2 += __ldcg(p + b + 64);
            s3 += __ldcg(p + b + 96);
        }
        for (; b < tpp; b += 32) s0 += __ldcg(p + b);
        const double sm_ = warp_sum((s0 + s1) + (s2 + s3));
        if (lane == 0) a.out_sum[pl] = sm_;
    }
    int fl = 0;
    for (int b = tid; b < (int)gridDim.x; b += SW) fl |= __ldcg(a.pflag + b);
    const int ff = block_or(fl, reinterpret_cast<int *>(red));
    if (tid == 0) *a.out_flags = ff;
}

template <int K>
static int launch_ssim_stream(const SsimArgs &a, int grid, cudaStream_t st) {
    ssim_stream_kernel<K><<<grid, SW, 0, st>>>(a);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

// ---------------------------------------------------------------------------------------------------
// Streaming variant 2 (no SSIM map; statistics, optionally the loss-coefficient planes): the same row
// ring and rolling vertical accumulators, but
//   * one thread owns TWO adjacent output columns of a 256-column strip: the K + 1 input values it needs per row and
//     array come from (K + 1) / 2 conflict-free LDS.64, the x^2 / y^2 / xy products of a position are formed once and
//     feed both columns, and the per-row bookkeeping (ring refill, barrier, address arithmetic) is paid per two
//     elements;
//   * without the per-image variance clamp (torchmetrics / kornia policies) sigma_x^2 and sigma_y^2 only enter SSIM
//     through their sum, so FOUR moment maps are filtered (x, y, x^2 + y^2, xy) instead of five: 88 instead of 110
//     multiply-adds per element.  MAPS5 keeps the five maps for `clamp_var` (ssim_linear), bit-identical to
//     ssim_stream_kernel.
// 312 -> ~170 instructions per element at K = 11.
constexpr int S2W = 256;
constexpr int S2T = 128;

template <int K, bool MAPS5, bool COEF = false, bool DUAL = false>
__global__ void __launch_bounds__(S2T, 3) ssim_stream2_kernel(const SsimArgs a) {
    static_assert(!DUAL || (MAPS5 && !COEF), "the second policy is evaluated from the five-map statistics");
    constexpr int P = K / 2;
    constexpr int RW = S2W + 2 * P;  // even: row starts stay 8-byte aligned
    constexpr int M = MAPS5 ? 5 : 4;
    static_assert(2 * P <= S2T && (RW % 2) == 0, "halo slots are loaded by the first 2P threads");
    __shared__ __align__(16) float sX[SD][RW], sY[SD][RW];
    __shared__ double red[32];
    const llie_ssim_desc &d = a.d;
    const int H = d.h, W = d.w, tid = threadIdx.x;
    const int tpp = a.tiles_x * a.tiles_y;
    const int plane = blockIdx.x / tpp;
    const int trem = blockIdx.x - plane * tpp;
    const int seg = trem / a.tiles_x, strip = trem - seg * a.tiles_x;
    const int x0 = strip * S2W, y0 = seg * a.seg_rows;
    const int rows_out = min(a.seg_rows, H - y0);
    const int nrows_in = rows_out + 2 * P;
    const long long plane_elems = (long long)H * W;
    const float *xb = a.x + (long long)plane * plane_elems;
    const float *yb = a.y + (long long)plane * plane_elems;

    float g[K];
#pragma unroll
    for (int j = 0; j < K; ++j) g[j] = __ldg(a.win + j);

    // ring slots of this thread: tid, S2T + tid, and 2 * S2T + tid for the right halo (tid < 2P)
    const bool has_ext = tid < 2 * P;
    int csrc[3];
    unsigned cbytes[3];
    const float *px[3], *py[3];
#pragma unroll
    for (int s = 0; s < 3; ++s) {
        const int gin = x0 - P + s * S2T + tid;
        const int c = ((s < 2 || has_ext) && gin <= W - 1 + P) ? pad_index(gin, W, d.pad_mode) : -1;
        csrc[s] = c;
        cbytes[s] = c >= 0 ? 4u : 0u;
        px[s] = xb + max(c, 0);
        py[s] = yb + max(c, 0);
    }
    const unsigned sxa = (unsigned)__cvta_generic_to_shared(&sX[0][tid]);
    const unsigned sya = (unsigned)__cvta_generic_to_shared(&sY[0][tid]);

    auto issue_row = [&](int i) {  // i is uniform over the CTA; one commit group per call
        if (i < nrows_in) {
            const int sy = pad_index(y0 - P + i, H, d.pad_mode);
            const long long ro = (long long)max(sy, 0) * W;
            const unsigned so = (unsigned)((i & (SD - 1)) * RW * 4);
            const unsigned ok = sy >= 0 ? 4u : 0u;
#pragma unroll
            for (int s = 0; s < 2; ++s) {
                cp_async4(sxa + so + s * S2T * 4, px[s] + ro, min(cbytes[s], ok));
                cp_async4(sya + so + s * S2T * 4, py[s] + ro, min(cbytes[s], ok));
            }
            if (has_ext) {
                cp_async4(sxa + so + 2 * S2T * 4, px[2] + ro, min(cbytes[2], ok));
                cp_async4(sya + so + 2 * S2T * 4, py[2] + ro, min(cbytes[2], ok));
            }
        }
        cp_async_commit();
    };

    const int gx = x0 + 2 * tid;  // first of the two output columns
    const int lo_y = d.interior ? P : 0, hi_y = d.interior ? H - P : H;
    bool col_ok[2];
#pragma unroll
    for (int c = 0; c < 2; ++c) col_ok[c] = gx + c < W && (!d.interior || (gx + c >= P && gx + c < W - P));
    float acc[M][2][K];
#pragma unroll
    for (int q = 0; q < M; ++q)
#pragma unroll
        for (int c = 0; c < 2; ++c)
#pragma unroll
            for (int j = 0; j < K; ++j) acc[q][c][j] = 0.f;
    float facc = 0.f, chk_x = 0.f, chk_y = 0.f;
    float facc2 = 0.f;  // DUAL: the second (interior-crop) policy, see llie_ssim_stats_dual
    const long long cstride = (long long)d.n * a.co * plane_elems;
    long long off = ((long long)plane * H + y0) * W + gx;  // COEF: first element of the next completed row
    const bool vec2 = (W & 1) == 0;                         // even rows keep the column pair 8-byte aligned

#pragma unroll
    for (int i = 0; i < SD - 1; ++i) issue_row(i);
    for (int base = 0; base < nrows_in; base += K) {
#pragma unroll
        for (int ii = 0; ii < K; ++ii) {
            const int i = base + ii;
            if (i >= nrows_in) break;  // uniform
            float *rx = sX[i & (SD - 1)], *ry = sY[i & (SD - 1)];
            cp_async_wait<SD - 2>();  // this thread's copies of row i have landed
#pragma unroll
            for (int s = 0; s < 3; ++s) {
                if (s == 2 && !has_ext) break;
                const float vx = rx[s * S2T + tid], vy = ry[s * S2T + tid];
                chk_x = fmaf(vx, 0.f, chk_x);
                chk_y = fmaf(vy, 0.f, chk_y);
                if (d.clamp_in) {
                    rx[s * S2T + tid] = fminf(fmaxf(vx, 0.f), 1.f);
                    ry[s * S2T + tid] = fminf(fmaxf(vy, 0.f), 1.f);
                }
            }
            __syncthreads();        // row i visible to all; everyone is done with row i - 1
            issue_row(i + SD - 1);  // refills the slot row i - 1 occupied
            float h[M][2];
#pragma unroll
            for (int q = 0; q < M; ++q) h[q][0] = h[q][1] = 0.f;
            const float2 *rx2 = reinterpret_cast<const float2 *>(rx) + tid;
            const float2 *ry2 = reinterpret_cast<const float2 *>(ry) + tid;
#pragma unroll
            for (int k = 0; k < (K + 1) / 2; ++k) {
                const float2 xv = rx2[k], yv = ry2[k];
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int p = 2 * k + e;  // position p feeds column 0 with tap p and column 1 with tap p - 1
                    const float xe = e ? xv.y : xv.x, ye = e ? yv.y : yv.x;
                    float m[M];
                    m[0] = xe;
                    m[1] = ye;
                    if (MAPS5) {
                        m[2] = xe * xe;
                        m[3] = ye * ye;
                        m[4] = xe * ye;
                    } else {
                        m[2] = fmaf(xe, xe, ye * ye);
                        m[3] = xe * ye;
                    }
#pragma unroll
                    for (int q = 0; q < M; ++q) {
                        if (p <= K - 1) h[q][0] = fmaf(g[p <= K - 1 ? p : 0], m[q], h[q][0]);
                        if (p >= 1) h[q][1] = fmaf(g[p >= 1 ? p - 1 : 0], m[q], h[q][1]);
                    }
                }
            }
            // input row i feeds output rows o = i - j with tap j; slot(o) = o mod K = (ii - j) mod K
#pragma unroll
            for (int j = 0; j < K; ++j) {
                const int sl = (ii - j + K) % K;
#pragma unroll
                for (int q = 0; q < M; ++q)
#pragma unroll
                    for (int c = 0; c < 2; ++c)
                        acc[q][c][sl] = (j == 0) ? g[0] * h[q][c] : fmaf(g[j], h[q][c], acc[q][c][sl]);
            }
            if (i >= K - 1) {  // output row o = i - (K - 1) is complete
                const int sl = (ii + 1) % K;
                const int gy = y0 + i - (K - 1);
                const bool row_ok = gy >= lo_y && gy < hi_y;
                float cf[3][2];
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    const float mx = acc[0][c][sl], my = acc[1][c][sl];
                    const float mx2 = mx * mx, my2 = my * my, mxy = mx * my;
                    float vsum, sxy;
                    if (MAPS5) {
                        float sx = acc[2][c][sl] - mx2, sy = acc[3][c][sl] - my2;
                        if (d.clamp_var) {
                            sx = fmaxf(sx, 0.f);
                            sy = fmaxf(sy, 0.f);
                        }
                        if (DUAL) {  // same closed form as ssim_ring_kernel<.., DUAL>: M' = r M, var' = r (var - eps mu^2)
                            const float su = (acc[2][c][sl] - mx2) + (acc[3][c][sl] - my2), B = mx2 + my2;
                            const float cxy = acc[4][c][sl] - mxy;
                            const float n1b = fmaf(mxy, a.dual_2r2, a.dual_c1), d1b = fmaf(B, a.dual_r2, a.dual_c1);
                            const float n2b = fmaf(fmaf(mxy, a.dual_neg_eps, cxy), a.dual_2r, a.dual_c2);
                            const float d2b = fmaf(fmaf(B, a.dual_neg_eps, su), a.dual_r, a.dual_c2);
                            const float s2 = (n1b * n2b) / (d1b * d2b);
                            if (gx + c < W && gx + c >= P && gx + c < W - P && gy >= P && gy < H - P) facc2 += s2;
                        }
                        vsum = sx + sy;
                        sxy = acc[4][c][sl] - mxy;
                    } else {
                        vsum = (acc[2][c][sl] - mx2) - my2;
                        sxy = acc[3][c][sl] - mxy;
                    }
                    const float num = (2.f * mxy + d.c1) * (2.f * sxy + d.c2);
                    const float den = (mx2 + my2 + d.c1) * (vsum + d.c2) + d.eps;
                    float sv = num / den;
                    if (COEF) {  // dLoss/dS times the sensitivities of S to mu_x, E[x^2], E[xy] (ssim_stream_kernel)
                        const float l = (1.f - sv) * 0.5f;
                        const float up =
                            (!d.loss_map) ? a.coef_scale : ((l >= 0.f && l <= 1.f) ? -0.5f * a.coef_scale : 0.f);
                        const float inv = 1.f / den;
                        const float n1 = 2.f * mxy + d.c1, n2 = 2.f * sxy + d.c2, d1 = mx2 + my2 + d.c1,
                                    d2 = vsum + d.c2;
                        const float a3 = 2.f * n1 * inv;
                        const float a2 = -sv * d1 * inv;
                        const float dmu = 2.f * (my * n2 - sv * mx * d2) * inv;
                        const float a1 = dmu - 2.f * mx * a2 - my * a3;
                        cf[0][c] = up * a1;
                        cf[1][c] = up * a2;
                        cf[2][c] = up * a3;
                    }
                    if (d.loss_map) sv = fminf(fmaxf((1.f - sv) * 0.5f, 0.f), 1.f);
                    if (col_ok[c] && row_ok) facc += sv;
                }
                if (COEF) {
                    float *cp = a.coef + off;
#pragma unroll
                    for (int q = 0; q < 3; ++q) {
                        if (vec2 && gx + 1 < W) {
                            *reinterpret_cast<float2 *>(cp + q * cstride) = make_float2(cf[q][0], cf[q][1]);
                        } else {
                            if (gx < W) cp[q * cstride] = cf[q][0];
                            if (gx + 1 < W) cp[q * cstride + 1] = cf[q][1];
                        }
                    }
                    off += W;
                }
            }
        }
    }
    cp_async_wait<0>();
    const double bs = block_sum<double>((double)facc, red);
    if (DUAL) {
        const double bs2 = block_sum<double>((double)facc2, red);
        if (tid == 0) a.part2[blockIdx.x] = bs2;
    }
    int flag = 0;
    if (!(chk_x == 0.f)) flag |= 1;
    if (!(chk_y == 0.f)) flag |= 2;
    const int f = block_or(flag, reinterpret_cast<int *>(red));
    if (tid == 0) {
        a.part[blockIdx.x] = bs;
        a.pflag[blockIdx.x] = f;
    }
    if (!last_block_done(a.ticket, gridDim.x)) return;
    const int lane = tid & 31, wrp = tid >> 5;
    const int planes = d.n * a.co;
    if (DUAL) {
        for (int pl = wrp; pl < planes; pl += S2T / 32) {
            const double *p2 = a.part2 + (size_t)pl * tpp;
            double t0 = 0.0;
            for (int b2 = lane; b2 < tpp; b2 += 32) t0 += __ldcg(p2 + b2);
            t0 = warp_sum(t0);
            if (lane == 0) a.out_sum2[pl] = t0;
        }
    }
    for (int pl = wrp; pl < planes; pl += S2T / 32) {
        const double *p = a.part + (size_t)pl * tpp;
        double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
        int b = lane;
        for (; b + 96 < tpp; b += 128) {
            s0 += __ldcg(p + b);
            s1 += __ldcg(p + b + 32);
            s2 += __ldcg(p + b + 64);
            s3 += __ldcg(p + b + 96);
        }
        for (; b < tpp; b += 32) s0 += __ldcg(p + b);
        const double sm_ = warp_sum((s0 + s1) + (s2 + s3));
        if (lane == 0) a.out_sum[pl] = sm_;
    }
    int fl = 0;
    for (int b = tid; b < (int)gridDim.x; b += S2T) fl |= __ldcg(a.pflag + b);
    const int ff = block_or(fl, reinterpret_cast<int *>(red));
    if (tid == 0) *a.out_flags = ff;
}

template <int K>
static int launch_ssim_stream2(const SsimArgs &a, int grid, cudaStream_t st) {
    if (a.coef != nullptr)
        ssim_stream2_kernel<K, false, true><<<grid, S2T, 0, st>>>(a);
    else if (a.d.clamp_var)
        ssim_stream2_kernel<K, true><<<grid, S2T, 0, st>>>(a);
    else
        ssim_stream2_kernel<K, false><<<grid, S2T, 0, st>>>(a);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

static std::atomic<int> g_ssim_path_a{0};      // 0 auto: two-column streaming kernel where eligible, else streaming | 1 tile kernel | 2 one-column streaming
static std::atomic<int> g_ssim_seg_rows_a{0};  // 0 = heuristic

// ---------------------------------------------------------------------------------------------------
// Backward of the (D)SSIM sum w.r.t. x (kornia policy: reflect border, "same" size):
//   dL/dx(q) = mask(q) * g_up * [ (G^T c1)(q) + 2 x(q) (G^T c2)(q) + y(q) (G^T c3)(q) ]
// where c1..c3 are the coefficient maps the forward kernel stored and G^T is the adjoint of
// reflect-pad o separable correlation: a zero-extended correlation t(j) on j in [-P, W+P) whose out-of-image
// taps fold back, out(q) = t(q) + [1 <= q <= P] t(-q) + [W-1-P <= q <= W-2] t(2W-2-q), per axis.
// Same tiling as the forward kernel: (A) coefficient tile with halo (zeros outside the image) -> smem,
// (B) horizontal pass + column folds, (C) vertical pass + row folds, combine with x, y (8 B/elem read).

__device__ __forceinline__ float fold_extra(int P, const float *line, int stride, int q, int o0, int L,
                                            const float *sg) {
    // line[(p - o0) * stride] holds the zero-extended sequence at image coordinate p
    float e = 0.f;
    if (q >= 1 && q <= P)
        for (int p = 0; p <= P - q && p < L; ++p) e = fmaf(sg[p + q + P], line[(p - o0) * stride], e);
    if (q >= L - 1 - P && q <= L - 2) {
        const int qq = 2 * L - 2 - q;
        for (int p = max(qq - P, 0); p <= L - 1; ++p) e = fmaf(sg[p - qq + P], line[(p - o0) * stride], e);
    }
    return e;
}

template <int K>
__global__ void __launch_bounds__(SNT, K ? 2 : 1) ssim_loss_bwd_kernel(const SsimBwdArgs a) {
    const int KK = K ? K : a.ksize;   // K == 0: runtime window (see ssim_stats_kernel)
    const int P = KK / 2;
    const int XR = STH + 2 * P;
    const int XP = (STW + 2 * P) | 1;
    constexpr int HP = STW + 1;
    extern __shared__ __align__(16) float sm[];
    float *sC = sm;                   // [3][XR][XP]
    float *sH = sC + 3 * XR * XP;     // [3][XR][HP]
    __shared__ float sg[K ? K : LLIE_SSIM_MAX_K];
    const int H = a.h, W = a.w, tid = threadIdx.x;
    const int tpp = a.tiles_x * a.tiles_y;
    const int plane = blockIdx.x / tpp;
    const int trem = blockIdx.x - plane * tpp;
    const int ty = trem / a.tiles_x, tx = trem - ty * a.tiles_x;
    const int y0 = ty * STH, x0 = tx * STW;
    const long long plane_elems = (long long)H * W;
    const long long stride = (long long)a.n * a.c * plane_elems;
    const float *cb = a.coef + (long long)plane * plane_elems;
    const float *gm = a.gout_is_map ? a.gout + (long long)plane * plane_elems : nullptr;
    const float gs = a.gout_is_map ? 1.f : __ldg(a.gout);

    float g[K ? K : 1];
#pragma unroll
    for (int j = 0; j < K; ++j) g[j] = __ldg(a.win + j);
    if (tid < KK) sg[tid] = __ldg(a.win + tid);

    // ---- (A) coefficient tiles, zero outside the image
    for (int idx = tid; idx < XR * (STW + 2 * P); idx += SNT) {
        const int r = idx / (STW + 2 * P), cc = idx - r * (STW + 2 * P);
        const int sy = y0 - P + r, sx = x0 - P + cc;
        float v1 = 0.f, v2 = 0.f, v3 = 0.f;
        if (sy >= 0 && sy < H && sx >= 0 && sx < W) {
            const long long off = (long long)sy * W + sx;
            const float up = gm ? __ldg(gm + off) : gs;
            v1 = up * __ldg(cb + off);
            v2 = up * __ldg(cb + stride + off);
            v3 = up * __ldg(cb + 2 * stride + off);
        }
        sC[r * XP + cc] = v1;
        sC[XR * XP + r * XP + cc] = v2;
        sC[2 * XR * XP + r * XP + cc] = v3;
    }
    __syncthreads();

    // ---- (B) horizontal pass (+ column folds on the tiles that touch the left / right image border)
    const bool fold_x = (x0 <= P) || (x0 + STW > W - 1 - P);
    for (int t = tid; t < XR * (STW / 4); t += SNT) {
        const int cg = t / XR, r = t - cg * XR;
        float acc[3][4];
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            const float *px = sC + q * XR * XP + r * XP + cg * 4;
#pragma unroll
            for (int o = 0; o < 4; ++o) acc[q][o] = 0.f;
#pragma unroll(K ? K + 3 : 1)
            for (int u = 0; u < KK + 3; ++u) {
                const float v = px[u];
#pragma unroll
                for (int o = 0; o < 4; ++o) {
                    const int j = u - o;
                    if (j >= 0 && j < KK) acc[q][o] = fmaf(SSIM_TAP(j), v, acc[q][o]);
                }
            }
        }
        if (fold_x) {
#pragma unroll
            for (int o = 0; o < 4; ++o) {
                const int qx = x0 + cg * 4 + o;
                if (qx < W && (qx <= P || qx >= W - 1 - P)) {
#pragma unroll
                    for (int q = 0; q < 3; ++q)
                        acc[q][o] += fold_extra(P, sC + q * XR * XP + r * XP, 1, qx, x0 - P, W, sg);
                }
            }
        }
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            float *h = sH + q * XR * HP + r * HP + cg * 4;
#pragma unroll
            for (int o = 0; o < 4; ++o) h[o] = acc[q][o];
        }
    }
    __syncthreads();

    // ---- (C) vertical pass (+ row folds), combine
    const int col = tid & (STW - 1), rg = tid / STW;
    const int gx = x0 + col;
    float m[3][8];
#pragma unroll
    for (int q = 0; q < 3; ++q) {
        const float *h = sH + q * XR * HP + (rg * 8) * HP + col;
#pragma unroll
        for (int o = 0; o < 8; ++o) m[q][o] = 0.f;
#pragma unroll(K ? K + 7 : 1)
        for (int u = 0; u < KK + 7; ++u) {
            const float v = h[u * HP];
#pragma unroll
            for (int o = 0; o < 8; ++o) {
                const int j = u - o;
                if (j >= 0 && j < KK) m[q][o] = fmaf(SSIM_TAP(j), v, m[q][o]);
            }
        }
    }
    const bool fold_y = (y0 <= P) || (y0 + STH > H - 1 - P);
    if (gx >= W) return;
    const float *xb = a.x + (long long)plane * plane_elems, *yb = a.y + (long long)plane * plane_elems;
    float *ob = a.out_grad + (long long)plane * plane_elems;
#pragma unroll
    for (int o = 0; o < 8; ++o) {
        const int gy = y0 + rg * 8 + o;
        if (gy >= H) break;
        float g1 = m[0][o], g2 = m[1][o], g3 = m[2][o];
        if (fold_y && (gy <= P || gy >= H - 1 - P)) {
            g1 += fold_extra(P, sH + col, HP, gy, y0 - P, H, sg);
            g2 += fold_extra(P, sH + XR * HP + col, HP, gy, y0 - P, H, sg);
            g3 += fold_extra(P, sH + 2 * XR * HP + col, HP, gy, y0 - P, H, sg);
        }
        const long long off = (long long)gy * W + gx;
        const float xr = __ldg(xb + off);
        float xv = xr, yv = __ldg(yb + off);
        bool pass = true;
        if (a.clamp_in) {
            pass = (xr >= 0.f && xr <= 1.f);
            xv = fminf(fmaxf(xv, 0.f), 1.f);
            yv = fminf(fmaxf(yv, 0.f), 1.f);
        }
        const float gr = g1 + 2.f * xv * g2 + yv * g3;
        const float res = pass ? gr : 0.f;
        ob[off] = a.accumulate ? ob[off] + res : res;
    }
}

// Streaming backward (default when the upstream gradient is a scalar): same structure as ssim_stream_kernel —
// one thread per column of a 128-column strip, coefficient rows (zero outside the image) streamed through a
// cp.async ring, K-tap horizontal pass from shared memory, K rolling vertical accumulators per map in registers.
// The mirror folds of the reflect adjoint are folded into the WEIGHTS: a border column adds the taps of its
// mirrored partner to its own per-thread horizontal taps (computed once), and the vertical scatter weight of (input row p, output row y) gains
// g[p + y + P] at the top (1 <= y <= P, p + y <= P) and g[p - (2H - 2 - y) + P] at the bottom; those tests are
// uniform over the CTA and only compiled into the path taken by segments that touch the top / bottom rows.
template <int K>
__global__ void __launch_bounds__(SW) ssim_bwd_stream_kernel(const SsimBwdArgs a, int seg_rows) {
    constexpr int P = K / 2;
    constexpr int RW = SW + 2 * P;
    __shared__ float sC[3][SD][RW];
    __shared__ float sg[K];
    const int H = a.h, W = a.w, tid = threadIdx.x;
    const int tpp = a.tiles_x * a.tiles_y;
    const int plane = blockIdx.x / tpp;
    const int trem = blockIdx.x - plane * tpp;
    const int seg = trem / a.tiles_x, strip = trem - seg * a.tiles_x;
    const int x0 = strip * SW, y0 = seg * seg_rows;
    const int rows_out = min(seg_rows, H - y0);
    const int nrows_in = rows_out + 2 * P;
    const long long plane_elems = (long long)H * W;
    const long long cstride = (long long)a.n * a.c * plane_elems;
    const float *cb = a.coef + (long long)plane * plane_elems;
    const float gs = __ldg(a.gout);

    float g[K];
#pragma unroll
    for (int j = 0; j < K; ++j) g[j] = __ldg(a.win + j);
    if (tid < K) sg[tid] = __ldg(a.win + tid);

    const int gin0 = x0 - P + tid, gin1 = x0 - P + SW + tid;
    const bool has_ext = tid < 2 * P;
    const unsigned bm = (gin0 >= 0 && gin0 < W) ? 4u : 0u;
    const unsigned be = (has_ext && gin1 >= 0 && gin1 < W) ? 4u : 0u;
    const float *pm = cb + min(max(gin0, 0), W - 1), *pe = cb + min(max(gin1, 0), W - 1);
    const unsigned sca = (unsigned)__cvta_generic_to_shared(&sC[0][0][tid]);

    auto issue_row = [&](int i) {
        if (i < nrows_in) {
            const int sy = y0 - P + i;
            const unsigned ok = (sy >= 0 && sy < H) ? 4u : 0u;
            const long long ro = (long long)min(max(sy, 0), H - 1) * W;
            const unsigned so = (unsigned)((i & (SD - 1)) * RW * 4);
#pragma unroll
            for (int m = 0; m < 3; ++m) {
                cp_async4(sca + m * SD * RW * 4 + so, pm + m * cstride + ro, min(bm, ok));
                if (has_ext) cp_async4(sca + m * SD * RW * 4 + so + SW * 4, pe + m * cstride + ro, min(be, ok));
            }
        }
        cp_async_commit();
    };

    const int gx = x0 + tid;
    const bool col_in = gx < W;
    const bool seg_fold = (y0 <= P) || (y0 + rows_out > H - 1 - P);
    const float *xb = a.x + (long long)plane * plane_elems + gx, *yb = a.y + (long long)plane * plane_elems + gx;
    float *ob = a.out_grad + (long long)plane * plane_elems + gx;
    float acc[3][K];
#pragma unroll
    for (int m = 0; m < 3; ++m)
#pragma unroll
        for (int j = 0; j < K; ++j) acc[m][j] = 0.f;

#pragma unroll
    for (int i = 0; i < SD - 1; ++i) issue_row(i);
    __syncthreads();  // sg visible
    // horizontal taps of THIS column with the mirror folds added: tap j reads image column x = gx - P + j; a left
    // border column q in [1, P] also collects t(-q), i.e. + g[j + 2q]; a right border column + g[j + 2q - 2W + 2]
    float gh[K];
#pragma unroll
    for (int j = 0; j < K; ++j) {
        float wj = g[j];
        const int x = gx - P + j;
        if (col_in && gx >= 1 && gx <= P && x >= 0 && j + 2 * gx <= 2 * P) wj += sg[j + 2 * gx];
        if (col_in && gx >= W - 1 - P && gx <= W - 2 && x <= W - 1 && j + 2 * gx - 2 * W + 2 >= 0)
            wj += sg[j + 2 * gx - 2 * W + 2];
        gh[j] = wj;
    }
    for (int base = 0; base < nrows_in; base += K) {
#pragma unroll
        for (int ii = 0; ii < K; ++ii) {
            const int i = base + ii;
            if (i >= nrows_in) break;  // uniform
            const int slot = i & (SD - 1);
            cp_async_wait<SD - 2>();
            __syncthreads();
            issue_row(i + SD - 1);
            const int p = y0 - P + i;
            const bool vfold = seg_fold && (p < P || p >= H - P);  // uniform; only <= 2P rows of the image
            // x, y of the output row this iteration completes (latency overlaps the arithmetic below)
            const int oy = y0 + i - (K - 1);
            float xr = 0.f, yr = 0.f;
            if (i >= K - 1 && col_in) {
                xr = __ldg(xb + (long long)oy * W);
                yr = __ldg(yb + (long long)oy * W);
            }
            float h[3];
#pragma unroll
            for (int m = 0; m < 3; ++m) {
                const float *row = &sC[m][slot][0];
                float v = 0.f;
#pragma unroll
                for (int j = 0; j < K; ++j) v = fmaf(gh[j], row[tid + j], v);
                h[m] = v;
            }
            if (!vfold) {
#pragma unroll
                for (int j = 0; j < K; ++j) {
                    const int sl = (ii - j + K) % K;
#pragma unroll
                    for (int m = 0; m < 3; ++m) acc[m][sl] = (j == 0) ? g[0] * h[m] : fmaf(g[j], h[m], acc[m][sl]);
                }
            } else {
#pragma unroll
                for (int j = 0; j < K; ++j) {
                    const int sl = (ii - j + K) % K;
                    const int y = y0 + i - j;
                    float wj = g[j];
                    if (y >= 1 && y <= P && p >= 0 && p + y <= P) wj += sg[p + y + P];
                    if (y >= H - 1 - P && y <= H - 2 && p <= H - 1 && p >= 2 * H - 2 - y - P)
                        wj += sg[p - (2 * H - 2 - y) + P];
#pragma unroll
                    for (int m = 0; m < 3; ++m) acc[m][sl] = (j == 0) ? wj * h[m] : fmaf(wj, h[m], acc[m][sl]);
                }
            }
            if (i >= K - 1 && col_in) {
                const int sl = (ii + 1) % K;
                float xv = xr, yv = yr;
                bool pass = true;
                if (a.clamp_in) {
                    pass = (xr >= 0.f && xr <= 1.f);
                    xv = fminf(fmaxf(xv, 0.f), 1.f);
                    yv = fminf(fmaxf(yv, 0.f), 1.f);
                }
                const float gr = gs * (acc[0][sl] + 2.f * xv * acc[1][sl] + yv * acc[2][sl]);
                const float res = pass ? gr : 0.f;
                float *o = ob + (long long)oy * W;
                *o = a.accumulate ? *o + res : res;
            }
        }
    }
    cp_async_wait<0>();
}

template <int K>
static int launch_ssim_bwd_stream(const SsimBwdArgs &a, int grid, int seg_rows, cudaStream_t st) {
    ssim_bwd_stream_kernel<K><<<grid, SW, 0, st>>>(a, seg_rows);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

// Two-column streaming backward (default for the 11-tap window): the structure of ssim_bwd_stream_kernel with the
// column pairing of ssim_stream2_kernel -- 256-column strips, one thread = two adjacent gradient columns, the K + 1
// coefficient values per row and plane from (K + 1) / 2 LDS.64, per-row bookkeeping paid per two elements.  Each of the
// two columns carries its own horizontal taps (mirror folds added), the vertical fold weights are shared.
template <int K>
__global__ void __launch_bounds__(S2T, 3) ssim_bwd_stream2_kernel(const SsimBwdArgs a, int seg_rows) {
    constexpr int P = K / 2;
    constexpr int RW = S2W + 2 * P;
    static_assert(2 * P <= S2T && (RW % 2) == 0, "halo slots are loaded by the first 2P threads");
    __shared__ __align__(16) float sC[3][SD][RW];
    __shared__ __align__(16) float sIO[3][SD][S2W];  // x, y and (accumulate) the old gradient of the row a ring entry completes
    __shared__ float sg[K];
    const int H = a.h, W = a.w, tid = threadIdx.x;
    const int tpp = a.tiles_x * a.tiles_y;
    const int plane = blockIdx.x / tpp;
    const int trem = blockIdx.x - plane * tpp;
    const int seg = trem / a.tiles_x, strip = trem - seg * a.tiles_x;
    const int x0 = strip * S2W, y0 = seg * seg_rows;
    const int rows_out = min(seg_rows, H - y0);
    const int nrows_in = rows_out + 2 * P;
    const long long plane_elems = (long long)H * W;
    const long long cstride = (long long)a.n * a.c * plane_elems;
    const float *cb = a.coef + (long long)plane * plane_elems;
    const float gs = __ldg(a.gout);

    float g[K];
#pragma unroll
    for (int j = 0; j < K; ++j) g[j] = __ldg(a.win + j);
    if (tid < K) sg[tid] = __ldg(a.win + tid);

    // ring slots of this thread: tid, S2T + tid, 2 * S2T + tid (right halo, tid < 2P); zero outside the image
    const bool has_ext = tid < 2 * P;
    unsigned cbytes[3];
    const float *pc[3];
#pragma unroll
    for (int s = 0; s < 3; ++s) {
        const int gin = x0 - P + s * S2T + tid;
        cbytes[s] = ((s < 2 || has_ext) && gin >= 0 && gin < W) ? 4u : 0u;
        pc[s] = cb + min(max(gin, 0), W - 1);
    }
    const unsigned sca = (unsigned)__cvta_generic_to_shared(&sC[0][0][tid]);
    const unsigned sio = (unsigned)__cvta_generic_to_shared(&sIO[0][0][2 * tid]);
    const int gx = x0 + 2 * tid;
    const bool in0 = gx < W, in1 = gx + 1 < W;
    const bool vec2 = (W & 1) == 0;  // even rows keep the column pair 8-byte aligned
    const float *xb = a.x + (long long)plane * plane_elems + gx, *yb = a.y + (long long)plane * plane_elems + gx;
    float *ob = a.out_grad + (long long)plane * plane_elems + gx;
    const float *iosrc[3] = {xb, yb, ob};
    const int nio = a.accumulate ? 3 : 2;

    auto issue_row = [&](int i) {
        if (i < nrows_in) {
            if (i >= K - 1) {
                // x / y / old gradient of output row oy = y0 + i - (K - 1), which ring entry i completes: they travel
                // with the coefficient row (three rows ahead) instead of stalling the row on global loads
                const long long oo = (long long)(y0 + i - (K - 1)) * W;
                const unsigned so2 = (unsigned)((i & (SD - 1)) * S2W * 4);
                for (int m = 0; m < nio; ++m) {
                    const unsigned dst = sio + m * SD * S2W * 4 + so2;
                    if (vec2) {
                        asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(dst),
                                     "l"(in0 ? iosrc[m] + oo : a.x), "r"(in0 ? 8u : 0u)
                                     : "memory");
                    } else {
                        cp_async4(dst, in0 ? iosrc[m] + oo : a.x, in0 ? 4u : 0u);
                        cp_async4(dst + 4, in1 ? iosrc[m] + oo + 1 : a.x, in1 ? 4u : 0u);
                    }
                }
            }
            const int sy = y0 - P + i;
            const unsigned ok = (sy >= 0 && sy < H) ? 4u : 0u;
            const long long ro = (long long)min(max(sy, 0), H - 1) * W;
            const unsigned so = (unsigned)((i & (SD - 1)) * RW * 4);
#pragma unroll
            for (int m = 0; m < 3; ++m) {
#pragma unroll
                for (int s = 0; s < 2; ++s)
                    cp_async4(sca + m * SD * RW * 4 + so + s * S2T * 4, pc[s] + m * cstride + ro, min(cbytes[s], ok));
                if (has_ext)
                    cp_async4(sca + m * SD * RW * 4 + so + 2 * S2T * 4, pc[2] + m * cstride + ro, min(cbytes[2], ok));
            }
        }
        cp_async_commit();
    };

    const bool seg_fold = (y0 <= P) || (y0 + rows_out > H - 1 - P);
    float acc[3][2][K];
#pragma unroll
    for (int m = 0; m < 3; ++m)
#pragma unroll
        for (int c = 0; c < 2; ++c)
#pragma unroll
            for (int j = 0; j < K; ++j) acc[m][c][j] = 0.f;

#pragma unroll
    for (int i = 0; i < SD - 1; ++i) issue_row(i);
    __syncthreads();  // sg visible
    // horizontal taps of the two columns with the mirror folds added (see ssim_bwd_stream_kernel)
    float gh[2][K];
#pragma unroll
    for (int c = 0; c < 2; ++c) {
        const int gc = gx + c;
        const bool cin = gc < W;
#pragma unroll
        for (int j = 0; j < K; ++j) {
            float wj = g[j];
            const int x = gc - P + j;
            if (cin && gc >= 1 && gc <= P && x >= 0 && j + 2 * gc <= 2 * P) wj += sg[j + 2 * gc];
            if (cin && gc >= W - 1 - P && gc <= W - 2 && x <= W - 1 && j + 2 * gc - 2 * W + 2 >= 0)
                wj += sg[j + 2 * gc - 2 * W + 2];
            gh[c][j] = wj;
        }
    }
    for (int base = 0; base < nrows_in; base += K) {
#pragma unroll
        for (int ii = 0; ii < K; ++ii) {
            const int i = base + ii;
            if (i >= nrows_in) break;  // uniform
            const int slot = i & (SD - 1);
            cp_async_wait<SD - 2>();
            __syncthreads();
            issue_row(i + SD - 1);
            const int p = y0 - P + i;
            const bool vfold = seg_fold && (p < P || p >= H - P);  // uniform; only <= 2P rows of the image
            const int oy = y0 + i - (K - 1);
            // x, y (and the old gradient) of the output row this iteration completes: this thread's own ring words
            const float2 xq = *reinterpret_cast<const float2 *>(&sIO[0][slot][2 * tid]);
            const float2 yq = *reinterpret_cast<const float2 *>(&sIO[1][slot][2 * tid]);
            const float2 oq = a.accumulate ? *reinterpret_cast<const float2 *>(&sIO[2][slot][2 * tid])
                                           : make_float2(0.f, 0.f);
            const float xr[2] = {xq.x, xq.y}, yr[2] = {yq.x, yq.y};
            float h[3][2];
#pragma unroll
            for (int m = 0; m < 3; ++m) {
                const float2 *row2 = reinterpret_cast<const float2 *>(&sC[m][slot][0]) + tid;
                float v0 = 0.f, v1 = 0.f;
#pragma unroll
                for (int k = 0; k < (K + 1) / 2; ++k) {
                    const float2 cv = row2[k];
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const int q = 2 * k + e;  // position q: tap q of column 0, tap q - 1 of column 1
                        const float ce = e ? cv.y : cv.x;
                        if (q <= K - 1) v0 = fmaf(gh[0][q <= K - 1 ? q : 0], ce, v0);
                        if (q >= 1) v1 = fmaf(gh[1][q >= 1 ? q - 1 : 0], ce, v1);
                    }
                }
                h[m][0] = v0;
                h[m][1] = v1;
            }
            if (!vfold) {
#pragma unroll
                for (int j = 0; j < K; ++j) {
                    const int sl = (ii - j + K) % K;
#pragma unroll
                    for (int m = 0; m < 3; ++m)
#pragma unroll
                        for (int c = 0; c < 2; ++c)
                            acc[m][c][sl] = (j == 0) ? g[0] * h[m][c] : fmaf(g[j], h[m][c], acc[m][c][sl]);
                }
            } else {
#pragma unroll
                for (int j = 0; j < K; ++j) {
                    const int sl = (ii - j + K) % K;
                    const int y = y0 + i - j;
                    float wj = g[j];
                    if (y >= 1 && y <= P && p >= 0 && p + y <= P) wj += sg[p + y + P];
                    if (y >= H - 1 - P && y <= H - 2 && p <= H - 1 && p >= 2 * H - 2 - y - P)
                        wj += sg[p - (2 * H - 2 - y) + P];
#pragma unroll
                    for (int m = 0; m < 3; ++m)
#pragma unroll
                        for (int c = 0; c < 2; ++c)
                            acc[m][c][sl] = (j == 0) ? wj * h[m][c] : fmaf(wj, h[m][c], acc[m][c][sl]);
                }
            }
            if (i >= K - 1 && in0) {
                const int sl = (ii + 1) % K;
                float res[2];
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    float xv = xr[c], yv = yr[c];
                    bool pass = true;
                    if (a.clamp_in) {
                        pass = (xr[c] >= 0.f && xr[c] <= 1.f);
                        xv = fminf(fmaxf(xv, 0.f), 1.f);
                        yv = fminf(fmaxf(yv, 0.f), 1.f);
                    }
                    const float gr = gs * (acc[0][c][sl] + 2.f * xv * acc[1][c][sl] + yv * acc[2][c][sl]);
                    res[c] = pass ? gr : 0.f;
                }
                float *o = ob + (long long)oy * W;
                if (vec2) {
                    *reinterpret_cast<float2 *>(o) = make_float2(oq.x + res[0], oq.y + res[1]);
                } else {
                    o[0] = oq.x + res[0];
                    if (in1) o[1] = oq.y + res[1];
                }
            }
        }
    }
    cp_async_wait<0>();
}

template <int K>
static int launch_ssim_bwd_stream2(const SsimBwdArgs &a, int grid, int seg_rows, cudaStream_t st) {
    ssim_bwd_stream2_kernel<K><<<grid, S2T, 0, st>>>(a, seg_rows);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

template <int K>
static int launch_ssim_bwd(const SsimBwdArgs &a, int grid, cudaStream_t st) {
    const int P = (K ? K : a.ksize) / 2, XR = STH + 2 * P, XP = (STW + 2 * P) | 1, HP = STW + 1;
    const size_t smem = (size_t)(3 * XR * XP + 3 * XR * HP) * sizeof(float);
    LLIE_CUDA_OK(cudaFuncSetAttribute(ssim_loss_bwd_kernel<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ssim_loss_bwd_kernel<K><<<grid, SNT, smem, st>>>(a);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

}  // namespace llie

using namespace llie;

// Output rows per CTA of the streaming kernels.  Every column strip is cut into the same number of row segments; the
// grid then runs in waves of `slots` resident CTAs, and a wave that is only partly filled costs a full segment time
// (e.g. 918 CTAs on 888 slots = 3 waves for 2.07 waves of work).  Pick, among 1..4 waves, the segment count whose
// (slot fill) x (useful rows / rows incl. the `halo` warm-up rows) is best, with at least `min_rows` rows per segment.
namespace llie {
int pick_seg_rows(long long cols, int h, int slots, int min_rows, int halo) {
    int best_rows = h;
    double best = -1.0;
    for (int waves = 1; waves <= 4; ++waves) {
        long long segs = (long long)waves * slots / cols;
        if (segs < 1) segs = 1;
        int rows = (int)((h + segs - 1) / segs);
        if (rows < min_rows) rows = min_rows < h ? min_rows : h;
        const long long tiles_y = (h + rows - 1) / rows;
        const long long grid = cols * tiles_y;
        const long long used_waves = (grid + slots - 1) / slots;
        const double eff = (double)grid / (double)(used_waves * slots) * rows / (rows + halo);
        if (eff > best + 1e-9) {
            best = eff;
            best_rows = rows;
        }
    }
    return best_rows;
}
}  // namespace llie

static int ssim_forward(const llie_ssim_desc *desc, const float *x, const float *y, const float *win1d,
                        const float *win1d_host, double *out_sum, float *out_map, float *coef, float coef_scale,
                        int32_t *out_flags, void *workspace, void *stream) {
    LLIE_REQUIRE(desc && x && y && win1d && out_sum && out_flags && workspace, "ssim_stats: null pointer");
    const llie_ssim_desc &d = *desc;
    LLIE_REQUIRE(d.n > 0 && d.c > 0 && d.h > 0 && d.w > 0, "ssim_stats: empty tensor");
    LLIE_REQUIRE(d.ksize >= 1 && (d.ksize & 1) && d.ksize <= LLIE_SSIM_MAX_K, "ssim_stats: window must be odd and <= %d",
                 LLIE_SSIM_MAX_K);
    LLIE_REQUIRE(d.ksize <= d.h && d.ksize <= d.w, "ssim_stats: window larger than the image");
    LLIE_REQUIRE(d.pad_mode >= LLIE_PAD_ZEROS && d.pad_mode <= LLIE_PAD_CIRCULAR, "ssim_stats: bad pad mode");
    LLIE_REQUIRE(!d.luma || d.c == 3, "ssim_stats: luma needs 3-channel inputs");
    LLIE_REQUIRE(!d.interior || (d.h > d.ksize - 1 && d.w > d.ksize - 1), "ssim_stats: interior crop leaves nothing");
    SsimArgs a{};
    a.d = d;
    a.x = x; a.y = y; a.win = win1d;
    a.out_sum = out_sum; a.out_map = out_map; a.out_flags = out_flags;
    a.coef = coef; a.coef_scale = coef_scale;
    a.co = d.luma ? 1 : d.c;
    Workspace ws(workspace);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const int g_ssim_path = g_ssim_path_a.load(std::memory_order_relaxed);  // diagnostic knobs, read once per call
    const int g_ssim_seg_rows = g_ssim_seg_rows_a.load(std::memory_order_relaxed);
    // default: TMA row ring + packed FP32 (ssim_ring.cu) for the 11-tap window of all three reference policies
    if (g_ssim_path == 0 && ssim_ring_eligible(d, x, y, win1d_host, out_map != nullptr, coef))
        return ssim_ring_forward(a, win1d_host, workspace, st);
    long long grid;
    const bool stream_ok = g_ssim_path != 1 && !d.luma && d.ksize <= 15;
    // statistics only (no map / coefficient planes), the 11-tap window of all three reference policies
    // (the loss-coefficient planes are written by the four-map instantiation: no variance clamp on that path)
    const bool stream2_ok = stream_ok && (g_ssim_path == 0 || g_ssim_path == 3) && d.ksize == 11 && out_map == nullptr &&
                            (coef == nullptr || !d.clamp_var);
    if (stream2_ok) {
        // strips of 256 columns x row segments sized for ~2 resident waves (148 SMs x 3 CTAs), >= 64 rows each
        a.tiles_x = div_up(d.w, S2W);
        const long long cols = (long long)d.n * a.co * a.tiles_x;
        const int rows = g_ssim_seg_rows > 0 ? (g_ssim_seg_rows < d.h ? g_ssim_seg_rows : d.h)
                                             : pick_seg_rows(cols, d.h, kNumSM * 3, 64, d.ksize - 1);
        a.seg_rows = rows;
        a.tiles_y = div_up(d.h, rows);
        grid = cols * a.tiles_y;
    } else if (stream_ok) {
        // strips of 128 columns x row segments sized for ~3 resident waves (148 SMs x 5 CTAs), >= 48 rows each
        a.tiles_x = div_up(d.w, SW);
        const long long cols = (long long)d.n * a.co * a.tiles_x;
        const int rows = g_ssim_seg_rows > 0 ? (g_ssim_seg_rows < d.h ? g_ssim_seg_rows : d.h)
                                             : pick_seg_rows(cols, d.h, kNumSM * 4, 48, d.ksize - 1);
        a.seg_rows = rows;
        a.tiles_y = div_up(d.h, rows);
        grid = cols * a.tiles_y;
    } else {
        a.tiles_x = div_up(d.w, STW);
        a.tiles_y = div_up(d.h, STH);
        grid = (long long)d.n * a.co * a.tiles_x * a.tiles_y;
    }
    LLIE_REQUIRE(grid <= 0x7fffffffLL && (size_t)grid * (sizeof(double) + sizeof(int)) <= ws.scratch_bytes,
                 "ssim_stats: workspace too small for %lld tiles", grid);
    a.part = reinterpret_cast<double *>(ws.scratch);
    a.pflag = reinterpret_cast<int *>(ws.scratch + (size_t)grid * sizeof(double));
    a.ticket = ws.tickets;
    if (stream2_ok) return launch_ssim_stream2<11>(a, (int)grid, st);
    if (stream_ok) {
        switch (d.ksize) {
            case 1: return launch_ssim_stream<1>(a, (int)grid, st);
            case 3: return launch_ssim_stream<3>(a, (int)grid, st);
            case 5: return launch_ssim_stream<5>(a, (int)grid, st);
            case 7: return launch_ssim_stream<7>(a, (int)grid, st);
            case 9: return launch_ssim_stream<9>(a, (int)grid, st);
            case 11: return launch_ssim_stream<11>(a, (int)grid, st);
            case 13: return launch_ssim_stream<13>(a, (int)grid, st);
            default: return launch_ssim_stream<15>(a, (int)grid, st);
        }
    }
    switch (d.ksize) {
        case 1: return launch_ssim<1>(a, (int)grid, st);
        case 3: return launch_ssim<3>(a, (int)grid, st);
        case 5: return launch_ssim<5>(a, (int)grid, st);
        case 7: return launch_ssim<7>(a, (int)grid, st);
        case 9: return launch_ssim<9>(a, (int)grid, st);
        case 11: return launch_ssim<11>(a, (int)grid, st);
        case 13: return launch_ssim<13>(a, (int)grid, st);
        case 15: return launch_ssim<15>(a, (int)grid, st);
        default: return launch_ssim<0>(a, (int)grid, st);  // 17..LLIE_SSIM_MAX_K taps: runtime-window instantiation
    }
}

namespace llie {
void ssim_tune(int key, int value) {
    if (key == LLIE_TUNE_SSIM_PATH) g_ssim_path_a.store(value, std::memory_order_relaxed);
    if (key == LLIE_TUNE_SSIM_SEG_ROWS) g_ssim_seg_rows_a.store(value, std::memory_order_relaxed);
}
}  // namespace llie

extern "C" int llie_ssim_stats(const llie_ssim_desc *desc, const float *x, const float *y, const float *win1d,
                               const float *win1d_host, double *out_sum, float *out_map, int32_t *out_flags,
                               void *workspace, void *stream) {
    return ssim_forward(desc, x, y, win1d, win1d_host, out_sum, out_map, nullptr, 0.f, out_flags, workspace, stream);
}

// pairs the ring kernel does not take (W % 4 != 0, unaligned planes, no host taps): the two-column cp.async kernel
static bool ssim_stream2_dual_eligible(const llie_ssim_desc &d) {
    return d.ksize == 11 && !d.luma && d.clamp_var && !d.clamp_in && !d.interior && !d.loss_map &&
           d.pad_mode == LLIE_PAD_REFLECT && d.h > 10 && d.w > 10;
}

extern "C" int llie_ssim_stats_dual_eligible(const llie_ssim_desc *desc, const float *x, const float *y,
                                            const float *win1d_host) {
    const int path = g_ssim_path_a.load(std::memory_order_relaxed);
    return desc != nullptr && ((path == 0 && ssim_ring_dual_eligible(*desc, x, y, win1d_host)) ||
                               ((path == 0 || path == 3) && ssim_stream2_dual_eligible(*desc)));
}

extern "C" int llie_ssim_stats_dual(const llie_ssim_desc *desc, float c1_b, float c2_b, double tap_sum_ratio,
                                    const float *x, const float *y, const float *win1d, const float *win1d_host,
                                    double *out_sum, double *out_sum_b, int32_t *out_flags, void *workspace,
                                    void *stream) {
    LLIE_REQUIRE(desc && x && y && win1d && out_sum && out_sum_b && out_flags && workspace, "ssim_stats_dual: null pointer");
    LLIE_REQUIRE(llie_ssim_stats_dual_eligible(desc, x, y, win1d_host),
                 "ssim_stats_dual: needs the 11-tap window, reflect padding, the variance clamp, no luma / input clamp / "
                 "interior flag and images larger than the window; make two llie_ssim_stats calls instead");
    LLIE_REQUIRE(tap_sum_ratio > 0.999 && tap_sum_ratio < 1.001, "ssim_stats_dual: the two windows must both be normalised");
    SsimArgs a{};
    a.d = *desc;
    a.x = x; a.y = y; a.win = win1d;
    a.out_sum = out_sum; a.out_sum2 = out_sum_b; a.out_flags = out_flags;
    a.co = desc->c;
    const double r = tap_sum_ratio * tap_sum_ratio;  // a moment is filtered along both axes
    a.dual_c1 = c1_b; a.dual_c2 = c2_b;
    a.dual_r = (float)r; a.dual_2r = (float)(2.0 * r); a.dual_r2 = (float)(r * r); a.dual_2r2 = (float)(2.0 * r * r);
    a.dual_neg_eps = (float)(1.0 - r);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (g_ssim_path_a.load(std::memory_order_relaxed) == 0 && ssim_ring_dual_eligible(*desc, x, y, win1d_host))
        return ssim_ring_forward(a, win1d_host, workspace, st);
    // two-column cp.async kernel, same geometry as the single-policy launch of ssim_forward
    const llie_ssim_desc &d = *desc;
    a.tiles_x = div_up(d.w, S2W);
    const long long cols = (long long)d.n * a.co * a.tiles_x;
    a.seg_rows = pick_seg_rows(cols, d.h, kNumSM * 3, 64, d.ksize - 1);
    a.tiles_y = div_up(d.h, a.seg_rows);
    const long long grid = cols * a.tiles_y;
    Workspace ws(workspace);
    LLIE_REQUIRE(grid <= 0x7fffffffLL && (size_t)grid * (2 * sizeof(double) + sizeof(int)) <= ws.scratch_bytes,
                 "ssim_stats_dual: workspace too small for %lld tiles", grid);
    a.part = reinterpret_cast<double *>(ws.scratch);
    a.part2 = a.part + grid;
    a.pflag = reinterpret_cast<int *>(ws.scratch + (size_t)grid * 2 * sizeof(double));
    a.ticket = ws.tickets;
    ssim_stream2_kernel<11, true, false, true><<<(int)grid, S2T, 0, st>>>(a);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

extern "C" int llie_ssim_loss_forward(const llie_ssim_desc *desc, const float *x, const float *y, const float *win1d,
                                      const float *win1d_host, float grad_scale, double *out_sum, float *out_map,
                                      float *out_coef, int32_t *out_flags, void *workspace, void *stream) {
    LLIE_REQUIRE(desc && out_coef, "ssim_loss_forward: null pointer");
    LLIE_REQUIRE(desc->pad_mode == LLIE_PAD_REFLECT && !desc->luma && !desc->clamp_var && !desc->interior,
                 "ssim_loss_forward: the differentiable path is the kornia policy (reflect, same size, no clamps)");
    LLIE_REQUIRE(desc->ksize / 2 < desc->h && desc->ksize / 2 < desc->w, "ssim_loss_forward: reflect pad >= image");
    return ssim_forward(desc, x, y, win1d, win1d_host, out_sum, out_map, out_coef, grad_scale, out_flags, workspace,
                        stream);
}

extern "C" int llie_ssim_loss_backward(const llie_ssim_desc *desc, const float *x, const float *y, const float *win1d,
                                       const float *win1d_host, const float *coef, const float *grad_out, int32_t grad_out_is_map,
                                       int32_t accumulate, float *out_grad, void *stream) {
    LLIE_REQUIRE(desc && x && y && win1d && coef && grad_out && out_grad, "ssim_loss_backward: null pointer");
    const llie_ssim_desc &d = *desc;
    LLIE_REQUIRE(d.n > 0 && d.c > 0 && d.h > 0 && d.w > 0, "ssim_loss_backward: empty tensor");
    LLIE_REQUIRE(d.ksize >= 1 && (d.ksize & 1) && d.ksize <= LLIE_SSIM_MAX_K, "ssim_loss_backward: bad window");
    LLIE_REQUIRE(d.pad_mode == LLIE_PAD_REFLECT && !d.luma && !d.clamp_var && !d.interior,
                 "ssim_loss_backward: kornia policy only");
    LLIE_REQUIRE(d.ksize / 2 < d.h && d.ksize / 2 < d.w && d.ksize <= d.h && d.ksize <= d.w,
                 "ssim_loss_backward: window larger than the image");
    SsimBwdArgs a{};
    a.n = d.n; a.c = d.c; a.h = d.h; a.w = d.w; a.ksize = d.ksize; a.clamp_in = d.clamp_in; a.gout_is_map = grad_out_is_map; a.accumulate = accumulate;
    a.x = x; a.y = y; a.win = win1d; a.coef = coef; a.gout = grad_out; a.out_grad = out_grad;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const int g_ssim_path = g_ssim_path_a.load(std::memory_order_relaxed);
    const int g_ssim_seg_rows = g_ssim_seg_rows_a.load(std::memory_order_relaxed);
    // default: TMA row ring + packed FP32 (ssim_ring.cu)
    if (g_ssim_path == 0 && ssim_ring_bwd_eligible(a, win1d_host)) return ssim_ring_backward(a, win1d_host, st);
    if ((g_ssim_path == 0 || g_ssim_path == 3) && !grad_out_is_map && d.ksize == 11) {
        a.tiles_x = div_up(d.w, S2W);
        const long long cols = (long long)d.n * d.c * a.tiles_x;
        const int rows = g_ssim_seg_rows > 0 ? (g_ssim_seg_rows < d.h ? g_ssim_seg_rows : d.h)
                                             : pick_seg_rows(cols, d.h, kNumSM * 3, 64, d.ksize - 1);
        a.tiles_y = div_up(d.h, rows);
        const long long grid = cols * a.tiles_y;
        LLIE_REQUIRE(grid <= 0x7fffffffLL, "ssim_loss_backward: too many tiles");
        return launch_ssim_bwd_stream2<11>(a, (int)grid, rows, st);
    }
    if (g_ssim_path != 1 && !grad_out_is_map && d.ksize <= 15) {
        a.tiles_x = div_up(d.w, SW);
        const long long cols = (long long)d.n * d.c * a.tiles_x;
        int segs = (int)((148LL * 5 * 3 + cols - 1) / cols);
        int rows = g_ssim_seg_rows > 0 ? g_ssim_seg_rows : (d.h + segs - 1) / segs;
        if (g_ssim_seg_rows <= 0 && rows < 48) rows = 48;
        if (rows > d.h) rows = d.h;
        a.tiles_y = div_up(d.h, rows);
        const long long grid = cols * a.tiles_y;
        LLIE_REQUIRE(grid <= 0x7fffffffLL, "ssim_loss_backward: too many tiles");
        switch (d.ksize) {
            case 1: return launch_ssim_bwd_stream<1>(a, (int)grid, rows, st);
            case 3: return launch_ssim_bwd_stream<3>(a, (int)grid, rows, st);
            case 5: return launch_ssim_bwd_stream<5>(a, (int)grid, rows, st);
            case 7: return launch_ssim_bwd_stream<7>(a, (int)grid, rows, st);
            case 9: return launch_ssim_bwd_stream<9>(a, (int)grid, rows, st);
            case 11: return launch_ssim_bwd_stream<11>(a, (int)grid, rows, st);
            case 13: return launch_ssim_bwd_stream<13>(a, (int)grid, rows, st);
            default: return launch_ssim_bwd_stream<15>(a, (int)grid, rows, st);
        }
    }
    a.tiles_x = div_up(d.w, STW);
    a.tiles_y = div_up(d.h, STH);
    const long long grid = (long long)d.n * d.c * a.tiles_x * a.tiles_y;
    LLIE_REQUIRE(grid <= 0x7fffffffLL, "ssim_loss_backward: too many tiles");
    switch (d.ksize) {
        case 1: return launch_ssim_bwd<1>(a, (int)grid, st);
        case 3: return launch_ssim_bwd<3>(a, (int)grid, st);
        case 5: return launch_ssim_bwd<5>(a, (int)grid, st);
        case 7: return launch_ssim_bwd<7>(a, (int)grid, st);
        case 9: return launch_ssim_bwd<9>(a, (int)grid, st);
        case 11: return launch_ssim_bwd<11>(a, (int)grid, st);
        case 13: return launch_ssim_bwd<13>(a, (int)grid, st);
        case 15: return launch_ssim_bwd<15>(a, (int)grid, st);
        default: return launch_ssim_bwd<0>(a, (int)grid, st);
    }
}
