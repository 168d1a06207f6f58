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

namespace llie {

constexpr int STH = 32, STW = 64, SNT = 256;

struct SsimArgs {
    llie_ssim_desc d;
    const float *x, *y, *win;
    double *out_sum;
    float *out_map;
    int *out_flags;
    double *part;  // [planes][tiles_per_plane]
    int *pflag;    // [gridDim.x]
    unsigned *ticket;
    int tiles_x, tiles_y, co;
};

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

template <int K>
__global__ void __launch_bounds__(SNT, 2) ssim_stats_kernel(const SsimArgs a) {
    constexpr int P = K / 2;
    constexpr int XR = STH + 2 * P;           // rows of the input tile / of the horizontal-pass output
    constexpr int XP = (STW + 2 * P) | 1;     // odd pitches: lanes walk rows without bank conflicts
    constexpr int HP = STW + 1;
    extern __shared__ __align__(16) float sm[];
    float *sX = sm, *sY = sX + XR * XP;
    float *sH = sY + XR * XP;  // [5][XR][HP]
    __shared__ double red[32];
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

    float g[K];
#pragma unroll
    for (int j = 0; j < K; ++j) g[j] = __ldg(a.win + j);

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
#pragma unroll
        for (int u = 0; u < K + 3; ++u) {
            const float xv = px[u], yv = py[u];
            const float xx = xv * xv, yy = yv * yv, xy = xv * yv;
#pragma unroll
            for (int o = 0; o < 4; ++o) {
                const int j = u - o;
                if (j >= 0 && j < K) {
                    ax[o] = fmaf(g[j], xv, ax[o]);
                    ay[o] = fmaf(g[j], yv, ay[o]);
                    axx[o] = fmaf(g[j], xx, axx[o]);
                    ayy[o] = fmaf(g[j], yy, ayy[o]);
                    axy[o] = fmaf(g[j], xy, axy[o]);
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
#pragma unroll
        for (int u = 0; u < K + 7; ++u) {
            const float v = h[u * HP];
#pragma unroll
            for (int o = 0; o < 8; ++o) {
                const int j = u - o;
                if (j >= 0 && j < K) m[q][o] = fmaf(g[j], v, m[q][o]);
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
    constexpr int P = K / 2, XR = STH + 2 * P, XP = (STW + 2 * P) | 1, HP = STW + 1;
    const size_t smem = (size_t)(2 * XR * XP + 5 * XR * HP) * sizeof(float);
    LLIE_CUDA_OK(cudaFuncSetAttribute(ssim_stats_kernel<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ssim_stats_kernel<K><<<grid, SNT, smem, st>>>(a);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

}  // namespace llie

using namespace llie;

extern "C" int llie_ssim_stats(const llie_ssim_desc *desc, const float *x, const float *y, const float *win1d,
                               double *out_sum, float *out_map, int32_t *out_flags, void *workspace, void *stream) {
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
    a.co = d.luma ? 1 : d.c;
    a.tiles_x = div_up(d.w, STW);
    a.tiles_y = div_up(d.h, STH);
    const long long grid = (long long)d.n * a.co * a.tiles_x * a.tiles_y;
    Workspace ws(workspace);
    LLIE_REQUIRE(grid <= 0x7fffffffLL && (size_t)grid * (sizeof(double) + sizeof(int)) <= ws.scratch_bytes,
                 "ssim_stats: workspace too small for %lld tiles", grid);
    a.part = reinterpret_cast<double *>(ws.scratch);
    a.pflag = reinterpret_cast<int *>(ws.scratch + (size_t)grid * sizeof(double));
    a.ticket = ws.tickets;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    switch (d.ksize) {
        case 1: return launch_ssim<1>(a, (int)grid, st);
        case 3: return launch_ssim<3>(a, (int)grid, st);
        case 5: return launch_ssim<5>(a, (int)grid, st);
        case 7: return launch_ssim<7>(a, (int)grid, st);
        case 9: return launch_ssim<9>(a, (int)grid, st);
        case 11: return launch_ssim<11>(a, (int)grid, st);
        case 13: return launch_ssim<13>(a, (int)grid, st);
        default: return launch_ssim<15>(a, (int)grid, st);
    }
}
