// Physics-consistency projector, TMA tile path (depthwise 3x3 PSF, W % 4 == 0, 16-byte aligned planes).
//
// Persistent CTAs walk the tile list round-robin.  Each tile's pred halo block (128 x (TH+2) floats,
// image columns x0-4 .. x0+123) and obs block are fetched by ONE thread with cp.async.bulk.tensor
// (TMA, out-of-image elements arrive as zeros) into a 3-stage shared-memory ring guarded by
// mbarriers, so the loads of tiles k+1, k+2 are in flight while all warps run the stencil on tile k
// with conflict-free 128-bit LDS and a rolling 3-row register window.  reflect / replicate padding
// is a fix-up of the out-of-image halo cells on image-border tiles only.
// Reductions: one float64 partial per (tile, warp) -> fixed-order finalize in the last CTA.
#include "phys_common.cuh"
#include "tma.cuh"

namespace llie {

constexpr int TWO = 120;  // output columns per tile
constexpr int TWS = 128;  // shared row pitch of halo blocks (floats)
constexpr int NSTAGE = 3;
constexpr int MAXC = 16;  // channels supported by the tile path

template <int TH>
struct TileCfg {
    static constexpr int NT = 8 * TH;  // one warp per 4 output rows
    static constexpr int NW = NT / 32;
};

struct TileCoord {
    int plane, n, c, x0, y0;
};
__device__ __forceinline__ TileCoord decode_tile(int t, int tiles_x, int tiles_y, int C, int th) {
    TileCoord tc;
    const int tpp = tiles_x * tiles_y;
    tc.plane = t / tpp;
    const int r = t - tc.plane * tpp;
    const int ty = r / tiles_x;
    tc.x0 = (r - ty * tiles_x) * TWO;
    tc.y0 = ty * th;
    tc.n = tc.plane / C;
    tc.c = tc.plane - tc.n * C;
    return tc;
}

// rewrite out-of-image halo cells of a 128-wide block with the reflect / replicate source (which is
// always an in-image cell of the same block for the 1-2 pixel halos used here)
__device__ __forceinline__ void fixup_halo(float *sp, int rows, int y_first, int x_first, int H, int W, int mode,
                                           int nthreads) {
    for (int idx = threadIdx.x; idx < rows * TWS; idx += nthreads) {
        const int r = idx >> 7, s = idx & 127;
        const int y = y_first + r, x = x_first + s;
        if (y < 0 || y >= H || x < 0 || x >= W) {
            const int sy = pad_index(y, H, mode), sx = pad_index(x, W, mode);
            const int rr = sy - y_first, ss = sx - x_first;
            if (sy >= 0 && sx >= 0 && rr >= 0 && rr < rows && ss >= 0 && ss < TWS) sp[idx] = sp[rr * TWS + ss];
        }
    }
}

struct Row6 {
    float v[6];  // block columns 4L-1 .. 4L+4 (left neighbour, own four, right neighbour)
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

// ------------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------------
struct TmaFwdExtra {
    int total_tiles, tiles_per_sample;
    double *part;  // [total_tiles][NW]
    int *pflag;    // [gridDim.x]
};

template <int TH>
__global__ void __launch_bounds__(TileCfg<TH>::NT) phys_fwd_tma(const __grid_constant__ CUtensorMap map_pred,
                                                                const __grid_constant__ CUtensorMap map_obs,
                                                                const PhysArgs a, const TmaFwdExtra e) {
    constexpr int NT = TileCfg<TH>::NT, NW = TileCfg<TH>::NW;
    constexpr int PRED_F = (TH + 2) * TWS, OBS_F = TH * TWO, STAGE_F = PRED_F + OBS_F;
    constexpr unsigned STAGE_BYTES = STAGE_F * 4;
    extern __shared__ __align__(128) unsigned char smraw[];
    float *ring = reinterpret_cast<float *>(smraw);
    __shared__ uint64_t full[NSTAGE];
    __shared__ float sw[MAXC * 9];
    __shared__ double red[32];

    const llie_phys_desc &d = a.d;
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    const int H = d.h, W = d.w, G = gridDim.x;
    const int my_tiles = ((int)blockIdx.x < e.total_tiles) ? (e.total_tiles - 1 - (int)blockIdx.x) / G + 1 : 0;

    for (int i = threadIdx.x; i < d.c_out * 9; i += NT) sw[i] = __ldg(a.psf + i);
    if (threadIdx.x == 0) {
        for (int s = 0; s < NSTAGE; ++s) mbar_init(&full[s], 1);
        mbar_fence_init();
        tma_prefetch_desc(&map_pred);
        tma_prefetch_desc(&map_obs);
    }
    __syncthreads();

    auto issue = [&](int k) {
        if (k >= my_tiles) return;
        const TileCoord tc = decode_tile(blockIdx.x + k * G, a.tiles_x, a.tiles_y, d.c_out, TH);
        const int s = k % NSTAGE;
        float *sp = ring + s * STAGE_F;
        mbar_expect_tx(&full[s], STAGE_BYTES);
        tma_load_3d(sp, &map_pred, tc.x0 - 4, tc.y0 - 1, tc.plane, &full[s]);
        tma_load_3d(sp + PRED_F, &map_obs, tc.x0, tc.y0, tc.plane, &full[s]);
    };
    if (threadIdx.x == 0)
        for (int k = 0; k < NSTAGE - 1; ++k) issue(k);

    const bool map_vec = a.out_map != nullptr && (a.wo & 3) == 0 && a.x_lo == 0;
    float chk_p = 0.f, chk_o = 0.f;  // x*0 accumulators: become NaN iff a non-finite value was seen

    for (int k = 0; k < my_tiles; ++k) {
        __syncthreads();  // every warp is done with tile k-1 -> its stage may be refilled
        if (threadIdx.x == 0) {
            fence_proxy_async();
            issue(k + NSTAGE - 1);
        }
        const int t = blockIdx.x + k * G;
        const TileCoord tc = decode_tile(t, a.tiles_x, a.tiles_y, d.c_out, TH);
        const float *sp = ring + (k % NSTAGE) * STAGE_F;
        const float *so = sp + PRED_F;
        mbar_wait(&full[k % NSTAGE], (k / NSTAGE) & 1);

        const int x0 = tc.x0, y0 = tc.y0;
        if (d.pad_mode != LLIE_PAD_ZEROS && (y0 == 0 || y0 + TH + 1 > H || x0 == 0 || x0 + TWO + 1 > W)) {
            fixup_halo(const_cast<float *>(sp), TH + 2, y0 - 1, x0 - 4, H, W, d.pad_mode, NT);
            __syncthreads();
        }
        float w[9];
#pragma unroll
        for (int q = 0; q < 9; ++q) w[q] = sw[tc.c * 9 + q];
        const float rho_nc = (a.rho == nullptr)
                                 ? d.rho_scalar
                                 : __ldg(a.rho + (long long)tc.n * d.rho_sn + (long long)tc.c * d.rho_sc);
        const bool full_tile = (x0 >= a.x_lo) && (x0 + TWO <= a.x_lo + a.wo) && (y0 >= a.y_lo) &&
                               (y0 + TH <= a.y_lo + a.ho);
        float acc = 0.f;
        const int xx = x0 + 4 * (lane - 1);
        if (lane >= 1 && lane <= 30) {
            const int rbase = 4 * wrp;
            Row6 r0 = lds_row6(sp + (rbase + 0) * TWS, lane);
            Row6 r1 = lds_row6(sp + (rbase + 1) * TWS, lane);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const Row6 r2 = lds_row6(sp + (rbase + j + 2) * TWS, lane);
                const int yy = y0 + rbase + j;
                float cv[4];
                conv3_rows(r0, r1, r2, w, cv);
                const float4 ob = reinterpret_cast<const float4 *>(so + (rbase + j) * TWO)[lane - 1];
                const float av[4] = {ob.x, ob.y, ob.z, ob.w};
                float mv[4];
                if (full_tile) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        chk_o = fmaf(av[i], 0.f, chk_o);
                        chk_p = fmaf(r1.v[i + 1], 0.f, chk_p);
                        float r;
                        acc += residual(d, cv[i], av[i], rho_nc, r);
                        mv[i] = fabsf(r);
                    }
                    if (a.out_map) {
                        float *mrow = a.out_map + (((long long)tc.plane) * a.ho + (yy - a.y_lo)) * a.wo;
                        if (map_vec) {
                            *reinterpret_cast<float4 *>(mrow + xx) = make_float4(mv[0], mv[1], mv[2], mv[3]);
                        } else {
#pragma unroll
                            for (int i = 0; i < 4; ++i) mrow[xx + i - a.x_lo] = mv[i];
                        }
                    }
                } else {
                    const bool row_in = (yy >= a.y_lo) && (yy < a.y_lo + a.ho);
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        chk_o = fmaf(av[i], 0.f, chk_o);
                        chk_p = fmaf(r1.v[i + 1], 0.f, chk_p);
                        const int x = xx + i;
                        const bool in_dom = row_in && (x >= a.x_lo) && (x < a.x_lo + a.wo);
                        float r;
                        const float l = residual(d, cv[i], av[i], rho_nc, r);
                        mv[i] = fabsf(r);
                        if (in_dom) acc += l;
                    }
                    if (a.out_map && row_in) {
                        float *mrow = a.out_map + (((long long)tc.plane) * a.ho + (yy - a.y_lo)) * a.wo;
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const int xo = xx + i - a.x_lo;
                            if (xo >= 0 && xo < a.wo) mrow[xo] = mv[i];
                        }
                    }
                }
                r0 = r1;
                r1 = r2;
            }
        }
        const double ws = warp_sum((double)acc);
        if (lane == 0) e.part[(size_t)t * NW + wrp] = ws;
    }

    int flag = 0;
    if (!(chk_p == 0.f)) flag |= 1;
    if (!(chk_o == 0.f)) flag |= 2;
    const int f = block_or(flag, reinterpret_cast<int *>(red));
    if (threadIdx.x == 0) e.pflag[blockIdx.x] = f;
    if (!last_block_done(a.ticket, gridDim.x)) return;

    // ---- finalize (last CTA): fixed-order sums of the per-(tile, warp) partials of each sample
    const int N = d.n;
    const size_t per_sample = (size_t)e.tiles_per_sample * NW;
    double batch = 0.0, total = 0.0;
    for (int n = 0; n < N; ++n) {
        double s = 0.0;
        const double *p = e.part + (size_t)n * per_sample;
        for (size_t b = threadIdx.x; b < per_sample; b += NT) s += __ldcg(p + b);
        s = block_sum<double, true>(s, red);
        if (threadIdx.x == 0) {
            const float m = (float)(s * a.inv_count);
            a.out_mean[n] = m;
            if (a.out_sum) a.out_sum[n] = s;
            batch += (double)m;
            total += s;
        }
    }
    int fl = 0;
    for (int b = threadIdx.x; b < G; b += NT) fl |= __ldcg(e.pflag + b);
    const int ff = block_or(fl, reinterpret_cast<int *>(red));
    if (threadIdx.x == 0) {
        a.out_mean[N] = (float)(batch / N);
        a.out_mean[N + 1] = (float)batch;
        if (a.out_sum) a.out_sum[N] = total;
        *a.out_flags = ff;
    }
}

// ------------------------------------------------------------------------------------------------
// backward: pred block (halo 2) + obs block (halo 1) -> q = dLoss/dconv on the tile + 1-pixel halo
// (zero outside the compared domain) in shared memory -> grad = flipped-K stencil over q, plus the
// mirror / edge fold of reflect / replicate padding on image-border tiles.
// ------------------------------------------------------------------------------------------------
template <int TH>
__global__ void __launch_bounds__(TileCfg<TH>::NT) phys_bwd_tma(const __grid_constant__ CUtensorMap map_pred,
                                                                const __grid_constant__ CUtensorMap map_obs,
                                                                const PhysArgs a, const int total_tiles) {
    constexpr int NT = TileCfg<TH>::NT, NW = TileCfg<TH>::NW;
    constexpr int PRED_F = (TH + 4) * TWS, OBS_F = (TH + 2) * TWS, STAGE_F = PRED_F + OBS_F;
    constexpr unsigned STAGE_BYTES = STAGE_F * 4;
    extern __shared__ __align__(128) unsigned char smraw[];
    float *ring = reinterpret_cast<float *>(smraw);
    float *sq = ring + NSTAGE * STAGE_F;
    __shared__ uint64_t full[NSTAGE];
    __shared__ float sw[MAXC * 9];

    const llie_phys_desc &d = a.d;
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    const int H = d.h, W = d.w, G = gridDim.x;
    const int my_tiles = ((int)blockIdx.x < total_tiles) ? (total_tiles - 1 - (int)blockIdx.x) / G + 1 : 0;

    for (int i = threadIdx.x; i < d.c_out * 9; i += NT) sw[i] = __ldg(a.psf + i);
    if (threadIdx.x == 0) {
        for (int s = 0; s < NSTAGE; ++s) mbar_init(&full[s], 1);
        mbar_fence_init();
        tma_prefetch_desc(&map_pred);
        tma_prefetch_desc(&map_obs);
    }
    __syncthreads();

    auto issue = [&](int k) {
        if (k >= my_tiles) return;
        const TileCoord tc = decode_tile(blockIdx.x + k * G, a.tiles_x, a.tiles_y, d.c_out, TH);
        const int s = k % NSTAGE;
        float *sp = ring + s * STAGE_F;
        mbar_expect_tx(&full[s], STAGE_BYTES);
        tma_load_3d(sp, &map_pred, tc.x0 - 4, tc.y0 - 2, tc.plane, &full[s]);
        tma_load_3d(sp + PRED_F, &map_obs, tc.x0 - 4, tc.y0 - 1, tc.plane, &full[s]);
    };
    if (threadIdx.x == 0)
        for (int k = 0; k < NSTAGE - 1; ++k) issue(k);

    for (int k = 0; k < my_tiles; ++k) {
        __syncthreads();  // tile k-1 fully consumed (its stage and sq)
        if (threadIdx.x == 0) {
            fence_proxy_async();
            issue(k + NSTAGE - 1);
        }
        const TileCoord tc = decode_tile(blockIdx.x + k * G, a.tiles_x, a.tiles_y, d.c_out, TH);
        float *sp = ring + (k % NSTAGE) * STAGE_F;
        const float *so = sp + PRED_F;
        mbar_wait(&full[k % NSTAGE], (k / NSTAGE) & 1);
        const int x0 = tc.x0, y0 = tc.y0;
        const bool border = (y0 == 0 || y0 + TH + 2 > H || x0 == 0 || x0 + TWO + 2 > W);
        if (d.pad_mode != LLIE_PAD_ZEROS && border) {
            fixup_halo(sp, TH + 4, y0 - 2, x0 - 4, H, W, d.pad_mode, NT);
            __syncthreads();
        }
        float w[9];
#pragma unroll
        for (int q = 0; q < 9; ++q) w[q] = sw[tc.c * 9 + q];
        const float rho_nc = (a.rho == nullptr)
                                 ? d.rho_scalar
                                 : __ldg(a.rho + (long long)tc.n * d.rho_sn + (long long)tc.c * d.rho_sc);
        const float gn = __ldg(a.gout + a.gout_stride * tc.n) * a.gscale;

        // ---- phase 2: q rows 0 .. TH+1 (image rows y0-1 .. y0+TH), all 128 block columns
        {
            constexpr int QR = TH + 2;
            constexpr int PER = (QR + NW - 1) / NW;
            const int qa = wrp * PER;
            const int qb = min(qa + PER, QR);
            const bool edge_lane = (lane == 0) || (lane == 31);
            const int xb = x0 - 4 + 4 * lane;  // image column of this lane's first element
            if (qa < qb && !edge_lane) {
                Row6 r0 = lds_row6(sp + (qa + 0) * TWS, lane);
                Row6 r1 = lds_row6(sp + (qa + 1) * TWS, lane);
                for (int qr = qa; qr < qb; ++qr) {
                    const Row6 r2 = lds_row6(sp + (qr + 2) * TWS, lane);
                    const int yy = y0 - 1 + qr;
                    float qv[4] = {0.f, 0.f, 0.f, 0.f};
                    if (yy >= a.y_lo && yy < a.y_lo + a.ho) {
                        float cv[4];
                        conv3_rows(r0, r1, r2, w, cv);
                        const float4 ob = reinterpret_cast<const float4 *>(so + qr * TWS)[lane];
                        const float av[4] = {ob.x, ob.y, ob.z, ob.w};
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const int x = xb + i;
                            if (x >= a.x_lo && x < a.x_lo + a.wo) qv[i] = gn * dresidual(d, cv[i], av[i], rho_nc);
                        }
                    }
                    reinterpret_cast<float4 *>(sq + qr * TWS)[lane] = make_float4(qv[0], qv[1], qv[2], qv[3]);
                    r0 = r1;
                    r1 = r2;
                }
            } else if (qa < qb) {
                // lanes 0 / 31 own only the halo column x0-1 (block col 3) / x0+120 (block col 124)
                const int i = (lane == 0) ? 3 : 0;
                const int sc = 4 * lane + i;
                const int x = xb + i;
                for (int qr = qa; qr < qb; ++qr) {
                    const int yy = y0 - 1 + qr;
                    float qv[4] = {0.f, 0.f, 0.f, 0.f};
                    if (yy >= a.y_lo && yy < a.y_lo + a.ho && x >= a.x_lo && x < a.x_lo + a.wo) {
                        float s = 0.f;
#pragma unroll
                        for (int ky = 0; ky < 3; ++ky) {
                            const float *sr = sp + (qr + ky) * TWS;
#pragma unroll
                            for (int kx = 0; kx < 3; ++kx) s = fmaf(w[ky * 3 + kx], sr[sc + kx - 1], s);
                        }
                        qv[i] = gn * dresidual(d, s, so[qr * TWS + sc], rho_nc);
                    }
                    reinterpret_cast<float4 *>(sq + qr * TWS)[lane] = make_float4(qv[0], qv[1], qv[2], qv[3]);
                }
            }
        }
        __syncthreads();

        // ---- phase 3: grad(y,x) = sum_k K[k] q(y - ky + 1, x - kx + 1)
        const int xx = x0 + 4 * (lane - 1);
        if (lane >= 1 && lane <= 30 && xx < W) {
            const int rbase = 4 * wrp;
            float wf[9];
#pragma unroll
            for (int q = 0; q < 9; ++q) wf[q] = w[8 - q];
            float *gplane = a.grad + (long long)tc.plane * H * W;
            Row6 r0 = lds_row6(sq + (rbase + 0) * TWS, lane);
            Row6 r1 = lds_row6(sq + (rbase + 1) * TWS, lane);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const Row6 r2 = lds_row6(sq + (rbase + j + 2) * TWS, lane);
                const int yy = y0 + rbase + j;
                if (yy < H) {
                    float g[4];
                    conv3_rows(r0, r1, r2, wf, g);
                    if (border && d.pad_mode != LLIE_PAD_ZEROS) {
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
                                    if (iy == 0 && ix == 0) continue;  // the pixel itself is already in g
                                    for (int ky = 0; ky < 3; ++ky) {
                                        const int oy = py[iy] - ky + 1;
                                        if (oy < 0 || oy >= H) continue;
                                        for (int kx = 0; kx < 3; ++kx) {
                                            const int ox = px[ix] - kx + 1;
                                            if (ox < 0 || ox >= W) continue;
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
// host
// ------------------------------------------------------------------------------------------------
constexpr int kTH = 16;

bool phys_tma_eligible(const PhysArgs &a) {
    const llie_phys_desc &d = a.d;
    if (d.psf_layout != LLIE_PSF_DEPTHWISE || d.kh != 3 || d.kw != 3) return false;
    if (d.c_out > MAXC) return false;
    if ((d.w & 3) != 0 || d.w < 4 || d.h < 4) return false;
    if (!aligned16(a.pred) || !aligned16(a.obs)) return false;
    if (a.rho != nullptr && (d.rho_sh != 0 || d.rho_sw != 0)) return false;
    const long long tiles = (long long)d.n * d.c_out * div_up(d.w, TWO) * div_up(d.h, kTH);
    if (tiles * TileCfg<kTH>::NW * sizeof(double) + 4096 * sizeof(int) > kWsBytes - kWsHeaderBytes) return false;
    return true;
}

template <typename K>
static int set_smem(K kernel, size_t bytes) {
    // per (function, device) attribute; cheap enough to set on every launch (multi-GPU safe)
    LLIE_CUDA_OK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    return LLIE_OK;
}

int launch_phys_fwd_tma(PhysArgs &a, void *workspace, cudaStream_t st) {
    const llie_phys_desc &d = a.d;
    constexpr int TH = kTH;
    a.tiles_x = div_up(d.w, TWO);
    a.tiles_y = div_up(d.h, TH);
    CUtensorMap mp, mo;
    const long long planes = (long long)d.n * d.c_out;
    if (int rc = make_plane_map(&mp, a.pred, planes, d.h, d.w, TWS, TH + 2)) return rc;
    if (int rc = make_plane_map(&mo, a.obs, planes, d.h, d.w, TWO, TH)) return rc;
    TmaFwdExtra e;
    e.tiles_per_sample = d.c_out * a.tiles_x * a.tiles_y;
    e.total_tiles = d.n * e.tiles_per_sample;
    Workspace ws(workspace);
    e.part = reinterpret_cast<double *>(ws.scratch);
    e.pflag = reinterpret_cast<int *>(ws.scratch + (size_t)e.total_tiles * TileCfg<TH>::NW * sizeof(double));
    a.ticket = ws.tickets;
    const size_t smem = (size_t)NSTAGE * ((TH + 2) * TWS + TH * TWO) * 4;
    if (int rc = set_smem(phys_fwd_tma<TH>, smem)) return rc;
    int grid = kNumSM * 4;
    if (grid > e.total_tiles) grid = e.total_tiles;
    phys_fwd_tma<TH><<<grid, TileCfg<TH>::NT, smem, st>>>(mp, mo, a, e);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

int launch_phys_bwd_tma(PhysArgs &a, void *, cudaStream_t st) {
    const llie_phys_desc &d = a.d;
    constexpr int TH = kTH;
    a.tiles_x = div_up(d.w, TWO);
    a.tiles_y = div_up(d.h, TH);
    CUtensorMap mp, mo;
    const long long planes = (long long)d.n * d.c_out;
    if (int rc = make_plane_map(&mp, a.pred, planes, d.h, d.w, TWS, TH + 4)) return rc;
    if (int rc = make_plane_map(&mo, a.obs, planes, d.h, d.w, TWS, TH + 2)) return rc;
    const int total = d.n * d.c_out * a.tiles_x * a.tiles_y;
    const size_t smem = ((size_t)NSTAGE * ((TH + 4) * TWS + (TH + 2) * TWS) + (TH + 2) * TWS) * 4;
    if (int rc = set_smem(phys_bwd_tma<TH>, smem)) return rc;
    int grid = kNumSM * 3;
    if (grid > total) grid = total;
    phys_bwd_tma<TH><<<grid, TileCfg<TH>::NT, smem, st>>>(mp, mo, a, total);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

}  // namespace llie
