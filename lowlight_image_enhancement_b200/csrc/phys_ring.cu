// Physics-consistency projector, TMA row-ring path (depthwise 3x3 PSF, W % 4 == 0, 16-byte aligned
// planes) -- the HBM-bound production path for the reference's CrosstalkPSF shapes.
//
// Persistent CTAs of 4 consumer warps + 1 producer warp.  A work item is a chunk of rows of a
// 512-column strip of one (n, c) plane.  The producer's elected lane streams the item row by row
// with 1-D TMA bulk copies (cp.async.bulk global -> shared, 2 KB per row and tensor, completion on
// an mbarrier) into a ring of kRing row slots, so ~45 KB per CTA (180 KB per SM at 4 CTAs) are in
// flight without holding a single register; consumers wait on a slot's full barrier, pull their 4
// pixels (+ neighbours) into registers with LDS, release the slot through its empty barrier and
// run the 3x3 stencil in rolling ("scatter") form on registers.  No CTA-wide barrier in the loop.
// Padding: zero rows are synthesised by the consumer, reflect / replicate rows are simply fetched
// from the mapped source row by the producer, and the image-edge lanes select their outside
// neighbour from their own pixels.
//
//   forward   r = clipY(sY*(K(*)pred)) - clipA(sA*obs), |r| or Charbonnier, fp64 partial per warp
//             and item, fixed-order finalize in the last CTA (bitwise reproducible).
//   backward  q = dLoss/dconv recomputed for the thread's 4 columns and its two outside neighbours,
//             flipped-kernel stencil in the same rolling form; the adjoint of the padding operator
//             (mirror / edge fold) is applied in-register on the image-edge lanes and through two
//             virtual rows on the first / last chunk of a plane.
#include "phys_common.cuh"
#include "tma.cuh"

namespace llie {

constexpr int kRing = 12;            // row slots per CTA
constexpr int kCons = 128;           // consumer threads (4 warps side by side = 512 columns)
constexpr int kRingThreads = kCons + 32;
constexpr int kCols = kCons * 4;     // 512
constexpr int kRowF = kCols + 8;     // floats per halo'd row segment: columns x0-4 .. x0+515
constexpr int kSlotF = 2 * kRowF;    // pred segment + obs segment
constexpr int kCtasPerSM = 4;

struct RingGeom {
    int groups, chunks;  // column groups of 512 px, row chunks per plane (even split of H)
    int items, items_per_sample;
    double *part;  // [items][4] per-warp partial sums
    double *samp;  // [n] per-sample sums (finalize scratch)
    int *pflag;    // [gridDim.x]
};

struct Item {
    int plane, n, c, x0, ya, yb;
    bool first, last;
};
__device__ __forceinline__ Item decode_item(int item, const RingGeom &g, int C, int H) {
    Item it;
    const int per_plane = g.groups * g.chunks;
    it.plane = item / per_plane;
    const int rem = item - it.plane * per_plane;
    const int group = rem / g.chunks;
    const int chunk = rem - group * g.chunks;
    it.n = it.plane / C;
    it.c = it.plane - it.n * C;
    it.x0 = group * kCols;
    it.ya = (int)(((long long)chunk * H) / g.chunks);
    it.yb = (int)(((long long)(chunk + 1) * H) / g.chunks);
    it.first = chunk == 0;
    it.last = chunk == g.chunks - 1;
    return it;
}

__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gsrc, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// source row of a (possibly padded) row index; -1 = the row is all zeros (zero padding or far
// outside, where it only feeds discarded outputs)
__device__ __forceinline__ int src_row(int r, int H, int mode) {
    if (r >= 0 && r < H) return r;
    if (mode == LLIE_PAD_REFLECT) {
        const int s = r < 0 ? -r : 2 * (H - 1) - r;
        return (s >= 0 && s < H) ? s : -1;
    }
    if (mode == LLIE_PAD_REPLICATE) return r < 0 ? 0 : H - 1;
    return -1;
}

// One ring entry = (pred row pr, obs row orow), either may be absent (-1).  Producer side.
__device__ __forceinline__ void produce(float *slot, uint64_t *full, const float *pred_plane, const float *obs_plane,
                                        int pr, int orow, int x0, int W) {
    const int cs = x0 - 4 < 0 ? 0 : x0 - 4;
    const int ce = x0 + kCols + 4 > W ? W : x0 + kCols + 4;
    const uint32_t bytes = (uint32_t)(ce - cs) * 4u;
    const int doff = cs - (x0 - 4);
    const uint32_t total = (pr >= 0 ? bytes : 0u) + (orow >= 0 ? bytes : 0u);
    mbar_expect_tx(full, total);
    if (pr >= 0) bulk_g2s(slot + doff, pred_plane + (long long)pr * W + cs, bytes, full);
    if (orow >= 0) bulk_g2s(slot + kRowF + doff, obs_plane + (long long)orow * W + cs, bytes, full);
}

template <bool ROBUST, bool CLIPY, bool CLIPA>
__device__ __forceinline__ float loss_r(float conv, float a, float ry, float ra, float eps2) {
    float y = conv * ry, t = a * ra;
    if (CLIPY) y = fminf(fmaxf(y, 0.f), 1.f);
    if (CLIPA) t = fminf(fmaxf(t, 0.f), 1.f);
    const float r = y - t;
    return ROBUST ? sqrtf(fmaf(r, r, eps2)) : fabsf(r);
}
template <bool CLIPY, bool CLIPA>
__device__ __forceinline__ float absres_r(float conv, float a, float ry, float ra) {
    float y = conv * ry, t = a * ra;
    if (CLIPY) y = fminf(fmaxf(y, 0.f), 1.f);
    if (CLIPA) t = fminf(fmaxf(t, 0.f), 1.f);
    return fabsf(y - t);
}
// upstream-weighted derivative of the loss w.r.t. conv; gs = g_n * (rho if rho scales the conv side)
template <bool ROBUST, bool CLIPY, bool CLIPA>
__device__ __forceinline__ float dloss_r(float conv, float a, float ry, float ra, float eps2, float gs) {
    float y = conv * ry, t = a * ra;
    bool pass = true;
    if (CLIPY) {
        pass = (y >= 0.f) && (y <= 1.f);  // torch.clamp backward: gradient where lo <= y <= hi
        y = fminf(fmaxf(y, 0.f), 1.f);
    }
    if (CLIPA) t = fminf(fmaxf(t, 0.f), 1.f);
    const float r = y - t;
    float q;
    if (ROBUST)
        q = gs * (r / sqrtf(fmaf(r, r, eps2)));
    else
        q = (r > 0.f) ? gs : ((r < 0.f) ? -gs : 0.f);
    return pass ? q : 0.f;
}

// Rolling ("scatter") form of the 9-tap chain over NC output columns; r has NC + 2 entries (one
// neighbour on each side).  When input row rho arrives it completes output row rho-1 (kernel row
// 2), extends output row rho (kernel row 1) and opens output row rho+1 (kernel row 0).  The
// summation order per output is row-0 taps, row-1 taps, row-2 taps, left to right.
template <int NC>
__device__ __forceinline__ void roll3(const float (&r)[NC + 2], const float (&w)[9], float (&A)[NC], float (&B)[NC],
                                      float (&out)[NC]) {
#pragma unroll
    for (int i = 0; i < NC; ++i) {
        float o = fmaf(w[6], r[i], A[i]);
        o = fmaf(w[7], r[i + 1], o);
        o = fmaf(w[8], r[i + 2], o);
        out[i] = o;
        float t = fmaf(w[3], r[i], B[i]);
        t = fmaf(w[4], r[i + 1], t);
        t = fmaf(w[5], r[i + 2], t);
        A[i] = t;
        float u = w[0] * r[i];
        u = fmaf(w[1], r[i + 1], u);
        u = fmaf(w[2], r[i + 2], u);
        B[i] = u;
    }
}

struct RingState {
    int slot;
    uint32_t phase;
    __device__ __forceinline__ void advance() {
        if (++slot == kRing) {
            slot = 0;
            phase ^= 1u;
        }
    }
};

// Thread 0 publishes the CTA's results; returns true in ALL threads of the last CTA to arrive.
__device__ __forceinline__ bool ring_last_block(unsigned *ticket, unsigned total_blocks) {
    __shared__ bool s_last;
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned t = atomicAdd(ticket, 1u);
        s_last = (t == total_blocks - 1);
        if (s_last) {
            *ticket = 0u;  // self-reset: the workspace is zero again when the kernel ends
            __threadfence();
        }
    }
    __syncthreads();
    return s_last;
}

// last CTA: per-sample sums by one warp per sample (all loads of a pass in flight together),
// fixed summation order -> bitwise reproducible
__device__ void ring_finalize(const PhysArgs &a, const RingGeom &g, int *red_i) {
    const int N = a.d.n;
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const int per = g.items_per_sample * 4;
    for (int n = wrp; n < N; n += nw) {
        const double *p = g.part + (size_t)n * per;
        double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
        int b = lane;
        for (; b + 96 < per; b += 128) {
            s0 += __ldcg(p + b);
            s1 += __ldcg(p + b + 32);
            s2 += __ldcg(p + b + 64);
            s3 += __ldcg(p + b + 96);
        }
        for (; b < per; b += 32) s0 += __ldcg(p + b);
        const double s = warp_sum((s0 + s1) + (s2 + s3));
        if (lane == 0) g.samp[n] = s;
    }
    int fl = 0;
    for (int b = threadIdx.x; b < (int)gridDim.x; b += blockDim.x) fl |= __ldcg(g.pflag + b);
    const int ff = block_or(fl, red_i);  // contains the __syncthreads that publishes samp[]
    if (threadIdx.x == 0) {
        double batch = 0.0, total = 0.0;
        for (int n = 0; n < N; ++n) {
            const double s = g.samp[n];
            const float m = (float)(s * a.inv_count);
            a.out_mean[n] = m;
            if (a.out_sum) a.out_sum[n] = s;
            batch += (double)m;
            total += s;
        }
        a.out_mean[N] = (float)(batch / N);
        a.out_mean[N + 1] = (float)batch;
        if (a.out_sum) a.out_sum[N] = total;
        *a.out_flags = ff;
    }
}

// ------------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------------
template <bool ROBUST, bool CLIPY, bool CLIPA>
__global__ void __launch_bounds__(kRingThreads, kCtasPerSM) phys_fwd_ring(const PhysArgs a, const RingGeom g) {
    extern __shared__ __align__(128) unsigned char smraw[];
    float *ring = reinterpret_cast<float *>(smraw);
    __shared__ uint64_t full[kRing], empty[kRing];
    __shared__ int red_i[32];
    const llie_phys_desc &d = a.d;
    const int H = d.h, W = d.w, mode = d.pad_mode;
    const int tid = threadIdx.x, lane = tid & 31, wrp = tid >> 5;

    if (tid == 0) {
        for (int s = 0; s < kRing; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], kCons / 32);
        }
        mbar_fence_init();
    }
    __syncthreads();

    float chk_p = 0.f, chk_o = 0.f;  // x*0 accumulators: NaN iff a non-finite value was read
    RingState rs{0, 0u};

    if (wrp == kCons / 32) {
        // ================= producer (one elected lane) =================
        if (lane == 0) {
            for (int item = blockIdx.x; item < g.items; item += gridDim.x) {
                const Item it = decode_item(item, g, d.c_out, H);
                const float *pplane = a.pred + (long long)it.plane * H * W;
                const float *oplane = a.obs + (long long)it.plane * H * W;
                // entry e: pred row ya-1+e (mapped), obs row ya-2+e (for e >= 2)
                const int E = it.yb - it.ya + 2;
                for (int e = 0; e < E; ++e) {
                    mbar_wait(&empty[rs.slot], rs.phase ^ 1u);
                    produce(ring + rs.slot * kSlotF, &full[rs.slot], pplane, oplane, src_row(it.ya - 1 + e, H, mode),
                            e >= 2 ? it.ya - 2 + e : -1, it.x0, W);
                    rs.advance();
                }
            }
        }
    } else {
        // ================= consumers =================
        for (int item = blockIdx.x; item < g.items; item += gridDim.x) {
            const Item it = decode_item(item, g, d.c_out, H);
            const int xr = it.x0 + tid * 4;
            const bool active = xr < W;
            const int x = active ? xr : W - 4;
            const bool is_left = xr == 0, is_right = xr + 4 == W;
            float w[9];
#pragma unroll
            for (int q = 0; q < 9; ++q) w[q] = __ldg(a.psf + it.c * 9 + q);
            const float rho = (a.rho == nullptr)
                                  ? d.rho_scalar
                                  : __ldg(a.rho + (long long)it.n * d.rho_sn + (long long)it.c * d.rho_sc);
            const float ry = d.rho_side == 0 ? rho : 1.f, ra = d.rho_side == 1 ? rho : 1.f;
            const float eps2 = d.eps * d.eps;
            float cm[4];  // 1 where the column belongs to the compared domain of an active lane
#pragma unroll
            for (int i = 0; i < 4; ++i) cm[i] = (active && xr + i >= a.x_lo && xr + i < a.x_lo + a.wo) ? 1.f : 0.f;
            const bool map_vec = (a.wo & 3) == 0 && a.x_lo == 0;
            float sA[4] = {0.f, 0.f, 0.f, 0.f}, sB[4] = {0.f, 0.f, 0.f, 0.f};
            float acc = 0.f;
            const int E = it.yb - it.ya + 2;
            for (int e = 0; e < E; ++e) {
                const int prow = it.ya - 1 + e;  // pred row of this entry; completes output row prow-1
                const float *sp = ring + rs.slot * kSlotF;
                mbar_wait(&full[rs.slot], rs.phase);
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f), ob = v;
                float l = 0.f, r = 0.f;
                if (src_row(prow, H, mode) >= 0) {  // CTA-uniform; absent rows are zeros
                    v = reinterpret_cast<const float4 *>(sp)[tid + 1];
                    l = sp[4 * tid + 3];
                    r = sp[4 * tid + 8];
                }
                if (e >= 2) ob = reinterpret_cast<const float4 *>(sp + kRowF)[tid + 1];
                __syncwarp();
                if (lane == 0) mbar_arrive(&empty[rs.slot]);
                rs.advance();

                if (!active) {  // lanes beyond W read stale smem
                    v = make_float4(0.f, 0.f, 0.f, 0.f);
                    ob = v;
                    l = r = 0.f;
                }
                if (is_left) l = mode == LLIE_PAD_REFLECT ? v.y : (mode == LLIE_PAD_REPLICATE ? v.x : 0.f);
                if (is_right) r = mode == LLIE_PAD_REFLECT ? v.z : (mode == LLIE_PAD_REPLICATE ? v.w : 0.f);
                const float r6[6] = {l, v.x, v.y, v.z, v.w, r};
                chk_p = fmaf(v.x, 0.f, chk_p);
                chk_p = fmaf(v.y, 0.f, chk_p);
                chk_p = fmaf(v.z, 0.f, chk_p);
                chk_p = fmaf(v.w, 0.f, chk_p);
                float cv[4];
                roll3<4>(r6, w, sA, sB, cv);
                if (e >= 2) {  // CTA-uniform
                    const int y = prow - 1;
                    const float av[4] = {ob.x, ob.y, ob.z, ob.w};
#pragma unroll
                    for (int i = 0; i < 4; ++i) chk_o = fmaf(av[i], 0.f, chk_o);
                    if (y >= a.y_lo && y < a.y_lo + a.ho) {  // CTA-uniform
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            acc = fmaf(loss_r<ROBUST, CLIPY, CLIPA>(cv[i], av[i], ry, ra, eps2), cm[i], acc);
                        if (a.out_map != nullptr && active) {
                            float mv[4];
#pragma unroll
                            for (int i = 0; i < 4; ++i) mv[i] = absres_r<CLIPY, CLIPA>(cv[i], av[i], ry, ra);
                            float *mrow = a.out_map + (((long long)it.plane) * a.ho + (y - a.y_lo)) * a.wo;
                            if (map_vec) {
                                *reinterpret_cast<float4 *>(mrow + x) = make_float4(mv[0], mv[1], mv[2], mv[3]);
                            } else {
#pragma unroll
                                for (int i = 0; i < 4; ++i)
                                    if (cm[i] != 0.f) mrow[x + i - a.x_lo] = mv[i];
                            }
                        }
                    }
                }
            }
            const double ws = warp_sum((double)acc);
            if (lane == 0) g.part[(size_t)item * 4 + wrp] = ws;
        }
    }

    int flag = 0;
    if (!(chk_p == 0.f)) flag |= 1;
    if (!(chk_o == 0.f)) flag |= 2;
    const int f = block_or(flag, red_i);
    if (tid == 0) g.pflag[blockIdx.x] = f;
    if (ring_last_block(a.ticket, gridDim.x)) ring_finalize(a, g, red_i);
}

// ------------------------------------------------------------------------------------------------
// backward: grad = padY^T K^T q,  q = g_n * dLoss/dconv on the compared domain (0 elsewhere)
// ------------------------------------------------------------------------------------------------
template <bool ROBUST, bool CLIPY, bool CLIPA>
__global__ void __launch_bounds__(kRingThreads, kCtasPerSM) phys_bwd_ring(const PhysArgs a, const RingGeom g) {
    extern __shared__ __align__(128) unsigned char smraw[];
    float *ring = reinterpret_cast<float *>(smraw);
    __shared__ uint64_t full[kRing], empty[kRing];
    const llie_phys_desc &d = a.d;
    const int H = d.h, W = d.w, mode = d.pad_mode;
    const int tid = threadIdx.x, lane = tid & 31, wrp = tid >> 5;
    const bool fold = mode != LLIE_PAD_ZEROS;
    const int fold_top = mode == LLIE_PAD_REFLECT ? 1 : 0;          // image row that padded row -1 reads
    const int fold_bot = mode == LLIE_PAD_REFLECT ? H - 2 : H - 1;  // image row that padded row H reads

    if (tid == 0) {
        for (int s = 0; s < kRing; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], kCons / 32);
        }
        mbar_fence_init();
    }
    __syncthreads();
    RingState rs{0, 0u};

    // An item emits the (virtual) grad rows ja .. jb-1; the first / last chunk of a plane also produce
    // the padded rows -1 / H.  Entry j (j = ja-4 .. jb-1) carries pred row j+2 and obs row j+1; the
    // first two entries only warm up the conv partial sums, the next two the grad partial sums.
    if (wrp == kCons / 32) {
        if (lane == 0) {
            for (int item = blockIdx.x; item < g.items; item += gridDim.x) {
                const Item it = decode_item(item, g, d.c_out, H);
                const float *pplane = a.pred + (long long)it.plane * H * W;
                const float *oplane = a.obs + (long long)it.plane * H * W;
                const int ja = (it.first && fold) ? -1 : it.ya;
                const int jb = (it.last && fold) ? H + 1 : it.yb;
                for (int j = ja - 4; j < jb; ++j) {
                    const int qr = j + 1;  // q row this entry completes
                    const bool need_obs = j >= ja - 2 && qr >= a.y_lo && qr < a.y_lo + a.ho;
                    mbar_wait(&empty[rs.slot], rs.phase ^ 1u);
                    produce(ring + rs.slot * kSlotF, &full[rs.slot], pplane, oplane, src_row(j + 2, H, mode),
                            need_obs ? qr : -1, it.x0, W);
                    rs.advance();
                }
            }
        }
        return;
    }

    for (int item = blockIdx.x; item < g.items; item += gridDim.x) {
        const Item it = decode_item(item, g, d.c_out, H);
        const int xr = it.x0 + tid * 4;
        const bool active = xr < W;
        const int x = active ? xr : W - 4;
        const bool is_left = xr == 0, is_right = xr + 4 == W;
        const int ja = (it.first && fold) ? -1 : it.ya;
        const int jb = (it.last && fold) ? H + 1 : it.yb;
        float *gp = a.grad + (long long)it.plane * H * W + x;
        float w[9], wf[9];
#pragma unroll
        for (int q = 0; q < 9; ++q) w[q] = __ldg(a.psf + it.c * 9 + q);
#pragma unroll
        for (int q = 0; q < 9; ++q) wf[q] = w[8 - q];
        const float rho = (a.rho == nullptr)
                              ? d.rho_scalar
                              : __ldg(a.rho + (long long)it.n * d.rho_sn + (long long)it.c * d.rho_sc);
        const float ry = d.rho_side == 0 ? rho : 1.f, ra = d.rho_side == 1 ? rho : 1.f;
        const float eps2 = d.eps * d.eps;
        const float gs = __ldg(a.gout + a.gout_stride * it.n) * a.gscale * ry;
        bool colok[6];  // columns x-1 .. x+4 inside the compared domain
#pragma unroll
        for (int i = 0; i < 6; ++i) colok[i] = active && (xr - 1 + i >= a.x_lo) && (xr - 1 + i < a.x_lo + a.wo);

        float cA[6], cB[6];
#pragma unroll
        for (int i = 0; i < 6; ++i) cA[i] = cB[i] = 0.f;
        float gA[4] = {0.f, 0.f, 0.f, 0.f}, gB[4] = {0.f, 0.f, 0.f, 0.f};
        float eA = 0.f, eB = 0.f;  // partial sums of the folded padded column (image-edge lanes only)
        float carry[4] = {0.f, 0.f, 0.f, 0.f};

        for (int j = ja - 4; j < jb; ++j) {
            const int prow = j + 2, qr = j + 1;
            const bool have_q = j >= ja - 2 && qr >= a.y_lo && qr < a.y_lo + a.ho;  // CTA-uniform
            const float *sp = ring + rs.slot * kSlotF;
            mbar_wait(&full[rs.slot], rs.phase);
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f), ob = v;
            float2 pl = make_float2(0.f, 0.f), pr = pl;
            float ol = 0.f, orr = 0.f;
            if (src_row(prow, H, mode) >= 0) {
                v = reinterpret_cast<const float4 *>(sp)[tid + 1];
                pl = reinterpret_cast<const float2 *>(sp)[2 * tid + 1];
                pr = reinterpret_cast<const float2 *>(sp)[2 * tid + 4];
            }
            if (have_q) {
                const float *so = sp + kRowF;
                ob = reinterpret_cast<const float4 *>(so)[tid + 1];
                ol = so[4 * tid + 3];
                orr = so[4 * tid + 8];
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[rs.slot]);
            rs.advance();

            if (!active) {  // lanes beyond W read stale smem
                v = make_float4(0.f, 0.f, 0.f, 0.f);
                pl = pr = make_float2(0.f, 0.f);
            }
            float l0 = pl.x, l1 = pl.y, r0 = pr.x, r1 = pr.y;
            if (is_left) {  // column -1 is padded; column -2 only feeds a discarded q
                l1 = mode == LLIE_PAD_REFLECT ? v.y : (mode == LLIE_PAD_REPLICATE ? v.x : 0.f);
                l0 = 0.f;
            }
            if (is_right) {
                r0 = mode == LLIE_PAD_REFLECT ? v.z : (mode == LLIE_PAD_REPLICATE ? v.w : 0.f);
                r1 = 0.f;
            }
            const float r8[8] = {l0, l1, v.x, v.y, v.z, v.w, r0, r1};
            float cv[6];
            roll3<6>(r8, w, cA, cB, cv);  // conv row qr over columns x-1 .. x+4
            if (j < ja - 2) continue;     // conv warm-up entries (CTA-uniform)
            float qc[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            if (have_q) {
                const float av[6] = {ol, ob.x, ob.y, ob.z, ob.w, orr};
#pragma unroll
                for (int i = 0; i < 6; ++i)
                    qc[i] = colok[i] ? dloss_r<ROBUST, CLIPY, CLIPA>(cv[i], av[i], ry, ra, eps2, gs) : 0.f;
            }
            // ---- grad (virtual) row j: q row j+1 completes it (flipped kernel, rolling form)
            float gr[4];
            roll3<4>(qc, wf, gA, gB, gr);
            if (fold && (is_left || is_right)) {
                // padded column -1 (W) = sum_ky K[ky][0] q(j-ky+1, 0)  (K[ky][2] q(.., W-1)), rolling too
                const float qe = is_left ? qc[1] : qc[4];
                const float k0 = is_left ? w[0] : w[2], k1 = is_left ? w[3] : w[5], k2 = is_left ? w[6] : w[8];
                const float ecol = fmaf(k0, qe, eA);
                eA = fmaf(k1, qe, eB);
                eB = k2 * qe;
                // folds onto column 1 / 0 (left: reflect / replicate) or W-2 / W-1 (right)
                const int tgt = is_left ? (mode == LLIE_PAD_REFLECT ? 1 : 0) : (mode == LLIE_PAD_REFLECT ? 2 : 3);
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    if (i == tgt) gr[i] += ecol;
            }
            if (j < ja) continue;  // grad warm-up entries (CTA-uniform)
            if (j == -1) {
#pragma unroll
                for (int i = 0; i < 4; ++i) carry[i] = gr[i];
            } else if (j == H) {
                if (active) {
                    float4 *dst = reinterpret_cast<float4 *>(gp + (long long)fold_bot * W);
                    float4 t = *dst;  // written by this very thread two / one iterations ago
                    t.x += gr[0]; t.y += gr[1]; t.z += gr[2]; t.w += gr[3];
                    *dst = t;
                }
            } else {
                if (it.first && fold && j == fold_top) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) gr[i] += carry[i];
                }
                if (active)
                    *reinterpret_cast<float4 *>(gp + (long long)j * W) = make_float4(gr[0], gr[1], gr[2], gr[3]);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// host
// ------------------------------------------------------------------------------------------------
static int g_ring_rows = 0;  // 0 = heuristic; set through llie_tune(LLIE_TUNE_PHYS_ROWS, r)
void phys_ring_set_rows(int r) { g_ring_rows = r; }

static RingGeom make_geom(const llie_phys_desc &d, int &grid) {
    RingGeom g{};
    g.groups = div_up(d.w, kCols);
    const long long cols = (long long)d.n * d.c_out * g.groups;  // independent column strips
    const int slots = kNumSM * kCtasPerSM;
    int chunks;
    if (g_ring_rows > 0) {
        chunks = d.h / g_ring_rows;
    } else {
        // items = k * (resident CTA slots) for the smallest k that keeps chunks >= ~24 rows: every
        // persistent CTA gets the same number of items (no tail wave), rows split evenly
        long long k = ((long long)d.h * cols + (long long)slots * 24 - 1) / ((long long)slots * 24);
        if (k < 1) k = 1;
        chunks = (int)((k * slots) / cols);
    }
    if (chunks > d.h / 2) chunks = d.h / 2;  // >= 2 rows per chunk (padding-fold contract)
    if (chunks < 1) chunks = 1;
    g.chunks = chunks;
    g.items_per_sample = d.c_out * g.groups * chunks;
    const long long items = (long long)d.n * g.items_per_sample;
    g.items = (int)items;
    grid = items < slots ? (int)items : slots;
    return g;
}

bool phys_ring_eligible(const PhysArgs &a) {
    const llie_phys_desc &d = a.d;
    if (d.psf_layout != LLIE_PSF_DEPTHWISE || d.kh != 3 || d.kw != 3) return false;
    if ((d.w & 3) != 0 || d.w < 8 || d.h < 4) return false;
    if (!aligned16(a.pred) || !aligned16(a.obs)) return false;
    if (a.rho != nullptr && (d.rho_sh != 0 || d.rho_sw != 0)) return false;
    const long long items = (long long)d.n * d.c_out * div_up(d.w, kCols) * (d.h / 2);
    if (items > 0x3fffffffLL) return false;
    int grid;
    const RingGeom g = make_geom(d, grid);
    const size_t need = (size_t)g.items * 4 * sizeof(double) + (size_t)d.n * sizeof(double) + (size_t)grid * sizeof(int);
    return need <= kWsBytes - kWsHeaderBytes;
}

#define LLIE_RING_DISPATCH(KERNEL, d, GRID, SMEM, ST, ...)                                                  \
    do {                                                                                                    \
        const int key = ((d).robust ? 4 : 0) | ((d).clip_y ? 2 : 0) | ((d).clip_a ? 1 : 0);                 \
        switch (key) {                                                                                      \
            case 0: LLIE_RING_ONE((KERNEL<false, false, false>), GRID, SMEM, ST, __VA_ARGS__); break;       \
            case 1: LLIE_RING_ONE((KERNEL<false, false, true>), GRID, SMEM, ST, __VA_ARGS__); break;        \
            case 2: LLIE_RING_ONE((KERNEL<false, true, false>), GRID, SMEM, ST, __VA_ARGS__); break;        \
            case 3: LLIE_RING_ONE((KERNEL<false, true, true>), GRID, SMEM, ST, __VA_ARGS__); break;         \
            case 4: LLIE_RING_ONE((KERNEL<true, false, false>), GRID, SMEM, ST, __VA_ARGS__); break;        \
            case 5: LLIE_RING_ONE((KERNEL<true, false, true>), GRID, SMEM, ST, __VA_ARGS__); break;         \
            case 6: LLIE_RING_ONE((KERNEL<true, true, false>), GRID, SMEM, ST, __VA_ARGS__); break;         \
            default: LLIE_RING_ONE((KERNEL<true, true, true>), GRID, SMEM, ST, __VA_ARGS__); break;         \
        }                                                                                                   \
    } while (0)
#define LLIE_RING_ONE(K, GRID, SMEM, ST, ...)                                                               \
    do {                                                                                                    \
        auto kern = K;                                                                                      \
        LLIE_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(SMEM))); \
        kern<<<GRID, kRingThreads, SMEM, ST>>>(__VA_ARGS__);                                                \
    } while (0)

constexpr size_t kRingSmem = (size_t)kRing * kSlotF * sizeof(float);

int launch_phys_fwd_ring(PhysArgs &a, void *workspace, cudaStream_t st) {
    int grid;
    RingGeom g = make_geom(a.d, grid);
    Workspace ws(workspace);
    g.part = reinterpret_cast<double *>(ws.scratch);
    g.samp = g.part + (size_t)g.items * 4;
    g.pflag = reinterpret_cast<int *>(g.samp + a.d.n);
    a.ticket = ws.tickets;
    LLIE_RING_DISPATCH(phys_fwd_ring, a.d, grid, kRingSmem, st, a, g);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

int launch_phys_bwd_ring(PhysArgs &a, cudaStream_t st) {
    int grid;
    const RingGeom g = make_geom(a.d, grid);
    LLIE_RING_DISPATCH(phys_bwd_ring, a.d, grid, kRingSmem, st, a, g);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

}  // namespace llie
