// Physics-consistency projector, TMA row-ring path (depthwise 3x3 PSF, W % 4 == 0, 16-byte aligned
// planes) -- the HBM-bound production path for the reference's CrosstalkPSF shapes.
//
// Persistent CTAs of 4 consumer warps + 1 producer warp.  A work item is a chunk of rows of a
// 512-column strip of one (n, c) plane.  The producer's elected lane streams the item with 1-D TMA
// bulk copies (cp.async.bulk global -> shared, one 2 KB row segment per copy, completion on an
// mbarrier) into a ring of kSlots slots of kG rows each, so the bytes in flight (~130 KB per SM)
// live in shared memory and not in registers; consumers wait on a slot's full barrier, run kG rows
// of the 3x3 stencil in rolling ("scatter") form on registers with straight-line code, and release
// the slot through its empty barrier.  No CTA-wide barrier in the loop.
//
// The per-row consumer code is branch-free:
//   * padded rows: reflect / replicate rows are fetched from the mapped source row by the producer;
//     zero rows are read from a permanently-zero row buffer (pointer select);
//   * padded columns: the two image-edge lanes read their outside neighbour from a per-thread
//     precomputed shared-memory index (own pixel for reflect / replicate, a zero word for zeros);
//   * crop, warm-up rows and lanes beyond W are float masks folded into the accumulation;
//   * rho on the conv side is folded into the 9 weights once per item.
//
//   forward   r = clipY(sY*(K(*)pred)) - clipA(sA*obs), |r| or Charbonnier, fp64 partial per warp
//             and item, fixed-order finalize in the last CTA (bitwise reproducible).
//   backward  q = dLoss/dconv recomputed for the thread's 4 columns and its two outside neighbours,
//             flipped-kernel stencil in the same rolling form; the adjoint of the padding operator
//             (mirror / edge fold) is applied in-register on the image-edge lanes and through two
//             virtual rows on the first / last chunk of a plane.
#include "phys_common.cuh"
#include <atomic>
#include <type_traits>

#include "tma.cuh"

namespace llie {

constexpr int kG = 4;                // rows per ring slot (one mbarrier pair per slot)
constexpr int kSlots = 3;            // ring depth
constexpr int kCons = 128;           // consumer threads (4 warps side by side = 512 columns)
constexpr int kRingThreads = kCons + 32;
constexpr int kCols = kCons * 4;     // 512
constexpr int kRowD = kCols + 8;     // data floats per halo'd row segment: columns x0-4 .. x0+515
constexpr int kRowF = kRowD + 4;     // + 4 always-zero floats (index kRowD = "the zero word")
constexpr int kPairF = 2 * kRowF;    // pred row + obs row
constexpr int kSlotF = kG * kPairF;
constexpr int kCtasPerSM = 4;
constexpr size_t kRingSmem = ((size_t)kSlots * kSlotF + kRowF) * sizeof(float);  // ring + zero row

constexpr int kSlotInterior = 0, kSlotGeneral = 1, kSlotWarm = 2;  // consumer code paths per ring slot
template <int K>
using SlotKind = std::integral_constant<int, K>;

// Timeline tracing for tools/micro/ring_trace.cu (compile with -DLLIE_RING_TRACE; never in the product build)
#ifdef LLIE_RING_TRACE
__device__ unsigned long long g_ring_trace[1024 * 16];
__device__ __forceinline__ void ring_stamp(int k) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    if (blockIdx.x < 1024) g_ring_trace[blockIdx.x * 16 + k] = t;
}
#define RING_STAMP(cond, k) do { if (cond) ring_stamp(k); } while (0)
#else
#define RING_STAMP(cond, k) do { } while (0)
#endif

// Row-loop unrolling of the backward slot bodies.  The fused (value + gradient) instantiation carries three slot bodies
// of four unrolled rows each plus the reduction epilogue: ~100 KB of SASS against 56 KB for the gradient-only kernel, and
// ncu showed 1.06 `no_instruction` stall cycles per issued instruction on it (none on the gradient-only kernel).  With
// the GENERAL body rolled and the interior body unrolled by two, the fused launch at C2 goes 25.2 -> 23.1 us
// stream-launched (tools/micro/ring_trace.cu) and 23.5 -> 21.6 us in bench.py's graphs; rolling the interior body
// completely changes nothing more, and sending the warm-up slot through the GENERAL body costs 0.7 us.  The same change
// costs the gradient-only kernel 0.8 us, so it keeps the fully unrolled bodies.
#ifndef LLIE_RING_UNROLL_G
#define LLIE_RING_UNROLL_G 1
#endif
#ifndef LLIE_RING_UNROLL_I
#define LLIE_RING_UNROLL_I 2
#endif
constexpr int kUnrollG = LLIE_RING_UNROLL_G, kUnrollI = LLIE_RING_UNROLL_I;

struct RingGeom {
    int groups, chunks;  // column groups of 512 px, row chunks per plane (even split of H)
    int items, items_per_sample;
    double *part;  // [items] per-item partial sums
    double *samp;  // [n] per-sample sums (finalize scratch)
    int *pflag;    // [gridDim.x]
    int early;     // stream-overlap contract (llie_phys_desc.stream_overlap): only the workspace depends on the previous kernel
    int prefetch;  // rows per tensor of the CTA's first item prefetched into L2 before griddepcontrol.wait (LLIE_TUNE_PHYS_PREFETCH)
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

// Packed fp32 pairs (sm_100 FFMA2 / FMUL2: one issue slot for two IEEE fp32 operations; a pair whose halves are the
// same register is encoded as a scalar broadcast operand, so broadcast weights cost nothing).
typedef unsigned long long f2_t;
__device__ __forceinline__ f2_t pack2(float lo, float hi) {
    f2_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(f2_t v, float &lo, float &hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
#ifndef LLIE_ROLL3_SCALAR
__device__ __forceinline__ f2_t fma2(f2_t a, f2_t b, f2_t c) {
    f2_t d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ f2_t mul2(f2_t a, f2_t b) {
    f2_t d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
#else  // A/B build (tools/micro): the same chains as scalar FFMA / FMUL
__device__ __forceinline__ f2_t fma2(f2_t a, f2_t b, f2_t c) {
    float al, ah, bl, bh, cl, ch;
    unpack2(a, al, ah); unpack2(b, bl, bh); unpack2(c, cl, ch);
    return pack2(fmaf(al, bl, cl), fmaf(ah, bh, ch));
}
__device__ __forceinline__ f2_t mul2(f2_t a, f2_t b) {
    float al, ah, bl, bh;
    unpack2(a, al, ah); unpack2(b, bl, bh);
    return pack2(al * bl, ah * bh);
}
#endif

// Rolling ("scatter") form of the 9-tap chain over NC output columns; r has NC + 2 entries (one
// neighbour on each side).  When input row rho arrives it completes output row rho-1 (kernel row
// 2), extends output row rho (kernel row 1) and opens output row rho+1 (kernel row 0).  The
// summation order per output is row-0 taps, row-1 taps, row-2 taps, left to right.
// Adjacent columns 2p, 2p + 1 share one packed accumulator (A[p], B[p] hold the pair).  The outer taps (dx = 0, 2) read
// the register-aligned data pairs (r[2p], r[2p+1]) / (r[2p+2], r[2p+3]) and are one FFMA2 each; the middle tap's pair
// (r[2p+1], r[2p+2]) straddles two aligned pairs, so it stays two scalar FFMAs on the halves of the accumulator (no
// repacking moves).  Same operation order as the scalar chain: bit-identical, 4 issue slots per 6 multiply-adds.
__device__ __forceinline__ f2_t mid_tap(float wq, float ra, float rb, f2_t acc) {
    float lo, hi;
    unpack2(acc, lo, hi);
    lo = fmaf(wq, ra, lo);
    hi = fmaf(wq, rb, hi);
    return pack2(lo, hi);
}
template <int NC>
__device__ __forceinline__ void roll3(const float (&r)[NC + 2], const float (&w)[9], f2_t (&A)[NC / 2],
                                      f2_t (&B)[NC / 2], float (&out)[NC]) {
    static_assert(NC % 2 == 0, "roll3 pairs columns");
    constexpr int HP = NC / 2;
    f2_t ev[HP + 1];  // (r0,r1) (r2,r3) ...
#pragma unroll
    for (int p = 0; p <= HP; ++p) ev[p] = pack2(r[2 * p], r[2 * p + 1]);
#pragma unroll
    for (int p = 0; p < HP; ++p) {
        const float ra = r[2 * p + 1], rb = r[2 * p + 2];
        f2_t o = fma2(pack2(w[6], w[6]), ev[p], A[p]);
        o = mid_tap(w[7], ra, rb, o);
        o = fma2(pack2(w[8], w[8]), ev[p + 1], o);
        unpack2(o, out[2 * p], out[2 * p + 1]);
        f2_t t = fma2(pack2(w[3], w[3]), ev[p], B[p]);
        t = mid_tap(w[4], ra, rb, t);
        t = fma2(pack2(w[5], w[5]), ev[p + 1], t);
        A[p] = t;
        f2_t u = mul2(pack2(w[0], w[0]), ev[p]);
        u = mid_tap(w[1], ra, rb, u);
        u = fma2(pack2(w[2], w[2]), ev[p + 1], u);
        B[p] = u;
    }
}

// sqrt(x) and 1 / sqrt(x) for the Charbonnier penalty and its derivative, x = r^2 + eps^2.  For x in [1e-30, 1e30] (every
// sane configuration: eps >= 1e-15, |r| <= 1e15): one MUFU.RSQ, a Newton step on the root (<= 1 ulp) and one on the
// reciprocal root (<= 1 ulp), 7 instructions without a slow path; the IEEE sqrtf + division pair costs ~20 issue slots
// per element in a kernel that is co-limited by instruction issue.  Outside that range: the IEEE forms.
__device__ __forceinline__ void sqrt_rsqrt_pos(float x, float &sq, float &rs) {
    if (x >= 1e-30f && x <= 1e30f) {
        float r0;
        asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r0) : "f"(x));
        const float s0 = x * r0;
        const float h = 0.5f * r0;
        const float e = fmaf(-s0, s0, x);        // x - s0^2
        sq = fmaf(e, h, s0);                     // s0 + e / (2 s0)
        rs = fmaf(r0, fmaf(-s0, h, 0.5f), r0);   // r0 (1.5 - 0.5 x r0^2)
    } else {
        sq = sqrtf(x);
        rs = 1.f / sq;
    }
}

// fminf / fmaxf drop NaN: non-finite values are caught by the chk_* accumulators before the clip
__device__ __forceinline__ float clamp01(float v) { return fminf(fmaxf(v, 0.f), 1.f); }

struct RingState {
    int slot;
    uint32_t phase;
    __device__ __forceinline__ void advance() {
        if (++slot == kSlots) {
            slot = 0;
            phase ^= 1u;
        }
    }
};

// common prologue: init the barriers and zero what the consumers may read without a TMA write: the zero row, the
// zero words and the 4-column halos of every row segment (never written in the first / last strip of a plane).
// Strips that the image only partly covers (W % kCols != 0) have idle lanes reading unwritten columns: those launches
// zero the whole ring (stale words must be finite, they meet 0-weights).
__device__ __forceinline__ void ring_setup(float *ring, uint64_t *full, uint64_t *empty, bool zero_all) {
    float4 *r4 = reinterpret_cast<float4 *>(ring);
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    if (zero_all) {
        const int n4 = (int)(kRingSmem / 16);
        for (int i = threadIdx.x; i < n4; i += kRingThreads) r4[i] = z;
    } else {
        constexpr int kSegs = kSlots * kG * 2, kRow4 = kRowF / 4;
        for (int i = threadIdx.x; i < kSegs * 3; i += kRingThreads) {
            const int seg = i / 3, part = i - seg * 3;
            r4[seg * kRow4 + (part == 0 ? 0 : (part == 1 ? kRowD / 4 - 1 : kRowD / 4))] = z;
        }
        for (int i = threadIdx.x; i < kRow4; i += kRingThreads) r4[kSegs * kRow4 + i] = z;  // the zero row
    }
    RING_STAMP(threadIdx.x == 0, 8);
    if (threadIdx.x == 0) {
        for (int s = 0; s < kSlots; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], kCons / 32);
        }
        mbar_fence_init();
    }
    RING_STAMP(threadIdx.x == 0, 9);
    fence_proxy_async();  // generic-proxy zero fill before the async-proxy (TMA) writes
    __syncthreads();
}

// Producer: fill one slot.  rows[u] = (pred source row or -1, obs row or -1) for u < kG.
struct SlotRows {
    int pr[kG], orow[kG];
};
__device__ __forceinline__ void produce_slot(float *slot, uint64_t *full, const float *pred_plane,
                                             const float *obs_plane, const SlotRows &sr, int x0, int W) {
    const int cs = x0 - 4 < 0 ? 0 : x0 - 4;
    const int ce = x0 + kCols + 4 > W ? W : x0 + kCols + 4;
    const uint32_t bytes = (uint32_t)(ce - cs) * 4u;
    const int doff = cs - (x0 - 4);
    uint32_t total = 0;
#pragma unroll
    for (int u = 0; u < kG; ++u) total += (sr.pr[u] >= 0 ? bytes : 0u) + (sr.orow[u] >= 0 ? bytes : 0u);
    mbar_expect_tx(full, total);
#pragma unroll
    for (int u = 0; u < kG; ++u) {
        float *dst = slot + u * kPairF + doff;
        if (sr.pr[u] >= 0) bulk_g2s(dst, pred_plane + (long long)sr.pr[u] * W + cs, bytes, full);
        if (sr.orow[u] >= 0) bulk_g2s(dst + kRowF, obs_plane + (long long)sr.orow[u] * W + cs, bytes, full);
    }
}

// L2 prefetch of the first rows of this CTA's first item, issued BEFORE griddepcontrol.wait.  A prefetch has no
// architectural effect (L2 is the point of coherence: whatever the previous kernel still writes to these lines lands in
// the same L2 lines), so it is legal under the default, fully serialised contract -- unlike the early loads of the
// stream-overlap mode.  Under programmatic dependent launch the CTAs of launch i + 1 become resident while launch i
// drains, and their opening rows stream into L2 behind the previous launch's tail.  Measured at C2 (graphs of 200
// serialised launches, final round-2 kernels): fused value + gradient 22.3 -> 21.5 us with 8 rows per tensor, gradient-only
// 17.5 -> 16.4, forward 14.6 -> 14.3; 16 or more rows per CTA compete with the previous launch's stragglers for DRAM
// and give the gain back (24 rows: 22.8 / 17.6 / 15.5).  Default 8 rows (two ring slots).
__device__ __forceinline__ void l2_prefetch(const void *p, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}
__device__ __forceinline__ void prefetch_first_item(const PhysArgs &a, const RingGeom &g, int halo, int lane) {
    if ((int)blockIdx.x >= g.items) return;
    const llie_phys_desc &d = a.d;
    const Item it = decode_item(blockIdx.x, g, d.c_out, d.h);
    const int cs = it.x0 - 4 < 0 ? 0 : it.x0 - 4;
    const int ce = it.x0 + kCols + 4 > d.w ? d.w : it.x0 + kCols + 4;
    const uint32_t bytes = (uint32_t)(ce - cs) * 4u;
    const int r0 = max(it.ya - halo, 0), r1 = min(min(it.yb + halo, d.h), r0 + g.prefetch);
    const float *pp = a.pred + (long long)it.plane * d.h * d.w + cs;
    const float *op = a.obs + (long long)it.plane * d.h * d.w + cs;
    for (int r = r0 + lane; r < r1; r += 32) {
        l2_prefetch(pp + (long long)r * d.w, bytes);
        l2_prefetch(op + (long long)r * d.w, bytes);
    }
}

// CTA epilogue: OR of the per-thread flags (one barrier), thread 0 publishes the CTA's flag word and takes a ticket;
// returns true in ALL threads of the last CTA to arrive.  `early`: thread 0 first waits for the previous grid (the
// workspace is shared with it; a no-op when it already waited for this CTA's per-item partial).
__device__ __forceinline__ bool ring_publish(int flag, int *pflag, unsigned *ticket, unsigned total_blocks, bool early) {
    __shared__ bool s_last;
    const int f = __syncthreads_or(flag);  // also orders thread 0's partial stores of the item loop before the fence
    if (threadIdx.x == 0) {
        if (early) pdl_wait();
        pflag[blockIdx.x] = f;
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

// last CTA: fixed-order per-sample sums of the per-item partials.  The partials (and flags) are fetched with ALL
// loads of a thread in flight together - one L2 round trip instead of one per 32 partials - and staged in the (now
// idle) ring shared memory; the per-sample sums then read shared memory, one warp per sample.  The summation order
// only depends on the launch geometry.
constexpr int kStageMax = (int)(kRingSmem / sizeof(double));
__device__ void ring_finalize(const PhysArgs &a, const RingGeom &g, int *red_i, double *red_d, double *stage) {
    constexpr int kSmemSamples = 64;
    __shared__ double s_samp[kSmemSamples];
    const int N = a.d.n;
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const int NT = blockDim.x;
    const int per = g.items_per_sample;
    double *samp = N <= kSmemSamples ? s_samp : g.samp;
    int fl = 0;
    const bool staged = g.items <= kStageMax;
    {
        // the first batch of partial loads is issued BEFORE the flag words are consumed: partials and flags then share
        // one L2 round trip instead of two on the serial tail of the launch
        double v[8];
        const int bp = threadIdx.x;
        if (staged) {
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = bp + k * NT < g.items ? __ldcg(g.part + bp + k * NT) : 0.0;
        }
        int fv[4];
        for (int b0 = threadIdx.x; b0 < (int)gridDim.x; b0 += 4 * NT) {
#pragma unroll
            for (int k = 0; k < 4; ++k) fv[k] = b0 + k * NT < (int)gridDim.x ? __ldcg(g.pflag + b0 + k * NT) : 0;
            fl |= (fv[0] | fv[1]) | (fv[2] | fv[3]);
        }
        if (staged) {
#pragma unroll
            for (int k = 0; k < 8; ++k)
                if (bp + k * NT < g.items) stage[bp + k * NT] = v[k];
            for (int b0 = threadIdx.x + 8 * NT; b0 < g.items; b0 += 8 * NT) {
#pragma unroll
                for (int k = 0; k < 8; ++k) v[k] = b0 + k * NT < g.items ? __ldcg(g.part + b0 + k * NT) : 0.0;
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    if (b0 + k * NT < g.items) stage[b0 + k * NT] = v[k];
            }
            __syncthreads();
        }
    }
    RING_STAMP(threadIdx.x == 0, 10);
    const double *src = staged ? stage : g.part;
    if (staged || N >= nw) {
        for (int n = wrp; n < N; n += nw) {
            const double *p = src + (size_t)n * per;
            double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
            int b = lane;
            for (; b + 96 < per; b += 128) {
                s0 += p[b];
                s1 += p[b + 32];
                s2 += p[b + 64];
                s3 += p[b + 96];
            }
            for (; b < per; b += 32) s0 += p[b];
            const double s = warp_sum((s0 + s1) + (s2 + s3));
            if (lane == 0) samp[n] = s;
        }
    } else {
        for (int n = 0; n < N; ++n) {
            const double *p = g.part + (size_t)n * per;
            double acc[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[k] = 0.0;
            int b = threadIdx.x;
            for (; b + 7 * NT < per; b += 8 * NT) {
#pragma unroll
                for (int k = 0; k < 8; ++k) acc[k] += __ldcg(p + b + k * NT);
            }
            for (; b < per; b += NT) acc[0] += __ldcg(p + b);
            const double s = block_sum<double, false>(((acc[0] + acc[1]) + (acc[2] + acc[3])) +
                                                          ((acc[4] + acc[5]) + (acc[6] + acc[7])), red_d);
            if (threadIdx.x == 0) samp[n] = s;
        }
    }
    const int ff = block_or(fl, red_i);  // contains the __syncthreads that publishes samp[]
    RING_STAMP(threadIdx.x == 0, 11);
    if (threadIdx.x == 0) {
        double batch = 0.0, total = 0.0;
        int f = ff;
        for (int n = 0; n < N; ++n) {
            const double s = samp[n];
            // an un-clipped non-finite input surfaces as a non-finite sum (source unknown: bit 2)
            if ((__double2hiint(s) & 0x7ff00000) == 0x7ff00000) f |= 4;
            const float m = (float)(s * a.inv_count);
            a.out_mean[n] = m;
            if (a.out_sum) a.out_sum[n] = s;
            batch += (double)m;
            total += s;
        }
        a.out_mean[N] = (float)(batch / N);
        a.out_mean[N + 1] = (float)batch;
        if (a.out_sum) a.out_sum[N] = total;
        *a.out_flags = f;
    }
}

// per-thread shared-memory indices of the horizontal neighbours (see header comment)
__device__ __forceinline__ void edge_indices(int tid, bool is_left, bool is_right, int mode, int &lidx, int &ridx) {
    lidx = 4 * tid + 3;
    ridx = 4 * tid + 8;
    if (is_left) lidx = mode == LLIE_PAD_REFLECT ? 4 * tid + 5 : (mode == LLIE_PAD_REPLICATE ? 4 * tid + 4 : kRowD);
    if (is_right) ridx = mode == LLIE_PAD_REFLECT ? 4 * tid + 6 : (mode == LLIE_PAD_REPLICATE ? 4 * tid + 7 : kRowD);
}

// ------------------------------------------------------------------------------------------------
// forward.  Entry e of an item: pred row ya-1+e (mapped), obs row ya-2+e (for e >= 2); it completes
// output row ya-2+e.  E = rows + 2 entries, kG entries per slot.
// ------------------------------------------------------------------------------------------------
template <bool ROBUST, bool CLIPY, bool CLIPA>
__global__ void __launch_bounds__(kRingThreads, kCtasPerSM) phys_fwd_ring(const PhysArgs a, const RingGeom g) {
    extern __shared__ __align__(128) unsigned char smraw[];
    float *ring = reinterpret_cast<float *>(smraw);
    const float *zrow = ring + kSlots * kSlotF;
    __shared__ uint64_t full[kSlots], empty[kSlots];
    __shared__ int red_i[32];
    __shared__ double red_d[32];
    __shared__ double red_w[8];
    const llie_phys_desc &d = a.d;
    const int H = d.h, W = d.w, mode = d.pad_mode;
    const int tid = threadIdx.x, lane = tid & 31, wrp = tid >> 5;
    pdl_launch_dependents();  // the next kernel's CTAs may run their prologue while this grid drains
    if (!g.early && g.prefetch && wrp == kCons / 32) prefetch_first_item(a, g, 2, lane);
    ring_setup(ring, full, empty, (W % kCols) != 0);
    if (!g.early) pdl_wait();  // nothing before this line reads or writes global memory (the L2 prefetch has no effect on it)
    int iter = 0;

    float chk_p = 0.f, chk_o = 0.f;  // x*0 accumulators: NaN iff a non-finite value went through a clip
    RingState rs{0, 0u};

    if (wrp == kCons / 32) {
        // ================= producer (one elected lane) =================
        if (lane == 0) {
            for (int item = blockIdx.x; item < g.items; item += gridDim.x) {
                const Item it = decode_item(item, g, d.c_out, H);
                const float *pplane = a.pred + (long long)it.plane * H * W;
                const float *oplane = a.obs + (long long)it.plane * H * W;
                const int E = it.yb - it.ya + 2;
                for (int e0 = 0; e0 < E; e0 += kG) {
                    SlotRows sr;
#pragma unroll
                    for (int u = 0; u < kG; ++u) {
                        const int e = e0 + u;
                        sr.pr[u] = e < E ? src_row(it.ya - 1 + e, H, mode) : -1;
                        sr.orow[u] = (e < E && e >= 2) ? it.ya - 2 + e : -1;
                    }
                    mbar_wait(&empty[rs.slot], rs.phase ^ 1u);
                    produce_slot(ring + rs.slot * kSlotF, &full[rs.slot], pplane, oplane, sr, it.x0, W);
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
            int lidx, ridx;
            edge_indices(tid, xr == 0, xr + 4 == W, mode, lidx, ridx);
            const float rho = (a.rho == nullptr)
                                  ? d.rho_scalar
                                  : __ldg(a.rho + (long long)it.n * d.rho_sn + (long long)it.c * d.rho_sc);
            const float ry = d.rho_side == 0 ? rho : 1.f, ra = d.rho_side == 1 ? rho : 1.f;
            float w[9];  // rho on the conv side is folded into the weights
#pragma unroll
            for (int q = 0; q < 9; ++q) w[q] = __ldg(a.psf + it.c * 9 + q) * ry;
            const float eps2 = d.eps * d.eps;
            float cm[4];  // 1 where the column belongs to the compared domain of an active lane
#pragma unroll
            for (int i = 0; i < 4; ++i) cm[i] = (active && xr + i >= a.x_lo && xr + i < a.x_lo + a.wo) ? 1.f : 0.f;
            const bool map_on = a.out_map != nullptr && active;
            const bool map_vec = (a.wo & 3) == 0 && a.x_lo == 0;
            f2_t sA[2] = {0ull, 0ull}, sB[2] = {0ull, 0ull};  // packed rolling sums (columns 0|1, 2|3)
            float acc = 0.f;
            const int E = it.yb - it.ya + 2;
            // one slot = kG entries.  GENERAL handles partial slots, zero-padding rows and cropped rows; the
            // interior slots of an item (all rows real, in-domain) take the lean path; WARM is the first slot
            // of an item that does not touch the plane's top edge: its first two entries only feed the
            // rolling conv sums (no residual), the other two are interior entries.
            auto do_slot = [&](auto kind_tag, int e0, const float *sp) {
                constexpr int KIND = decltype(kind_tag)::value;
                constexpr bool GENERAL = KIND == kSlotGeneral, WARM = KIND == kSlotWarm;
#pragma unroll
                for (int u = 0; u < kG; ++u) {
                    const int e = e0 + u;
                    if (GENERAL && e >= E) break;  // CTA-uniform (only the last slot of an item is partial)
                    const int y = it.ya - 2 + e;   // output row completed by this entry (pred row y+1)
                    const float *rowp = sp + u * kPairF;
                    const float *pz = rowp;
                    float rowmask = 1.f;
                    if (GENERAL) {
                        if (src_row(y + 1, H, mode) < 0) pz = zrow;
                        rowmask = (e >= 2 && y >= a.y_lo && y < a.y_lo + a.ho) ? 1.f : 0.f;
                    }
                    const float4 v = reinterpret_cast<const float4 *>(pz)[tid + 1];
                    const float r6[6] = {pz[lidx], v.x, v.y, v.z, v.w, pz[ridx]};
                    const float *so = (GENERAL && e < 2) ? zrow : rowp + kRowF;  // no obs row travels with warm-up entries
                    const float4 ob = reinterpret_cast<const float4 *>(so)[tid + 1];
                    float cv[4];
                    roll3<4>(r6, w, sA, sB, cv);
                    if (WARM && u < 2) continue;  // compile-time after unrolling: cv and the obs row are dead
                    const float av[4] = {ob.x, ob.y, ob.z, ob.w};
                    float mv[4], racc = 0.f;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        float yv = cv[i], t = av[i] * ra;
                        if (CLIPY) {
                            chk_p = fmaf(yv, 0.f, chk_p);
                            yv = clamp01(yv);
                        }
                        if (CLIPA) {
                            chk_o = fmaf(t, 0.f, chk_o);
                            t = clamp01(t);
                        }
                        const float r = yv - t;
                        mv[i] = fabsf(r);
                        float pen = mv[i];
                        if (ROBUST) {
                            float rs_unused;
                            sqrt_rsqrt_pos(fmaf(r, r, eps2), pen, rs_unused);
                        }
                        racc = fmaf(pen, cm[i], racc);
                    }
                    acc = GENERAL ? fmaf(racc, rowmask, acc) : acc + racc;
                    if (map_on && rowmask != 0.f) {
                        float *mrow = a.out_map + (((long long)it.plane) * a.ho + (y - a.y_lo)) * a.wo;
                        if (map_vec) {
                            *reinterpret_cast<float4 *>(mrow + xr) = make_float4(mv[0], mv[1], mv[2], mv[3]);
                        } else {
#pragma unroll
                            for (int i = 0; i < 4; ++i)
                                if (cm[i] != 0.f) mrow[xr + i - a.x_lo] = mv[i];
                        }
                    }
                }
            };
            // entries [e_lo, e_hi) are interior: real pred rows, e >= 2, output row inside the domain
            const int y_first = it.ya > a.y_lo ? it.ya : a.y_lo;  // first compared output row of the item
            int y_end = it.yb < a.y_lo + a.ho ? it.yb : a.y_lo + a.ho;
            if (y_end > H - 1) y_end = H - 1;                      // pred row y+1 must be a real row
            const int e_lo = y_first - it.ya + 2, e_hi = y_end - it.ya + 2;
            // warm-up slot: pred rows ya-1 .. ya+2 real, output rows ya, ya+1 compared
            const bool warm_ok = it.ya >= 1 && it.ya + 2 <= H - 1 && it.ya >= a.y_lo && it.ya + 1 < a.y_lo + a.ho;
            for (int e0 = 0; e0 < E; e0 += kG) {
                const float *sp = ring + rs.slot * kSlotF;
                mbar_wait(&full[rs.slot], rs.phase);
                if (e0 >= e_lo && e0 + kG <= e_hi)
                    do_slot(SlotKind<kSlotInterior>{}, e0, sp);
                else if (e0 == 0 && warm_ok)
                    do_slot(SlotKind<kSlotWarm>{}, e0, sp);
                else
                    do_slot(SlotKind<kSlotGeneral>{}, e0, sp);
                __syncwarp();
                if (lane == 0) mbar_arrive(&empty[rs.slot]);
                rs.advance();
            }
            // one partial per item: the 4 consumer warps meet on named barrier 1 (red_w is double-
            // buffered by item parity, so one barrier per item suffices)
            const double ws = warp_sum((double)acc);
            double *rw = red_w + (iter & 1) * 4;
            if (lane == 0) rw[wrp] = ws;
            asm volatile("bar.sync 1, %0;" ::"n"(kCons) : "memory");
            if (tid == 0) {
                if (g.early && iter == 0) pdl_wait();  // the workspace is shared with the previous launch
                g.part[item] = (rw[0] + rw[1]) + (rw[2] + rw[3]);
            }
            ++iter;
        }
    }

    int flag = 0;
    if (!(chk_p == 0.f)) flag |= 1;
    if (!(chk_o == 0.f)) flag |= 2;
    if (ring_publish(flag, g.pflag, a.ticket, gridDim.x, g.early != 0))
        ring_finalize(a, g, red_i, red_d, reinterpret_cast<double *>(smraw));
}

// ------------------------------------------------------------------------------------------------
// backward: grad = padY^T K^T q,  q = g_n * dLoss/dconv on the compared domain (0 elsewhere).
// An item emits the (virtual) grad rows ja .. jb-1; the first / last chunk of a plane also produce
// the padded rows -1 / H.  Entry k (k = 0 .. jb-ja+3) is j = ja-4+k: it carries pred row j+2 and obs
// row j+1, completes conv / q row j+1 and grad row j.  The first two entries only warm up the conv
// partial sums (q forced to 0 by the row mask), the next two the grad partial sums.
// ------------------------------------------------------------------------------------------------
template <bool ROBUST, bool CLIPY, bool CLIPA, bool FUSED = false>
__global__ void __launch_bounds__(kRingThreads, kCtasPerSM) phys_bwd_ring(const PhysArgs a, const RingGeom g) {
    extern __shared__ __align__(128) unsigned char smraw[];
    float *ring = reinterpret_cast<float *>(smraw);
    const float *zrow = ring + kSlots * kSlotF;
    __shared__ uint64_t full[kSlots], empty[kSlots];
    // FUSED (value + gradient in one pass): the loss of the residual rows an item OWNS (qr in [ya, yb), centre
    // four columns of each lane) is reduced exactly like in phys_fwd_ring
    __shared__ int red_i[FUSED ? 32 : 1];
    __shared__ double red_d[FUSED ? 32 : 1];
    __shared__ double red_w[FUSED ? 8 : 1];
    int iter = 0;
    float chk_p = 0.f, chk_o = 0.f;
    const llie_phys_desc &d = a.d;
    const int H = d.h, W = d.w, mode = d.pad_mode;
    const int tid = threadIdx.x, lane = tid & 31, wrp = tid >> 5;
    const bool fold = mode != LLIE_PAD_ZEROS;
    const int fold_top = mode == LLIE_PAD_REFLECT ? 1 : 0;          // image row that padded row -1 reads
    const int fold_bot = mode == LLIE_PAD_REFLECT ? H - 2 : H - 1;  // image row that padded row H reads
    RING_STAMP(threadIdx.x == 0, 0);
    pdl_launch_dependents();
    if (!g.early && g.prefetch && wrp == kCons / 32) prefetch_first_item(a, g, 4, lane);
    ring_setup(ring, full, empty, (W % kCols) != 0);
    RING_STAMP(threadIdx.x == 0, 1);
    if (!g.early) pdl_wait();  // nothing before this line reads or writes global memory (the L2 prefetch has no effect on it)
    RING_STAMP(threadIdx.x == 0, 2);
    RingState rs{0, 0u};

    if (wrp == kCons / 32) {
        if (lane == 0) {
            for (int item = blockIdx.x; item < g.items; item += gridDim.x) {
                const Item it = decode_item(item, g, d.c_out, H);
                const float *pplane = a.pred + (long long)it.plane * H * W;
                const float *oplane = a.obs + (long long)it.plane * H * W;
                const int ja = (it.first && fold) ? -1 : it.ya;
                const int jb = (it.last && fold) ? H + 1 : it.yb;
                const int E = jb - ja + 4;
                for (int k0 = 0; k0 < E; k0 += kG) {
                    SlotRows sr;
#pragma unroll
                    for (int u = 0; u < kG; ++u) {
                        const int k = k0 + u, j = ja - 4 + k, qr = j + 1;
                        sr.pr[u] = k < E ? src_row(j + 2, H, mode) : -1;
                        sr.orow[u] = (k < E && k >= 2 && qr >= a.y_lo && qr < a.y_lo + a.ho) ? qr : -1;
                    }
                    mbar_wait(&empty[rs.slot], rs.phase ^ 1u);
                    produce_slot(ring + rs.slot * kSlotF, &full[rs.slot], pplane, oplane, sr, it.x0, W);
                    rs.advance();
                }
            }
        }
        if (!FUSED) return;
    }

    for (int item = blockIdx.x; wrp < kCons / 32 && item < g.items; item += gridDim.x) {
        const Item it = decode_item(item, g, d.c_out, H);
        const int xr = it.x0 + tid * 4;
        const bool active = xr < W;
        const int x = active ? xr : W - 4;
        const bool is_left = xr == 0, is_right = xr + 4 == W;
        int lidx, ridx;  // smem index of column x-1 / x+4 (own pixel or the zero word on image-edge lanes)
        edge_indices(tid, is_left, is_right, mode, lidx, ridx);
        const int ja = (it.first && fold) ? -1 : it.ya;
        const int jb = (it.last && fold) ? H + 1 : it.yb;
        float *gp = a.grad + (long long)it.plane * H * W + x;
        const float rho = (a.rho == nullptr)
                              ? d.rho_scalar
                              : __ldg(a.rho + (long long)it.n * d.rho_sn + (long long)it.c * d.rho_sc);
        const float ry = d.rho_side == 0 ? rho : 1.f, ra = d.rho_side == 1 ? rho : 1.f;
        float w[9], wf[9];  // rho on the conv side is folded into the weights (forward and adjoint)
#pragma unroll
        for (int q = 0; q < 9; ++q) w[q] = __ldg(a.psf + it.c * 9 + q) * ry;
#pragma unroll
        for (int q = 0; q < 9; ++q) wf[q] = w[8 - q];
        const float eps2 = d.eps * d.eps;
        const float gs = __ldg(a.gout + a.gout_stride * it.n) * a.gscale;
        float qm[6];  // gs where column x-1+i is inside the compared domain, else 0
#pragma unroll
        for (int i = 0; i < 6; ++i)
            qm[i] = (active && (xr - 1 + i >= a.x_lo) && (xr - 1 + i < a.x_lo + a.wo)) ? gs : 0.f;
        float dm[4];  // FUSED: 1 where centre column x+i is inside the compared domain (independent of gs)
#pragma unroll
        for (int i = 0; i < 4; ++i) dm[i] = (active && (xr + i >= a.x_lo) && (xr + i < a.x_lo + a.wo)) ? 1.f : 0.f;
        float lacc = 0.f;
        // edge-fold weights (image-edge lanes): column -1 gets K[.][0] q(., 0), column W gets K[.][2] q(., W-1)
        const float k0 = is_left ? w[0] : w[2], k1 = is_left ? w[3] : w[5], k2 = is_left ? w[6] : w[8];
        const int tgt = is_left ? (mode == LLIE_PAD_REFLECT ? 1 : 0) : (mode == LLIE_PAD_REFLECT ? 2 : 3);
        const bool edge_fold = fold && (is_left || is_right);

        f2_t cA[3] = {0ull, 0ull, 0ull}, cB[3] = {0ull, 0ull, 0ull};  // packed rolling sums (columns 0|1, 2|3, 4|5)
        f2_t gA[2] = {0ull, 0ull}, gB[2] = {0ull, 0ull};
        float eA = 0.f, eB = 0.f;  // partial sums of the folded padded column (image-edge lanes only)
        float carry[4] = {0.f, 0.f, 0.f, 0.f};
        const int E = jb - ja + 4;

        // GENERAL slots: partial slots, zero-padding rows, warm-up entries, cropped q rows, virtual rows;
        // interior slots (all pred rows real, q row compared, grad row a plain image row) run lean.
        // WARM: the first slot of an item away from the plane's top edge (entries 0..3 = exactly the warm-up):
        // entries 0, 1 only feed the rolling conv sums, entries 2, 3 the rolling grad sums; nothing is stored.
        auto do_slot = [&](auto kind_tag, int kb, const float *sp) {
            constexpr int KIND = decltype(kind_tag)::value;
            constexpr bool GENERAL = KIND == kSlotGeneral, WARM = KIND == kSlotWarm;
            float *gslot = gp + (long long)(ja - 4 + kb) * W;  // grad row of entry kb; one 64-bit product per slot
#pragma unroll(!FUSED || WARM ? kG : (GENERAL ? kUnrollG : kUnrollI))
            for (int u = 0; u < kG; ++u, gslot += W) {
                const int k = kb + u;
                if (GENERAL && k >= E) break;  // CTA-uniform
                const int j = ja - 4 + k, qr = j + 1;
                const float *rowp = sp + u * kPairF;
                const float *pz = rowp;
                float rowm = 1.f;
                if (GENERAL) {
                    if (src_row(j + 2, H, mode) < 0) pz = zrow;
                    rowm = (k >= 2 && qr >= a.y_lo && qr < a.y_lo + a.ho) ? 1.f : 0.f;
                }
                const float4 v = reinterpret_cast<const float4 *>(pz)[tid + 1];
                const float r8[8] = {pz[4 * tid + 2], pz[lidx], v.x, v.y, v.z, v.w, pz[ridx], pz[4 * tid + 9]};
                const float *so = (GENERAL && rowm == 0.f) ? zrow : rowp + kRowF;  // obs travels with compared rows only
                const float4 ob = reinterpret_cast<const float4 *>(so)[tid + 1];
                const float av[6] = {so[4 * tid + 3], ob.x, ob.y, ob.z, ob.w, so[4 * tid + 8]};
                float cv[6];
                roll3<6>(r8, w, cA, cB, cv);  // conv row qr over columns x-1 .. x+4
                if (WARM && u < 2) continue;  // compile-time after unrolling: cv and the obs row are dead
                float qc[6];
                float lrow = 0.f;
#pragma unroll
                for (int i = 0; i < 6; ++i) {
                    float yv = cv[i], t = av[i] * ra;
                    float m = GENERAL ? qm[i] * rowm : qm[i];
                    if (CLIPY) {
                        if (FUSED && i >= 1 && i <= 4) chk_p = fmaf(yv, 0.f, chk_p);
                        m = (yv >= 0.f && yv <= 1.f) ? m : 0.f;  // torch.clamp backward passes lo <= y <= hi
                        yv = clamp01(yv);
                    }
                    if (CLIPA) {
                        if (FUSED && i >= 1 && i <= 4) chk_o = fmaf(t, 0.f, chk_o);
                        t = clamp01(t);
                    }
                    const float r = yv - t;
                    float pen = fabsf(r), rsq = 0.f;
                    if (ROBUST) sqrt_rsqrt_pos(fmaf(r, r, eps2), pen, rsq);
                    if (FUSED && !(WARM && u == 2) && i >= 1 && i <= 4) lrow = fmaf(pen, dm[i - 1], lrow);
                    if (ROBUST) {
                        qc[i] = (m != 0.f) ? m * (r * rsq) : 0.f;
                    } else {
                        // m * sign(r), sign(0) = 0 (stale rows hold finite values; m == 0 masks them)
                        qc[i] = (r != 0.f) ? copysignf(m, r) : 0.f;
                    }
                }
                if (FUSED) {  // rows outside [ya, yb) are recomputed halo rows owned by the neighbouring item
                    if (WARM) {
                        if (u == 3) lacc += lrow;  // q row ya; entry 2 is row ya - 1 of the item above
                    } else {
                        lacc = fmaf(lrow, (rowm != 0.f && qr >= it.ya && qr < it.yb) ? 1.f : 0.f, lacc);
                    }
                }
                // ---- grad (virtual) row j: q row j+1 completes it (flipped kernel, rolling form)
                float gr[4];
                roll3<4>(qc, wf, gA, gB, gr);
                if (edge_fold) {
                    const float qe = is_left ? qc[1] : qc[4];
                    const float ecol = fmaf(k0, qe, eA);
                    eA = fmaf(k1, qe, eB);
                    eB = k2 * qe;
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        if (i == tgt) gr[i] += ecol;
                }
                if (KIND == kSlotInterior) {
                    if (active) *reinterpret_cast<float4 *>(gslot) = make_float4(gr[0], gr[1], gr[2], gr[3]);
                } else if (GENERAL && k >= 4) {  // CTA-uniform: entries 0..3 are warm-up
                    if (j == -1) {
#pragma unroll
                        for (int i = 0; i < 4; ++i) carry[i] = gr[i];
                    } else if (j == H) {
                        if (active) {
                            float4 *dst = reinterpret_cast<float4 *>(gp + (long long)fold_bot * W);
                            float4 t = *dst;  // written by this very thread two / one entries ago
                            t.x += gr[0]; t.y += gr[1]; t.z += gr[2]; t.w += gr[3];
                            *dst = t;
                        }
                    } else {
                        if (it.first && fold && j == fold_top) {
#pragma unroll
                            for (int i = 0; i < 4; ++i) gr[i] += carry[i];
                        }
                        if (active)
                            *reinterpret_cast<float4 *>(gp + (long long)j * W) =
                                make_float4(gr[0], gr[1], gr[2], gr[3]);
                    }
                }
            }
        };
        // interior entries: k >= 4, grad row j in [max(ja, fold_top + 1), ...), q row j+1 compared and
        // pred row j+2 a real image row:  j >= max(0, y_lo - 1, fold_top + 1 if first),  j + 2 <= H - 1,
        // j + 1 < y_lo + ho
        int j_lo = ja + 0 > 0 ? ja : 0;
        if (j_lo < a.y_lo - 1) j_lo = a.y_lo - 1;
        if (it.first && fold && j_lo < fold_top + 1) j_lo = fold_top + 1;
        int j_hi = jb < H - 2 ? jb : H - 2;  // exclusive
        if (j_hi > a.y_lo + a.ho - 1) j_hi = a.y_lo + a.ho - 1;
        const int k_lo = (j_lo - ja + 4) > 4 ? (j_lo - ja + 4) : 4, k_hi = j_hi - ja + 4;

        // warm-up slot: pred rows ja-2 .. ja+1 real (so ja == ya: not the folded first chunk), q rows ja-1, ja compared
        const bool warm_ok = ja >= 2 && ja + 1 <= H - 1 && ja - 1 >= a.y_lo && ja < a.y_lo + a.ho;
        for (int kb = 0; kb < E; kb += kG) {
            const float *sp = ring + rs.slot * kSlotF;
            mbar_wait(&full[rs.slot], rs.phase);
            RING_STAMP(tid == 0 && kb == 0 && item == (int)blockIdx.x, 3);
            if (kb >= k_lo && kb + kG <= k_hi)
                do_slot(SlotKind<kSlotInterior>{}, kb, sp);
            else if (kb == 0 && warm_ok)
                do_slot(SlotKind<kSlotWarm>{}, kb, sp);
            else
                do_slot(SlotKind<kSlotGeneral>{}, kb, sp);
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[rs.slot]);
            rs.advance();
        }
        RING_STAMP(tid == 0, 4);
        if (FUSED) {
            const double ws = warp_sum((double)lacc);
            double *rw = red_w + (iter & 1) * 4;
            if (lane == 0) rw[wrp] = ws;
            asm volatile("bar.sync 1, %0;" ::"n"(kCons) : "memory");
            if (tid == 0) {
                if (g.early && iter == 0) pdl_wait();  // the workspace is shared with the previous launch
                g.part[item] = (rw[0] + rw[1]) + (rw[2] + rw[3]);
            }
            ++iter;
        }
    }
    // gradient-only launches touch no workspace; the wait keeps grid completion transitively ordered
    if (!FUSED && g.early && tid == 0) pdl_wait();
    if (FUSED) {
        int flag = 0;
        if (!(chk_p == 0.f)) flag |= 1;
        if (!(chk_o == 0.f)) flag |= 2;
        RING_STAMP(tid == 0, 5);
        const bool last = ring_publish(flag, g.pflag, a.ticket, gridDim.x, g.early != 0);
        RING_STAMP(tid == 0, 6);
        if (last) ring_finalize(a, g, red_i, red_d, reinterpret_cast<double *>(smraw));
        RING_STAMP(tid == 0 && last, 7);
    }
}

// ------------------------------------------------------------------------------------------------
// host
// ------------------------------------------------------------------------------------------------
// Diagnostic knobs (llie_tune): relaxed atomics, so that a sweep thread and a launching thread never race on them.
static std::atomic<int> g_ring_rows{0};  // 0 = heuristic (LLIE_TUNE_PHYS_ROWS)
void phys_ring_set_rows(int r) { g_ring_rows.store(r, std::memory_order_relaxed); }
static std::atomic<int> g_ring_pdl{1};   // programmatic dependent launch of the ring kernels (LLIE_TUNE_PHYS_PDL)
void phys_ring_set_pdl(int on) { g_ring_pdl.store(on ? 1 : 0, std::memory_order_relaxed); }
static std::atomic<int> g_ring_prefetch{8};  // pre-wait L2 prefetch (LLIE_TUNE_PHYS_PREFETCH)
void phys_ring_set_prefetch(int rows) { g_ring_prefetch.store(rows < 0 ? 0 : rows, std::memory_order_relaxed); }
// Stream-overlap contract (llie_phys_desc.stream_overlap, a PER-CALL promise): no buffer of this ring launch except the
// workspace is read or written by the kernel launched right before it on the same stream.  The kernels then wait for
// the previous grid only where they touch the workspace, so the CTAs of launch i + 1 stream and compute on the SMs
// that the early finishers of launch i have left (and under its last-CTA finalize).

static RingGeom make_geom(const llie_phys_desc &d, int &grid) {
    RingGeom g{};
    g.groups = div_up(d.w, kCols);
    const long long cols = (long long)d.n * d.c_out * g.groups;  // independent column strips
    const int slots = kNumSM * kCtasPerSM;
    int chunks;
    const int tune_rows = g_ring_rows.load(std::memory_order_relaxed);
    const bool overlap = g_ring_pdl.load(std::memory_order_relaxed) && d.stream_overlap != 0;
    if (tune_rows > 0) {
        chunks = d.h / tune_rows;
    } else if (overlap) {
        // stream-overlap loops: several launches are resident at once (launch i + 1 starts as soon as every CTA of
        // launch i has started), so the machine is filled ACROSS launches and a CTA should live long: ~128-row items
        // halve the warm-up rows and amortise the per-CTA ramp (setup, first TMA round trip, epilogue ~ 5 us) that a
        // one-item CTA of ~21 rows pays once per 13 us of work (C2 sweep: 17.7 us per fused launch at 21 rows,
        // 13.5 at 64, 12.9 at 128..256, 13.9 at 512)
        chunks = (d.h + 64) / 128;
    } else {
        // items = k * (resident CTA slots): every persistent CTA gets the same number of items (no
        // tail wave) and the rows of a plane are split evenly; k aims at ~48-row chunks (warm-up rows
        // cost 2 (forward) / 4 (backward) extra row reads per chunk), but never leaves slots idle
        const long long per_k = (long long)slots * 48;
        long long k = ((long long)d.h * cols + per_k / 2) / per_k;
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
    g.early = overlap ? 1 : 0;
    g.prefetch = g_ring_pdl.load(std::memory_order_relaxed) ? g_ring_prefetch.load(std::memory_order_relaxed) : 0;
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
    const size_t need = (size_t)g.items * sizeof(double) + (size_t)d.n * sizeof(double) + (size_t)grid * sizeof(int);
    return need <= kWsBytes - kWsHeaderBytes;
}

#define LLIE_RING_ONE(K, GRID, ST, ...)                                                                          \
    do {                                                                                                         \
        auto kern = K;                                                                                           \
        LLIE_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kRingSmem));   \
        cudaLaunchConfig_t cfg = {};                                                                             \
        cfg.gridDim = dim3(GRID);                                                                                \
        cfg.blockDim = dim3(kRingThreads);                                                                       \
        cfg.dynamicSmemBytes = kRingSmem;                                                                        \
        cfg.stream = ST;                                                                                         \
        cudaLaunchAttribute attr[1];                                                                             \
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;                                         \
        attr[0].val.programmaticStreamSerializationAllowed = g_ring_pdl.load(std::memory_order_relaxed);                                         \
        cfg.attrs = attr;                                                                                        \
        cfg.numAttrs = 1;                                                                                        \
        LLIE_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, __VA_ARGS__));                                               \
    } while (0)
#define LLIE_RING_DISPATCH(KERNEL, d, GRID, ST, ...)                                               \
    do {                                                                                           \
        const int key = ((d).robust ? 4 : 0) | ((d).clip_y ? 2 : 0) | ((d).clip_a ? 1 : 0);        \
        switch (key) {                                                                             \
            case 0: LLIE_RING_ONE((KERNEL<false, false, false>), GRID, ST, __VA_ARGS__); break;    \
            case 1: LLIE_RING_ONE((KERNEL<false, false, true>), GRID, ST, __VA_ARGS__); break;     \
            case 2: LLIE_RING_ONE((KERNEL<false, true, false>), GRID, ST, __VA_ARGS__); break;     \
            case 3: LLIE_RING_ONE((KERNEL<false, true, true>), GRID, ST, __VA_ARGS__); break;      \
            case 4: LLIE_RING_ONE((KERNEL<true, false, false>), GRID, ST, __VA_ARGS__); break;     \
            case 5: LLIE_RING_ONE((KERNEL<true, false, true>), GRID, ST, __VA_ARGS__); break;      \
            case 6: LLIE_RING_ONE((KERNEL<true, true, false>), GRID, ST, __VA_ARGS__); break;      \
            default: LLIE_RING_ONE((KERNEL<true, true, true>), GRID, ST, __VA_ARGS__); break;      \
        }                                                                                          \
    } while (0)

int launch_phys_fwd_ring(PhysArgs &a, void *workspace, cudaStream_t st) {
    int grid;
    RingGeom g = make_geom(a.d, grid);
    Workspace ws(workspace);
    g.part = reinterpret_cast<double *>(ws.scratch);
    g.samp = g.part + (size_t)g.items;
    g.pflag = reinterpret_cast<int *>(g.samp + a.d.n);
    a.ticket = ws.tickets;
    LLIE_RING_DISPATCH(phys_fwd_ring, a.d, grid, st, a, g);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

#define LLIE_RING_DISPATCH_FUSED(d, GRID, ST, ...)                                                            \
    do {                                                                                                      \
        const int key = ((d).robust ? 4 : 0) | ((d).clip_y ? 2 : 0) | ((d).clip_a ? 1 : 0);                   \
        switch (key) {                                                                                        \
            case 0: LLIE_RING_ONE((phys_bwd_ring<false, false, false, true>), GRID, ST, __VA_ARGS__); break;  \
            case 1: LLIE_RING_ONE((phys_bwd_ring<false, false, true, true>), GRID, ST, __VA_ARGS__); break;   \
            case 2: LLIE_RING_ONE((phys_bwd_ring<false, true, false, true>), GRID, ST, __VA_ARGS__); break;   \
            case 3: LLIE_RING_ONE((phys_bwd_ring<false, true, true, true>), GRID, ST, __VA_ARGS__); break;    \
            case 4: LLIE_RING_ONE((phys_bwd_ring<true, false, false, true>), GRID, ST, __VA_ARGS__); break;   \
            case 5: LLIE_RING_ONE((phys_bwd_ring<true, false, true, true>), GRID, ST, __VA_ARGS__); break;    \
            case 6: LLIE_RING_ONE((phys_bwd_ring<true, true, false, true>), GRID, ST, __VA_ARGS__); break;    \
            default: LLIE_RING_ONE((phys_bwd_ring<true, true, true, true>), GRID, ST, __VA_ARGS__); break;    \
        }                                                                                                     \
    } while (0)

// loss value (per-sample means / batch mean / sum, flags) AND gradient from one pass over pred / obs
int launch_phys_fused_ring(PhysArgs &a, void *workspace, cudaStream_t st) {
    int grid;
    RingGeom g = make_geom(a.d, grid);
    Workspace ws(workspace);
    g.part = reinterpret_cast<double *>(ws.scratch);
    g.samp = g.part + (size_t)g.items;
    g.pflag = reinterpret_cast<int *>(g.samp + a.d.n);
    a.ticket = ws.tickets;
    LLIE_RING_DISPATCH_FUSED(a.d, grid, st, a, g);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

int launch_phys_bwd_ring(PhysArgs &a, cudaStream_t st) {
    int grid;
    const RingGeom g = make_geom(a.d, grid);
    LLIE_RING_DISPATCH(phys_bwd_ring, a.d, grid, st, a, g);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

#ifdef LLIE_RING_TRACE
extern "C" int llie_debug_ring_trace(unsigned long long *host_dst) {
    return (int)cudaMemcpyFromSymbol(host_dst, g_ring_trace, sizeof(unsigned long long) * 1024 * 16);
}
#endif

}  // namespace llie
