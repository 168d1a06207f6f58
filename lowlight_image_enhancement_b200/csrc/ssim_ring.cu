// 11-tap separable SSIM statistics on a TMA-fed row ring with packed FP32 arithmetic (sm_100a) — the default
// kernel behind every statistics-only SSIM entry point and the DSSIM loss forward (reference functions: see ssim.cu).
//
// Shape of the computation (unchanged from ssim_stream2_kernel, ssim.cu): a CTA owns a 256-column strip x a row
// segment of one (n, c) plane; each of its 128 threads owns TWO adjacent output columns and marches down the segment;
// per input row it runs the horizontal pass of the moment maps (x, y, x^2 + y^2, xy; five maps with the variance clamp
// of ssim_linear) and scatters the results into 11 rolling vertical accumulators per map held in registers.
// What is new here:
//   * Packed FP32 (fma.rn.f32x2 / mul.rn.f32x2, SASS FFMA2 / FMUL2): the two columns of a thread share one 64-bit
//     accumulator.  Vertical pass: acc(c0|c1) += g[j] * h(c0|c1) with a broadcast weight; horizontal pass: the value
//     at window position p is a broadcast operand and the weight is the pair (g[p-1], g[p-2]) - position p is tap
//     p - 1 of the left column and tap p - 2 of the right one.  88 scalar multiply-adds per element become 46 issue
//     slots; the x^2 + y^2 / xy products are formed on the register pairs the 64-bit shared-memory loads deliver.
//   * Window taps travel in the kernel parameters (constant bank operands): no registers, no loads.
//   * Rows arrive by TMA: one lane issues two cp.async.bulk copies per row (x and y segments, 16-byte aligned
//     supersets of the strip) that complete on an mbarrier; the <= 5 + 5 padded columns of a border strip are fetched
//     by per-lane 4-byte cp.async that report to the SAME mbarrier (cp.async.mbarrier.arrive.noinc), so reflect /
//     replicate / circular padding costs the consumers nothing and zero padding is a ring that starts zeroed.  The
//     producing duty rotates over the four warps (row r is issued six rows ahead by warp r mod 4); slots are handed
//     back through a second mbarrier per slot.  No CTA-wide barrier in the row loop (`clamp_in` clamps the 14 + 14
//     loaded values in registers: an in-place clamp of the landed row needed a __syncthreads per row and ran the
//     DSSIM forward at C3 at 179 us instead of the ~120 us the unclamped kernel's rate predicts).
// Same operation order per output as ssim_stream2_kernel -> equal to it to rounding (tests: <= 1e-6 relative).
#include "ssim_common.cuh"
#include "tma.cuh"

namespace llie {
namespace {

typedef unsigned long long f2_t;
__device__ __forceinline__ f2_t pack2(float lo, float hi) {
    f2_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(f2_t v, float &lo, float &hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f2_t bc2(float v) { return pack2(v, v); }
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
__device__ __forceinline__ f2_t add2(f2_t a, f2_t b) {
    f2_t d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ f2_t ld2(const float *p) {  // 8-byte aligned shared-memory pair
    const float2 v = *reinterpret_cast<const float2 *>(p);
    return pack2(v.x, v.y);
}

// n / d for a finite d > 0 (the SSIM denominator is >= c1 * c2 > 0), branch-free: SFU reciprocal refined by one Newton
// step, quotient corrected with the exact FMA remainder (Markstein) - the correctly rounded quotient except for rare
// 1-ulp cases, against ~14 instructions + a guarded slow-path call for the IEEE division instruction sequence.
__device__ __forceinline__ float rcp_pos(float d) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(d));
    return fmaf(fmaf(-d, r, 1.f), r, r);
}
__device__ __forceinline__ float div_pos(float n, float d, float r) {
    const float q = n * r;
    return fmaf(fmaf(-d, q, n), r, q);
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
__device__ __forceinline__ void cp_async4_to(uint32_t dst, const float *src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}
// the executing thread's earlier cp.async copies arrive on `bar` when they have landed (count pre-accounted at init)
__device__ __forceinline__ void cp_async_arrive_noinc(uint64_t *bar) {
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

constexpr int RK = 11, RP = 5;
constexpr int RT = 128;                  // 4 warps, all consumers; the producing duty rotates
constexpr int RCOLS = 256;               // output columns per strip (two per thread)
constexpr int RLEAD = 8;                 // ring row index 0 <-> image column x0 - 8 (16-byte aligned in global memory)
constexpr int RROW = RCOLS + 2 * RLEAD;  // floats per ring row
constexpr int RD = 8;                    // ring depth (rows)
constexpr int RPF = 6;                   // rows issued ahead of the row being consumed
constexpr int NPOS = 14;                 // window positions of a thread: image columns gx - 6 .. gx + 7 (7 aligned pairs)
constexpr int HALO0 = 8;                 // lanes HALO0 .. HALO0 + 19 of the producing warp fetch the padded columns

struct RingTaps {       // by value in the kernel parameters -> constant-bank operands
    float2 wp[NPOS];    // position p: (tap p - 1 for column gx, tap p - 2 for column gx + 1), zero outside 0..10
    float g[RK];        // vertical taps
};

// Vertical scatter of one input row whose index is II modulo 11: the row feeds output rows o = i - j with tap j, and the
// accumulator slot of output row o is o mod 11 = (II - j) mod 11 - static here, so the 11 slots per map stay in
// registers.  The slot of the row this input completes, (II + 1) mod 11, is handed out in `done`.
// The row loop is NOT unrolled by 11 (that body is 64 KB of SASS, more than the instruction cache holds: ncu showed
// 2.6 `no_instruction` stall cycles per issued instruction, 43 % issue-active); only this scatter exists in 11 copies,
// picked by a CTA-uniform switch (0.27 `no_instruction`, 56 % issue-active, 333 -> 311 us per 3 x 2848 x 4256 pair).
template <int II, int M>
__device__ __forceinline__ void scatter_row(f2_t (&acc)[M][RK], const f2_t (&h)[M], f2_t (&done)[M], const RingTaps &tp) {
#pragma unroll
    for (int j = 0; j < RK; ++j) {
        const int sl = (II - j + RK) % RK;
        const f2_t gj = bc2(tp.g[j]);
#pragma unroll
        for (int q = 0; q < M; ++q) acc[q][sl] = (j == 0) ? mul2(gj, h[q]) : fma2(gj, h[q], acc[q][sl]);
    }
#pragma unroll
    for (int q = 0; q < M; ++q) done[q] = acc[q][(II + 1) % RK];
}

template <bool MAPS5, bool COEF, bool CLAMP, bool DUAL = false>
__global__ void __launch_bounds__(RT, MAPS5 ? 3 : 4) ssim_ring_kernel(const SsimArgs a, const RingTaps tp) {
    static_assert(!DUAL || (MAPS5 && !COEF), "the second policy is evaluated from the five-map statistics");
    constexpr int M = MAPS5 ? 5 : 4;
    __shared__ __align__(128) float ring[RD][2][RROW];
    __shared__ __align__(8) uint64_t full[RD], empty[RD];
    __shared__ double red[32];
    const llie_ssim_desc &d = a.d;
    const int H = d.h, W = d.w, tid = threadIdx.x, lane = tid & 31, wrp = tid >> 5;
    const int tpp = a.tiles_x * a.tiles_y;
    const int plane = blockIdx.x / tpp;
    const int trem = blockIdx.x - plane * tpp;
    const int seg = trem / a.tiles_x, strip = trem - seg * a.tiles_x;
    const int x0 = strip * RCOLS, y0 = seg * a.seg_rows;
    const int rows_out = min(a.seg_rows, H - y0);
    const int nrows_in = rows_out + 2 * RP;
    const long long plane_elems = (long long)H * W;
    const float *xb = a.x + (long long)plane * plane_elems;
    const float *yb = a.y + (long long)plane * plane_elems;

    // ---- producer geometry: the aligned column range of a row this strip reads, and this lane's padded column
    const int cx0 = max(x0 - RLEAD, 0), cx1 = min(x0 + RCOLS + RLEAD, W);
    const uint32_t row_bytes = (uint32_t)(cx1 - cx0) * 4u;
    const int doff = cx0 - (x0 - RLEAD);
    int hsrc = -1, hdst = 0;
    const int hl = lane - HALO0;
    const int htensor = hl >= 10 ? 1 : 0;
    if (hl >= 0 && hl < 20) {
        const int h = hl - 10 * htensor;  // 0..4 left of the image, 5..9 right of it
        const int col = h < 5 ? -1 - h : W + (h - 5);
        const bool need = h < 5 ? (x0 == 0) : (col <= x0 + RCOLS - 1 + RP);
        hsrc = need ? pad_index(col, W, d.pad_mode) : -1;
        hdst = col - (x0 - RLEAD);
    }
    const bool halo_lane = hsrc >= 0;
    const int nhalo = __popc(__ballot_sync(0xffffffffu, halo_lane));
    const float *hbase = (htensor ? yb : xb) + max(hsrc, 0);

    {   // the ring starts zeroed: columns no copy ever writes (outside the image under zero padding, beyond the
        // right edge of a partial strip) read as zero for the whole kernel
        float4 *r4 = reinterpret_cast<float4 *>(&ring[0][0][0]);
        for (int i = tid; i < RD * 2 * RROW / 4; i += RT) r4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (tid == 0) {
#pragma unroll
            for (int s = 0; s < RD; ++s) {
                mbar_init(&full[s], 1 + nhalo);
                mbar_init(&empty[s], RT / 32);
            }
            mbar_fence_init();
        }
        fence_proxy_async();  // generic-proxy zero fill before the async-proxy (TMA) writes to the same bytes
        __syncthreads();
    }

    auto issue_row = [&](int r) {  // executed by ONE whole warp
        const int slot = r & (RD - 1), use = r / RD;
        if (use > 0) mbar_wait(&empty[slot], (uint32_t)(use - 1) & 1u);
        const int sy = pad_index(y0 - RP + r, H, d.pad_mode);
        float *dst = &ring[slot][0][0];
        if (sy >= 0) {
            const long long ro = (long long)sy * W;
            if (lane == 0) {
                mbar_expect_tx(&full[slot], 2u * row_bytes);
                bulk_g2s(dst + doff, xb + ro + cx0, row_bytes, &full[slot]);
                bulk_g2s(dst + RROW + doff, yb + ro + cx0, row_bytes, &full[slot]);
            }
            if (halo_lane) {
                cp_async4_to(smem_u32(dst + htensor * RROW + hdst), hbase + ro);
                cp_async_arrive_noinc(&full[slot]);
            }
        } else {  // a row of zero padding
            float4 *r4 = reinterpret_cast<float4 *>(dst);
            for (int i = lane; i < 2 * RROW / 4; i += 32) r4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            fence_proxy_async();
            __syncwarp();
            if (lane == 0 || halo_lane) mbar_arrive(&full[slot]);
        }
    };

    const int gx = x0 + 2 * tid;  // first of the two output columns (even)
    const int lo_y = d.interior ? RP : 0, hi_y = d.interior ? H - RP : H;
    float cm[2];
#pragma unroll
    for (int c = 0; c < 2; ++c) cm[c] = (gx + c < W && (!d.interior || (gx + c >= RP && gx + c < W - RP))) ? 1.f : 0.f;
    const f2_t CM = pack2(cm[0], cm[1]);
    // DUAL: column mask of the interior crop the second policy averages over
    const f2_t CM2 = pack2((gx >= RP && gx < W - RP) ? 1.f : 0.f, (gx + 1 >= RP && gx + 1 < W - RP) ? 1.f : 0.f);
    f2_t facc2 = 0ull;
    const f2_t C1 = bc2(d.c1), C2 = bc2(d.c2), EPS = bc2(d.eps), TWO = bc2(2.f), NEG1 = bc2(-1.f);
    f2_t acc[M][RK];
#pragma unroll
    for (int q = 0; q < M; ++q)
#pragma unroll
        for (int j = 0; j < RK; ++j) acc[q][j] = 0ull;
    f2_t facc = 0ull, chkx = 0ull, chky = 0ull;
    const long long cstride = (long long)d.n * a.co * plane_elems;
    long long off = ((long long)plane * H + y0) * W + gx;  // COEF: first element of the next completed row

#pragma unroll
    for (int r = 0; r < RPF; ++r)
        if (r < nrows_in && (r & 3) == wrp) issue_row(r);

    int ii = 0;  // i mod 11
#pragma unroll 1
    for (int i = 0; i < nrows_in; ++i) {
        {
            if ((i & 3) == wrp && i + RPF < nrows_in) issue_row(i + RPF);
            const int slot = i & (RD - 1);
            mbar_wait(&full[slot], (uint32_t)(i / RD) & 1u);
            float *rx = &ring[slot][0][0], *ry = rx + RROW;
            // ---- horizontal pass: 7 aligned pairs per tensor = window positions 0..13 (image columns gx-6 .. gx+7)
            const float *wx = rx + 2 * tid + 2, *wy = ry + 2 * tid + 2;
            f2_t h[M];
#pragma unroll
            for (int q = 0; q < M; ++q) h[q] = 0ull;
#pragma unroll
            for (int k = 0; k < NPOS / 2; ++k) {
                f2_t X = ld2(wx + 2 * k), Y = ld2(wy + 2 * k);
                if (k == 3) {  // positions 6, 7 are the thread's own columns: finite check on the raw values
                    chkx = fma2(X, 0ull, chkx);
                    chky = fma2(Y, 0ull, chky);
                }
                if (CLAMP) {  // clamp(v, 0, 1) in registers (one FADD.SAT per value; NaN -> 0, flagged above)
                    float xl, xh, yl, yh;
                    unpack2(X, xl, xh);
                    unpack2(Y, yl, yh);
                    X = pack2(__saturatef(xl), __saturatef(xh));
                    Y = pack2(__saturatef(yl), __saturatef(yh));
                }
                // the products as SCALAR multiplies: every value enters the horizontal pass broadcast to both lanes
                // of an FFMA2 (a register operand form, no move), and ptxas turns a packed product that is only ever
                // read one lane at a time into one half-rate FMUL2 per lane (36 per row instead of 21); 36 full-rate
                // FMULs are half the pipe time of those
                float v[M][2];
                unpack2(X, v[0][0], v[0][1]);
                unpack2(Y, v[1][0], v[1][1]);
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    if (MAPS5) {
                        v[2][e] = __fmul_rn(v[0][e], v[0][e]);
                        v[3][e] = __fmul_rn(v[1][e], v[1][e]);
                        v[4][e] = __fmul_rn(v[0][e], v[1][e]);
                    } else {
                        v[2][e] = __fmaf_rn(v[0][e], v[0][e], __fmul_rn(v[1][e], v[1][e]));
                        v[3][e] = __fmul_rn(v[0][e], v[1][e]);
                    }
                }
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int p = 2 * k + e;
                    if (p < 1 || p > RK + 1) continue;  // positions 0 and 13 carry no tap
                    const f2_t w = pack2(tp.wp[p].x, tp.wp[p].y);
#pragma unroll
                    for (int q = 0; q < M; ++q) h[q] = fma2(bc2(v[q][e]), w, h[q]);
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[slot]);  // this warp has consumed row i
            // ---- vertical scatter (11 static variants, CTA-uniform switch)
            f2_t done[M];
            switch (ii) {
                case 0: scatter_row<0, M>(acc, h, done, tp); break;
                case 1: scatter_row<1, M>(acc, h, done, tp); break;
                case 2: scatter_row<2, M>(acc, h, done, tp); break;
                case 3: scatter_row<3, M>(acc, h, done, tp); break;
                case 4: scatter_row<4, M>(acc, h, done, tp); break;
                case 5: scatter_row<5, M>(acc, h, done, tp); break;
                case 6: scatter_row<6, M>(acc, h, done, tp); break;
                case 7: scatter_row<7, M>(acc, h, done, tp); break;
                case 8: scatter_row<8, M>(acc, h, done, tp); break;
                case 9: scatter_row<9, M>(acc, h, done, tp); break;
                default: scatter_row<10, M>(acc, h, done, tp); break;
            }
            ii = ii == RK - 1 ? 0 : ii + 1;
            if (i >= RK - 1) {  // output row o = i - 10 is complete
                const int gy = y0 + i - (RK - 1);
                const f2_t MX = done[0], MY = done[1];
                const f2_t mx2 = mul2(MX, MX), my2 = mul2(MY, MY), mxy = mul2(MX, MY);
                f2_t vsum, sxy;
                if (MAPS5) {
                    f2_t sx = fma2(mx2, NEG1, done[2]), sy = fma2(my2, NEG1, done[3]);
                    if (DUAL) {
                        // second policy: its taps are this window's up to <= 5 ulp per tap and a different tap SUM, and
                        // only the sum matters: M' = r M, var' = r (var - eps mu^2), no clamp, no eps
                        const f2_t NE = bc2(a.dual_neg_eps), B = add2(mx2, my2);
                        const f2_t n1b = fma2(mxy, bc2(a.dual_2r2), bc2(a.dual_c1));
                        const f2_t d1b = fma2(B, bc2(a.dual_r2), bc2(a.dual_c1));
                        const f2_t n2b = fma2(fma2(mxy, NE, fma2(mxy, NEG1, done[4])), bc2(a.dual_2r), bc2(a.dual_c2));
                        const f2_t d2b = fma2(fma2(B, NE, add2(sx, sy)), bc2(a.dual_r), bc2(a.dual_c2));
                        float nb[2], db[2];
                        unpack2(mul2(n1b, n2b), nb[0], nb[1]);
                        unpack2(mul2(d1b, d2b), db[0], db[1]);
                        const float s0 = div_pos(nb[0], db[0], rcp_pos(db[0])), s1 = div_pos(nb[1], db[1], rcp_pos(db[1]));
                        if (gy >= RP && gy < H - RP) facc2 = fma2(pack2(s0, s1), CM2, facc2);
                    }
                    if (d.clamp_var) {
                        float a0, a1, b0, b1;
                        unpack2(sx, a0, a1);
                        unpack2(sy, b0, b1);
                        sx = pack2(fmaxf(a0, 0.f), fmaxf(a1, 0.f));
                        sy = pack2(fmaxf(b0, 0.f), fmaxf(b1, 0.f));
                    }
                    vsum = add2(sx, sy);
                    sxy = fma2(mxy, NEG1, done[4]);
                } else {
                    vsum = fma2(my2, NEG1, fma2(mx2, NEG1, done[2]));
                    sxy = fma2(mxy, NEG1, done[3]);
                }
                const f2_t n1 = fma2(mxy, TWO, C1), n2 = fma2(sxy, TWO, C2);
                const f2_t d1 = add2(add2(mx2, my2), C1), d2 = add2(vsum, C2);
                const f2_t NUM = mul2(n1, n2), DEN = fma2(d1, d2, EPS);
                float nu[2], de[2], sv[2], rinv[2];
                unpack2(NUM, nu[0], nu[1]);
                unpack2(DEN, de[0], de[1]);
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    rinv[c] = rcp_pos(de[c]);
                    sv[c] = div_pos(nu[c], de[c], rinv[c]);
                }
                if (COEF) {  // dLoss/dS times the sensitivities of S to mu_x, E[x^2], E[xy] (ssim_stream_kernel)
                    float mx[2], my[2], n1v[2], n2v[2], d1v[2], d2v[2], cf[3][2];
                    unpack2(MX, mx[0], mx[1]);
                    unpack2(MY, my[0], my[1]);
                    unpack2(n1, n1v[0], n1v[1]);
                    unpack2(n2, n2v[0], n2v[1]);
                    unpack2(d1, d1v[0], d1v[1]);
                    unpack2(d2, d2v[0], d2v[1]);
#pragma unroll
                    for (int c = 0; c < 2; ++c) {
                        const float l = (1.f - sv[c]) * 0.5f;
                        const float up =
                            (!d.loss_map) ? a.coef_scale : ((l >= 0.f && l <= 1.f) ? -0.5f * a.coef_scale : 0.f);
                        const float inv = rinv[c];
                        const float a3 = 2.f * n1v[c] * inv;
                        const float a2 = -sv[c] * d1v[c] * inv;
                        const float dmu = 2.f * (my[c] * n2v[c] - sv[c] * mx[c] * d2v[c]) * inv;
                        const float a1 = dmu - 2.f * mx[c] * a2 - my[c] * a3;
                        cf[0][c] = up * a1;
                        cf[1][c] = up * a2;
                        cf[2][c] = up * a3;
                    }
                    if (gx < W) {  // W is even and gx is even: both columns are inside, the pair is 8-byte aligned
                        float *cp = a.coef + off;
#pragma unroll
                        for (int q = 0; q < 3; ++q)
                            *reinterpret_cast<float2 *>(cp + q * cstride) = make_float2(cf[q][0], cf[q][1]);
                    }
                    off += W;
                }
                if (d.loss_map) {
#pragma unroll
                    for (int c = 0; c < 2; ++c) sv[c] = fminf(fmaxf((1.f - sv[c]) * 0.5f, 0.f), 1.f);
                }
                if (gy >= lo_y && gy < hi_y) facc = fma2(pack2(sv[0], sv[1]), CM, facc);
            }
        }
    }
    float f0, f1, cx0f, cx1f, cy0f, cy1f;
    unpack2(facc, f0, f1);
    unpack2(chkx, cx0f, cx1f);
    unpack2(chky, cy0f, cy1f);
    const double bs = block_sum<double>((double)f0 + (double)f1, red);
    if (DUAL) {
        float g0, g1;
        unpack2(facc2, g0, g1);
        const double bs2 = block_sum<double>((double)g0 + (double)g1, red);
        if (tid == 0) a.part2[blockIdx.x] = bs2;
    }
    int flag = 0;
    if (!(cx0f == 0.f) || !(cx1f == 0.f)) flag |= 1;
    if (!(cy0f == 0.f) || !(cy1f == 0.f)) flag |= 2;
    const int f = block_or(flag, reinterpret_cast<int *>(red));
    if (tid == 0) {
        a.part[blockIdx.x] = bs;
        a.pflag[blockIdx.x] = f;
    }
    if (!last_block_done(a.ticket, gridDim.x)) return;
    const int planes = d.n * a.co;
    for (int pl = wrp; pl < planes; pl += RT / 32) {
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
        if (DUAL) {
            const double *p2 = a.part2 + (size_t)pl * tpp;
            double t0 = 0.0;
            for (int b2 = lane; b2 < tpp; b2 += 32) t0 += __ldcg(p2 + b2);
            t0 = warp_sum(t0);
            if (lane == 0) a.out_sum2[pl] = t0;
        }
    }
    int fl = 0;
    for (int b = tid; b < (int)gridDim.x; b += RT) fl |= __ldcg(a.pflag + b);
    const int ff = block_or(fl, reinterpret_cast<int *>(red));
    if (tid == 0) *a.out_flags = ff;
}


// ------------------------------------------------------------------------------------------------------------------
// DSSIM backward on the same skeleton (replaces ssim_bwd_stream2_kernel of ssim.cu as the default for a scalar upstream
// gradient and the 11-tap window): grad_x = pad^T G^T (A1) + 2 x pad^T G^T (A2) + y pad^T G^T (A3), the three
// coefficient planes written by the forward kernel filtered with the ADJOINT of the reflect-padded separable window.
//   * Ring slot = one row of the three coefficient planes (zero outside the image: the ring starts zeroed, rows above /
//     below the image are zero-filled by the issuing warp) + the x / y (/ old gradient) rows of the output row that
//     ring entry completes, all by cp.async.bulk on one mbarrier per slot; no halo lanes (zero extension).
//   * The mirror folds of the adjoint live in the WEIGHTS, as in ssim_bwd_stream2_kernel: a border column's horizontal
//     taps gain the taps of its mirrored partner (per-thread tap pairs in registers), and the <= 2 x 5 input rows next
//     to the top / bottom edge of the image scatter with row-dependent vertical weights (CTA-uniform, rare).
//   * Horizontal and vertical passes in packed FP32 exactly like the forward kernel: 69 FFMA2 per thread-row for two
//     columns instead of 138 FFMA.
// Same operation order per output as ssim_bwd_stream2_kernel -> equal to it to rounding.
struct BwdTaps {
    float g[RK];
};

template <int II, class WF>
__device__ __forceinline__ void scatter_row3(f2_t (&acc)[3][RK], const f2_t (&h)[3], f2_t (&done)[3], WF wf) {
#pragma unroll
    for (int j = 0; j < RK; ++j) {
        const int sl = (II - j + RK) % RK;
        const f2_t gj = bc2(wf(j));
#pragma unroll
        for (int q = 0; q < 3; ++q) acc[q][sl] = (j == 0) ? mul2(gj, h[q]) : fma2(gj, h[q], acc[q][sl]);
    }
#pragma unroll
    for (int q = 0; q < 3; ++q) done[q] = acc[q][(II + 1) % RK];
}

constexpr size_t ring_bwd_smem(bool accumulate) {
    return (size_t)RD * (3 * RROW + (accumulate ? 3 : 2) * RCOLS) * sizeof(float);
}

template <bool ACC>
__global__ void __launch_bounds__(RT, ACC ? 3 : 4) ssim_ring_bwd_kernel(const SsimBwdArgs a, const BwdTaps tp, int seg_rows) {
    constexpr int NIO = ACC ? 3 : 2;
    extern __shared__ __align__(128) unsigned char smraw[];
    float (*ringc)[3][RROW] = reinterpret_cast<float (*)[3][RROW]>(smraw);
    float (*ringio)[NIO][RCOLS] = reinterpret_cast<float (*)[NIO][RCOLS]>(smraw + (size_t)RD * 3 * RROW * sizeof(float));
    __shared__ __align__(8) uint64_t full[RD], empty[RD];
    const int H = a.h, W = a.w, tid = threadIdx.x, lane = tid & 31, wrp = tid >> 5;
    const int tpp = a.tiles_x * a.tiles_y;
    const int plane = blockIdx.x / tpp;
    const int trem = blockIdx.x - plane * tpp;
    const int seg = trem / a.tiles_x, strip = trem - seg * a.tiles_x;
    const int x0 = strip * RCOLS, y0 = seg * seg_rows;
    const int rows_out = min(seg_rows, H - y0);
    const int nrows_in = rows_out + 2 * RP;
    const long long plane_elems = (long long)H * W;
    const long long cstride = (long long)a.n * a.c * plane_elems;
    const float *cb = a.coef + (long long)plane * plane_elems;
    const float *xb = a.x + (long long)plane * plane_elems;
    const float *yb = a.y + (long long)plane * plane_elems;
    float *ob = a.out_grad + (long long)plane * plane_elems;

    const int cx0 = max(x0 - RLEAD, 0), cx1 = min(x0 + RCOLS + RLEAD, W);
    const uint32_t row_bytes = (uint32_t)(cx1 - cx0) * 4u;
    const int doff = cx0 - (x0 - RLEAD);
    const uint32_t io_bytes = (uint32_t)(min(x0 + RCOLS, W) - x0) * 4u;

    {
        float4 *r4 = reinterpret_cast<float4 *>(smraw);
        for (int i = tid; i < (int)(ring_bwd_smem(ACC) / 16); i += RT) r4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (tid == 0) {
#pragma unroll
            for (int s = 0; s < RD; ++s) {
                mbar_init(&full[s], 1);
                mbar_init(&empty[s], RT / 32);
            }
            mbar_fence_init();
        }
        fence_proxy_async();
        __syncthreads();
    }

    auto issue_row = [&](int r) {  // executed by ONE whole warp
        const int slot = r & (RD - 1), use = r / RD;
        if (use > 0) mbar_wait(&empty[slot], (uint32_t)(use - 1) & 1u);
        const int sy = y0 - RP + r;
        const bool in_img = sy >= 0 && sy < H;
        const bool has_io = r >= RK - 1;  // ring entry r completes output row y0 + r - 10
        float *dst = &ringc[slot][0][0];
        if (!in_img) {  // the coefficient planes are zero outside the image
            float4 *r4 = reinterpret_cast<float4 *>(dst);
            for (int i = lane; i < 3 * RROW / 4; i += 32) r4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            fence_proxy_async();
            __syncwarp();
        }
        if (lane == 0) {
            const uint32_t bytes = (in_img ? 3u * row_bytes : 0u) + (has_io ? (uint32_t)NIO * io_bytes : 0u);
            if (bytes == 0u) {
                mbar_arrive(&full[slot]);
            } else {
                mbar_expect_tx(&full[slot], bytes);
                if (in_img) {
                    const float *src = cb + (long long)sy * W + cx0;
#pragma unroll
                    for (int m = 0; m < 3; ++m) bulk_g2s(dst + m * RROW + doff, src + m * cstride, row_bytes, &full[slot]);
                }
                if (has_io) {
                    const long long oo = (long long)(y0 + r - (RK - 1)) * W + x0;
                    float *io = &ringio[slot][0][0];
                    bulk_g2s(io, xb + oo, io_bytes, &full[slot]);
                    bulk_g2s(io + RCOLS, yb + oo, io_bytes, &full[slot]);
                    if (ACC) bulk_g2s(io + 2 * RCOLS, ob + oo, io_bytes, &full[slot]);
                }
            }
        }
    };

    const int gx = x0 + 2 * tid;
    // horizontal tap pairs of this thread's two columns with the mirror folds added: position p (image column
    // gx - 6 + p) carries tap p - 1 of column gx and tap p - 2 of column gx + 1
    f2_t WP[NPOS];
    {
        float gh[2][RK];
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            const int gc = gx + c;
            const bool cin = gc < W;
#pragma unroll
            for (int j = 0; j < RK; ++j) {
                float wj = tp.g[j];
                const int x = gc - RP + j;
                if (cin && gc >= 1 && gc <= RP && x >= 0 && j + 2 * gc <= 2 * RP) wj += tp.g[j + 2 * gc];
                if (cin && gc >= W - 1 - RP && gc <= W - 2 && x <= W - 1 && j + 2 * gc - 2 * W + 2 >= 0)
                    wj += tp.g[j + 2 * gc - 2 * W + 2];
                gh[c][j] = wj;
            }
        }
#pragma unroll
        for (int p = 0; p < NPOS; ++p)
            WP[p] = pack2((p >= 1 && p <= RK) ? gh[0][(p >= 1 && p <= RK) ? p - 1 : 0] : 0.f,
                          (p >= 2 && p <= RK + 1) ? gh[1][(p >= 2 && p <= RK + 1) ? p - 2 : 0] : 0.f);
    }
    const bool seg_fold = (y0 <= RP) || (y0 + rows_out > H - 1 - RP);
    const f2_t GS = bc2(__ldg(a.gout)), TWO = bc2(2.f);
    f2_t acc[3][RK];
#pragma unroll
    for (int q = 0; q < 3; ++q)
#pragma unroll
        for (int j = 0; j < RK; ++j) acc[q][j] = 0ull;

#pragma unroll
    for (int r = 0; r < RPF; ++r)
        if (r < nrows_in && (r & 3) == wrp) issue_row(r);

    int ii = 0;
#pragma unroll 1
    for (int i = 0; i < nrows_in; ++i) {
        if ((i & 3) == wrp && i + RPF < nrows_in) issue_row(i + RPF);
        const int slot = i & (RD - 1);
        mbar_wait(&full[slot], (uint32_t)(i / RD) & 1u);
        f2_t h[3];
#pragma unroll
        for (int m = 0; m < 3; ++m) {
            const float *wc = &ringc[slot][m][0] + 2 * tid + 2;
            f2_t v = 0ull;
#pragma unroll
            for (int k = 0; k < NPOS / 2; ++k) {
                float lo, hi;
                unpack2(ld2(wc + 2 * k), lo, hi);
                if (2 * k >= 1) v = fma2(bc2(lo), WP[2 * k], v);           // position 0 carries no tap
                if (2 * k + 1 <= RK + 1) v = fma2(bc2(hi), WP[2 * k + 1], v);  // position 13 neither
            }
            h[m] = v;
        }
        f2_t X = 0ull, Y = 0ull, O = 0ull;
        if (i >= RK - 1) {
            X = ld2(&ringio[slot][0][0] + 2 * tid);
            Y = ld2(&ringio[slot][1][0] + 2 * tid);
            if (ACC) O = ld2(&ringio[slot][2][0] + 2 * tid);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[slot]);
        const int p = y0 - RP + i;
        const bool vfold = seg_fold && (p < RP || p >= H - RP);  // CTA-uniform; only <= 2 x 5 rows of the image
        f2_t done[3];
        if (!vfold) {
            auto wf = [&](int j) { return tp.g[j]; };
            switch (ii) {
                case 0: scatter_row3<0>(acc, h, done, wf); break;
                case 1: scatter_row3<1>(acc, h, done, wf); break;
                case 2: scatter_row3<2>(acc, h, done, wf); break;
                case 3: scatter_row3<3>(acc, h, done, wf); break;
                case 4: scatter_row3<4>(acc, h, done, wf); break;
                case 5: scatter_row3<5>(acc, h, done, wf); break;
                case 6: scatter_row3<6>(acc, h, done, wf); break;
                case 7: scatter_row3<7>(acc, h, done, wf); break;
                case 8: scatter_row3<8>(acc, h, done, wf); break;
                case 9: scatter_row3<9>(acc, h, done, wf); break;
                default: scatter_row3<10>(acc, h, done, wf); break;
            }
        } else {
            float wv[RK];
#pragma unroll
            for (int j = 0; j < RK; ++j) {
                const int y = y0 + i - j;  // output row that tap j of this input row feeds
                float wj = tp.g[j];
                if (y >= 1 && y <= RP && p >= 0 && p + y <= RP) wj += tp.g[p + y + RP];
                if (y >= H - 1 - RP && y <= H - 2 && p <= H - 1 && p >= 2 * H - 2 - y - RP)
                    wj += tp.g[p - (2 * H - 2 - y) + RP];
                wv[j] = wj;
            }
            auto wf = [&](int j) { return wv[j]; };
            switch (ii) {
                case 0: scatter_row3<0>(acc, h, done, wf); break;
                case 1: scatter_row3<1>(acc, h, done, wf); break;
                case 2: scatter_row3<2>(acc, h, done, wf); break;
                case 3: scatter_row3<3>(acc, h, done, wf); break;
                case 4: scatter_row3<4>(acc, h, done, wf); break;
                case 5: scatter_row3<5>(acc, h, done, wf); break;
                case 6: scatter_row3<6>(acc, h, done, wf); break;
                case 7: scatter_row3<7>(acc, h, done, wf); break;
                case 8: scatter_row3<8>(acc, h, done, wf); break;
                case 9: scatter_row3<9>(acc, h, done, wf); break;
                default: scatter_row3<10>(acc, h, done, wf); break;
            }
        }
        ii = ii == RK - 1 ? 0 : ii + 1;
        if (i >= RK - 1 && gx < W) {  // W is a multiple of 4 and gx is even: both columns are inside
            const int oy = y0 + i - (RK - 1);
            float xr[2], yr[2];
            unpack2(X, xr[0], xr[1]);
            unpack2(Y, yr[0], yr[1]);
            f2_t XV = X, YV = Y;
            if (a.clamp_in) {
                XV = pack2(fminf(fmaxf(xr[0], 0.f), 1.f), fminf(fmaxf(xr[1], 0.f), 1.f));
                YV = pack2(fminf(fmaxf(yr[0], 0.f), 1.f), fminf(fmaxf(yr[1], 0.f), 1.f));
            }
            const f2_t t = fma2(YV, done[2], fma2(mul2(TWO, XV), done[1], done[0]));
            float r0, r1, o0, o1;
            unpack2(mul2(GS, t), r0, r1);
            unpack2(O, o0, o1);
            if (a.clamp_in) {
                r0 = (xr[0] >= 0.f && xr[0] <= 1.f) ? r0 : 0.f;
                r1 = (xr[1] >= 0.f && xr[1] <= 1.f) ? r1 : 0.f;
            }
            *reinterpret_cast<float2 *>(ob + (long long)oy * W + gx) = make_float2(o0 + r0, o1 + r1);
        }
    }
}

}  // namespace

bool ssim_ring_eligible(const llie_ssim_desc &d, const float *x, const float *y, const float *win_host, bool want_map,
                        const float *coef) {
    return win_host != nullptr && d.ksize == RK && !d.luma && !want_map && (d.w & 3) == 0 && aligned16(x) &&
           aligned16(y) && (coef == nullptr || (!d.clamp_var && (reinterpret_cast<uintptr_t>(coef) & 7u) == 0));
}

bool ssim_ring_dual_eligible(const llie_ssim_desc &d, const float *x, const float *y, const float *win_host) {
    return ssim_ring_eligible(d, x, y, win_host, false, nullptr) && d.clamp_var && !d.clamp_in && !d.interior &&
           !d.loss_map && d.pad_mode == LLIE_PAD_REFLECT && d.h > 2 * RP && d.w > 2 * RP;
}

int ssim_ring_forward(SsimArgs &a, const float *win_host, void *workspace, cudaStream_t st) {
    const llie_ssim_desc &d = a.d;
    RingTaps tp;
    for (int p = 0; p < NPOS; ++p) {
        tp.wp[p].x = (p - 1 >= 0 && p - 1 < RK) ? win_host[p - 1] : 0.f;
        tp.wp[p].y = (p - 2 >= 0 && p - 2 < RK) ? win_host[p - 2] : 0.f;
    }
    for (int j = 0; j < RK; ++j) tp.g[j] = win_host[j];
    a.tiles_x = div_up(d.w, RCOLS);
    const long long cols = (long long)d.n * a.co * a.tiles_x;
    // resident CTAs per SM: 4 at 128 registers (a few spilled words, 2-4 % faster than 3 at 148) except for the five-map
    // variant, whose 168 registers would spill 128 bytes (362 instead of 322 us per C4 pair)
    a.seg_rows = pick_seg_rows(cols, d.h, kNumSM * (d.clamp_var ? 3 : 4), 64, RK - 1);
    a.tiles_y = div_up(d.h, a.seg_rows);
    const long long grid = cols * a.tiles_y;
    Workspace ws(workspace);
    LLIE_REQUIRE(grid <= 0x7fffffffLL && (size_t)grid * (2 * sizeof(double) + sizeof(int)) <= ws.scratch_bytes,
                 "ssim_stats: workspace too small for %lld tiles", grid);
    a.part = reinterpret_cast<double *>(ws.scratch);
    a.part2 = a.part + grid;
    a.pflag = reinterpret_cast<int *>(ws.scratch + (size_t)grid * 2 * sizeof(double));
    a.ticket = ws.tickets;
    const bool clamp = d.clamp_in != 0;
    if (a.out_sum2 != nullptr) {
        ssim_ring_kernel<true, false, false, true><<<(int)grid, RT, 0, st>>>(a, tp);
    } else if (a.coef != nullptr) {
        if (clamp) ssim_ring_kernel<false, true, true><<<(int)grid, RT, 0, st>>>(a, tp);
        else ssim_ring_kernel<false, true, false><<<(int)grid, RT, 0, st>>>(a, tp);
    } else if (d.clamp_var) {
        if (clamp) ssim_ring_kernel<true, false, true><<<(int)grid, RT, 0, st>>>(a, tp);
        else ssim_ring_kernel<true, false, false><<<(int)grid, RT, 0, st>>>(a, tp);
    } else {
        if (clamp) ssim_ring_kernel<false, false, true><<<(int)grid, RT, 0, st>>>(a, tp);
        else ssim_ring_kernel<false, false, false><<<(int)grid, RT, 0, st>>>(a, tp);
    }
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

bool ssim_ring_bwd_eligible(const SsimBwdArgs &a, const float *win_host) {
    return win_host != nullptr && a.ksize == RK && !a.gout_is_map && (a.w & 3) == 0 && aligned16(a.x) &&
           aligned16(a.y) && aligned16(a.coef) && aligned16(a.out_grad);
}

int ssim_ring_backward(SsimBwdArgs &a, const float *win_host, cudaStream_t st) {
    BwdTaps tp;
    for (int j = 0; j < RK; ++j) tp.g[j] = win_host[j];
    a.tiles_x = div_up(a.w, RCOLS);
    const long long cols = (long long)a.n * a.c * a.tiles_x;
    // 4 resident CTAs per SM at 128 registers, 3 for the accumulating variant (its spills cost more than the warps gain)
    const int rows = pick_seg_rows(cols, a.h, kNumSM * (a.accumulate ? 3 : 4), 64, RK - 1);
    a.tiles_y = div_up(a.h, rows);
    const long long grid = cols * a.tiles_y;
    LLIE_REQUIRE(grid <= 0x7fffffffLL, "ssim_loss_backward: too many tiles");
    if (a.accumulate) {
        LLIE_CUDA_OK(cudaFuncSetAttribute(ssim_ring_bwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)ring_bwd_smem(true)));
        ssim_ring_bwd_kernel<true><<<(int)grid, RT, ring_bwd_smem(true), st>>>(a, tp, rows);
    } else {
        LLIE_CUDA_OK(cudaFuncSetAttribute(ssim_ring_bwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)ring_bwd_smem(false)));
        ssim_ring_bwd_kernel<false><<<(int)grid, RT, ring_bwd_smem(false), st>>>(a, tp, rows);
    }
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

}  // namespace llie
