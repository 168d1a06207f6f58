// Physics-consistency projector, register-rolling path (depthwise 3x3 PSF, W % 4 == 0, 16-byte
// aligned planes) -- the HBM-bound production path for the reference's CrosstalkPSF shapes.
//
// A CTA of 4 warps owns 512 adjacent columns of a row chunk of one (n, c) plane; every thread holds 4
// adjacent pixels (one coalesced 128-bit load per row and tensor) and marches down the chunk.  The
// 3x3 stencil runs in "scatter" form -- an arriving input row completes one output row, extends the
// next and opens a third -- so only two partial-sum rows live in registers, and rows are software-
// prefetched PF iterations ahead.  No shared memory, no block barrier and no shuffles in the main
// loop: horizontal neighbours are read with (L1-hit) scalar loads at fixed offsets from the row
// pointer, everything that depends on the row is CTA-uniform (uniform datapath), and padding
// (zeros / reflect / replicate) is a select on the row index and on the two image-edge lanes.
//
//   forward   r = clipY(sY*(K(*)pred)) - clipA(sA*obs), |r| or Charbonnier, fp64 partial per CTA.
//   backward  q = dLoss/dconv is recomputed for the thread's 4 columns and its two outside
//             neighbours (6 columns), then the flipped-kernel stencil runs in the same rolling form;
//             the adjoint of the padding operator (mirror / edge fold) is applied in-register on the
//             image-edge lanes and through two virtual rows on the first / last chunk.
//
// Reductions: one float64 partial per CTA -> fixed-order finalize in the last CTA (bitwise
// reproducible, no float atomics).
#include "phys_common.cuh"

namespace llie {

#ifndef LLIE_ROLL_PF
#define LLIE_ROLL_PF 3
#endif
#ifndef LLIE_ROLL_MINB
#define LLIE_ROLL_MINB 5
#endif
constexpr int PF = LLIE_ROLL_PF;   // rows prefetched ahead per thread
constexpr int kRollThreads = 128;  // 4 warps side by side = 512 columns
constexpr int kRollCols = kRollThreads * 4;

struct RollGeom {
    int groups, chunks, R;  // column groups of 512 px; R rows per chunk (last chunk absorbs the remainder)
    int ctas_per_sample;
    double *part;  // [gridDim.x]
    int *pflag;    // [gridDim.x]
};

__device__ __forceinline__ float4 ldg4(const float *p) { return __ldg(reinterpret_cast<const float4 *>(p)); }
__device__ __forceinline__ float2 ldg2(const float *p) { return __ldg(reinterpret_cast<const float2 *>(p)); }

template <bool ROBUST, bool CLIPY, bool CLIPA>
__device__ __forceinline__ float loss_t(float conv, float a, float ry, float ra, float eps2) {
    float y = conv * ry, t = a * ra;
    if (CLIPY) y = fminf(fmaxf(y, 0.f), 1.f);
    if (CLIPA) t = fminf(fmaxf(t, 0.f), 1.f);
    const float r = y - t;
    return ROBUST ? sqrtf(fmaf(r, r, eps2)) : fabsf(r);
}
template <bool CLIPY, bool CLIPA>
__device__ __forceinline__ float absres_t(float conv, float a, float ry, float ra) {
    float y = conv * ry, t = a * ra;
    if (CLIPY) y = fminf(fmaxf(y, 0.f), 1.f);
    if (CLIPA) t = fminf(fmaxf(t, 0.f), 1.f);
    return fabsf(y - t);
}

// upstream-weighted derivative of the loss w.r.t. conv; gs = g_n * (rho if rho scales the conv side)
template <bool ROBUST, bool CLIPY, bool CLIPA>
__device__ __forceinline__ float dloss_t(float conv, float a, float ry, float ra, float eps2, float gs) {
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

// source row of a (possibly one-row-padded) row index; CTA-uniform.  zero is set for zero padding.
__device__ __forceinline__ int map_row(int r, int H, int mode, bool &zero) {
    zero = false;
    if (r >= 0 && r < H) return r;
    if (mode == LLIE_PAD_ZEROS) {
        zero = true;
        return r < 0 ? 0 : H - 1;
    }
    if (mode == LLIE_PAD_REFLECT) {
        const int s = r < 0 ? -r : 2 * (H - 1) - r;
        return s < 0 ? 0 : (s >= H ? H - 1 : s);  // rows beyond one reflection only feed discarded outputs
    }
    return r < 0 ? 0 : H - 1;
}

struct BlockCoord {
    int plane, n, c, group, chunk;
};
__device__ __forceinline__ BlockCoord decode_block(const RollGeom &g, int C) {
    BlockCoord bc;
    const int b = blockIdx.x;
    const int per_plane = g.groups * g.chunks;
    bc.plane = b / per_plane;
    const int rem = b - bc.plane * per_plane;
    bc.group = rem / g.chunks;
    bc.chunk = rem - bc.group * g.chunks;
    bc.n = bc.plane / C;
    bc.c = bc.plane - bc.n * C;
    return bc;
}

// Thread 0 must be the only writer of data the finalizing CTA reads.  Returns true in ALL threads
// of the last CTA to arrive.
__device__ __forceinline__ bool last_block_done_t0(unsigned *ticket, unsigned total_blocks) {
    __shared__ bool s_last;
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

// ------------------------------------------------------------------------------------------------
// finalize (last CTA): fixed-order sums of the per-CTA partials of each sample
// ------------------------------------------------------------------------------------------------
__device__ void roll_finalize(const PhysArgs &a, const RollGeom &g, double *red) {
    const int N = a.d.n, NT = blockDim.x;
    double batch = 0.0, total = 0.0;
    for (int n = 0; n < N; ++n) {
        double s = 0.0;
        const double *p = g.part + (size_t)n * g.ctas_per_sample;
        for (int b = threadIdx.x; b < g.ctas_per_sample; b += NT) s += __ldcg(p + b);
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
    for (int b = threadIdx.x; b < (int)gridDim.x; b += NT) fl |= __ldcg(g.pflag + b);
    const int ff = block_or(fl, reinterpret_cast<int *>(red));
    if (threadIdx.x == 0) {
        a.out_mean[N] = (float)(batch / N);
        a.out_mean[N + 1] = (float)batch;
        if (a.out_sum) a.out_sum[N] = total;
        *a.out_flags = ff;
    }
}

// ------------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------------
struct FwdRow {
    float4 p, o;  // pred row y+1 (pad-mapped), obs row y
    float l, r;   // pred neighbours at columns x-1 / x+4 of the same row
};

template <bool ROBUST, bool CLIPY, bool CLIPA>
__global__ void __launch_bounds__(kRollThreads, LLIE_ROLL_MINB) phys_fwd_roll(const PhysArgs a, const RollGeom g) {
    __shared__ double red[32];
    const llie_phys_desc &d = a.d;
    const int H = d.h, W = d.w, mode = d.pad_mode;
    const BlockCoord bc = decode_block(g, d.c_out);
    const int xr = bc.group * kRollCols + threadIdx.x * 4;
    const bool active = xr < W;
    const int x = active ? xr : W - 4;  // inactive lanes shadow the last valid quad (loads stay in bounds)
    const int ya = bc.chunk * g.R;
    const int yb = (bc.chunk == g.chunks - 1) ? H : ya + g.R;
    const float *pp = a.pred + (long long)bc.plane * H * W + x;
    const float *op = a.obs + (long long)bc.plane * H * W + x;
    const bool is_left = x == 0, is_right = x + 4 == W;
    const int offl = is_left ? 0 : -1, offr = is_right ? 3 : 4;  // in-bounds neighbour addresses
    float w[9];
#pragma unroll
    for (int q = 0; q < 9; ++q) w[q] = __ldg(a.psf + bc.c * 9 + q);
    const float rho = (a.rho == nullptr) ? d.rho_scalar
                                         : __ldg(a.rho + (long long)bc.n * d.rho_sn + (long long)bc.c * d.rho_sc);
    const float ry = d.rho_side == 0 ? rho : 1.f, ra = d.rho_side == 1 ? rho : 1.f;
    const float eps2 = d.eps * d.eps;
    float cm[4];  // 1 where the column belongs to the compared domain of an active lane
#pragma unroll
    for (int i = 0; i < 4; ++i) cm[i] = (active && x + i >= a.x_lo && x + i < a.x_lo + a.wo) ? 1.f : 0.f;
    const bool map_vec = (a.wo & 3) == 0 && a.x_lo == 0;
    float chk_p = 0.f, chk_o = 0.f;  // x*0 accumulators: NaN iff a non-finite value was loaded

    auto load_pred = [&](int r, float4 &v, float &l, float &rr) {
        bool zero;
        const float *row = pp + (long long)map_row(r, H, mode, zero) * W;
        v = ldg4(row);
        l = __ldg(row + offl);
        rr = __ldg(row + offr);
        // NOTE: the loaded registers are not touched here (a select on them would stall the thread
        // until the load lands and defeat the prefetch); zero padding is applied in expand()
    };
    auto expand = [&](int r, float4 v, float l, float rr, float (&o6)[6]) {
        if (mode == LLIE_PAD_ZEROS && (r < 0 || r >= H)) {  // CTA-uniform
            v = make_float4(0.f, 0.f, 0.f, 0.f);
            l = rr = 0.f;
        }
        if (is_left) l = mode == LLIE_PAD_REFLECT ? v.y : (mode == LLIE_PAD_REPLICATE ? v.x : 0.f);
        if (is_right) rr = mode == LLIE_PAD_REFLECT ? v.z : (mode == LLIE_PAD_REPLICATE ? v.w : 0.f);
        o6[0] = l; o6[1] = v.x; o6[2] = v.y; o6[3] = v.z; o6[4] = v.w; o6[5] = rr;
        chk_p = fmaf(v.x, 0.f, chk_p);
        chk_p = fmaf(v.y, 0.f, chk_p);
        chk_p = fmaf(v.z, 0.f, chk_p);
        chk_p = fmaf(v.w, 0.f, chk_p);
    };

    float sA[4] = {0.f, 0.f, 0.f, 0.f}, sB[4] = {0.f, 0.f, 0.f, 0.f};
    FwdRow buf[PF];
    {   // warm-up rows ya-1 and ya open the partial sums of output row ya; all loads issued up front
        float4 v0, v1;
        float l0, r0, l1, r1;
        load_pred(ya - 1, v0, l0, r0);
        load_pred(ya, v1, l1, r1);
#pragma unroll
        for (int u = 0; u < PF; ++u) {
            const int y = ya + u;
            load_pred(y + 1, buf[u].p, buf[u].l, buf[u].r);
            buf[u].o = ldg4(op + (long long)(y < H ? y : H - 1) * W);
        }
        float r6[6], dump[4];
        expand(ya - 1, v0, l0, r0, r6);
        roll3<4>(r6, w, sA, sB, dump);
        expand(ya, v1, l1, r1, r6);
        roll3<4>(r6, w, sA, sB, dump);
    }
    float acc = 0.f;
    for (int y0 = ya; y0 < yb; y0 += PF) {
#pragma unroll
        for (int u = 0; u < PF; ++u) {
            const int y = y0 + u;
            if (y < yb) {  // CTA-uniform
                const FwdRow cur = buf[u];
                const int yn = y + PF;
                if (yn < yb) {  // prefetch the rows of iteration y + PF
                    load_pred(yn + 1, buf[u].p, buf[u].l, buf[u].r);
                    buf[u].o = ldg4(op + (long long)yn * W);
                }
                float r6[6], cv[4];
                expand(y + 1, cur.p, cur.l, cur.r, r6);
                roll3<4>(r6, w, sA, sB, cv);
                const float av[4] = {cur.o.x, cur.o.y, cur.o.z, cur.o.w};
#pragma unroll
                for (int i = 0; i < 4; ++i) chk_o = fmaf(av[i], 0.f, chk_o);
                if (y >= a.y_lo && y < a.y_lo + a.ho) {  // CTA-uniform
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        acc = fmaf(loss_t<ROBUST, CLIPY, CLIPA>(cv[i], av[i], ry, ra, eps2), cm[i], acc);
                    if (a.out_map != nullptr && active) {
                        float mv[4];
#pragma unroll
                        for (int i = 0; i < 4; ++i) mv[i] = absres_t<CLIPY, CLIPA>(cv[i], av[i], ry, ra);
                        float *mrow = a.out_map + (((long long)bc.plane) * a.ho + (y - a.y_lo)) * a.wo;
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
    }
    // ---- CTA partial (fixed order) + flags
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    const double ws = warp_sum((double)acc);
    int flag = 0;
    if (!(chk_p == 0.f)) flag |= 1;
    if (!(chk_o == 0.f)) flag |= 2;
    flag = __reduce_or_sync(0xffffffffu, (unsigned)flag);
    __shared__ int sflag[kRollThreads / 32];
    if (lane == 0) {
        red[wrp] = ws;
        sflag[wrp] = flag;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = red[0];
        int f = sflag[0];
#pragma unroll
        for (int k = 1; k < kRollThreads / 32; ++k) {
            s += red[k];
            f |= sflag[k];
        }
        g.part[blockIdx.x] = s;
        g.pflag[blockIdx.x] = f;
    }
    if (last_block_done_t0(a.ticket, gridDim.x)) roll_finalize(a, g, red);
}

// ------------------------------------------------------------------------------------------------
// backward: grad = padY^T K^T q,  q = g_n * dLoss/dconv on the compared domain (0 elsewhere)
// ------------------------------------------------------------------------------------------------
struct BwdRow {
    float4 p, o;    // pred row j+2 (pad-mapped), obs row j+1, columns x .. x+3
    float2 pl, pr;  // pred columns x-2, x-1 / x+4, x+5
    float ol, orr;  // obs columns x-1 / x+4
};

template <bool ROBUST, bool CLIPY, bool CLIPA>
__global__ void __launch_bounds__(kRollThreads, LLIE_ROLL_MINB) phys_bwd_roll(const PhysArgs a, const RollGeom g) {
    const llie_phys_desc &d = a.d;
    const int H = d.h, W = d.w, mode = d.pad_mode;
    const BlockCoord bc = decode_block(g, d.c_out);
    const int xr = bc.group * kRollCols + threadIdx.x * 4;
    const bool active = xr < W;
    const int x = active ? xr : W - 4;
    const bool first = bc.chunk == 0, last = bc.chunk == g.chunks - 1;
    const int ya = bc.chunk * g.R;
    const int yb = last ? H : ya + g.R;
    const bool fold = mode != LLIE_PAD_ZEROS;
    // emitted (virtual) rows: the first / last chunk also produce the padded rows -1 / H
    const int ja = (first && fold) ? -1 : ya;
    const int jb = (last && fold) ? H + 1 : yb;
    const int fold_top = mode == LLIE_PAD_REFLECT ? 1 : 0;          // image row that padded row -1 reads
    const int fold_bot = mode == LLIE_PAD_REFLECT ? H - 2 : H - 1;  // image row that padded row H reads
    const bool is_left = x == 0, is_right = x + 4 == W;
    const int offl = is_left ? 0 : -2, offr = is_right ? 2 : 4;  // in-bounds float2 neighbour addresses

    const float *pp = a.pred + (long long)bc.plane * H * W + x;
    const float *op = a.obs + (long long)bc.plane * H * W + x;
    float *gp = a.grad + (long long)bc.plane * H * W + x;
    float w[9], wf[9];
#pragma unroll
    for (int q = 0; q < 9; ++q) w[q] = __ldg(a.psf + bc.c * 9 + q);
#pragma unroll
    for (int q = 0; q < 9; ++q) wf[q] = w[8 - q];
    const float rho = (a.rho == nullptr) ? d.rho_scalar
                                         : __ldg(a.rho + (long long)bc.n * d.rho_sn + (long long)bc.c * d.rho_sc);
    const float ry = d.rho_side == 0 ? rho : 1.f, ra = d.rho_side == 1 ? rho : 1.f;
    const float eps2 = d.eps * d.eps;
    const float gs = __ldg(a.gout + a.gout_stride * bc.n) * a.gscale * ry;
    bool colok[6];  // columns x-1 .. x+4 inside the compared domain
#pragma unroll
    for (int i = 0; i < 6; ++i) colok[i] = (x - 1 + i >= a.x_lo) && (x - 1 + i < a.x_lo + a.wo);

    auto load_row = [&](int j, BwdRow &b) {  // pred row j+2, obs row j+1
        bool zero;
        const float *prow = pp + (long long)map_row(j + 2, H, mode, zero) * W;
        b.p = ldg4(prow);
        b.pl = ldg2(prow + offl);
        b.pr = ldg2(prow + offr);
        // (zero padding is applied at consumption time, see expand8)
        const int orow_i = j + 1 < 0 ? 0 : (j + 1 >= H ? H - 1 : j + 1);
        const float *orow = op + (long long)orow_i * W;
        b.o = ldg4(orow);
        b.ol = __ldg(orow + (is_left ? 0 : -1));
        b.orr = __ldg(orow + (is_right ? 3 : 4));
    };
    auto expand8 = [&](int r, BwdRow b, float (&o8)[8]) {  // pred row r, columns x-2 .. x+5
        if (mode == LLIE_PAD_ZEROS && (r < 0 || r >= H)) {  // CTA-uniform
            b.p = make_float4(0.f, 0.f, 0.f, 0.f);
            b.pl = b.pr = make_float2(0.f, 0.f);
        }
        float l0 = b.pl.x, l1 = b.pl.y, r0 = b.pr.x, r1 = b.pr.y;
        if (is_left) {  // column -1 is padded; column -2 only feeds a discarded q
            l1 = mode == LLIE_PAD_REFLECT ? b.p.y : (mode == LLIE_PAD_REPLICATE ? b.p.x : 0.f);
            l0 = 0.f;
        }
        if (is_right) {
            r0 = mode == LLIE_PAD_REFLECT ? b.p.z : (mode == LLIE_PAD_REPLICATE ? b.p.w : 0.f);
            r1 = 0.f;
        }
        o8[0] = l0; o8[1] = l1; o8[2] = b.p.x; o8[3] = b.p.y; o8[4] = b.p.z; o8[5] = b.p.w; o8[6] = r0; o8[7] = r1;
    };

    // rolling partial sums: conv rows over 6 columns (from pred rows), grad rows over 4 columns (from q rows)
    float cA[6], cB[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) cA[i] = cB[i] = 0.f;
    float gA[4] = {0.f, 0.f, 0.f, 0.f}, gB[4] = {0.f, 0.f, 0.f, 0.f};
    float eA = 0.f, eB = 0.f;  // partial sums of the folded padded column (image-edge lanes only)
    const int j0 = ja - 2;     // two warm-up iterations fill the partial sums
    BwdRow buf[PF];
    {
        BwdRow w0, w1;
        load_row(j0 - 2, w0);  // pred row j0
        load_row(j0 - 1, w1);  // pred row j0+1
#pragma unroll
        for (int u = 0; u < PF; ++u) load_row(j0 + u, buf[u]);
        float r8[8], dump[6];
        expand8(j0, w0, r8);
        roll3<6>(r8, w, cA, cB, dump);
        expand8(j0 + 1, w1, r8);
        roll3<6>(r8, w, cA, cB, dump);
    }
    float carry[4] = {0.f, 0.f, 0.f, 0.f};

    for (int jj = j0; jj < jb; jj += PF) {
#pragma unroll
        for (int u = 0; u < PF; ++u) {
            const int j = jj + u;
            if (j < jb) {  // CTA-uniform
                const BwdRow cur = buf[u];
                if (j + PF < jb) load_row(j + PF, buf[u]);
                // ---- q row j+1 over columns x-1 .. x+4
                float r8[8], cv[6];
                expand8(j + 2, cur, r8);
                roll3<6>(r8, w, cA, cB, cv);
                const int qr = j + 1;
                float qc[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                if (qr >= a.y_lo && qr < a.y_lo + a.ho) {  // CTA-uniform
                    const float av[6] = {cur.ol, cur.o.x, cur.o.y, cur.o.z, cur.o.w, cur.orr};
#pragma unroll
                    for (int i = 0; i < 6; ++i)
                        qc[i] = colok[i] ? dloss_t<ROBUST, CLIPY, CLIPA>(cv[i], av[i], ry, ra, eps2, gs) : 0.f;
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
                if (j >= ja) {  // CTA-uniform
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
                        if (first && fold && j == fold_top) {
#pragma unroll
                            for (int i = 0; i < 4; ++i) gr[i] += carry[i];
                        }
                        if (active)
                            *reinterpret_cast<float4 *>(gp + (long long)j * W) = make_float4(gr[0], gr[1], gr[2], gr[3]);
                    }
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// host
// ------------------------------------------------------------------------------------------------
static int g_roll_rows = 0;  // 0 = heuristic; set through llie_tune(LLIE_TUNE_PHYS_ROWS, r)

void phys_roll_set_rows(int r) { g_roll_rows = r; }

static RollGeom make_geom(const llie_phys_desc &d) {
    RollGeom g{};
    g.groups = div_up(d.w, kRollCols);
    const long long planes = (long long)d.n * d.c_out;
    int R = g_roll_rows;
    if (R <= 0) {
        // enough CTAs to give every SM ~5 resident CTAs (20 warps) in a single wave when the problem
        // is small; at most 32 rows per chunk (2-4 halo rows per chunk are L2 re-reads)
        const long long want = (long long)kNumSM * 5;
        const long long cols = planes * g.groups;
        long long r = ((long long)d.h * cols) / want;
        R = (int)(r < 8 ? 8 : (r > 32 ? 32 : r));
    }
    if (R < 2) R = 2;
    if (R > d.h) R = d.h;
    g.R = R;
    g.chunks = d.h / R;  // floor: the last chunk absorbs the remainder (always >= 2 rows)
    if (g.chunks < 1) g.chunks = 1;
    g.ctas_per_sample = d.c_out * g.groups * g.chunks;
    return g;
}

bool phys_roll_eligible(const PhysArgs &a) {
    const llie_phys_desc &d = a.d;
    if (d.psf_layout != LLIE_PSF_DEPTHWISE || d.kh != 3 || d.kw != 3) return false;
    if ((d.w & 3) != 0 || d.w < 8 || d.h < 4) return false;
    if (!aligned16(a.pred) || !aligned16(a.obs)) return false;
    if (a.rho != nullptr && (d.rho_sh != 0 || d.rho_sw != 0)) return false;
    const RollGeom g = make_geom(d);
    const long long ctas = (long long)d.n * g.ctas_per_sample;
    if (ctas > 0x7fffffffLL) return false;
    if (ctas * (long long)(sizeof(double) + sizeof(int)) > (long long)(kWsBytes - kWsHeaderBytes)) return false;
    return true;
}

#define LLIE_ROLL_DISPATCH(KERNEL, d, ...)                                                      \
    do {                                                                                        \
        const int key = ((d).robust ? 4 : 0) | ((d).clip_y ? 2 : 0) | ((d).clip_a ? 1 : 0);     \
        switch (key) {                                                                          \
            case 0: KERNEL<false, false, false> __VA_ARGS__; break;                             \
            case 1: KERNEL<false, false, true> __VA_ARGS__; break;                              \
            case 2: KERNEL<false, true, false> __VA_ARGS__; break;                              \
            case 3: KERNEL<false, true, true> __VA_ARGS__; break;                               \
            case 4: KERNEL<true, false, false> __VA_ARGS__; break;                              \
            case 5: KERNEL<true, false, true> __VA_ARGS__; break;                               \
            case 6: KERNEL<true, true, false> __VA_ARGS__; break;                               \
            default: KERNEL<true, true, true> __VA_ARGS__; break;                               \
        }                                                                                       \
    } while (0)

int launch_phys_fwd_roll(PhysArgs &a, void *workspace, cudaStream_t st) {
    RollGeom g = make_geom(a.d);
    Workspace ws(workspace);
    const int grid = a.d.n * g.ctas_per_sample;
    g.part = reinterpret_cast<double *>(ws.scratch);
    g.pflag = reinterpret_cast<int *>(ws.scratch + (size_t)grid * sizeof(double));
    a.ticket = ws.tickets;
    LLIE_ROLL_DISPATCH(phys_fwd_roll, a.d, <<<grid, kRollThreads, 0, st>>>(a, g));
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

int launch_phys_bwd_roll(PhysArgs &a, cudaStream_t st) {
    const RollGeom g = make_geom(a.d);
    const int grid = a.d.n * g.ctas_per_sample;
    LLIE_ROLL_DISPATCH(phys_bwd_roll, a.d, <<<grid, kRollThreads, 0, st>>>(a, g));
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

}  // namespace llie
