// Exact order statistics / torch.quantile(..., interpolation='linear') on the device by a 3-pass radix
// select over the float bit patterns (11 + 11 + 10 bits), replacing the full sorts the reference
// pays for the edge threshold, the edge p95 and the dE00 percentiles
// (metrics/color_error.py:213-231, 334-342).  Optionally the population is gated by a second array
// (`gate[i] >= gate_thr[image]`, the Sobel edge mask of color_error.py:336-339), in which case the
// first pass also yields the gated count (the bit-exact integer of the edge mask) and the gated sum.
//
// torch semantics reproduced exactly: rank = float(q) * float(n-1) in FLOAT32, lo = floor, hi = ceil,
// result = lerp(v[lo], v[hi], rank - lo) with torch.lerp's two-sided formula.
// All counting is integer (atomics on integers are order-independent) -> bit-reproducible.
#include "common.cuh"

namespace llie {

constexpr int kBins = 2048;
constexpr int kMaxQ = 2;            // quantiles per call
constexpr int kMaxR = 2 * kMaxQ;    // ranks tracked (lo / hi of every quantile)

struct SelState {                    // per segment, lives in the workspace
    unsigned long long count;        // population size
    unsigned long long k[kMaxR];     // remaining rank inside the current prefix
    unsigned prefix[kMaxR];          // decided high bits (monotone-mapped)
    float w[kMaxQ];                  // interpolation weights
};

struct SelArgs {
    llie_select_desc d;
    const float *values, *gate, *gate_thr;
    float q[kMaxQ];
    float *out_q;                    // [segments][nq]
    long long *out_count;            // [images] gated (or plain) element count per image
    double *out_sum;                 // [segments] sum of the participating values (or NULL)
    unsigned long long *hist;        // [segments][kMaxR][kBins]
    unsigned long long *img_count;   // [images]
    double *psum;                    // [images][gridDim.x]
    SelState *state;                 // [segments]
    int pass;                        // 0, 1, 2
    unsigned *ticket;                // last-block ticket of the histogram kernel (fold)
    int fold;                        // 1: one segment -- the last histogram block also runs the scan (no second launch)
};

// monotone map float bits -> unsigned (total order incl. negatives)
__device__ __forceinline__ unsigned fkey(float v) {
    const unsigned b = __float_as_uint(v);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float funkey(unsigned k) {
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}
__device__ __forceinline__ int pass_shift(int pass) { return pass == 0 ? 21 : (pass == 1 ? 10 : 0); }
__device__ __forceinline__ unsigned pass_bins(int pass) { return pass == 2 ? 1024u : 2048u; }
__device__ __forceinline__ unsigned pass_himask(int pass) {  // bits already decided before this pass
    return pass == 0 ? 0u : (pass == 1 ? 0xffe00000u : 0xfffffc00u);
}

__device__ void select_scan_body(const SelArgs &a, int images, int blocks_x, int seg);

// Fast path of the histogram pass: at most one histogram row, 16-byte aligned rows.  GATE: elements take part only
// where gate >= thr; PASS0: count the participating elements (+ SUM: their float64 sum, in a fixed per-thread order) and
// every participating element enters the histogram (no prefix yet); later passes only histogram the elements whose
// decided high bits equal the prefix.  Four independent 128-bit loads per tensor in flight per thread: with one, 2048
// resident threads per SM keep 4.8 MB in flight chip-wide and the pass ran at 2.1 TB/s (latency-bound).
template <bool GATE, bool PASS0, bool SUM>
__device__ __forceinline__ void hist_fast(const float *__restrict__ v, const float *__restrict__ gt, float thr,
                                          long long nv, unsigned *sh, bool want_hist, unsigned himask, unsigned pre0,
                                          int shift, unsigned bmask, unsigned &c32, double &sum) {
    auto elem = [&](float val, float gv) {
        if (GATE && !(gv >= thr)) return;
        const unsigned b = __float_as_uint(val);
        const unsigned key = b ^ ((unsigned)((int)b >> 31) | 0x80000000u);  // == fkey(val)
        if (PASS0) {
            ++c32;
            if (SUM) sum += (double)val;
            if (want_hist) atomicAdd(&sh[key >> shift], 1u);               // pass 0: shift = 21, every key is in range
        } else if ((key & himask) == pre0) {
            atomicAdd(&sh[(key >> shift) & bmask], 1u);
        }
    };
    const long long stride = (long long)gridDim.x * blockDim.x;
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 3 * stride < nv; i += 4 * stride) {
        float4 x[4], gg[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) x[u] = ldg_stream4(v + 4 * (i + u * stride));
        if (GATE) {
#pragma unroll
            for (int u = 0; u < 4; ++u) gg[u] = ldg_stream4(gt + 4 * (i + u * stride));
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            elem(x[u].x, GATE ? gg[u].x : 0.f);
            elem(x[u].y, GATE ? gg[u].y : 0.f);
            elem(x[u].z, GATE ? gg[u].z : 0.f);
            elem(x[u].w, GATE ? gg[u].w : 0.f);
        }
    }
    for (; i < nv; i += stride) {
        const float4 x = ldg_stream4(v + 4 * i);
        float4 gg = make_float4(0.f, 0.f, 0.f, 0.f);
        if (GATE) gg = ldg_stream4(gt + 4 * i);
        elem(x.x, gg.x);
        elem(x.y, gg.y);
        elem(x.z, gg.z);
        elem(x.w, gg.w);
    }
}

// grid (bx, images).  Segment = image, or 0 when d.joint.
__global__ void __launch_bounds__(256) select_hist_kernel(const SelArgs a) {
    extern __shared__ unsigned sh[];  // [nr][bins]
    __shared__ double red[32];
    const llie_select_desc &d = a.d;
    const int img = blockIdx.y, seg = d.joint ? 0 : img;
    const int nr = 2 * d.nq;
    const unsigned bins = pass_bins(a.pass);
    for (unsigned i = threadIdx.x; i < nr * bins; i += blockDim.x) sh[i] = 0u;
    __syncthreads();
    const int shift = pass_shift(a.pass);
    const unsigned himask = pass_himask(a.pass);
    unsigned pre[kMaxR];
#pragma unroll
    for (int r = 0; r < kMaxR; ++r) pre[r] = (a.pass > 0 && r < nr) ? a.state[seg].prefix[r] : 0u;
    const bool same01 = nr >= 2 && pre[0] == pre[1], same23 = nr >= 4 && pre[2] == pre[3];
    const float thr = a.gate ? __ldg(a.gate_thr + img) : 0.f;
    const float *v = a.values + (long long)img * d.per_image;
    const float *gt = a.gate ? a.gate + (long long)img * d.per_image : nullptr;
    unsigned long long cnt = 0;
    double sum = 0.0;
    // Fast path (the common case): at most one histogram row (one quantile whose two ranks still share their
    // prefix, or the count/sum-only call) and 16-byte aligned rows -> 128-bit streaming loads, ~10 instructions
    // per element instead of ~65.
    const bool one_row = nr == 0 || nr == 1 || (nr == 2 && same01);
    const bool vec_ok = (d.per_image & 3) == 0 && (reinterpret_cast<uintptr_t>(v) & 15u) == 0 &&
                        (gt == nullptr || (reinterpret_cast<uintptr_t>(gt) & 15u) == 0);
    if (one_row && vec_ok) {
        const bool want_hist = nr >= 1, pass0 = a.pass == 0, want_sum = pass0 && a.out_sum != nullptr;
        // the run-time switches become template arguments: the pass is a streaming read that should be HBM-bound, and
        // with the switches tested per element it executed 37 instructions per element (62 % issue-active, 2 TB/s)
        const int variant = (gt ? 4 : 0) | (pass0 ? 2 : 0) | (want_sum ? 1 : 0);
        unsigned c32 = 0;
        switch (variant) {
            case 0: hist_fast<false, false, false>(v, gt, thr, d.per_image >> 2, sh, want_hist, himask, pre[0], shift, bins - 1u, c32, sum); break;
            case 2: hist_fast<false, true, false>(v, gt, thr, d.per_image >> 2, sh, want_hist, himask, pre[0], shift, bins - 1u, c32, sum); break;
            case 3: hist_fast<false, true, true>(v, gt, thr, d.per_image >> 2, sh, want_hist, himask, pre[0], shift, bins - 1u, c32, sum); break;
            case 4: hist_fast<true, false, false>(v, gt, thr, d.per_image >> 2, sh, want_hist, himask, pre[0], shift, bins - 1u, c32, sum); break;
            case 6: hist_fast<true, true, false>(v, gt, thr, d.per_image >> 2, sh, want_hist, himask, pre[0], shift, bins - 1u, c32, sum); break;
            default: hist_fast<true, true, true>(v, gt, thr, d.per_image >> 2, sh, want_hist, himask, pre[0], shift, bins - 1u, c32, sum); break;
        }
        cnt = pass0 ? c32 : 0ull;
    } else {
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < d.per_image;
             i += (long long)gridDim.x * blockDim.x) {
            if (gt && !(__ldg(gt + i) >= thr)) continue;
            const float val = __ldg(v + i);
            const unsigned key = fkey(val);
            if (a.pass == 0) {
                ++cnt;
                sum += (double)val;
            }
            const unsigned digit = (key >> shift) & (bins - 1u);
            const unsigned hi = key & himask;
            // ranks that share a prefix share a histogram row (the common case: lo and hi are adjacent ranks)
            if (nr >= 1 && hi == pre[0]) atomicAdd(&sh[digit], 1u);
            if (nr >= 2 && !same01 && hi == pre[1]) atomicAdd(&sh[bins + digit], 1u);
            if (nr >= 4 && hi == pre[2]) atomicAdd(&sh[2 * bins + digit], 1u);
            if (nr >= 4 && !same23 && hi == pre[3]) atomicAdd(&sh[3 * bins + digit], 1u);
        }
    }
    __syncthreads();
    unsigned long long *gh = a.hist + (size_t)seg * kMaxR * kBins;
    for (unsigned i = threadIdx.x; i < nr * bins; i += blockDim.x) {
        const unsigned c = sh[i];
        if (c) atomicAdd(&gh[(i / bins) * kBins + (i % bins)], (unsigned long long)c);
    }
    if (a.pass == 0) {
        // gated count per image (integer) and a deterministic per-block partial of the gated sum
        unsigned long long wc = cnt;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) wc += __shfl_xor_sync(0xffffffffu, wc, o);
        if ((threadIdx.x & 31) == 0 && wc) atomicAdd(&a.img_count[img], wc);
        const double bs = block_sum<double>(sum, red);
        if (threadIdx.x == 0) a.psum[(size_t)img * gridDim.x + blockIdx.x] = bs;
    }
    // one segment: the block that finishes last walks the histogram itself (saves the one-block scan launch per pass)
    if (a.fold && last_block_done(a.ticket, gridDim.x * gridDim.y)) select_scan_body(a, d.images, gridDim.x, 0);
}

// one block per segment: walk the histogram rows, fix the next digit of every rank
// (every value another block of the same launch may have written -- histogram rows, counts, partial sums -- is read
// with __ldcg: the folded form runs inside the histogram kernel)
__device__ void select_scan_body(const SelArgs &a, int images, int blocks_x, int seg) {
    __shared__ unsigned long long chunk[256];
    __shared__ unsigned long long newk[kMaxR];
    __shared__ unsigned newp[kMaxR];
    __shared__ double red[32];
    const llie_select_desc &d = a.d;
    const int nr = 2 * d.nq;
    SelState &st = a.state[seg];
    const unsigned bins = pass_bins(a.pass);
    const int shift = pass_shift(a.pass);
    unsigned long long *gh = a.hist + (size_t)seg * kMaxR * kBins;

    if (a.pass == 0) {
        // population, ranks and weights (torch.quantile float32 arithmetic)
        const int i0 = d.joint ? 0 : seg, i1 = d.joint ? images : seg + 1;
        if (threadIdx.x == 0) {
            unsigned long long c = 0;
            for (int i = i0; i < i1; ++i) {
                const unsigned long long ci = __ldcg(a.img_count + i);
                c += ci;
                a.out_count[i] = (long long)ci;
            }
            st.count = c;
            for (int qi = 0; qi < d.nq; ++qi) {
                const float rank = a.q[qi] * (float)((long long)c - 1);
                const float lo = floorf(rank), hi = ceilf(rank);
                st.k[2 * qi] = c ? (unsigned long long)lo : 0ull;
                st.k[2 * qi + 1] = c ? (unsigned long long)hi : 0ull;
                st.w[qi] = rank - lo;
                st.prefix[2 * qi] = st.prefix[2 * qi + 1] = 0u;
            }
        }
        if (a.out_sum) {  // fixed-order sum of the per-block partials
            double s = 0.0;
            const long long tot = (long long)(i1 - i0) * blocks_x;
            const double *p = a.psum + (size_t)i0 * blocks_x;
            for (long long b = threadIdx.x; b < tot; b += blockDim.x) s += __ldcg(p + b);
            s = block_sum<double>(s, red);
            if (threadIdx.x == 0) a.out_sum[seg] = s;
        }
        __syncthreads();
    }
    for (int r = 0; r < nr; ++r) {
        // rows are shared when the two ranks of a quantile had the same prefix going into this pass
        const bool shared_row = (r & 1) && st.prefix[r] == st.prefix[r - 1] && a.pass >= 0;
        const unsigned long long *row = gh + (size_t)(shared_row ? r - 1 : r) * kBins;
        const unsigned long long k = st.k[r];
        // 256 threads x (bins / 256) bins each: block-wide exclusive scan of the chunk sums, then the one thread
        // whose chunk holds rank k walks its own <= 8 bins (registers)
        const int per = bins / 256;
        unsigned long long bv[8], cs = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            bv[j] = j < per ? __ldcg(row + threadIdx.x * per + j) : 0ull;
            cs += bv[j];
        }
        const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
        unsigned long long incl = cs;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) chunk[wid] = incl;
        __syncthreads();
        unsigned long long base = 0;
        for (int w2 = 0; w2 < wid; ++w2) base += chunk[w2];
        const unsigned long long excl = base + incl - cs;
        bool sel = excl <= k && k < excl + cs;
        const int any = __syncthreads_or(sel ? 1 : 0);   // also orders the chunk[] reads before the writes below
        if (!any && threadIdx.x == 255) sel = true;       // rank beyond the population (empty set): last chunk
        if (sel) {
            unsigned long long cum = excl;
            int jb = 0;
#pragma unroll
            for (int j = 0; j < 7; ++j)
                if (jb == j && j < per - 1 && cum + bv[j] <= k) {
                    cum += bv[j];
                    jb = j + 1;
                }
            chunk[0] = (unsigned long long)(threadIdx.x * per + jb);
            chunk[1] = k - cum;
        }
        __syncthreads();
        const unsigned digit = (unsigned)chunk[0];
        const unsigned long long knew = chunk[1];
        __syncthreads();
        if (threadIdx.x == 0) {
            // NOTE: prefix[r] is updated only after a possible sharing partner (r+1) compared against
            // the OLD prefix[r]; handled by deferring the write through `newp`
            newp[r] = st.prefix[r] | (digit << shift);
            newk[r] = knew;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        for (int r = 0; r < nr; ++r) {
            st.prefix[r] = newp[r];
            st.k[r] = newk[r];
        }
    }
    __syncthreads();
    // clear the histogram rows for the next pass / next call
    for (unsigned i = threadIdx.x; i < (unsigned)kMaxR * kBins; i += blockDim.x) gh[i] = 0ull;
    if ((a.pass == 2 || d.nq == 0) && threadIdx.x == 0) {
        for (int qi = 0; qi < d.nq; ++qi) {
            float res = __int_as_float(0x7fc00000);  // empty population -> NaN (color_error.py:340-341)
            if (st.count) {
                const float lo = funkey(st.prefix[2 * qi]), hi = funkey(st.prefix[2 * qi + 1]);
                const float w = st.w[qi];
                // torch.lerp: w < 0.5 ? a + w (b - a) : b - (b - a) (1 - w)
                const float df = hi - lo;
                res = w < 0.5f ? lo + w * df : hi - df * (1.f - w);
            }
            a.out_q[seg * d.nq + qi] = res;
        }
        const int i0 = d.joint ? 0 : seg, i1 = d.joint ? images : seg + 1;
        for (int i = i0; i < i1; ++i) a.img_count[i] = 0ull;  // self-reset
    }
}

__global__ void __launch_bounds__(256) select_scan_kernel(const SelArgs a, int images, int blocks_x) {
    select_scan_body(a, images, blocks_x, blockIdx.x);
}

}  // namespace llie

using namespace llie;

extern "C" size_t llie_select_workspace_bytes(int32_t images, int32_t blocks_hint) {
    (void)blocks_hint;
    const size_t segs = (size_t)images;
    return segs * kMaxR * kBins * sizeof(unsigned long long) + segs * sizeof(unsigned long long) +
           segs * sizeof(SelState) + segs * (size_t)(kNumSM * 8) * sizeof(double) + 256 + 64;  // + ticket
}

extern "C" int llie_select_quantiles(const llie_select_desc *desc, const float *values, const float *gate,
                                     const float *gate_thr, const double *q, float *out_q, int64_t *out_count,
                                     double *out_sum, void *select_ws, void *stream) {
    LLIE_REQUIRE(desc && values && out_count && select_ws, "select: null pointer");
    LLIE_REQUIRE(desc->nq == 0 || (q && out_q), "select: q / out_q missing");
    const llie_select_desc &d = *desc;
    LLIE_REQUIRE(d.images > 0 && d.images <= 65535 && d.per_image > 0, "select: bad population");
    LLIE_REQUIRE(d.nq >= 0 && d.nq <= kMaxQ, "select: 0..%d quantiles per call", kMaxQ);
    LLIE_REQUIRE((gate == nullptr) == (gate_thr == nullptr), "select: gate and gate_thr go together");
    for (int i = 0; i < d.nq; ++i) LLIE_REQUIRE(q[i] >= 0.0 && q[i] <= 1.0, "select: q must be in [0,1]");
    SelArgs a{};
    a.d = d;
    a.values = values; a.gate = gate; a.gate_thr = gate_thr;
    for (int i = 0; i < d.nq; ++i) a.q[i] = (float)q[i];
    a.out_q = out_q; a.out_count = reinterpret_cast<long long *>(out_count); a.out_sum = out_sum;
    const int segs = d.joint ? 1 : d.images;
    static int occ_nq[8] = {0, 0, 0, 0, 0, 0, 0, 0};  // resident blocks per SM by nq (dynamic shared memory grows with it)
    const int oi = d.nq < 7 ? d.nq : 7;
    const int sel_threads = 256;
    if (occ_nq[oi] == 0 || oi == 7)
        occ_nq[oi] = resident_blocks(select_hist_kernel, sel_threads, (size_t)2 * d.nq * kBins * sizeof(unsigned));
    const int per_sm = occ_nq[oi] < 8 ? occ_nq[oi] : 8;  // the partial-sum workspace holds kNumSM * 8 slots per segment
    int bx = kNumSM * per_sm / d.images;
    const long long maxb = (d.per_image + 4 * sel_threads - 1) / (4 * sel_threads);
    if (bx > maxb) bx = (int)maxb;
    if (bx < 1) bx = 1;
    unsigned char *w = static_cast<unsigned char *>(select_ws);
    a.hist = reinterpret_cast<unsigned long long *>(w);
    w += (size_t)d.images * kMaxR * kBins * sizeof(unsigned long long);
    a.img_count = reinterpret_cast<unsigned long long *>(w);
    w += (size_t)d.images * sizeof(unsigned long long);
    a.state = reinterpret_cast<SelState *>(w);
    w += (size_t)d.images * sizeof(SelState);
    a.psum = reinterpret_cast<double *>(((uintptr_t)w + 7) & ~(uintptr_t)7);
    a.ticket = reinterpret_cast<unsigned *>(a.psum + (size_t)d.images * (kNumSM * 8));  // inside the trailing slack
    a.fold = segs == 1 ? 1 : 0;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const size_t smem = (size_t)2 * d.nq * kBins * sizeof(unsigned);
    const int passes = d.nq == 0 ? 1 : 3;  // nq == 0: gated count + sum only, one pass
    for (int pass = 0; pass < passes; ++pass) {
        a.pass = pass;
        select_hist_kernel<<<dim3(bx, d.images), sel_threads, smem, st>>>(a);
        if (!a.fold) select_scan_kernel<<<segs, 256, 0, st>>>(a, d.images, bx);
    }
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}
