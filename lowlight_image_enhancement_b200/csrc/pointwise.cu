// PSNR family (sm_100a): one streaming pass producing float64 sums of squared differences per (n,c)
// plane, global min/max and finiteness flags; the last block finalises every PSNR flavour the
// reference defines (metrics/linear.py:139-210, metrics/channelwise.py:113-222, metrics/psnr.py:39-88).
// HBM-bound: 8 B/element (f32) read once, 128-bit loads, fixed-order two-stage reduction.
#include <math.h>

#include "common.cuh"

// waves of resident blocks for the value + gradient passes: 1 / 4 / 12 -> 82 / 91 / 107 us per 36 M elements (the
// per-block reduction epilogue outweighs finer scheduling; tools/time_pix.py)
#ifndef LLIE_PIX_WAVES
#define LLIE_PIX_WAVES 1
#endif
#ifndef LLIE_PSNR_WAVES
#define LLIE_PSNR_WAVES 1
#endif
namespace llie {

struct PsnrArgs {
    llie_psnr_desc d;
    const void *x, *y;
    double *out;
    void *out_minmax;
    int *out_flags;
    double *part_sum;  // [n*c][G]
    double *part_min;  // [n*c][G]
    double *part_max;
    int *part_flag;
    unsigned *ticket;
    int vec_ok;
};

template <typename T>
struct Vec;
template <>
struct Vec<float> {
    using type = float4;
    static constexpr int N = 4;
};
template <>
struct Vec<double> {
    using type = double2;
    static constexpr int N = 2;
};

template <typename T>
__device__ __forceinline__ bool is_finite(T v);
template <>
__device__ __forceinline__ bool is_finite<float>(float v) { return finite_f(v); }
template <>
__device__ __forceinline__ bool is_finite<double>(double v) {
    return (__double2hiint(v) & 0x7ff00000) != 0x7ff00000;
}

template <typename T>
struct Acc {
    double s = 0.0;
    T mn, mx;
    int flag = 0;
    const llie_psnr_desc &d;
    T hi_t;
    __device__ Acc(const llie_psnr_desc &dd) : d(dd) {
        mn = (T)INFINITY;
        mx = (T)-INFINITY;
        hi_t = (T)dd.clamp_hi;
    }
    __device__ __forceinline__ void add(T xv, T yv) {
        if (!is_finite(xv)) flag |= 1;
        if (!is_finite(yv)) flag |= 2;
        mn = min(mn, min(xv, yv));
        mx = max(mx, max(xv, yv));
        double df;
        if (d.cast_first) {
            double a = (double)xv, b = (double)yv;
            if (d.clamp) {
                a = fmin(fmax(a, 0.0), d.clamp_hi);
                b = fmin(fmax(b, 0.0), d.clamp_hi);
            }
            df = a - b;
        } else {
            T a = xv, b = yv;
            if (d.clamp) {
                a = min(max(a, (T)0), hi_t);
                b = min(max(b, (T)0), hi_t);
            }
            df = (double)(a - b);
        }
        s = fma(df, df, s);
    }
};

__device__ __forceinline__ void unpack(const float4 &v, float o[4]) { o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w; }
__device__ __forceinline__ void unpack(const double2 &v, double o[2]) { o[0] = v.x; o[1] = v.y; }

__device__ double psnr_db(double range2, double mse, double eps) { return 10.0 * log10(range2 / fmax(mse, eps)); }

template <typename T>
__global__ void __launch_bounds__(256) psnr_stats_kernel(const PsnrArgs a) {
    __shared__ double red[32];
    using V = typename Vec<T>::type;
    constexpr int VN = Vec<T>::N;
    const llie_psnr_desc &d = a.d;
    const int nc = blockIdx.y;
    const int G = gridDim.x;
    const T *xp = static_cast<const T *>(a.x) + (long long)nc * d.plane;
    const T *yp = static_cast<const T *>(a.y) + (long long)nc * d.plane;
    Acc<T> acc(d);
    const long long stride = (long long)G * blockDim.x;
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (a.vec_ok) {
        const long long nv = d.plane / VN;
        const V *xv = reinterpret_cast<const V *>(xp);
        const V *yv = reinterpret_cast<const V *>(yp);
        // 2 independent 128-bit loads per tensor in flight per thread
        for (; i + stride < nv; i += 2 * stride) {
            const V x0 = __ldcs(xv + i), x1 = __ldcs(xv + i + stride);
            const V y0 = __ldcs(yv + i), y1 = __ldcs(yv + i + stride);
            T xa[VN], ya[VN];
            unpack(x0, xa); unpack(y0, ya);
#pragma unroll
            for (int k = 0; k < VN; ++k) acc.add(xa[k], ya[k]);
            unpack(x1, xa); unpack(y1, ya);
#pragma unroll
            for (int k = 0; k < VN; ++k) acc.add(xa[k], ya[k]);
        }
        for (; i < nv; i += stride) {
            T xa[VN], ya[VN];
            unpack(__ldcs(xv + i), xa); unpack(__ldcs(yv + i), ya);
#pragma unroll
            for (int k = 0; k < VN; ++k) acc.add(xa[k], ya[k]);
        }
        // tail (plane % VN elements)
        for (long long t = nv * VN + blockIdx.x * blockDim.x + threadIdx.x; t < d.plane; t += stride)
            acc.add(xp[t], yp[t]);
    } else {
        for (; i < d.plane; i += stride) acc.add(__ldcs(xp + i), __ldcs(yp + i));
    }
    // block reduce: sum, min, max, flags
    const double s = block_sum<double>(acc.s, red);
    double mn = (double)acc.mn, mx = (double)acc.mx;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    __shared__ double smn[8], smx[8];
    __syncthreads();
    if ((threadIdx.x & 31) == 0) { smn[threadIdx.x >> 5] = mn; smx[threadIdx.x >> 5] = mx; }
    const int f = block_or(acc.flag, reinterpret_cast<int *>(red));
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w) { mn = fmin(mn, smn[w]); mx = fmax(mx, smx[w]); }
        const size_t slot = (size_t)nc * G + blockIdx.x;
        a.part_sum[slot] = s;
        a.part_min[slot] = mn;
        a.part_max[slot] = mx;
        a.part_flag[slot] = f;
    }
    if (!last_block_done(a.ticket, gridDim.x * gridDim.y)) return;

    // ---------------- finalize (one block): per-(n,c) totals, then the PSNR table
    const int NC = d.n * d.c;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    double *tot = a.part_sum;  // reuse slot [row*G + 0] for the row total after reading the row
    double gmn = INFINITY, gmx = -INFINITY;
    int gf = 0;
    for (int row = wid; row < NC; row += 8) {
        double ss = 0.0;
        for (int b = lane; b < G; b += 32) {
            const size_t slot = (size_t)row * G + b;
            ss += __ldcg(a.part_sum + slot);
            gmn = fmin(gmn, __ldcg(a.part_min + slot));
            gmx = fmax(gmx, __ldcg(a.part_max + slot));
            gf |= __ldcg(a.part_flag + slot);
        }
        ss = warp_sum(ss);
        __syncwarp();
        if (lane == 0) tot[(size_t)row * G] = ss;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        gmn = fmin(gmn, __shfl_xor_sync(0xffffffffu, gmn, o));
        gmx = fmax(gmx, __shfl_xor_sync(0xffffffffu, gmx, o));
    }
    __syncthreads();
    if (lane == 0) { smn[wid] = gmn; smx[wid] = gmx; }
    const int ff = block_or(gf, reinterpret_cast<int *>(red));
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w) { gmn = fmin(gmn, smn[w]); gmx = fmax(gmx, smx[w]); }
        if (a.out_minmax) {
            if (d.dtype == 0) {
                static_cast<float *>(a.out_minmax)[0] = (float)gmn;
                static_cast<float *>(a.out_minmax)[1] = (float)gmx;
            } else {
                static_cast<double *>(a.out_minmax)[0] = gmn;
                static_cast<double *>(a.out_minmax)[1] = gmx;
            }
        }
        *a.out_flags = ff;
    }
    // table rows: one thread per sample
    const int C = d.c, cols = 2 * C + 3;
    const double r2 = d.data_range * d.data_range;
    for (int n = threadIdx.x; n < d.n; n += blockDim.x) {
        double *o = a.out + (size_t)n * cols;
        double msum = 0.0, psum = 0.0;
        for (int c = 0; c < C; ++c) {
            const double mse = tot[(size_t)(n * C + c) * G] / (double)d.plane;
            o[c] = mse;
            const double p = psnr_db(r2, mse, d.eps);
            o[C + c] = p;
            msum += mse;
            psum += p;
        }
        o[2 * C] = psum / C;
        o[2 * C + 1] = psnr_db(r2, msum / C, d.eps);
        const double mall = msum / C;  // equal plane sizes: mean over C*H*W == mean of channel means
        o[2 * C + 2] = (mall <= d.eps) ? INFINITY : psnr_db(r2, mall, d.eps);
    }
    __syncthreads();
    for (int col = threadIdx.x; col < cols; col += blockDim.x) {
        double sacc = 0.0;
        for (int n = 0; n < d.n; ++n) sacc += a.out[(size_t)n * cols + col];
        a.out[(size_t)d.n * cols + col] = sacc / d.n;
        a.out[(size_t)(d.n + 1) * cols + col] = sacc;
    }
}

}  // namespace llie

using namespace llie;

extern "C" int llie_psnr_stats(const llie_psnr_desc *desc, const void *x, const void *y, double *out,
                               void *out_minmax, int32_t *out_flags, void *workspace, void *stream) {
    LLIE_REQUIRE(desc && x && y && out && out_flags && workspace, "psnr_stats: null pointer");
    LLIE_REQUIRE(desc->n > 0 && desc->c > 0 && desc->plane > 0, "psnr_stats: empty tensor");
    LLIE_REQUIRE((long long)desc->n * desc->c <= 65535, "psnr_stats: n*c too large for grid.y");
    LLIE_REQUIRE(desc->dtype == 0 || desc->dtype == 1, "psnr_stats: dtype must be 0 (f32) or 1 (f64)");
    LLIE_REQUIRE(desc->data_range > 0 && desc->eps > 0, "psnr_stats: data_range and eps must be positive");
    PsnrArgs a{};
    a.d = *desc;
    a.x = x; a.y = y; a.out = out; a.out_minmax = out_minmax; a.out_flags = out_flags;
    const int nc = desc->n * desc->c;
    const int vn = desc->dtype ? 2 : 4;
    const size_t esz = desc->dtype ? 8 : 4;
    a.vec_ok = aligned16(x) && aligned16(y) && ((desc->plane * esz) % 16 == 0);
    long long per_block = 256LL * vn * 4;
    // exactly one wave of resident blocks (the register count, not the 2048-thread limit, decides how many fit)
    static int occ[2] = {0, 0};
    if (occ[desc->dtype] == 0) {
        int o = 0;
        if (desc->dtype == 0)
            LLIE_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, psnr_stats_kernel<float>, 256, 0));
        else
            LLIE_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, psnr_stats_kernel<double>, 256, 0));
        occ[desc->dtype] = o < 1 ? 1 : o;
    }
    int G = (kNumSM * occ[desc->dtype] * LLIE_PSNR_WAVES) / nc;
    long long maxg = (desc->plane + per_block - 1) / per_block;
    if (G > maxg) G = (int)maxg;
    if (G < 1) G = 1;
    Workspace ws(workspace);
    const size_t slots = (size_t)nc * G;
    LLIE_REQUIRE(slots * (3 * sizeof(double) + sizeof(int)) <= ws.scratch_bytes, "psnr_stats: workspace too small");
    a.part_sum = reinterpret_cast<double *>(ws.scratch);
    a.part_min = a.part_sum + slots;
    a.part_max = a.part_min + slots;
    a.part_flag = reinterpret_cast<int *>(a.part_max + slots);
    a.ticket = ws.tickets;
    dim3 grid(G, nc);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (desc->dtype == 0)
        psnr_stats_kernel<float><<<grid, 256, 0, st>>>(a);
    else
        psnr_stats_kernel<double><<<grid, 256, 0, st>>>(a);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

// ---------------------------------------------------------------------------------------------------
// llie_psnr_stats_static: the same kernel and outputs as llie_psnr_stats, for callers that can neither allocate nor
// carry a workspace (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  The scratch the kernel
// needs -- the "last block" ticket and the per-block partial records -- lives in ONE file-scope device array per GPU
// that loads the library (LLIE_PSNR_STATIC_SCRATCH_BYTES, held until the process exits).  No counterpart in the
// reference.
//   * In host code `g_psnr_static_scratch` is only a stand-in for the symbol: its device address is looked up once per
//     device, with that device current, by cudaGetSymbolAddress and cached; the kernel receives that pointer.
//   * The entry point calls no cudaMalloc / cudaFree / allocating routine; cudaMemsetAsync on the caller's stream
//     clears the ticket header before every launch (the array keeps the previous call's contents; the partial
//     records are fully rewritten by the launch that reads them).
//   * Every call on a device shares the array: calls that overlap on the GPU (different streams, or graphs replayed
//     concurrently) RACE on it.  Serialise them on one stream, or use llie_psnr_stats with one workspace per stream.
namespace llie {
constexpr size_t kPsnrStaticSlots = LLIE_PSNR_STATIC_MAX_SLOTS;
constexpr size_t kPsnrStaticBytes = LLIE_PSNR_STATIC_SCRATCH_BYTES;
static_assert(kPsnrStaticBytes >= kWsHeaderBytes + kPsnrStaticSlots * (3 * sizeof(double) + sizeof(int)),
              "static scratch too small for its slot count");
static __device__ __align__(256) unsigned char g_psnr_static_scratch[kPsnrStaticBytes];
constexpr int kMaxDevices = 64;
static unsigned char *g_psnr_static_ptr[kMaxDevices];  // device address of the array, per device (0 = not looked up)
}  // namespace llie

extern "C" int llie_psnr_stats_static(const llie_psnr_desc *desc, const void *x, const void *y, double *out,
                                      void *out_minmax, int32_t *out_flags, void *stream) {
    LLIE_REQUIRE(desc && x && y && out && out_flags, "psnr_stats_static: null pointer");
    LLIE_REQUIRE(desc->n > 0 && desc->c > 0 && desc->plane > 0, "psnr_stats_static: empty tensor");
    LLIE_REQUIRE(desc->dtype == 0 || desc->dtype == 1, "psnr_stats_static: dtype must be 0 (f32) or 1 (f64)");
    LLIE_REQUIRE(desc->data_range > 0 && desc->eps > 0, "psnr_stats_static: data_range and eps must be positive");
    const long long nc = (long long)desc->n * desc->c;
    LLIE_REQUIRE(nc <= (long long)kPsnrStaticSlots,
                 "psnr_stats_static: n*c = %lld planes exceed the static scratch (LLIE_PSNR_STATIC_MAX_SLOTS = %d "
                 "partial records, %d bytes per device); use llie_psnr_stats with a caller-owned workspace",
                 nc, (int)kPsnrStaticSlots, (int)kPsnrStaticBytes);
    int dev = -1;
    LLIE_CUDA_OK(cudaGetDevice(&dev));
    LLIE_REQUIRE(dev >= 0 && dev < kMaxDevices, "psnr_stats_static: device index %d out of range", dev);
    unsigned char *base = g_psnr_static_ptr[dev];
    if (base == nullptr) {  // once per device, that device current (benign race: every thread stores the same value)
        void *p = nullptr;
        LLIE_CUDA_OK(cudaGetSymbolAddress(&p, g_psnr_static_scratch));
        LLIE_REQUIRE(p != nullptr, "psnr_stats_static: cudaGetSymbolAddress returned a null address");
        base = g_psnr_static_ptr[dev] = static_cast<unsigned char *>(p);
    }
    PsnrArgs a{};
    a.d = *desc;
    a.x = x; a.y = y; a.out = out; a.out_minmax = out_minmax; a.out_flags = out_flags;
    const int vn = desc->dtype ? 2 : 4;
    const size_t esz = desc->dtype ? 8 : 4;
    a.vec_ok = aligned16(x) && aligned16(y) && ((desc->plane * esz) % 16 == 0);
    const long long per_block = 256LL * vn * 4;
    static int occ[2] = {0, 0};
    if (occ[desc->dtype] == 0) {
        int o = 0;
        if (desc->dtype == 0)
            LLIE_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, psnr_stats_kernel<float>, 256, 0));
        else
            LLIE_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, psnr_stats_kernel<double>, 256, 0));
        occ[desc->dtype] = o < 1 ? 1 : o;
    }
    // the grid of llie_psnr_stats (one wave of resident blocks), capped by the slots the static array holds
    long long G = ((long long)kNumSM * occ[desc->dtype] * LLIE_PSNR_WAVES) / nc;
    const long long maxg = (desc->plane + per_block - 1) / per_block;
    if (G > maxg) G = maxg;
    if (G * nc > (long long)kPsnrStaticSlots) G = (long long)kPsnrStaticSlots / nc;
    if (G < 1) G = 1;
    const size_t slots = (size_t)nc * (size_t)G;
    unsigned char *scratch = base + kWsHeaderBytes;
    a.part_sum = reinterpret_cast<double *>(scratch);
    a.part_min = a.part_sum + slots;
    a.part_max = a.part_min + slots;
    a.part_flag = reinterpret_cast<int *>(a.part_max + slots);
    a.ticket = reinterpret_cast<unsigned *>(base);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    LLIE_CUDA_OK(cudaMemsetAsync(base, 0, kWsHeaderBytes, st));  // tickets must start at zero; no allocation
    dim3 grid((unsigned)G, (unsigned)nc);
    if (desc->dtype == 0)
        psnr_stats_kernel<float><<<grid, 256, 0, st>>>(a);
    else
        psnr_stats_kernel<double><<<grid, 256, 0, st>>>(a);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}


// ---------------------------------------------------------------------------------------------------
// nn.L1Loss (global mean) value + gradient in one streaming pass (NewBP_model/losses.py:249, 332):
//   out_sum = sum |x - y| (float64),  grad = scale * sign(x - y)  [ + grad when accumulate ]
// 8 B/element read (+ 4 read when accumulating), 4 written; sign(0) = 0 as in ATen.
namespace llie {

struct L1Args {
    const float *x, *y;
    float *grad;
    double *out_sum;
    int *out_flags;
    double *part;
    int *pflag;
    unsigned *ticket;
    long long n;
    float scale;
    int accumulate, vec_ok;
};

__device__ __forceinline__ float l1_term(float xv, float yv, float scale, float &acc, float &chk_x, float &chk_y) {
    chk_x = fmaf(xv, 0.f, chk_x);
    chk_y = fmaf(yv, 0.f, chk_y);
    const float dlt = xv - yv;
    acc += fabsf(dlt);
    return dlt > 0.f ? scale : (dlt < 0.f ? -scale : 0.f);
}

__global__ void __launch_bounds__(256) l1_grad_kernel(const L1Args a) {
    __shared__ double red[32];
    float acc = 0.f, chk_x = 0.f, chk_y = 0.f;
    double dacc = 0.0;
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, nth = (long long)gridDim.x * blockDim.x;
    if (a.vec_ok) {
        const long long nv = a.n >> 2;
        float4 *g4 = reinterpret_cast<float4 *>(a.grad);
        int it = 0;
        for (long long i = tid; i < nv; i += nth) {
            const float4 xv = ldg_stream4(a.x + 4 * i), yv = ldg_stream4(a.y + 4 * i);
            float4 g;
            g.x = l1_term(xv.x, yv.x, a.scale, acc, chk_x, chk_y);
            g.y = l1_term(xv.y, yv.y, a.scale, acc, chk_x, chk_y);
            g.z = l1_term(xv.z, yv.z, a.scale, acc, chk_x, chk_y);
            g.w = l1_term(xv.w, yv.w, a.scale, acc, chk_x, chk_y);
            if (g4) {
                if (a.accumulate) {
                    const float4 o = g4[i];
                    g.x += o.x; g.y += o.y; g.z += o.z; g.w += o.w;
                }
                g4[i] = g;
            }
            if (++it == 64) {  // bound the float32 run length: 256 terms per flush
                dacc += (double)acc;
                acc = 0.f;
                it = 0;
            }
        }
    } else {
        int it = 0;
        for (long long i = tid; i < a.n; i += nth) {
            float g = l1_term(__ldg(a.x + i), __ldg(a.y + i), a.scale, acc, chk_x, chk_y);
            if (a.grad) {
                if (a.accumulate) g += a.grad[i];
                a.grad[i] = g;
            }
            if (++it == 256) {
                dacc += (double)acc;
                acc = 0.f;
                it = 0;
            }
        }
    }
    dacc += (double)acc;
    const double bs = block_sum<double>(dacc, red);
    int flag = 0;
    if (!(chk_x == 0.f)) flag |= 1;
    if (!(chk_y == 0.f)) flag |= 2;
    const int f = block_or(flag, reinterpret_cast<int *>(red));
    if (threadIdx.x == 0) {
        a.part[blockIdx.x] = bs;
        a.pflag[blockIdx.x] = f;
    }
    if (!last_block_done(a.ticket, gridDim.x)) return;
    double s0 = 0.0;
    int fl = 0;
    for (int b = threadIdx.x; b < (int)gridDim.x; b += blockDim.x) {
        s0 += __ldcg(a.part + b);
        fl |= __ldcg(a.pflag + b);
    }
    const double t = block_sum<double>(s0, red);
    const int ff = block_or(fl, reinterpret_cast<int *>(red));
    if (threadIdx.x == 0) {
        *a.out_sum = t;
        *a.out_flags = ff;
    }
}

}  // namespace llie

extern "C" int llie_l1_grad(const float *x, const float *y, int64_t n, float grad_scale, int32_t accumulate,
                            double *out_sum, float *out_grad, int32_t *out_flags, void *workspace, void *stream) {
    LLIE_REQUIRE(x && y && out_sum && out_flags && workspace && n > 0, "l1_grad: bad arguments");
    llie::L1Args a{};
    a.x = x; a.y = y; a.grad = out_grad; a.out_sum = out_sum; a.out_flags = out_flags;
    a.n = n; a.scale = grad_scale; a.accumulate = accumulate;
    a.vec_ok = llie::aligned16(x) && llie::aligned16(y) && (!out_grad || llie::aligned16(out_grad)) && (n % 4 == 0);
    long long blocks = (n / 4 + 255) / 256;
    static const int occ = llie::resident_blocks(llie::l1_grad_kernel, 256);
    const long long cap = llie::kNumSM * occ * LLIE_PIX_WAVES;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    llie::Workspace ws(workspace);
    a.part = reinterpret_cast<double *>(ws.scratch);
    a.pflag = reinterpret_cast<int *>(ws.scratch + (size_t)blocks * sizeof(double));
    a.ticket = ws.tickets;
    llie::l1_grad_kernel<<<(int)blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(a);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}


// ---------------------------------------------------------------------------------------------------
// Overlapping-tile merge of the validation driver (image_restoration_model.py:215-241 `grids_inverse`):
//   out[c, y, x] = ( sum over tiles t covering (y, x), in tile order, of tiles[t, c, y - i_t, x - j_t] ) / count
// The reference does this on the HOST (zeros + slice adds + divide, then one upload); here every output
// pixel gathers its covering tiles on the device in the same order, so the fp32 result is identical.
namespace llie {

// A CTA owns an 8-row x 128-column rectangle of the output (all channels).  It first culls the tile list to the tiles
// that intersect its rectangle -- in ascending tile order, so that every pixel still adds its covering tiles in the
// reference's order -- and every pixel then walks that short list (<= 4 entries on a regular grid) instead of all T
// tiles (54 tiles of 512^2 over a 2848 x 4256 image: 2.04 ms -> see DESIGN.md for the measured time).
constexpr int kMergeRows = 8, kMergeCols = 128, kMergeList = 64;

__global__ void __launch_bounds__(256) tile_merge_kernel(const float *tiles, const int *idx, int T, int C, int th,
                                                         int tw, int H, int W, float *out) {
    __shared__ int s_list[kMergeList];
    __shared__ int s_warp[8];
    __shared__ int s_count;
    const int x0 = blockIdx.x * kMergeCols, y0 = blockIdx.y * kMergeRows;
    const int x1 = min(x0 + kMergeCols, W), y1 = min(y0 + kMergeRows, H);
    const int tid = threadIdx.x, lane = tid & 31, wrp = tid >> 5;
    if (tid == 0) s_count = 0;
    __syncthreads();
    for (int t0 = 0; t0 < T; t0 += 256) {  // ordered compaction, 256 tiles per round
        const int t = t0 + tid;
        bool hit = false;
        if (t < T) {
            const int i = __ldg(idx + 2 * t), j = __ldg(idx + 2 * t + 1);
            hit = i < y1 && i + th > y0 && j < x1 && j + tw > x0;
        }
        const unsigned m = __ballot_sync(0xffffffffu, hit);
        if (lane == 0) s_warp[wrp] = __popc(m);
        __syncthreads();
        int base = s_count;
        for (int w2 = 0; w2 < wrp; ++w2) base += s_warp[w2];
        const int pos = base + __popc(m & ((1u << lane) - 1u));
        if (hit && pos < kMergeList) s_list[pos] = t;
        __syncthreads();
        if (tid == 0) {
            int tot = 0;
            for (int w2 = 0; w2 < 8; ++w2) tot += s_warp[w2];
            s_count += tot;
        }
        __syncthreads();
    }
    const int count = s_count;
    const bool culled = count <= kMergeList;   // otherwise (pathologically overlapping tiles) scan all T
    const int L = culled ? count : T;
    const int y = y0 + (tid >> 5);
    if (y >= H) return;
    for (int c = 0; c < C; ++c) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int x = x0 + lane + 32 * q;   // consecutive lanes -> consecutive columns
            if (x >= W) continue;
            float acc = 0.f, cnt = 0.f;
            for (int k = 0; k < L; ++k) {
                const int t = culled ? s_list[k] : k;
                const int i = __ldg(idx + 2 * t), j = __ldg(idx + 2 * t + 1);
                if (y >= i && y < i + th && x >= j && x < j + tw) {
                    acc += __ldg(tiles + (((long long)t * C + c) * th + (y - i)) * tw + (x - j));
                    cnt += 1.f;
                }
            }
            out[((long long)c * H + y) * W + x] = acc / cnt;  // an uncovered pixel is 0 / 0 = NaN, as in the reference
        }
    }
}

}  // namespace llie

extern "C" int llie_tile_merge(const float *tiles, const int32_t *tile_ij, int32_t n_tiles, int32_t c, int32_t tile_h,
                               int32_t tile_w, int32_t h, int32_t w, float *out, void *stream) {
    LLIE_REQUIRE(tiles && tile_ij && out && n_tiles > 0 && c > 0 && tile_h > 0 && tile_w > 0 && h > 0 && w > 0,
                 "tile_merge: bad arguments");
    const dim3 grid((w + llie::kMergeCols - 1) / llie::kMergeCols, (h + llie::kMergeRows - 1) / llie::kMergeRows);
    LLIE_REQUIRE(grid.y <= 65535, "tile_merge: image too tall for the grid");
    llie::tile_merge_kernel<<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(tiles, tile_ij, n_tiles, c, tile_h,
                                                                                      tile_w, h, w, out);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}


// ---------------------------------------------------------------------------------------------------
// BasicSR pixel losses (NAFNet_base/basicsr/models/losses/losses.py:18-31, loss_util.py:31-60): value and gradient
// of  sum_e w_e psi(x_e - y_e)  in one streaming pass, psi = |d| (L1), d^2 (MSE) or sqrt(d^2 + eps) (Charbonnier;
// eps is NOT squared there), optional element weights w [N, 1|C, H, W] and an optional per-sample device
// coefficient (PSNRLoss backward).  grad = scale * coef[n] * w * psi'(d).
namespace llie {

struct PixArgs {
    const float *x, *y, *w, *coef;
    float *grad;
    double *out_sum;
    int *out_flags;
    double *part;
    int *pflag;
    unsigned *ticket;
    long long n_elem, plane, chw;
    int mode, w_per_channel, accumulate;
    float eps, scale;
};

__device__ __forceinline__ float pix_term(int mode, float eps, float d, float &dpsi) {
    if (mode == LLIE_PIX_L1) {
        dpsi = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
        return fabsf(d);
    }
    if (mode == LLIE_PIX_MSE) {
        dpsi = 2.f * d;
        return d * d;
    }
    const float r = sqrtf(fmaf(d, d, eps));
    dpsi = d / r;
    return r;
}

__global__ void __launch_bounds__(256) pixel_loss_kernel(const PixArgs a) {
    __shared__ double red[32];
    float acc = 0.f, chk_x = 0.f, chk_y = 0.f;
    double dacc = 0.0;
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, nth = (long long)gridDim.x * blockDim.x;
    const long long nv = a.n_elem >> 2;  // launcher guarantees plane % 4 == 0 and 16-byte alignment
    int it = 0;
    for (long long i = tid; i < nv; i += nth) {
        const long long e = 4 * i;
        const float4 xv = ldg_stream4(a.x + e), yv = ldg_stream4(a.y + e);
        float4 wv = make_float4(1.f, 1.f, 1.f, 1.f);
        const long long n = e / a.chw;
        if (a.w) {
            const long long we = a.w_per_channel ? e : n * a.plane + (e % a.plane);
            wv = ldg_stream4(a.w + we);
        }
        const float s = a.coef ? a.scale * __ldg(a.coef + n) : a.scale;
        const float xs[4] = {xv.x, xv.y, xv.z, xv.w}, ys[4] = {yv.x, yv.y, yv.z, yv.w};
        const float ws[4] = {wv.x, wv.y, wv.z, wv.w};
        float g[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            chk_x = fmaf(xs[k], 0.f, chk_x);
            chk_y = fmaf(ys[k], 0.f, chk_y);
            float dpsi;
            const float v = pix_term(a.mode, a.eps, xs[k] - ys[k], dpsi);
            acc = fmaf(ws[k], v, acc);
            g[k] = s * ws[k] * dpsi;
        }
        if (a.grad) {
            float4 *gp = reinterpret_cast<float4 *>(a.grad + e);
            float4 o = make_float4(g[0], g[1], g[2], g[3]);
            if (a.accumulate) {
                const float4 p = *gp;
                o.x += p.x; o.y += p.y; o.z += p.z; o.w += p.w;
            }
            *gp = o;
        }
        if (++it == 64) {
            dacc += (double)acc;
            acc = 0.f;
            it = 0;
        }
    }
    dacc += (double)acc;
    const double bs = block_sum<double>(dacc, red);
    int flag = 0;
    if (!(chk_x == 0.f)) flag |= 1;
    if (!(chk_y == 0.f)) flag |= 2;
    const int f = block_or(flag, reinterpret_cast<int *>(red));
    if (threadIdx.x == 0) {
        a.part[blockIdx.x] = bs;
        a.pflag[blockIdx.x] = f;
    }
    if (!last_block_done(a.ticket, gridDim.x)) return;
    double s0 = 0.0;
    int fl = 0;
    for (int b = threadIdx.x; b < (int)gridDim.x; b += blockDim.x) {
        s0 += __ldcg(a.part + b);
        fl |= __ldcg(a.pflag + b);
    }
    const double t = block_sum<double>(s0, red);
    const int ff = block_or(fl, reinterpret_cast<int *>(red));
    if (threadIdx.x == 0) {
        *a.out_sum = t;
        *a.out_flags = ff;
    }
}

}  // namespace llie

extern "C" int llie_pixel_loss_grad(int32_t mode, const float *x, const float *y, const float *weight,
                                    int32_t weight_per_channel, const float *sample_coef, int32_t n, int32_t c,
                                    int64_t plane, float eps, float grad_scale, int32_t accumulate, double *out_sum,
                                    float *out_grad, int32_t *out_flags, void *workspace, void *stream) {
    LLIE_REQUIRE(x && y && out_sum && out_flags && workspace && n > 0 && c > 0 && plane > 0, "pixel_loss: bad arguments");
    LLIE_REQUIRE(mode >= LLIE_PIX_L1 && mode <= LLIE_PIX_CHARBONNIER, "pixel_loss: unknown mode %d", mode);
    LLIE_REQUIRE(plane % 4 == 0 && llie::aligned16(x) && llie::aligned16(y) && (!weight || llie::aligned16(weight)) &&
                     (!out_grad || llie::aligned16(out_grad)),
                 "pixel_loss: H*W must be a multiple of 4 and the tensors 16-byte aligned");
    llie::PixArgs a{};
    a.x = x; a.y = y; a.w = weight; a.coef = sample_coef; a.grad = out_grad;
    a.out_sum = out_sum; a.out_flags = out_flags;
    a.plane = plane; a.chw = (long long)c * plane; a.n_elem = (long long)n * a.chw;
    a.mode = mode; a.w_per_channel = weight_per_channel; a.accumulate = accumulate;
    a.eps = eps; a.scale = grad_scale;
    long long blocks = (a.n_elem / 4 + 255) / 256;
    static const int occ = llie::resident_blocks(llie::pixel_loss_kernel, 256);
    const long long cap = llie::kNumSM * occ * LLIE_PIX_WAVES;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    llie::Workspace ws(workspace);
    a.part = reinterpret_cast<double *>(ws.scratch);
    a.pflag = reinterpret_cast<int *>(ws.scratch + (size_t)blocks * sizeof(double));
    a.ticket = ws.tickets;
    llie::pixel_loss_kernel<<<(int)blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(a);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}


// ---------------------------------------------------------------------------------------------------
// g *= *scale, skipped entirely when *scale == 1 (the upstream gradient of `loss.backward()`): the value + gradient
// kernels of the losses produce d loss / d x in their forward pass, and autograd's backward only has to apply the
// upstream scalar -- as a torch multiply that is a read + write of the whole gradient (34 us for a 50 MB tensor) for
// a factor that is almost always exactly one.
namespace llie {
__global__ void __launch_bounds__(256) scale_inplace_kernel(float *__restrict__ g, long long n,
                                                            const float *__restrict__ scale, bool vec) {
    const float s = __ldg(scale);
    if (s == 1.f) return;
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long t0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    long long done = 0;
    if (vec) {
        const long long n4 = n >> 2;
        float4 *g4 = reinterpret_cast<float4 *>(g);
        for (long long i = t0; i < n4; i += stride) {
            float4 v = g4[i];
            v.x *= s; v.y *= s; v.z *= s; v.w *= s;
            g4[i] = v;
        }
        done = n4 << 2;
    }
    for (long long i = done + t0; i < n; i += stride) g[i] *= s;
}
}  // namespace llie

namespace llie {
struct HybridCombineArgs {
    const double *l1_sum, *de_sums, *ssim_sums;
    const float *phys_value;
    double l1_count, de_count, ssim_count;
    int n_de, n_ssim;
    float w[4];
    float *out_terms, *out_total;
};

// one warp: term values from the raw device sums, weighted total in the order L1, dE, SSIM, Phys
__global__ void __launch_bounds__(32) hybrid_combine_kernel(const HybridCombineArgs a) {
    const int lane = threadIdx.x;
    double de = 0.0, ss = 0.0;
    if (a.de_sums)
        for (int i = lane; i < a.n_de; i += 32) de += a.de_sums[i];
    if (a.ssim_sums)
        for (int i = lane; i < a.n_ssim; i += 32) ss += a.ssim_sums[i];
    de = warp_sum(de);
    ss = warp_sum(ss);
    if (lane != 0) return;
    const float nanv = __int_as_float(0x7fc00000);
    float t[4];
    t[0] = a.l1_sum ? (float)(a.l1_sum[0] / a.l1_count) : nanv;
    t[1] = a.de_sums ? (float)(de / a.de_count) : nanv;
    t[2] = a.ssim_sums ? (float)(ss / a.ssim_count) : nanv;
    t[3] = a.phys_value ? a.phys_value[0] : nanv;
    const bool have[4] = {a.l1_sum != nullptr, a.de_sums != nullptr, a.ssim_sums != nullptr, a.phys_value != nullptr};
    float total = 0.f;
    bool first = true;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        a.out_terms[i] = t[i];
        if (!have[i]) continue;
        const float p = __fmul_rn(t[i], a.w[i]);
        total = first ? p : __fadd_rn(total, p);
        first = false;
    }
    a.out_total[0] = total;
}
}  // namespace llie

extern "C" int llie_hybrid_combine(const double *l1_sum, double l1_count, const double *de_sums, int32_t n_de,
                                   double de_count, const double *ssim_sums, int32_t n_ssim, double ssim_count,
                                   const float *phys_value, const float *weights, float *out_terms, float *out_total,
                                   void *stream) {
    LLIE_REQUIRE(weights && out_terms && out_total, "hybrid_combine: null pointer");
    LLIE_REQUIRE((!l1_sum || l1_count > 0) && (!de_sums || (n_de > 0 && de_count > 0)) &&
                     (!ssim_sums || (n_ssim > 0 && ssim_count > 0)),
                 "hybrid_combine: a present term needs a positive element count");
    llie::HybridCombineArgs a{};
    a.l1_sum = l1_sum; a.de_sums = de_sums; a.ssim_sums = ssim_sums; a.phys_value = phys_value;
    a.l1_count = l1_count; a.de_count = de_count; a.ssim_count = ssim_count;
    a.n_de = n_de; a.n_ssim = n_ssim;
    for (int i = 0; i < 4; ++i) a.w[i] = weights[i];
    a.out_terms = out_terms; a.out_total = out_total;
    llie::hybrid_combine_kernel<<<1, 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(a);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

extern "C" int llie_scale_inplace(float *g, int64_t n, const float *scale, void *stream) {
    LLIE_REQUIRE(g != nullptr && scale != nullptr && n >= 0, "scale_inplace: null pointer");
    if (n == 0) return LLIE_OK;
    long long blocks = ((n + 3) / 4 + 255) / 256;
    if (blocks > llie::kNumSM * 8) blocks = llie::kNumSM * 8;
    llie::scale_inplace_kernel<<<(int)blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(g, n, scale,
                                                                                               llie::aligned16(g));
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}
