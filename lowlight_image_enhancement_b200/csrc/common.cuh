// Shared device/host helpers for the sm_100a loss/metric kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/llie_b200.h"

namespace llie {

constexpr int kNumSM = 148;  // B200: 2 dies x 74 SMs (grids are sized in multiples of this)

// Blocks of `kernel` that fit on one SM at `threads` threads (registers and shared memory included): grid-stride
// kernels size their grid to ONE wave of kNumSM * resident_blocks(...) so that no partly filled second wave trails
// (8 blocks of 256 threads only fit below 33 registers).  Call sites cache the value in a function-local static.
template <typename K>
inline int resident_blocks(K kernel, int threads, size_t smem = 0) {
    int o = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, kernel, threads, smem) != cudaSuccess || o < 1) o = 1;
    return o;
}

// ---------------------------------------------------------------- error plumbing
void set_error(const char *fmt, ...);
int cuda_fail(cudaError_t e, const char *what);

#define LLIE_CUDA_OK(expr)                                  \
    do {                                                    \
        cudaError_t _e = (expr);                            \
        if (_e != cudaSuccess) return ::llie::cuda_fail(_e, #expr); \
    } while (0)

#define LLIE_REQUIRE(cond, ...)              \
    do {                                     \
        if (!(cond)) {                       \
            ::llie::set_error(__VA_ARGS__);  \
            return LLIE_EINVAL;              \
        }                                    \
    } while (0)

// ---------------------------------------------------------------- workspace layout
// [0]      unsigned ticket A (self-resetting "last block" election)
// [1..15]  spare tickets / counters (all zero between calls)
// [256 B]  start of scratch partials (no need to be zero between calls)
constexpr size_t kWsHeaderBytes = 256;
constexpr size_t kWsBytes = 4u << 20;  // 4 MiB: header + per-block partial records + histograms

struct Workspace {
    unsigned *tickets;
    unsigned char *scratch;
    size_t scratch_bytes;
    __host__ explicit Workspace(void *p)
        : tickets(reinterpret_cast<unsigned *>(p)),
          scratch(reinterpret_cast<unsigned char *>(p) + kWsHeaderBytes),
          scratch_bytes(kWsBytes - kWsHeaderBytes) {}
};

// ---------------------------------------------------------------- index helpers
// Source index for a padded coordinate; -1 means "reads zero".
__host__ __device__ __forceinline__ int pad_index(int i, int n, int mode) {
    if (i >= 0 && i < n) return i;
    switch (mode) {
        case LLIE_PAD_REFLECT:
            i = (i < 0) ? -i : 2 * (n - 1) - i;
            return (i >= 0 && i < n) ? i : -1;
        case LLIE_PAD_REPLICATE:
            return i < 0 ? 0 : n - 1;
        case LLIE_PAD_CIRCULAR:
            i = (i < 0) ? i + n : i - n;
            return (i >= 0 && i < n) ? i : -1;
        default:
            return -1;
    }
}

__device__ __forceinline__ bool finite_f(float v) { return (__float_as_uint(v) & 0x7f800000u) != 0x7f800000u; }

// ---------------------------------------------------------------- reductions (deterministic)
template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Sum over the block; result valid in thread 0 (and broadcast to all if BCAST).
// `red` must hold >= 32 elements of T.  Fixed shuffle/tree order -> bitwise reproducible.
template <typename T, bool BCAST = false>
__device__ __forceinline__ T block_sum(T v, T *red) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int nw = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();  // protect `red` against a previous use
    if (lane == 0) red[wid] = v;
    __syncthreads();
    if (wid == 0) {
        T t = lane < nw ? red[lane] : T(0);
        t = warp_sum(t);
        if (lane == 0) red[0] = t;
    }
    if (BCAST) {
        __syncthreads();
        v = red[0];
    } else {
        v = (threadIdx.x == 0) ? red[0] : T(0);
    }
    return v;
}

__device__ __forceinline__ int block_or(int v, int *red) {
    v = __reduce_or_sync(0xffffffffu, (unsigned)v);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int nw = (blockDim.x + 31) >> 5;
    __syncthreads();
    if (lane == 0) red[wid] = v;
    __syncthreads();
    int t = 0;
    if (wid == 0) {
        t = lane < nw ? red[lane] : 0;
        t = __reduce_or_sync(0xffffffffu, (unsigned)t);
    }
    return t;  // valid in warp 0
}

// Elect the last block of the grid to finish (threadFenceReduction pattern).  Every block must
// have made its global writes before calling.  Returns true in ALL threads of the last block.
__device__ __forceinline__ bool last_block_done(unsigned *ticket, unsigned total_blocks) {
    __shared__ bool s_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned t = atomicAdd(ticket, 1u);
        s_last = (t == total_blocks - 1);
        if (s_last) *ticket = 0u;  // self-reset: workspace is zero again when the kernel ends
    }
    __syncthreads();
    if (s_last) __threadfence();
    return s_last;
}

// streaming 128-bit loads (read-once data: bypass L1 allocation)
__device__ __forceinline__ float4 ldg_stream4(const float *p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ float ldg_stream(const float *p) {
    float r;
    asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(r) : "l"(p));
    return r;
}

inline bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

inline int div_up(int a, int b) { return (a + b - 1) / b; }

}  // namespace llie
