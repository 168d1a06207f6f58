// Microbenchmark: scalar FFMA vs packed FFMA2 (fma.rn.f32x2, sm_100+) issue throughput on one B200.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ffma2_bench tools/micro/ffma2_bench.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned long long ffma2(unsigned long long a, unsigned long long b, unsigned long long c) {
    unsigned long long d;
    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ float ffma1(float a, float b, float c) {
    float d;
    asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}

template <int ILP>
__global__ void k_scalar(float *out, float a, float b, int iters) {
    float acc[ILP];
#pragma unroll
    for (int i = 0; i < ILP; ++i) acc[i] = threadIdx.x * 1e-3f + i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) acc[i] = ffma1(acc[i], a, b);
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int ILP>
__global__ void k_packed(float *out, float a, float b, int iters) {
    unsigned long long acc[ILP];
    float2 av = make_float2(a, a), bv = make_float2(b, b);
    const unsigned long long A = *reinterpret_cast<unsigned long long *>(&av), B = *reinterpret_cast<unsigned long long *>(&bv);
#pragma unroll
    for (int i = 0; i < ILP; ++i) {
        float2 v = make_float2(threadIdx.x * 1e-3f + i, threadIdx.x * 2e-3f + i);
        acc[i] = *reinterpret_cast<unsigned long long *>(&v);
    }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) acc[i] = ffma2(acc[i], A, B);
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < ILP; ++i) {
        float2 v = *reinterpret_cast<float2 *>(&acc[i]);
        s += v.x + v.y;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F>
static float time_ms(F launch) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    launch();
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    launch();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    return ms;
}

int main() {
    const int blocks = 148 * 4, threads = 512, iters = 1 << 15;
    float *out;
    cudaMalloc(&out, sizeof(float) * blocks * threads);
    const double n_thr = (double)blocks * threads;
    {
        float ms = time_ms([&] { k_scalar<8><<<blocks, threads>>>(out, 0.999f, 1e-3f, iters); });
        printf("scalar FFMA  ILP8: %.3f ms  %.2f TFLOP/s  %.1f G warp-instr/s\n", ms, n_thr * iters * 8 * 2 / ms * 1e-9,
               n_thr / 32 * iters * 8 / ms * 1e-6);
    }
    {
        float ms = time_ms([&] { k_packed<8><<<blocks, threads>>>(out, 0.999f, 1e-3f, iters); });
        printf("packed FFMA2 ILP8: %.3f ms  %.2f TFLOP/s  %.1f G warp-instr/s\n", ms, n_thr * iters * 8 * 4 / ms * 1e-9,
               n_thr / 32 * iters * 8 / ms * 1e-6);
    }
    {
        float ms = time_ms([&] { k_packed<4><<<blocks, threads>>>(out, 0.999f, 1e-3f, iters); });
        printf("packed FFMA2 ILP4: %.3f ms  %.2f TFLOP/s\n", ms, n_thr * iters * 4 * 4 / ms * 1e-9);
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
