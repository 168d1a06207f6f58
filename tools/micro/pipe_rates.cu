// Microbenchmark: issue / pipe throughput of the instruction classes the SSIM and dE00 kernels are made of, on one B200.
// Output: warp-instructions per SMSP per clock (assuming the SM clock printed), for dependent chains at ILP 8.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build_micro/pipe_rates tools/micro/pipe_rates.cu
#include <cstdio>
#include <cuda_runtime.h>

typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float lo, float hi) {
    u64 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ float lo_of(u64 v) {
    float lo, hi;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
    return lo + hi;
}

struct OpFfma { static __device__ __forceinline__ float f(float a, float b, float c) { float d; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; } };
struct OpFmnmx { static __device__ __forceinline__ float f(float a, float b, float c) { float d; asm volatile("max.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b)); return d; } };
struct OpRcp { static __device__ __forceinline__ float f(float a, float b, float c) { float d; asm volatile("rcp.approx.ftz.f32 %0, %1;" : "=f"(d) : "f"(a)); return d; } };
struct OpRsq { static __device__ __forceinline__ float f(float a, float b, float c) { float d; asm volatile("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(d) : "f"(a)); return d; } };
struct OpSqrt { static __device__ __forceinline__ float f(float a, float b, float c) { float d; asm volatile("sqrt.approx.ftz.f32 %0, %1;" : "=f"(d) : "f"(a)); return d; } };
struct OpEx2 { static __device__ __forceinline__ float f(float a, float b, float c) { float d; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(d) : "f"(a)); return d; } };
struct OpLg2 { static __device__ __forceinline__ float f(float a, float b, float c) { float d; asm volatile("lg2.approx.ftz.f32 %0, %1;" : "=f"(d) : "f"(a)); return d; } };
struct OpSin { static __device__ __forceinline__ float f(float a, float b, float c) { float d; asm volatile("sin.approx.ftz.f32 %0, %1;" : "=f"(d) : "f"(a)); return d; } };
struct OpIadd { static __device__ __forceinline__ float f(float a, float b, float c) { int d; asm volatile("add.s32 %0, %1, %2;" : "=r"(d) : "r"(__float_as_int(a)), "r"(__float_as_int(b))); return __int_as_float(d); } };
struct OpSel { static __device__ __forceinline__ float f(float a, float b, float c) { float d; asm volatile("{ .reg .pred p; setp.gt.f32 p, %1, %2; selp.f32 %0, %1, %3, p; }" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; } };

struct Op2Ffma { static __device__ __forceinline__ u64 f(u64 a, u64 b, u64 c) { u64 d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; } };
struct Op2Fmul { static __device__ __forceinline__ u64 f(u64 a, u64 b, u64 c) { u64 d; asm volatile("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; } };
struct Op2Fadd { static __device__ __forceinline__ u64 f(u64 a, u64 b, u64 c) { u64 d; asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; } };

template <class OP, int ILP>
__global__ void k1(float *out, float a, float b, int iters) {
    float acc[ILP];
#pragma unroll
    for (int i = 0; i < ILP; ++i) acc[i] = 1.0f + threadIdx.x * 1e-3f + i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) acc[i] = OP::f(acc[i], a, b);
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <class OP, int ILP>
__global__ void k2(float *out, float a, float b, int iters) {
    u64 acc[ILP];
    const u64 A = pk(a, a), B = pk(b, b * 1.0001f);
#pragma unroll
    for (int i = 0; i < ILP; ++i) acc[i] = pk(1.0f + threadIdx.x * 1e-3f + i, 1.0f + threadIdx.x * 2e-3f + i);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) acc[i] = OP::f(acc[i], A, B);
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += lo_of(acc[i]);
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// 8 FFMA2 + NM MUFU (rsqrt) per iteration, independent chains: does the MUFU pipe overlap the packed FMA pipe?
template <int NM>
__global__ void kmix(float *out, float a, float b, int iters) {
    u64 acc[8];
    float m[NM > 0 ? NM : 1];
    const u64 A = pk(a, a), B = pk(b, b * 1.0001f);
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = pk(1.0f + threadIdx.x * 1e-3f + i, 1.0f + threadIdx.x * 2e-3f + i);
#pragma unroll
    for (int i = 0; i < NM; ++i) m[i] = 1.0f + threadIdx.x * 1e-3f + i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            acc[i] = Op2Ffma::f(acc[i], A, B);
            if (i < NM) m[i] = OpRsq::f(m[i], a, b);
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += lo_of(acc[i]);
#pragma unroll
    for (int i = 0; i < NM; ++i) s += m[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// 8 scalar FFMA + 8 FMNMX (alu pipe) per iteration: do the two pipes dual-issue from one scheduler slot stream?
__global__ void kfma_alu(float *out, float a, float b, int iters) {
    float acc[8], m[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { acc[i] = 1.0f + threadIdx.x * 1e-3f + i; m[i] = acc[i] * 0.5f; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            acc[i] = OpFfma::f(acc[i], a, b);
            m[i] = OpFmnmx::f(m[i], a, b);
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += acc[i] + m[i];
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
    const int blocks = 148 * 4, threads = 512, iters = 1 << 14;
    float *out;
    cudaMalloc(&out, sizeof(float) * blocks * threads);
    int clk_khz = 0;
    cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    const double smsp_clk = 148.0 * 4 * clk_khz * 1e3;  // SMSP-cycles per second at the nominal max clock
    const double warps = (double)blocks * threads / 32;
    printf("clock (attr) %.0f MHz; rates are warp-instructions per SMSP per clock at that clock\n", clk_khz * 1e-3);
#define RUN(name, kern, per_iter)                                                                     \
    {                                                                                                 \
        float ms = time_ms([&] { kern<<<blocks, threads>>>(out, 0.999f, 1e-3f, iters); });            \
        printf("%-28s %8.3f ms  %6.3f instr/SMSP/clk\n", name, ms, warps * iters * (per_iter) / (ms * 1e-3) / smsp_clk); \
    }
    RUN("FFMA  ILP8", (k1<OpFfma, 8>), 8)
    RUN("FFMA2 ILP8", (k2<Op2Ffma, 8>), 8)
    RUN("FFMA2 ILP4", (k2<Op2Ffma, 4>), 4)
    RUN("FMUL2 ILP8", (k2<Op2Fmul, 8>), 8)
    RUN("FADD2 ILP8", (k2<Op2Fadd, 8>), 8)
    RUN("FMNMX ILP8", (k1<OpFmnmx, 8>), 8)
    RUN("IADD  ILP8", (k1<OpIadd, 8>), 8)
    RUN("SETP+SEL ILP8 (2 instr)", (k1<OpSel, 8>), 16)
    RUN("MUFU.RCP ILP8", (k1<OpRcp, 8>), 8)
    RUN("MUFU.RSQ ILP8", (k1<OpRsq, 8>), 8)
    RUN("sqrt.approx ILP8", (k1<OpSqrt, 8>), 8)
    RUN("MUFU.EX2 ILP8", (k1<OpEx2, 8>), 8)
    RUN("MUFU.LG2 ILP8", (k1<OpLg2, 8>), 8)
    RUN("MUFU.SIN ILP8", (k1<OpSin, 8>), 8)
    RUN("8 FFMA2 + 0 MUFU", kmix<0>, 8)
    RUN("8 FFMA2 + 1 MUFU", kmix<1>, 9)
    RUN("8 FFMA2 + 2 MUFU", kmix<2>, 10)
    RUN("8 FFMA2 + 4 MUFU", kmix<4>, 12)
    RUN("8 FFMA + 8 FMNMX", kfma_alu, 16)
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
