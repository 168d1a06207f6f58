// Shared declarations of the physics-consistency kernels (generic path in phys.cu, TMA row-ring path in
// phys_ring.cu).
#pragma once
#include "common.cuh"

namespace llie {

struct PhysArgs {
    llie_phys_desc d;
    const float *pred, *obs, *psf, *rho;
    float *out_mean;
    double *out_sum;
    float *out_map;
    int *out_flags;
    double *part_sum;  // [n][gridDim.x]
    int *part_flag;    // [n][gridDim.x]
    unsigned *ticket;
    int y_lo, x_lo, ho, wo;  // compared (output) domain
    int tiles_x, tiles_y;
    double inv_count;        // 1 / (c_out*ho*wo)
    // backward only
    const float *gout;
    long long gout_stride;
    float gscale;
    float *grad;
    float *scratch;
};

// residual and loss for one output element; returns loss value, sets r
__device__ __forceinline__ float residual(const llie_phys_desc &d, float conv, float a, float rho, float &r) {
    float y = d.rho_side == 0 ? conv * rho : conv;
    float t = d.rho_side == 1 ? a * rho : a;
    if (d.clip_y) y = fminf(fmaxf(y, 0.f), 1.f);
    if (d.clip_a) t = fminf(fmaxf(t, 0.f), 1.f);
    r = y - t;
    return d.robust ? sqrtf(r * r + d.eps * d.eps) : fabsf(r);
}

// d loss / d conv for one output element (without the upstream weight)
__device__ __forceinline__ float dresidual(const llie_phys_desc &d, float conv, float a, float rho) {
    float y = d.rho_side == 0 ? conv * rho : conv;
    float t = d.rho_side == 1 ? a * rho : a;
    float pass = 1.f;
    if (d.clip_y) {
        // torch.clamp backward passes the gradient where lo <= y <= hi
        pass = (y >= 0.f && y <= 1.f) ? 1.f : 0.f;
        y = fminf(fmaxf(y, 0.f), 1.f);
    }
    if (d.clip_a) t = fminf(fmaxf(t, 0.f), 1.f);
    float r = y - t;
    float psi;
    if (d.robust)
        psi = r / sqrtf(r * r + d.eps * d.eps);
    else
        psi = (r > 0.f) ? 1.f : ((r < 0.f) ? -1.f : 0.f);
    return psi * pass * (d.rho_side == 0 ? rho : 1.f);
}

__device__ __forceinline__ float rho_at(const PhysArgs &a, int n, int c, int y, int x) {
    if (a.rho == nullptr) return a.d.rho_scalar;
    return __ldg(a.rho + (long long)n * a.d.rho_sn + (long long)c * a.d.rho_sc + (long long)y * a.d.rho_sh +
                 (long long)x * a.d.rho_sw);
}


// 1-D set of padded coordinates that read image index i (the adjoint of the padding operator)
__device__ __forceinline__ int preimages(int i, int n, int p, int mode, int out[10]) {
    int k = 0;
    out[k++] = i;
    if (mode == LLIE_PAD_REFLECT) {
        if (i >= 1 && i <= p) out[k++] = -i;
        const int m = n - 1 - i;
        if (m >= 1 && m <= p) out[k++] = n - 1 + m;
    } else if (mode == LLIE_PAD_REPLICATE) {
        if (i == 0)
            for (int m = 1; m <= p; ++m) out[k++] = -m;
        if (i == n - 1)
            for (int m = 1; m <= p; ++m) out[k++] = n - 1 + m;
    }
    return k;
}

// TMA row-ring path (phys_ring.cu)
bool phys_ring_eligible(const PhysArgs &a);
int launch_phys_fwd_ring(PhysArgs &a, void *workspace, cudaStream_t st);
int launch_phys_bwd_ring(PhysArgs &a, cudaStream_t st);
int launch_phys_fused_ring(PhysArgs &a, void *workspace, cudaStream_t st);
void phys_ring_set_rows(int r);
void phys_ring_set_pdl(int on);
void phys_ring_set_prefetch(int on);

}  // namespace llie
