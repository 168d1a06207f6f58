// Argument blocks shared by the SSIM kernels of ssim.cu (tile / cp.async streaming kernels) and ssim_ring.cu (TMA row
// ring + packed-FP32 kernels).
#pragma once
#include "common.cuh"

namespace llie {

struct SsimArgs {
    llie_ssim_desc d;
    const float *x, *y, *win;
    double *out_sum;
    float *out_map;
    float *coef;       // NULL or [3][planes][H][W]: dLoss/dS * (A1, A2, A3) for ssim_loss_bwd_kernel
    float coef_scale;  // 1 / numel for reduction 'mean'
    int *out_flags;
    double *part;  // [planes][tiles_per_plane]
    int *pflag;    // [gridDim.x]
    unsigned *ticket;
    int tiles_x, tiles_y, co;
    int seg_rows;  // streaming kernel: output rows per CTA
    // second policy evaluated from the same five moments (llie_ssim_stats_dual): the interior-crop policy whose taps sum
    // to sqrt(r) times this window's, r = dual_r; moments scale by r, central moments by r after a -eps * mean^2
    // correction (eps = r - 1)
    double *out_sum2, *part2;
    float dual_c1, dual_c2, dual_r, dual_2r, dual_r2, dual_2r2, dual_neg_eps;
};

struct SsimBwdArgs {
    int n, c, h, w, ksize, clamp_in, gout_is_map, accumulate;
    const float *x, *y, *win, *coef;
    const float *gout;  // device scalar (upstream gradient of the reduced loss) or a [n,c,h,w] map
    float *out_grad;
    int tiles_x, tiles_y;
};

// launchers of ssim_ring.cu; `eligible` tells ssim.cu whether the ring kernels take this problem
bool ssim_ring_eligible(const llie_ssim_desc &d, const float *x, const float *y, const float *win_host, bool want_map,
                        const float *coef);
int ssim_ring_forward(SsimArgs &a, const float *win_host, void *workspace, cudaStream_t st);
bool ssim_ring_dual_eligible(const llie_ssim_desc &d, const float *x, const float *y, const float *win_host);
bool ssim_ring_bwd_eligible(const SsimBwdArgs &a, const float *win_host);
int ssim_ring_backward(SsimBwdArgs &a, const float *win_host, cudaStream_t st);
int pick_seg_rows(long long cols, int h, int slots, int min_rows, int halo);

}  // namespace llie
