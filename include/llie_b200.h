/*
 * llie_b200.h — C ABI of the B200-native loss/metric hot path
 * (drop-in compute boundary for RUA1027/Lowlight_Image_Enhancement).
 *
 * The reference has no FFI of its own: its boundary is Python call signatures
 * (SURVEY.md §8b).  Every entry point below names the reference function whose
 * arithmetic it replaces (paths relative to the reference repo root).  The
 * Python mirror in lowlight_image_enhancement_b200/ binds these with ctypes;
 * INTEGRATION.md shows the stub a reference maintainer would add.
 *
 * Conventions
 *   - all data pointers are DEVICE pointers to contiguous NCHW buffers on the
 *     current device; `stream` is a cudaStream_t passed as void*;
 *   - every call is asynchronous on `stream`, never synchronises, never
 *     allocates, and is CUDA-graph capturable;
 *   - the return value is 0 (LLIE_OK) or a negative LLIE_E* code; nothing
 *     throws across the ABI; llie_last_error() returns a static message;
 *   - `workspace` is a caller-owned device buffer of llie_workspace_bytes()
 *     bytes that must be ZERO before its first use and must not be shared by
 *     calls running concurrently on different streams; every call leaves it
 *     zeroed again (self-resetting tickets), so one cudaMemset at allocation
 *     time is enough;
 *   - reductions are deterministic (fixed-order two-stage, no float atomics).
 */
#ifndef LLIE_B200_H
#define LLIE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LLIE_OK 0
#define LLIE_EINVAL (-1)   /* bad argument / unsupported combination */
#define LLIE_ECUDA (-2)    /* a CUDA runtime call failed (see llie_last_error) */
#define LLIE_ENOSM100 (-3) /* device is not sm_100 */

/* padding modes (phys: zeros/reflect/replicate; ssim adds circular) */
#define LLIE_PAD_ZEROS 0
#define LLIE_PAD_REFLECT 1
#define LLIE_PAD_REPLICATE 2
#define LLIE_PAD_CIRCULAR 3

/* PSF layouts */
#define LLIE_PSF_DEPTHWISE 0 /* [C,1,kh,kw], output channel c reads input channel c */
#define LLIE_PSF_DENSE 1     /* [C_out,C_in,kh,kw] cross-channel */

int llie_abi_version(void);
const char *llie_last_error(void);
size_t llie_workspace_bytes(void);
/* 0 if the current device is compute capability 10.x, else LLIE_ENOSM100 */
int llie_check_device(void);

/* Diagnostic knobs for A/B measurements.  Process-wide (relaxed atomics: setting one while another thread launches is
 * not a data race, but it does change that thread's kernel choice); the defaults are the product and no product code
 * path sets them.  Anything a caller may legitimately vary per call is a descriptor field, not a knob
 * (llie_phys_desc.stream_overlap). */
#define LLIE_TUNE_PHYS_PATH 1 /* 0 TMA row ring (default) | 2 generic kernels only */
#define LLIE_TUNE_PHYS_ROWS 2 /* rows per item of the row-ring kernels; 0 = heuristic */
#define LLIE_TUNE_SSIM_PATH 3 /* 0 auto (TMA row ring + packed FP32 kernel for statistics-only / loss-forward 11-tap
                                 calls, else the one-column streaming kernel) | 1 shared-memory tile kernel |
                                 2 one-column streaming | 3 two-column cp.async streaming kernel (round-1 default) */
#define LLIE_TUNE_SSIM_SEG_ROWS 4 /* output rows per CTA of the streaming SSIM kernel; 0 = heuristic */
#define LLIE_TUNE_PHYS_PDL 5 /* 1 (default): ring kernels are launched with programmatic stream serialization, so the
                                prologue of a kernel overlaps the tail of the previous one | 0 plain launches */
/* key 6 (process-wide stream overlap, ABI 2) is gone: llie_phys_desc.stream_overlap */
#define LLIE_TUNE_PHYS_PREFETCH 7 /* rows per tensor of its first item that a ring CTA prefetches into L2 before it waits
                                     for the previous grid (no architectural effect, legal under serialised semantics);
                                     default 8, 0 = off */
int llie_tune(int key, int value);

/* ------------------------------------------------------------------------
 * Physics-consistency projector  r = clipY(sY * (K (*) padY(pred))) - clipA(sA * obs)
 *
 * replaces  metrics/phys_consistency.py:193-255  _phys_cons_core          (rho_side=0)
 *           NewBP_model/losses.py:173-192        PhysicsConsistencyLoss   (rho_side=1, replicate, same)
 *           NewBP_model/losses.py:217-220        PhysicalConsistencyLossSRGB (rho_side=1, zeros, same)
 * ------------------------------------------------------------------------ */
typedef struct llie_phys_desc {
    int32_t n, c_in, c_out, h, w;
    int32_t kh, kw;         /* odd */
    int32_t psf_layout;     /* LLIE_PSF_* (depthwise requires c_in == c_out) */
    int32_t pad_mode;       /* LLIE_PAD_ZEROS / REFLECT / REPLICATE */
    int32_t crop_valid;     /* 1: compare only [kh/2, H-kh/2) x [kw/2, W-kw/2) */
    int32_t robust;         /* 0: |r|   1: sqrt(r^2 + eps^2) */
    int32_t rho_side;       /* 0: rho multiplies K*pred (metric)  1: rho multiplies obs (loss) */
    int32_t clip_y;         /* clamp sY*(K*pred) to [0,1] */
    int32_t clip_a;         /* clamp sA*obs to [0,1] */
    int32_t stream_overlap; /* 0 (default) | 1: PER-CALL stream-overlap contract for loops of phys calls over independent
                               batches (dataset evaluation, benchmarks).  The caller guarantees that the kernel launched
                               right before this call on the same stream writes none of the call's buffers pred, obs,
                               psf, rho, gout, grad, map and reads none of grad / map (out_mean / out_sum / out_flags
                               are written after the wait and may be shared).  The row-ring kernels then wait for the
                               previous grid only before they touch the shared workspace, so this launch streams and
                               computes on the SMs the early finishers of the previous one have left.  Never set it
                               under autograd (gout comes from the previous kernel).  Ignored by the generic kernels. */
    float eps;
    float rho_scalar;       /* used when rho == NULL */
    /* element strides of the rho tensor for (n, c, y, x); 0 = broadcast */
    int64_t rho_sn, rho_sc, rho_sh, rho_sw;
} llie_phys_desc;

/* Forward.
 *   psf          prepared (normalised) weights, layout per desc
 *   rho          NULL (use rho_scalar) or device pointer with the strides in desc
 *   out_mean     [n + 2] floats: per-sample mean of the loss map, then mean over
 *                samples, then sum over samples (phys_consistency.py:250-251)
 *   out_sum      NULL or [n + 1] doubles: raw per-sample sums of the loss map, then the total
 *   out_map      NULL or [n, c_out, h', w'] |r| map (phys_consistency.py:244, return_map)
 *   out_flags    [1] int: bit0 non-finite value seen in pred, bit1 in obs, bit2 a per-sample sum
 *                is non-finite but the source tensor is unknown (un-clipped fast path: the caller
 *                resolves it on the error path); replaces the two isfinite().all() passes,
 *                phys_consistency.py:57-60
 */
int llie_phys_forward(const llie_phys_desc *desc, const float *pred, const float *obs,
                      const float *psf, const float *rho, float *out_mean, double *out_sum,
                      float *out_map, int32_t *out_flags, void *workspace, void *stream);

/* Backward (adjoint): grad_pred = padY^T K^T ( sY * 1[clipY passes] * psi(r) * g_n )
 *   where psi = sign (robust=0, sign(0)=0) or r/sqrt(r^2+eps^2), and
 *   g_n = gout[gout_stride * n] * gscale  (gscale folds 1/count and the batch reduction).
 * This is the conv_transpose2d dual of NewBP_model/newbp_layer.py:16-21 fused with the
 * residual derivative (autograd of losses.py:191, :220 in the reference).
 *   scratch      NULL, or [n, c_out, h, w] floats; REQUIRED for dense PSFs or kh,kw > 9
 */
int llie_phys_backward(const llie_phys_desc *desc, const float *pred, const float *obs,
                       const float *psf, const float *rho, const float *gout, int64_t gout_stride,
                       float gscale, float *grad_pred, float *scratch, void *stream);

/* 1 if llie_phys_backward needs `scratch` for this problem (generic path), 0 for the TMA tile path */
int llie_phys_backward_needs_scratch(const llie_phys_desc *desc, const float *pred, const float *obs,
                                     const float *rho, const float *grad_pred);

/* Loss value AND gradient from ONE pass over pred / obs (training use: 12 B/element instead of 8 + 12): the outputs
 * of llie_phys_forward (out_mean [n + 2], out_sum, out_flags; no map) and of llie_phys_backward (grad_pred, for the
 * given upstream `gout` / `gscale`) with identical values.  Shapes the fused kernel does not take run the two
 * single-purpose kernels back to back (then `scratch` follows llie_phys_backward_needs_scratch). */
int llie_phys_value_and_grad(const llie_phys_desc *desc, const float *pred, const float *obs, const float *psf,
                             const float *rho, const float *gout, int64_t gout_stride, float gscale, float *out_mean,
                             double *out_sum, int32_t *out_flags, float *grad_pred, float *scratch, void *workspace,
                             void *stream);

/* ------------------------------------------------------------------------
 * NewBP operator: y = conv2d(x, K, padding, groups) and its explicit adjoint
 * gx = conv_transpose2d(gy, K, padding, groups)
 * replaces NewBP_model/newbp_layer.py:9-21 (NewBPFunction) and :117-126 (CrosstalkPSF.forward).
 * groups must be 1 (dense K [c_out,c_in,kh,kw]) or c_in == c_out (depthwise K [c,1,kh,kw]).
 * dtype: 0 = float32, 1 = float64 (the reference's gradcheck tests run in double).
 * ------------------------------------------------------------------------ */
typedef struct llie_conv_desc {
    int32_t n, c_in, c_out, h, w; /* input spatial size of the FORWARD op */
    int32_t kh, kw;
    int32_t pad_h, pad_w;         /* zero padding */
    int32_t depthwise;            /* 1: groups == c_in == c_out */
    int32_t dtype;                /* 0 f32, 1 f64 */
} llie_conv_desc;

/* y: [n, c_out, h + 2*pad_h - kh + 1, w + 2*pad_w - kw + 1] */
int llie_newbp_forward(const llie_conv_desc *desc, const void *x, const void *kernel, void *y,
                       void *stream);
/* gy has the forward OUTPUT shape, gx the forward INPUT shape */
int llie_newbp_adjoint(const llie_conv_desc *desc, const void *gy, const void *kernel, void *gx,
                       void *stream);

/* ------------------------------------------------------------------------
 * PSNR family: per-(n,c) sums of squared differences in float64, finalised on the device.
 * replaces metrics/linear.py:139-210 (psnr_linear: subtract in the input dtype, THEN cast to f64),
 *          metrics/channelwise.py:113-222 (rgb_psnr / cpsnr_rgb: cast to f64, clamp, subtract),
 *          metrics/psnr.py:39-88 (calculate_psnr: f64 global mean, n=c=1 over the whole tensor).
 * ------------------------------------------------------------------------ */
typedef struct llie_psnr_desc {
    int32_t n, c;
    int64_t plane;        /* h*w elements per (n,c) plane */
    int32_t dtype;        /* 0 f32, 1 f64 inputs */
    int32_t cast_first;   /* 0: diff in input dtype then f64 (linear.py:197); 1: f64 first (channelwise.py:146-150) */
    int32_t clamp;        /* clamp both inputs to [0, clamp_hi] before the difference */
    double clamp_hi;
    double data_range;
    double eps;
} llie_psnr_desc;

/* columns of the [n + 2, LLIE_PSNR_COLS(c)] float64 result matrix (rows n / n+1: batch mean / sum) */
#define LLIE_PSNR_COLS(c) (2 * (c) + 3)
/* [0,c)      per-channel MSE
 * [c,2c)     per-channel PSNR = 10 log10(L^2 / max(mse_c, eps))           (channelwise.py:153-155)
 * 2c         mean over channels of the per-channel PSNR                     (channelwise.py:156)
 * 2c+1       CPSNR from the mean of channel MSEs, eps floor                  (channelwise.py:211-216)
 * 2c+2       PSNR of the whole sample, +inf where mse <= eps                 (linear.py:198-208, psnr.py:84-85)
 *   out_minmax  NULL or [2] floats (dtype f32) / doubles (f64): global min / max over BOTH tensors
 *               (metrics/ssim.py:62-71 auto data range)
 *   out_flags   [1] int: bit0 non-finite in x, bit1 non-finite in y */
int llie_psnr_stats(const llie_psnr_desc *desc, const void *x, const void *y, double *out,
                    void *out_minmax, int32_t *out_flags, void *workspace, void *stream);

/* llie_psnr_stats without a caller-owned workspace and without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same kernel, same outputs as llie_psnr_stats; its scratch (ticket +
 * per-block partial records) is a file-scope `static __device__` array of LLIE_PSNR_STATIC_SCRATCH_BYTES that every
 * GPU loading the library holds until exit.  Limits: n * c <= LLIE_PSNR_STATIC_MAX_SLOTS planes (LLIE_EINVAL with a
 * message beyond that); the grid is capped so that n * c * blocks-per-plane <= LLIE_PSNR_STATIC_MAX_SLOTS.  The call
 * issues one cudaMemsetAsync (ticket header) and one kernel on `stream`; it calls no cudaMalloc / cudaFree.  All calls
 * on a device SHARE the array: calls that overlap on the GPU (other streams, concurrently replayed graphs) race --
 * serialise them on one stream or use llie_psnr_stats with one workspace per stream. */
#define LLIE_PSNR_STATIC_MAX_SLOTS 8192
#define LLIE_PSNR_STATIC_SCRATCH_BYTES (256 + LLIE_PSNR_STATIC_MAX_SLOTS * 32) /* 262 400 bytes per device */
int llie_psnr_stats_static(const llie_psnr_desc *desc, const void *x, const void *y, double *out,
                           void *out_minmax, int32_t *out_flags, void *stream);

/* ------------------------------------------------------------------------
 * Gaussian / uniform-window SSIM statistics: per-(n, c) float64 sums of the SSIM map (or of the DSSIM
 * loss map) over the full image or its interior, one fused separable-filter kernel.
 * replaces metrics/linear.py:213-324 (ssim_linear), metrics/channelwise.py:226-302 (rgb_ssim),
 *          metrics/ssim.py:246-311 (SSIMEvaluator -> torchmetrics 1.2.0 _ssim_update),
 *          NewBP_model/losses.py:146-155 (SSIMLoss -> kornia 0.6.12 ssim_loss), forward value.
 * The K x K window of all three is the outer product of `win1d` (K normalised taps, device memory).
 * `win1d_host` is a HOST copy of the same K taps (or NULL): the default 11-tap kernels (TMA row ring + packed FP32,
 * csrc/ssim_ring.cu) take the taps by value in their kernel parameters; without it the call runs on the kernels that
 * read the device copy (same results to rounding).
 * ------------------------------------------------------------------------ */
#define LLIE_SSIM_MAX_K 63 /* 1..15 taps: unrolled kernels; 17..63: the runtime-window instantiation of the tile kernels */
typedef struct llie_ssim_desc {
    int32_t n, c, h, w;   /* c = channels of the input tensors */
    int32_t ksize;        /* odd, <= LLIE_SSIM_MAX_K, <= min(h, w) */
    int32_t pad_mode;     /* LLIE_PAD_* (ZEROS == 'constant') */
    int32_t luma;         /* 1: c == 3 and SSIM runs on 0.2989 R + 0.5870 G + 0.1140 B (ssim.py:116-127) */
    int32_t clamp_in;     /* clamp both inputs to [0,1] first (losses.py:155) */
    int32_t clamp_var;    /* sigma_x^2, sigma_y^2 clamped >= 0 (linear.py:307-308) */
    int32_t interior;     /* 1: reduce over [K/2, H-K/2) x [K/2, W-K/2) only (torchmetrics crop) */
    int32_t loss_map;     /* 1: reduce clamp((1 - S) / 2, 0, 1) instead of S (kornia ssim_loss) */
    float c1, c2, eps;    /* (k1 L)^2, (k2 L)^2, denominator eps (0 for torchmetrics) */
} llie_ssim_desc;
/*   out_sum   [n * co] doubles, co = luma ? 1 : c;  out_map NULL or [n, co, h, w] SSIM map
 *   out_flags [1] int: bit0 non-finite in x, bit1 in y */
int llie_ssim_stats(const llie_ssim_desc *desc, const float *x, const float *y, const float *win1d,
                    const float *win1d_host, double *out_sum, float *out_map, int32_t *out_flags, void *workspace,
                    void *stream);

/* Two SSIM policies of one image pair from ONE pass (the evaluation driver asks for both `linear_ssim`, torchmetrics
 * policy, and `rgb_ssim`, ssim_linear policy: metrics/ssim.py:188-538 and metrics/channelwise.py:226-302).
 * `desc` / `win1d` describe the first policy exactly as for llie_ssim_stats (reflect border, same size, variance
 * clamp); the second is the interior-crop policy (no clamp, no eps) with constants c1_b / c2_b whose normalised 1-D
 * window is the first one's up to rounding: tap_sum_ratio = sum(taps_b) / sum(taps_a) in double.  The two windows
 * differ by <= 5 ulp per tap; only the difference of their SUMS is visible in the score (1e-4 on flat images, where
 * E[x^2] - mu^2 cancels), and it is applied in closed form: moments scale by r = ratio^2, central moments become
 * r (var - (r - 1) mu^2).  Measured against a float64 evaluation with the second window: <= 3e-9 relative.
 *   out_sum    [n * c] doubles, first policy over h x w;  out_sum_b [n * c] doubles over (h - 10) x (w - 10)
 * llie_ssim_stats_dual_eligible: 1 when this pair of tensors can take the call (11 taps, reflect padding, clamp_var,
 * no input clamp / luma / interior, images larger than the window), else 0: make two llie_ssim_stats calls.  W % 4 == 0
 * with 16-byte aligned planes and a host copy of the taps runs on the TMA ring kernel, everything else on the
 * two-column cp.async kernel. */
int llie_ssim_stats_dual_eligible(const llie_ssim_desc *desc, const float *x, const float *y, const float *win1d_host);
int llie_ssim_stats_dual(const llie_ssim_desc *desc, float c1_b, float c2_b, double tap_sum_ratio, const float *x,
                         const float *y, const float *win1d, const float *win1d_host, double *out_sum,
                         double *out_sum_b, int32_t *out_flags, void *workspace, void *stream);

/* Differentiable (D)SSIM, kornia policy (reflect border, same size, no variance clamp): forward as
 * llie_ssim_stats plus out_coef [3][n*c][h][w] = grad_scale * dLossMap/dS * (A1, A2, A3), the per-window
 * sensitivities of S to (mu_x, E[x^2], E[xy]); backward folds them through the adjoint of reflect-pad o
 * separable correlation:  out_grad = mask_[0,1](x) * g * (G^T c1 + 2 x G^T c2 + y G^T c3).
 * `grad_out` is a DEVICE pointer: one float (gradient of the reduced loss) or, with grad_out_is_map, a
 * [n,c,h,w] map (reduction 'none'); with `accumulate` the result is added to out_grad (fused hybrid loss).
 * replaces kornia.losses.SSIMLoss autograd (losses.py:146-155). */
int llie_ssim_loss_forward(const llie_ssim_desc *desc, const float *x, const float *y, const float *win1d,
                           const float *win1d_host, float grad_scale, double *out_sum, float *out_map, float *out_coef,
                           int32_t *out_flags, void *workspace, void *stream);
int llie_ssim_loss_backward(const llie_ssim_desc *desc, const float *x, const float *y, const float *win1d,
                            const float *win1d_host, const float *coef, const float *grad_out, int32_t grad_out_is_map, int32_t accumulate,
                            float *out_grad, void *stream);

/* ------------------------------------------------------------------------
 * sRGB -> Lab(D65/2deg) -> CIEDE2000, per-image float64 sums (+ optional map / L plane of x).
 * replaces metrics/color_error.py:68-97,105-210,235-266 (variant 0, metric) and
 *          NewBP_model/losses.py:99-143 (variant 1, loss: eps 1e-6, inputs clamped to [0,1]).
 * ------------------------------------------------------------------------ */
typedef struct llie_deltae_desc {
    int32_t n, h, w;      /* inputs are [n, 3, h, w] */
    int32_t variant;      /* 0 metric (color_error.py) | 1 loss (losses.py) */
    int32_t input_lab;    /* 1: inputs already hold (L, a, b) planes (stage-level parity tests) */
    int32_t clamp_in;     /* clamp sRGB inputs to [0,1] first (losses.py:140-141) */
    float kL, kC, kH, eps;
} llie_deltae_desc;
/*   out_sum [n] doubles; out_map NULL or [n, h, w]; out_l NULL or [n, h, w] (L of x, for the edge mask) */
int llie_deltae(const llie_deltae_desc *desc, const float *x, const float *y, double *out_sum, float *out_map,
                float *out_l, int32_t *out_flags, void *workspace, void *stream);

/* llie_deltae plus out_sqerr[n] = sum over the image of (x - y)^2 in float64 on the RAW (unclamped) inputs, so
 * that the whole-image PSNR (metrics/psnr.py:39-88) rides on the colour pass instead of re-reading both images. */
int llie_deltae_sqerr(const llie_deltae_desc *desc, const float *x, const float *y, double *out_sum,
                      double *out_sqerr, float *out_map, float *out_l, int32_t *out_flags, void *workspace,
                      void *stream);
/* ... plus out_sqerr_rgb[n, 3] = the same float64 sums per channel (out_sqerr may be NULL): the CPSNR / RGB-PSNR family
 * (metrics/channelwise.py:113-222: float64 difference of the raw inputs, per-channel MSE) rides on the colour pass too,
 * and the dataset driver needs no llie_psnr_stats launch (one full read of both images less per pair). */
int llie_deltae_sqerr_rgb(const llie_deltae_desc *desc, const float *x, const float *y, double *out_sum,
                          double *out_sqerr, double *out_sqerr_rgb, float *out_map, float *out_l,
                          int32_t *out_flags, void *workspace, void *stream);

/* DeltaE00Loss forward + analytic backward in ONE pass (variant 1, sRGB inputs): per-image sums of dE00 and
 * out_grad[n,3,h,w] = grad_scale * d dE00(pixel) / d x (clamp pass-through mask included), i.e. the gradient
 * of `dE.mean()` when grad_scale = 1 / (n h w).  replaces losses.py:138-143 + its autograd graph. */
int llie_deltae_loss_grad(const llie_deltae_desc *desc, const float *x, const float *y, float grad_scale,
                          double *out_sum, float *out_grad, int32_t *out_flags, void *workspace, void *stream);

/* nn.L1Loss value + gradient in one pass: out_sum[1] = sum |x - y| (float64); out_grad (NULL = value only) =
 * grad_scale * sign(x - y), added to its previous content when `accumulate`.  n = element count.
 * replaces NewBP_model/losses.py:249,332 (nn.L1Loss on the RAW tensors) and its autograd. */
int llie_l1_grad(const float *x, const float *y, int64_t n, float grad_scale, int32_t accumulate, double *out_sum,
                 float *out_grad, int32_t *out_flags, void *workspace, void *stream);

/* BasicSR pixel losses, value + gradient in one pass:
 *   out_sum[1]  = sum_e w_e psi(x_e - y_e)                                  (float64)
 *   out_grad    = grad_scale * sample_coef[n] * w_e * psi'(x_e - y_e)       (NULL = value only; += when accumulate)
 * psi: LLIE_PIX_L1 |d|, LLIE_PIX_MSE d^2, LLIE_PIX_CHARBONNIER sqrt(d^2 + eps).  `weight` NULL or [n, c, plane]
 * (weight_per_channel = 1) / [n, 1, plane] (0); `sample_coef` NULL or [n] device floats.  plane = H*W, % 4 == 0.
 * replaces NAFNet_base/basicsr/models/losses/losses.py:18-31, 112-138 and loss_util.py:31-60 + autograd. */
#define LLIE_PIX_L1 0
#define LLIE_PIX_MSE 1
#define LLIE_PIX_CHARBONNIER 2
int llie_pixel_loss_grad(int32_t mode, const float *x, const float *y, const float *weight, int32_t weight_per_channel,
                         const float *sample_coef, int32_t n, int32_t c, int64_t plane, float eps, float grad_scale,
                         int32_t accumulate, double *out_sum, float *out_grad, int32_t *out_flags, void *workspace,
                         void *stream);

/* Epilogue of the fused hybrid loss (losses.py:249-345: `total = w_l1 L1 + w_de dE + w_ssim DSSIM + w_phys Phys`, the
 * `logs` values): the four term values from the raw device sums of the loss kernels and their weighted total in ONE
 * launch, in place of the dozen scalar tensor operations of the eager expression.
 *   l1_sum [1] f64 / l1_count; de_sums [n_de] f64 / de_count; ssim_sums [n_ssim] f64 / ssim_count: term = float(sum / count)
 *   phys_value: device float (batch mean from llie_phys_value_and_grad); a NULL pointer marks an absent term (NaN in
 *   out_terms, left out of the total).  weights: HOST float[4] in the order L1, dE, SSIM, Phys.
 *   out_terms [4] float, out_total [1] float (sum of float32 products in that order). */
int llie_hybrid_combine(const double *l1_sum, double l1_count, const double *de_sums, int32_t n_de, double de_count,
                        const double *ssim_sums, int32_t n_ssim, double ssim_count, const float *phys_value,
                        const float *weights, float *out_terms, float *out_total, void *stream);

/* g[0..n) *= *scale (device scalar), a no-op launch when *scale == 1: applies autograd's upstream scalar to a gradient
 * that a value + gradient kernel already produced in the forward pass.  replaces the `grad * grad_output` multiply of
 * the autograd graph behind NewBP_model/losses.py:92-372. */
int llie_scale_inplace(float *g, int64_t n, const float *scale, void *stream);

/* Merge of overlapping validation tiles: out[1, c, h, w] = average over the tiles covering each pixel, summed in
 * tile order (tiles [n_tiles, c, tile_h, tile_w], tile_ij [n_tiles][2] = top-left (row, col), int32, device).
 * replaces NAFNet_base/basicsr/models/image_restoration_model.py:215-241 (grids_inverse, done on the host there). */
int llie_tile_merge(const float *tiles, const int32_t *tile_ij, int32_t n_tiles, int32_t c, int32_t tile_h,
                    int32_t tile_w, int32_t h, int32_t w, float *out, void *stream);

/* ------------------------------------------------------------------------
 * Input decode / normalise step in front of the path.
 * src is the decoded 16-bit PNG, interleaved [n, h, w, c] uint16 on the device; outputs are planar float32
 * [n, c, crop_h, crop_w] (either may be NULL):
 *   out_raw     = float32(u) / 65535                      (one IEEE division)
 *   out_aligned = clip(out_raw * rho, 0, 1)               (rho: [n] device floats, or rho_scalar when NULL)
 * bit-exact with the host arithmetic it replaces: NAFNet_base/basicsr/data/sony_sid_lmdb_dataset.py:208-212
 * (astype(float32) / MAX_16BIT_VALUE, np.clip(short_raw * expo_ratio, 0, 1)), :161-194 (crop window),
 * :57 (cv2 BGR -> RGB, swap_rb = 1) and basicsr/utils/img_util.py:28-34 (HWC -> CHW float tensor).
 * ------------------------------------------------------------------------ */
typedef struct llie_decode_desc {
    int32_t n, h, w, c;          /* source images */
    int32_t top, left;           /* crop window origin inside the source */
    int32_t crop_h, crop_w;      /* output size (== h, w for no crop) */
    int32_t swap_rb;             /* 1: output channel k reads source channel 2 - k (c >= 3, first three channels) */
} llie_decode_desc;
int llie_decode_u16(const llie_decode_desc *desc, const uint16_t *src, const float *rho, float rho_scalar,
                    float *out_raw, float *out_aligned, void *stream);

/* linear RGB -> sRGB (IEC 61966-2-1) of the evaluation helpers, NaN-propagating like torch.clamp / torch.where:
 *   s = where(x' <= 0.0031308, 12.92 x', 1.055 max(x', 0.0031308)^(1/2.4) - 0.055),  x' = clamp(x, 0, 1)
 * out_srgb = s, out_lpips = 2 s - 1 (either may be NULL; both from one read of x).
 * replaces metrics/evaluation.py:27-38 (_linear_to_srgb) and :50-52 (_norm_to_lpips_range). */
int llie_linear_to_srgb(const float *x, int64_t n, float *out_srgb, float *out_lpips, void *stream);

/* BT.601 luma plane of [n, 3, plane] float32 images: y = 0.2989 r + 0.5870 g + 0.1140 b with every product and sum
 * rounded separately, the value of the eager expression in metrics/ssim.py:116-127 (_to_luma_bt601; color_space='y'
 * of SSIMEvaluator).  y: [n, 1, plane]. */
int llie_luma_bt601(const float *x, int32_t n, int64_t plane, float *y, void *stream);

/* ------------------------------------------------------------------------
 * Per-pair epilogue of the sharded evaluation driver: adapter values from the raw device sums, added to the rank's
 * running vector acc = [k metric sums, pair count, pairs with non-finite input] (float64), one launch, no host sync.
 * replaces the host arithmetic of NAFNet_base/basicsr/metrics/lowlight_metrics.py:211-272 (the five adapters) and
 * the `metric_results[name] += value` accumulation of image_restoration_model.py:420-428.
 *   sqerr        [n] float64 sums of (pred - target)^2 per image            (llie_deltae_sqerr)   -> linear_psnr
 *   ssim_sums    [n, ssim_c] float64 SSIM-map sums over ssim_count elements (llie_ssim_stats)     -> linear_ssim
 *   de_sums      [n] float64 dE00 sums over hw pixels                       (llie_deltae)         -> deltae2000_mean
 *   de_p95       [1] float                                                  (llie_select_quantiles) -> deltae2000_p95
 *   edge_sum     [1] float64, edge_counts [n] int64                         (gated select)        -> edge_deltae2000_mean
 *   flags0..3    NULL or the out_flags words of the kernels above
 * A slot of -1 skips that metric (its inputs may be NULL).
 * ------------------------------------------------------------------------ */
typedef struct llie_eval_desc {
    int32_t n, c;            /* images in this pair batch, channels */
    int64_t hw;              /* pixels per image */
    int32_t ssim_c;          /* columns of ssim_sums */
    int64_t ssim_count;      /* elements per image the SSIM sums cover */
    double data_range;       /* PSNR peak */
    int32_t k;               /* metric slots in acc */
    int32_t slot_psnr, slot_ssim, slot_de_mean, slot_de_p95, slot_edge;
    /* channel-wise family (metrics/channelwise.py:113-302), c == 3: */
    int32_t slot_cpsnr;        /* cpsnr_rgb(reduction='mean')                                   <- psnr_table */
    int32_t slot_rgb_psnr[4];  /* rgb_psnr(reduction='mean')['R' | 'G' | 'B' | 'mean']            <- psnr_table */
    int32_t slot_rgb_ssim[4];  /* rgb_ssim(reduction='mean')['R' | 'G' | 'B' | 'mean']            <- rgb_ssim_sums */
} llie_eval_desc;
/*   psnr_table     [n + 2, LLIE_PSNR_COLS(3)] float64 as written by llie_psnr_stats(cast_first = 1), or NULL with
 *   sqerr_rgb      [n, 3] float64 per-channel squared-error sums (llie_deltae_sqerr_rgb; data_range 1, eps 1e-12)
 *   rgb_ssim_sums  [n, 3] float64 SSIM-map sums over hw pixels of the ssim_linear policy (llie_ssim_stats, clamp_var) */
int llie_eval_accumulate(const llie_eval_desc *desc, const double *sqerr, const double *ssim_sums,
                         const double *de_sums, const float *de_p95, const double *edge_sum,
                         const int64_t *edge_counts, const double *psnr_table, const double *sqerr_rgb,
                         const double *rgb_ssim_sums, const int32_t *flags0, const int32_t *flags1, const int32_t *flags2, const int32_t *flags3,
                         const int32_t *flags4, const int32_t *flags5, double *acc, void *stream);

/* Accuracy self-test hook: out[i] = the kernels' sRGB -> linear transfer of in[i] (pow24 path), so that the
 * < 1 ulp claim can be checked against a float64 pow from the test-suite. */
int llie_selftest_srgb_to_linear(const float *in, float *out, int64_t n, void *stream);

/* Accuracy self-test hook: the kernels divide by their compile-time constants (which: 0 1.055, 1 12.92, 2 0.95047,
 * 3 1.08883, 4 25, 5 pi, 6 65535 of the uint16 decode) with a three-instruction FMA sequence that must equal the IEEE division; *out_mismatch (device)
 * = number of operands among the float bit patterns first_bits .. first_bits + count - 1 where it does not. */
int llie_selftest_div_const(int32_t which, uint32_t first_bits, uint64_t count, uint64_t *out_mismatch, void *stream);

/* Sobel magnitude sqrt(gx^2 + gy^2 + 1e-12) of [planes, h, w], zero padding (color_error.py:23-30, 296-301) */
int llie_sobel_magnitude(const float *l_planes, float *out_g, int64_t planes, int32_t h, int32_t w, void *stream);

/* ------------------------------------------------------------------------
 * Exact torch.quantile(..., 'linear') by radix select, optionally gated by `gate[i] >= gate_thr[image]`.
 * replaces the full sorts of metrics/color_error.py:213-231 (percentiles), :334-335 (edge threshold),
 * :336-342 (boolean gather + mean + p95 of the edge region).
 * ------------------------------------------------------------------------ */
typedef struct llie_select_desc {
    int32_t images;       /* values is [images, per_image] */
    int64_t per_image;
    int32_t nq;           /* 1 or 2 quantiles */
    int32_t joint;        /* 0: one population per image | 1: all images form ONE population */
} llie_select_desc;
/* bytes of the caller-owned `select_ws`: ZERO before first use, self-resetting like `workspace`, and to be
 * used with the SAME `images` on every call (its internal layout depends on it) */
size_t llie_select_workspace_bytes(int32_t images, int32_t reserved);
/*   q          HOST array of nq fractions in [0,1]
 *   out_q      [segments, nq] floats (segments = joint ? 1 : images); NaN for an empty population
 *   out_count  [images] int64: participating elements per image (the edge-mask pixel count)
 *   out_sum    NULL or [segments] doubles: sum of the participating values */
/* nq == 0 (q, out_q may be NULL): one pass that only yields the gated counts and sum (masked mean). */
int llie_select_quantiles(const llie_select_desc *desc, const float *values, const float *gate,
                          const float *gate_thr, const double *q, float *out_q, int64_t *out_count,
                          double *out_sum, void *select_ws, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* LLIE_B200_H */
