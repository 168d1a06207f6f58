// Physics-consistency projector, forward and adjoint backward (sm_100a).
//
//   r = clipY(sY * (K (*) padY(pred))) - clipA(sA * obs)      loss map = |r| or sqrt(r^2+eps^2)
//
// Replaces metrics/phys_consistency.py:193-255 (_phys_cons_core) and the physics loss terms
// NewBP_model/losses.py:173-192, :217-220 of the reference, fusing pad + conv + rho + clip + crop +
// residual + robust penalty + per-sample reduction (+ finiteness flags) into ONE pass over pred/obs,
// and the residual derivative + conv_transpose2d adjoint (newbp_layer.py:16-21) into one backward pass.
//
// Two code paths, both parity-tested:
//   * row ring     (phys_ring.cu) depthwise 3x3, W % 4 == 0, 16-byte aligned planes: persistent CTAs stream rows
//                  through a TMA ring in shared memory, HBM-bound.
//   * generic path (this file) any kh x kw, dense cross-channel PSFs, arbitrary W / alignment / rho strides.
#include "phys_common.cuh"
#include <atomic>

namespace llie {

// conv value at output (n, co, oy, ox), any layout / size / pad mode (generic path)
__device__ __forceinline__ float conv_at(const PhysArgs &a, int n, int co, int oy, int ox) {
    const llie_phys_desc &d = a.d;
    const int ph = d.kh >> 1, pw = d.kw >> 1;
    float acc = 0.f;
    const int ci0 = d.psf_layout == LLIE_PSF_DENSE ? 0 : co;
    const int ci1 = d.psf_layout == LLIE_PSF_DENSE ? d.c_in : co + 1;
    for (int ci = ci0; ci < ci1; ++ci) {
        const float *plane = a.pred + ((long long)n * d.c_in + ci) * d.h * d.w;
        const float *wk = a.psf + (d.psf_layout == LLIE_PSF_DENSE ? ((long long)co * d.c_in + ci) : co) * d.kh * d.kw;
        for (int ky = 0; ky < d.kh; ++ky) {
            int sy = pad_index(oy + ky - ph, d.h, d.pad_mode);
            if (sy < 0) continue;
            for (int kx = 0; kx < d.kw; ++kx) {
                int sx = pad_index(ox + kx - pw, d.w, d.pad_mode);
                if (sx < 0) continue;
                acc = fmaf(__ldg(wk + ky * d.kw + kx), __ldg(plane + (long long)sy * d.w + sx), acc);
            }
        }
    }
    return acc;
}

// ------------------------------------------------------------------------------------------------
// Finalize: last block sums the per-block partials in a fixed order and writes the outputs.
// ------------------------------------------------------------------------------------------------
__device__ void phys_finalize(const PhysArgs &a, double *red) {
    const int G = gridDim.x, N = a.d.n;
    double batch = 0.0, total = 0.0;
    int flags = 0;
    for (int n = 0; n < N; ++n) {
        double s = 0.0;
        for (int b = threadIdx.x; b < G; b += blockDim.x) {
            s += __ldcg(a.part_sum + (size_t)n * G + b);
            flags |= __ldcg(a.part_flag + (size_t)n * G + b);
        }
        s = block_sum<double, true>(s, red);
        if (threadIdx.x == 0) {
            float m = (float)(s * a.inv_count);
            a.out_mean[n] = m;
            if (a.out_sum) a.out_sum[n] = s;
            batch += (double)m;
            total += s;
        }
    }
    int f = block_or(flags, reinterpret_cast<int *>(red));
    if (threadIdx.x == 0) {
        a.out_mean[N] = (float)(batch / N);
        a.out_mean[N + 1] = (float)batch;
        if (a.out_sum) a.out_sum[N] = total;
        *a.out_flags = f;
    }
}

// ------------------------------------------------------------------------------------------------
// Generic forward: one thread per element of the full [c_out, H, W] grid of sample blockIdx.y.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) phys_fwd_generic(const PhysArgs a) {
    __shared__ double red[32];
    const llie_phys_desc &d = a.d;
    const int n = blockIdx.y;
    const long long plane = (long long)d.h * d.w;
    const long long total = plane * d.c_out;
    double acc = 0.0;
    int flag = 0;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        const int co = (int)(idx / plane);
        const int rem = (int)(idx - (long long)co * plane);
        const int y = rem / d.w, x = rem - y * d.w;
        const float av = __ldg(a.obs + ((long long)n * d.c_out + co) * plane + rem);
        if (!finite_f(av)) flag |= 2;
        const bool in_dom = (y >= a.y_lo) && (y < a.y_lo + a.ho) && (x >= a.x_lo) && (x < a.x_lo + a.wo);
        if (!in_dom) {
            // cropped border: only the finiteness contract of the reference applies here
            if (co < d.c_in && !finite_f(__ldg(a.pred + ((long long)n * d.c_in + co) * plane + rem))) flag |= 1;
            continue;
        }
        const float cv = conv_at(a, n, co, y, x);
        if (!finite_f(cv)) flag |= 1;
        float r;
        const float l = residual(d, cv, av, rho_at(a, n, co, y, x), r);
        acc += (double)l;
        if (a.out_map)
            a.out_map[(((long long)n * d.c_out + co) * a.ho + (y - a.y_lo)) * a.wo + (x - a.x_lo)] = fabsf(r);
    }
    acc = block_sum<double>(acc, red);
    const int f = block_or(flag, reinterpret_cast<int *>(red));
    if (threadIdx.x == 0) {
        a.part_sum[(size_t)n * gridDim.x + blockIdx.x] = acc;
        a.part_flag[(size_t)n * gridDim.x + blockIdx.x] = f;
    }
    if (last_block_done(a.ticket, gridDim.x * gridDim.y)) phys_finalize(a, red);
}

// ------------------------------------------------------------------------------------------------
// Backward, generic path: (1) q map into scratch, (2) transposed gather with padding fold.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) phys_bwd_q_generic(const PhysArgs a) {
    const llie_phys_desc &d = a.d;
    const int n = blockIdx.y;
    const long long plane = (long long)d.h * d.w;
    const long long total = plane * d.c_out;
    const float gn = __ldg(a.gout + a.gout_stride * n) * a.gscale;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        const int co = (int)(idx / plane);
        const int rem = (int)(idx - (long long)co * plane);
        const int y = rem / d.w, x = rem - y * d.w;
        float q = 0.f;
        if (y >= a.y_lo && y < a.y_lo + a.ho && x >= a.x_lo && x < a.x_lo + a.wo) {
            const float cv = conv_at(a, n, co, y, x);
            const float av = __ldg(a.obs + ((long long)n * d.c_out + co) * plane + rem);
            q = gn * dresidual(d, cv, av, rho_at(a, n, co, y, x));
        }
        a.scratch[((long long)n * d.c_out + co) * plane + rem] = q;
    }
}

__global__ void __launch_bounds__(256) phys_bwd_gather_generic(const PhysArgs a) {
    const llie_phys_desc &d = a.d;
    const int n = blockIdx.y;
    const int ph = d.kh >> 1, pw = d.kw >> 1;
    const long long plane = (long long)d.h * d.w;
    const long long total = plane * d.c_in;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        const int ci = (int)(idx / plane);
        const int rem = (int)(idx - (long long)ci * plane);
        const int y = rem / d.w, x = rem - y * d.w;
        float acc = 0.f;
        const int co0 = d.psf_layout == LLIE_PSF_DENSE ? 0 : ci;
        const int co1 = d.psf_layout == LLIE_PSF_DENSE ? d.c_out : ci + 1;
        // candidate padded coordinates that may alias onto (y, x): the pixel itself plus the
        // out-of-image bands [-p, 0) and [size, size+p)
        for (int ty = 0; ty <= 2 * ph; ++ty) {
            const int jy = (ty == 0) ? y : (ty <= ph ? -ty : d.h - 1 + (ty - ph));
            if (ty != 0 && pad_index(jy, d.h, d.pad_mode) != y) continue;
            for (int tx = 0; tx <= 2 * pw; ++tx) {
                const int jx = (tx == 0) ? x : (tx <= pw ? -tx : d.w - 1 + (tx - pw));
                if (tx != 0 && pad_index(jx, d.w, d.pad_mode) != x) continue;
                for (int co = co0; co < co1; ++co) {
                    const float *wk =
                        a.psf + (d.psf_layout == LLIE_PSF_DENSE ? ((long long)co * d.c_in + ci) : co) * d.kh * d.kw;
                    const float *qp = a.scratch + ((long long)n * d.c_out + co) * plane;
                    for (int ky = 0; ky < d.kh; ++ky) {
                        const int oy = jy - ky + ph;
                        if (oy < 0 || oy >= d.h) continue;
                        for (int kx = 0; kx < d.kw; ++kx) {
                            const int ox = jx - kx + pw;
                            if (ox < 0 || ox >= d.w) continue;
                            acc = fmaf(__ldg(wk + ky * d.kw + kx), qp[(long long)oy * d.w + ox], acc);
                        }
                    }
                }
            }
        }
        a.grad[((long long)n * d.c_in + ci) * plane + rem] = acc;
    }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static int validate(const llie_phys_desc *d) {
    LLIE_REQUIRE(d != nullptr, "phys: null descriptor");
    LLIE_REQUIRE(d->n > 0 && d->c_in > 0 && d->c_out > 0 && d->h > 0 && d->w > 0, "phys: empty tensor");
    LLIE_REQUIRE(d->kh >= 1 && d->kw >= 1 && (d->kh & 1) && (d->kw & 1), "phys: PSF must be odd-sized");
    LLIE_REQUIRE(d->psf_layout == LLIE_PSF_DENSE || d->c_in == d->c_out, "phys: depthwise PSF needs c_in == c_out");
    LLIE_REQUIRE(d->pad_mode >= LLIE_PAD_ZEROS && d->pad_mode <= LLIE_PAD_REPLICATE, "phys: bad pad mode");
    LLIE_REQUIRE(d->pad_mode != LLIE_PAD_REFLECT || (d->kh / 2 < d->h && d->kw / 2 < d->w),
                 "phys: reflect padding needs pad < size");
    LLIE_REQUIRE(d->n <= 65535, "phys: batch too large for grid.y");
    if (d->crop_valid) LLIE_REQUIRE(d->h > 2 * (d->kh / 2) && d->w > 2 * (d->kw / 2), "phys: valid crop leaves nothing");
    return LLIE_OK;
}

static void fill_domain(PhysArgs &a) {
    const llie_phys_desc &d = a.d;
    a.y_lo = d.crop_valid ? d.kh / 2 : 0;
    a.x_lo = d.crop_valid ? d.kw / 2 : 0;
    a.ho = d.h - 2 * a.y_lo;
    a.wo = d.w - 2 * a.x_lo;
    a.inv_count = 1.0 / ((double)d.c_out * a.ho * a.wo);
}

static int grid_x_for(int n, int work_items) {
    // persistent-style grid: ~8 resident CTAs of 256 threads per SM, split evenly over samples
    int per_sample = div_up(kNumSM * 8, n);
    if (per_sample > work_items) per_sample = work_items;
    if (per_sample < 1) per_sample = 1;
    return per_sample;
}

void ssim_tune(int key, int value);  // ssim.cu

static std::atomic<int> g_phys_path{0};  // 0 TMA row ring (default) | 2 generic kernels only (llie_tune, diagnostic)
static inline bool ring_enabled() { return g_phys_path.load(std::memory_order_relaxed) == 0; }

}  // namespace llie

using namespace llie;

extern "C" int llie_tune(int key, int value) {
    switch (key) {
        case LLIE_TUNE_PHYS_PATH:
            if (value != 0 && value != 2) {
                llie::set_error("llie_tune: LLIE_TUNE_PHYS_PATH takes 0 (row ring) or 2 (generic kernels), got %d", value);
                return LLIE_EINVAL;
            }
            g_phys_path.store(value, std::memory_order_relaxed);
            return LLIE_OK;
        case LLIE_TUNE_PHYS_ROWS: phys_ring_set_rows(value); return LLIE_OK;
        case LLIE_TUNE_PHYS_PDL: phys_ring_set_pdl(value); return LLIE_OK;
        case LLIE_TUNE_PHYS_PREFETCH: phys_ring_set_prefetch(value); return LLIE_OK;
        case 6:
            llie::set_error("llie_tune: key 6 (process-wide stream overlap) was removed in ABI 3; set "
                            "llie_phys_desc.stream_overlap on the calls that meet the contract");
            return LLIE_EINVAL;
        case LLIE_TUNE_SSIM_PATH:
        case LLIE_TUNE_SSIM_SEG_ROWS: ssim_tune(key, value); return LLIE_OK;
        default: break;
    }
    llie::set_error("llie_tune: unknown key %d", key);
    return LLIE_EINVAL;
}

extern "C" int llie_phys_forward(const llie_phys_desc *desc, const float *pred, const float *obs, const float *psf,
                                 const float *rho, float *out_mean, double *out_sum, float *out_map,
                                 int32_t *out_flags, void *workspace, void *stream) {
    if (int rc = validate(desc)) return rc;
    LLIE_REQUIRE(pred && obs && psf && out_mean && out_flags && workspace, "phys_forward: null pointer");
    PhysArgs a{};
    a.d = *desc;
    a.pred = pred; a.obs = obs; a.psf = psf; a.rho = rho;
    a.out_mean = out_mean; a.out_sum = out_sum; a.out_map = out_map; a.out_flags = out_flags;
    fill_domain(a);
    Workspace ws(workspace);
    a.ticket = ws.tickets;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (ring_enabled() && phys_ring_eligible(a)) return launch_phys_fwd_ring(a, workspace, st);
    long long items = (long long)desc->c_out * desc->h * desc->w;
    const int gx = grid_x_for(desc->n, (int)((items + 255) / 256));
    const size_t need = (size_t)desc->n * gx * (sizeof(double) + sizeof(int));
    LLIE_REQUIRE(need <= ws.scratch_bytes, "phys_forward: workspace too small");
    a.part_sum = reinterpret_cast<double *>(ws.scratch);
    a.part_flag = reinterpret_cast<int *>(ws.scratch + (size_t)desc->n * gx * sizeof(double));
    dim3 grid(gx, desc->n);
    phys_fwd_generic<<<grid, 256, 0, st>>>(a);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

extern "C" int llie_phys_backward_needs_scratch(const llie_phys_desc *desc, const float *pred, const float *obs,
                                               const float *rho, const float *grad_pred) {
    if (validate(desc)) return 1;
    PhysArgs a{};
    a.d = *desc;
    a.pred = pred; a.obs = obs; a.rho = rho;
    return (ring_enabled() && phys_ring_eligible(a) && aligned16(grad_pred)) ? 0 : 1;
}

extern "C" int llie_phys_backward(const llie_phys_desc *desc, const float *pred, const float *obs, const float *psf,
                                  const float *rho, const float *gout, int64_t gout_stride, float gscale,
                                  float *grad_pred, float *scratch, void *stream) {
    if (int rc = validate(desc)) return rc;
    LLIE_REQUIRE(pred && obs && psf && gout && grad_pred, "phys_backward: null pointer");
    PhysArgs a{};
    a.d = *desc;
    a.pred = pred; a.obs = obs; a.psf = psf; a.rho = rho;
    a.gout = gout; a.gout_stride = gout_stride; a.gscale = gscale;
    a.grad = grad_pred; a.scratch = scratch;
    fill_domain(a);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (ring_enabled() && phys_ring_eligible(a) && aligned16(grad_pred)) {
        return launch_phys_bwd_ring(a, st);
    } else {
        LLIE_REQUIRE(scratch != nullptr, "phys_backward: this PSF/shape needs the scratch buffer");
        LLIE_REQUIRE(desc->kh / 2 <= 16 && desc->kw / 2 <= 16, "phys_backward: PSF larger than 33x33 unsupported");
        long long items = (long long)desc->c_out * desc->h * desc->w;
        dim3 grid(grid_x_for(desc->n, (int)((items + 255) / 256)), desc->n);
        phys_bwd_q_generic<<<grid, 256, 0, st>>>(a);
        long long items2 = (long long)desc->c_in * desc->h * desc->w;
        dim3 grid2(grid_x_for(desc->n, (int)((items2 + 255) / 256)), desc->n);
        phys_bwd_gather_generic<<<grid2, 256, 0, st>>>(a);
    }
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}

extern "C" int llie_phys_value_and_grad(const llie_phys_desc *desc, const float *pred, const float *obs,
                                        const float *psf, const float *rho, const float *gout, int64_t gout_stride,
                                        float gscale, float *out_mean, double *out_sum, int32_t *out_flags,
                                        float *grad_pred, float *scratch, void *workspace, void *stream) {
    if (int rc = validate(desc)) return rc;
    LLIE_REQUIRE(pred && obs && psf && gout && grad_pred && out_mean && out_flags && workspace,
                 "phys_value_and_grad: null pointer");
    PhysArgs a{};
    a.d = *desc;
    a.pred = pred; a.obs = obs; a.psf = psf; a.rho = rho;
    a.gout = gout; a.gout_stride = gout_stride; a.gscale = gscale;
    a.grad = grad_pred; a.scratch = scratch;
    a.out_mean = out_mean; a.out_sum = out_sum; a.out_map = nullptr; a.out_flags = out_flags;
    fill_domain(a);
    if (ring_enabled() && phys_ring_eligible(a) && aligned16(grad_pred))
        return launch_phys_fused_ring(a, workspace, reinterpret_cast<cudaStream_t>(stream));
    // shapes the fused kernel does not take: the two single-purpose passes, same results
    if (int rc = llie_phys_forward(desc, pred, obs, psf, rho, out_mean, out_sum, nullptr, out_flags, workspace, stream))
        return rc;
    return llie_phys_backward(desc, pred, obs, psf, rho, gout, gout_stride, gscale, grad_pred, scratch, stream);
}
