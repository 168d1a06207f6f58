// Per-pair epilogue of the sharded evaluation driver: turns the raw device sums the metric kernels left behind into the
// adapter values (lowlight_metrics.py:211-272) and adds them to the rank's running float64 vector -- one single-thread
// launch instead of the ~30 tiny elementwise / reduce launches the same arithmetic costs as torch glue (10 % of the
// per-image time at 12 Mpx).  Follows the float32 / float64 roundings of the adapters step by step.
#include <math_constants.h>

#include "common.cuh"

namespace llie {

struct EvalAccArgs {
    llie_eval_desc d;
    const double *sqerr, *ssim_sums, *de_sums, *edge_sum;
    const float *de_p95;
    const long long *edge_counts;
    const double *psnr_table, *sqerr_rgb, *rgb_ssim_sums;
    const int *flags[6];
    double *acc;
};

__global__ void eval_accumulate_kernel(const EvalAccArgs a) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const llie_eval_desc &d = a.d;
    const int n = d.n;
    if (d.slot_psnr >= 0) {  // metrics/psnr.py:39-88 on the whole batch tensor
        double s = 0.0;
        for (int i = 0; i < n; ++i) s += a.sqerr[i];
        const double mse = s / ((double)n * (double)d.c * (double)d.hw);
        double v = 10.0 * log10(d.data_range * d.data_range / mse);
        if (fabs(mse) <= 1e-12) v = CUDART_INF;
        a.acc[d.slot_psnr] += v;
    }
    if (d.slot_ssim >= 0) {  // float32 per-image scores, float64 mean over the batch
        double m = 0.0;
        for (int i = 0; i < n; ++i) {
            double s = 0.0;
            for (int c = 0; c < d.ssim_c; ++c) s += a.ssim_sums[i * d.ssim_c + c];
            m += (double)(float)(s / (double)d.ssim_count);
        }
        a.acc[d.slot_ssim] += m / n;
    }
    if (d.slot_de_mean >= 0) {  // float32 per-image means, float32 mean over the batch
        float m = 0.f;
        for (int i = 0; i < n; ++i) m += (float)(a.de_sums[i] / (double)d.hw);
        a.acc[d.slot_de_mean] += (double)(n > 1 ? m / (float)n : m);
    }
    if (d.slot_de_p95 >= 0) a.acc[d.slot_de_p95] += (double)a.de_p95[0];
    if (d.slot_edge >= 0) {  // masked mean over the whole batch; 0 / 0 = NaN for an empty mask, as the reference
        long long cnt = 0;
        for (int i = 0; i < n; ++i) cnt += a.edge_counts[i];
        a.acc[d.slot_edge] += a.edge_sum[0] / (double)cnt;
    }
    if (a.psnr_table != nullptr) {  // channelwise.py:113-222, reduction 'mean' = table row n; columns see llie_b200.h
        const double *row = a.psnr_table + (size_t)n * 9;
        if (d.slot_cpsnr >= 0) a.acc[d.slot_cpsnr] += row[7];
        for (int c = 0; c < 3; ++c)
            if (d.slot_rgb_psnr[c] >= 0) a.acc[d.slot_rgb_psnr[c]] += row[3 + c];
        if (d.slot_rgb_psnr[3] >= 0) a.acc[d.slot_rgb_psnr[3]] += row[6];
    }
    if (a.psnr_table == nullptr && a.sqerr_rgb != nullptr) {
        // the same table row from the per-channel squared-error sums the dE00 pass leaves behind (llie_deltae_sqerr_rgb):
        // the arithmetic of psnr_stats_kernel's finalize (pointwise.cu) with data_range 1, eps 1e-12 -- the defaults of
        // channelwise.py:113-222, the only configuration this path is used for
        double pc[3] = {0.0, 0.0, 0.0}, pmean = 0.0, cps = 0.0;
        for (int i = 0; i < n; ++i) {
            double msum = 0.0, psum = 0.0;
            for (int c = 0; c < 3; ++c) {
                const double mse = a.sqerr_rgb[i * 3 + c] / (double)d.hw;
                const double p = 10.0 * log10(1.0 / fmax(mse, 1e-12));
                pc[c] += p;
                msum += mse;
                psum += p;
            }
            pmean += psum / 3;
            cps += 10.0 * log10(1.0 / fmax(msum / 3, 1e-12));
        }
        if (d.slot_cpsnr >= 0) a.acc[d.slot_cpsnr] += cps / n;
        for (int c = 0; c < 3; ++c)
            if (d.slot_rgb_psnr[c] >= 0) a.acc[d.slot_rgb_psnr[c]] += pc[c] / n;
        if (d.slot_rgb_psnr[3] >= 0) a.acc[d.slot_rgb_psnr[3]] += pmean / n;
    }
    if (a.rgb_ssim_sums != nullptr) {  // channelwise.py:262-295: float32 per-(image, channel) means, batch mean
        float ch[3] = {0.f, 0.f, 0.f}, mean = 0.f;
        for (int i = 0; i < n; ++i) {
            float m = 0.f;
            for (int c = 0; c < 3; ++c) {
                const float v = (float)(a.rgb_ssim_sums[i * 3 + c] / (double)d.hw);
                ch[c] += v;
                m += v;
            }
            mean += m / 3.f;
        }
        for (int c = 0; c < 3; ++c)
            if (d.slot_rgb_ssim[c] >= 0) a.acc[d.slot_rgb_ssim[c]] += (double)(n > 1 ? ch[c] / (float)n : ch[c]);
        if (d.slot_rgb_ssim[3] >= 0) a.acc[d.slot_rgb_ssim[3]] += (double)(n > 1 ? mean / (float)n : mean);
    }
    int fl = 0;
    for (int k = 0; k < 6; ++k)
        if (a.flags[k] != nullptr) fl |= *a.flags[k];
    a.acc[d.k] += 1.0;
    a.acc[d.k + 1] += fl != 0 ? 1.0 : 0.0;
}

}  // namespace llie

extern "C" int llie_eval_accumulate(const llie_eval_desc *desc, const double *sqerr, const double *ssim_sums,
                                    const double *de_sums, const float *de_p95, const double *edge_sum,
                                    const int64_t *edge_counts, const double *psnr_table, const double *sqerr_rgb,
                                    const double *rgb_ssim_sums, const int32_t *flags0, const int32_t *flags1,
                                    const int32_t *flags2, const int32_t *flags3, const int32_t *flags4,
                                    const int32_t *flags5, double *acc, void *stream) {
    LLIE_REQUIRE(desc != nullptr && acc != nullptr, "eval_accumulate: null pointer");
    const llie_eval_desc &d = *desc;
    LLIE_REQUIRE(d.n >= 1 && d.k >= 1, "eval_accumulate: empty batch / no metric slots");
    const int slots[14] = {d.slot_psnr, d.slot_ssim, d.slot_de_mean, d.slot_de_p95, d.slot_edge, d.slot_cpsnr,
                           d.slot_rgb_psnr[0], d.slot_rgb_psnr[1], d.slot_rgb_psnr[2], d.slot_rgb_psnr[3],
                           d.slot_rgb_ssim[0], d.slot_rgb_ssim[1], d.slot_rgb_ssim[2], d.slot_rgb_ssim[3]};
    bool want_tab = d.slot_cpsnr >= 0, want_rs = false;
    for (int q = 0; q < 4; ++q) {
        want_tab = want_tab || d.slot_rgb_psnr[q] >= 0;
        want_rs = want_rs || d.slot_rgb_ssim[q] >= 0;
    }
    LLIE_REQUIRE((!want_tab || ((psnr_table || sqerr_rgb) && d.c == 3)) && (!want_rs || (rgb_ssim_sums && d.c == 3)),
                 "eval_accumulate: a channel-wise metric needs its table / sums and 3-channel images");
    if (!want_tab) psnr_table = sqerr_rgb = nullptr;
    if (!want_rs) rgb_ssim_sums = nullptr;
    for (int s : slots) LLIE_REQUIRE(s >= -1 && s < d.k, "eval_accumulate: slot %d outside [0, %d)", s, d.k);
    LLIE_REQUIRE(d.slot_psnr < 0 || sqerr, "eval_accumulate: linear_psnr needs the squared-error sums");
    LLIE_REQUIRE(d.slot_ssim < 0 || (ssim_sums && d.ssim_c >= 1 && d.ssim_count >= 1),
                 "eval_accumulate: linear_ssim needs the SSIM sums");
    LLIE_REQUIRE((d.slot_de_mean < 0 || de_sums) && (d.slot_de_p95 < 0 || de_p95) &&
                     (d.slot_edge < 0 || (edge_sum && edge_counts)),
                 "eval_accumulate: a requested colour metric has no input");
    llie::EvalAccArgs a{d, sqerr, ssim_sums, de_sums, edge_sum, de_p95,
                        reinterpret_cast<const long long *>(edge_counts), psnr_table, sqerr_rgb, rgb_ssim_sums,
                        {flags0, flags1, flags2, flags3, flags4, flags5}, acc};
    llie::eval_accumulate_kernel<<<1, 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(a);
    LLIE_CUDA_OK(cudaGetLastError());
    return LLIE_OK;
}
