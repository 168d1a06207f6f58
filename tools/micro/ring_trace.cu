// Per-CTA timeline of the fused phys ring kernel at BASELINE configs[1] (8x3x512x512): builds on a library compiled
// with -DLLIE_RING_TRACE (tools/micro/build_trace.sh) and prints, relative to the earliest CTA start, the median /
// max over CTAs of: setup done, first slot ready, loop done, ticket taken, plus the last CTA's finalize end.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#include "llie_b200.h"

extern "C" int llie_debug_ring_trace(unsigned long long *host_dst);

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("cuda error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

int main(int argc, char **argv) {
    const int n = argc > 1 ? atoi(argv[1]) : 8, c = 3, h = 512, w = 512;
    const size_t el = (size_t)n * c * h * w;
    const int pool = 6;
    std::vector<float> hp(el), ho(el);
    for (size_t i = 0; i < el; ++i) { hp[i] = (float)((i * 2654435761u) % 1000) * 1e-5f; ho[i] = (float)((i * 40503u) % 1000) * 1e-3f; }
    float *pred[pool], *obs[pool], *grad[pool];
    for (int i = 0; i < pool; ++i) {
        CK(cudaMalloc(&pred[i], el * 4)); CK(cudaMalloc(&obs[i], el * 4)); CK(cudaMalloc(&grad[i], el * 4));
        CK(cudaMemcpy(pred[i], hp.data(), el * 4, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(obs[i], ho.data(), el * 4, cudaMemcpyHostToDevice));
    }
    float hk[27];
    for (int q = 0; q < 27; ++q) hk[q] = (q % 9 == 4) ? 0.6f : 0.05f;
    float hrho[8] = {100, 250, 300, 100, 250, 300, 100, 250};
    std::vector<float> hr(n);
    for (int i = 0; i < n; ++i) hr[i] = hrho[i % 8];
    float one = 1.f, *psf, *rho, *gout, *omean; double *osum; int *flags; void *ws;
    CK(cudaMalloc(&psf, sizeof hk)); CK(cudaMemcpy(psf, hk, sizeof hk, cudaMemcpyHostToDevice));
    CK(cudaMalloc(&rho, n * 4)); CK(cudaMemcpy(rho, hr.data(), n * 4, cudaMemcpyHostToDevice));
    CK(cudaMalloc(&gout, 4)); CK(cudaMemcpy(gout, &one, 4, cudaMemcpyHostToDevice));
    CK(cudaMalloc(&omean, (n + 2) * 4)); CK(cudaMalloc(&osum, (n + 1) * 8)); CK(cudaMalloc(&flags, 4));
    CK(cudaMalloc(&ws, llie_workspace_bytes())); CK(cudaMemset(ws, 0, llie_workspace_bytes()));
    llie_phys_desc d = {};
    d.n = n; d.c_in = c; d.c_out = c; d.h = h; d.w = w; d.kh = 3; d.kw = 3;
    d.psf_layout = LLIE_PSF_DEPTHWISE; d.pad_mode = LLIE_PAD_REFLECT; d.crop_valid = 1; d.robust = 0; d.rho_side = 0;
    d.clip_y = 0; d.clip_a = 0; d.eps = 1e-12f; d.rho_scalar = 1.f; d.rho_sn = 1;
    const float gscale = 1.f / ((float)n * c * (h - 2) * (w - 2));
    if (argc > 2) llie_tune(LLIE_TUNE_PHYS_PDL, atoi(argv[2]));
    cudaStream_t st; CK(cudaStreamCreate(&st));
    const int mode = argc > 3 ? atoi(argv[3]) : 0;  // 0 fused value + gradient, 1 gradient only, 2 forward only
    printf("mode %d (%s)\n", mode, mode == 0 ? "fused value + gradient" : (mode == 1 ? "gradient only" : "forward only"));
    auto step = [&](int i) {
        int rc;
        if (mode == 0)
            rc = llie_phys_value_and_grad(&d, pred[i % pool], obs[i % pool], psf, rho, gout, 0, gscale, omean, osum, flags,
                                          grad[i % pool], nullptr, ws, st);
        else if (mode == 1)
            rc = llie_phys_backward(&d, pred[i % pool], obs[i % pool], psf, rho, gout, 0, gscale, grad[i % pool], nullptr, st);
        else
            rc = llie_phys_forward(&d, pred[i % pool], obs[i % pool], psf, rho, omean, osum, nullptr, flags, ws, st);
        if (rc) { printf("rc %d: %s\n", rc, llie_last_error()); exit(1); }
    };
    for (int i = 0; i < 10; ++i) step(i);
    CK(cudaStreamSynchronize(st));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0, st);
    for (int i = 0; i < 50; ++i) step(i);
    cudaEventRecord(e1, st);
    CK(cudaStreamSynchronize(st));
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("stream-launched: %.2f us per launch\n", ms * 1e3f / 50);
    CK(cudaStreamSynchronize(st));
    step(3);  // isolated launch for the trace
    CK(cudaStreamSynchronize(st));
    std::vector<unsigned long long> tr(1024 * 16);
    if (llie_debug_ring_trace(tr.data())) { printf("trace copy failed\n"); return 1; }
    int ctas = 0;
    unsigned long long t0 = ~0ull;
    for (int b = 0; b < 1024; ++b) if (tr[b * 16]) { ++ctas; t0 = std::min(t0, tr[b * 16]); }
    const char *names[12] = {"cta start", "setup done", "pdl wait done", "first slot ready", "loop done", "flags published",
                             "ticket taken", "finalize done (last CTA)", "setup: zero fill done", "setup: barriers ready",
                             "finalize: partials staged", "finalize: sums + flags"};
    printf("ctas traced: %d\n", ctas);
    for (int k = 0; k < 12; ++k) {
        std::vector<double> v;
        for (int b = 0; b < 1024; ++b) if (tr[b * 16] && tr[b * 16 + k] >= t0 && (!(k == 7 || k >= 10) || tr[b * 16 + k] > tr[b * 16 + 6])) v.push_back((tr[b * 16 + k] - t0) * 1e-3);
        if (v.empty()) continue;
        std::sort(v.begin(), v.end());
        printf("%-26s n=%4zu  min %6.2f  p50 %6.2f  p90 %6.2f  max %6.2f us\n", names[k], v.size(), v.front(), v[v.size() / 2],
               v[v.size() * 9 / 10], v.back());
    }
    return 0;
}
