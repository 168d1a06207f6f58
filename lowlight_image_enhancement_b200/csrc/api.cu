// ABI bookkeeping: version, error string, workspace size, device check.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace llie {
static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
int cuda_fail(cudaError_t e, const char *what) {
    set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
    return LLIE_ECUDA;
}
}  // namespace llie

extern "C" int llie_abi_version(void) { return 3; }  // 2: SSIM entry points take a host copy of the window taps; 3: llie_eval_accumulate channel-wise slots
extern "C" const char *llie_last_error(void) { return llie::g_err; }
extern "C" size_t llie_workspace_bytes(void) { return llie::kWsBytes; }
extern "C" int llie_check_device(void) {
    int dev = 0;
    LLIE_CUDA_OK(cudaGetDevice(&dev));
    int major = 0;
    LLIE_CUDA_OK(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
    if (major != 10) {
        llie::set_error("device %d has compute capability %d.x; this library is sm_100a only", dev, major);
        return LLIE_ENOSM100;
    }
    return LLIE_OK;
}
