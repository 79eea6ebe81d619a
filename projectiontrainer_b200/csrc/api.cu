// Error plumbing and device checks behind the C ABI (include/pt_stage1.h).
#include "pt_common.cuh"
#include <stdarg.h>

namespace pt {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int check_launch(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("%s: launch failed: %s", what, cudaGetErrorString(e));
        return PT_ERR_CUDA;
    }
    return PT_OK;
}

int num_sms() {
    static int cached = 0;
    if (cached) return cached;
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) {
        (void)cudaGetLastError();
        return kNumSMsFallback;
    }
    cached = n;
    return cached;
}

static int g_pdl = 0;
bool pdl_enabled(int kind) { return (g_pdl & kind) != 0; }

}  // namespace pt

extern "C" int pt_set_dependent_launch(int on) {   // bit set of kPdl* kinds
    const int was = pt::g_pdl;
    pt::g_pdl = on;
    return was;
}

extern "C" const char* pt_last_error(void) { return pt::g_err; }

extern "C" int pt_version(void) { return 100; }

extern "C" int pt_device_check(void) {
    static int ok = 0;
    if (ok) return PT_OK;
    int dev = 0, major = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        pt::set_error("no CUDA device: %s (this library has no CPU path)", cudaGetErrorString(e));
        return PT_ERR_CUDA;
    }
    if (major != 10) {
        pt::set_error("device compute capability %d.x is not sm_100 (kernels are sm_100a only)", major);
        return PT_ERR_CUDA;
    }
    ok = 1;
    return PT_OK;
}
