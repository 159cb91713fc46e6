// core.cu — error reporting, device check, launch counter.
#include <stdarg.h>
#include <string.h>
#include "common.cuh"

namespace tb200 {

static thread_local char t_err[512] = "";
std::atomic<uint64_t> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(t_err, sizeof(t_err), fmt, ap);
    va_end(ap);
}

static int g_dev_state = 1;  // 1 = unknown, 0 = ok, <0 = error
static int g_sms = 0;

int check_device() {
    if (g_dev_state <= 0) {
        if (g_dev_state < 0) set_error("no sm_100 CUDA device available (libternary_b200 has no CPU fallback)");
        return g_dev_state;
    }
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    cudaDeviceProp p;
    if (e == cudaSuccess) e = cudaGetDeviceProperties(&p, dev);
    if (e != cudaSuccess) {
        cudaGetLastError();
        g_dev_state = TB200_ENODEV;
        set_error("no CUDA device: %s (libternary_b200 has no CPU fallback)", cudaGetErrorString(e));
        return g_dev_state;
    }
    if (p.major != 10) {
        g_dev_state = TB200_ENODEV;
        set_error("device '%s' is sm_%d%d; libternary_b200 is built for sm_100a only", p.name, p.major, p.minor);
        return g_dev_state;
    }
    g_sms = p.multiProcessorCount;
    g_dev_state = TB200_OK;
    return TB200_OK;
}

int sm_count() { return g_sms > 0 ? g_sms : 148; }

}  // namespace tb200

extern "C" {
int tb200_abi_version(void) { return TB200_ABI_VERSION; }
const char* tb200_last_error(void) { return tb200::t_err; }
uint64_t tb200_launch_count(void) { return tb200::g_launches.load(); }
}
