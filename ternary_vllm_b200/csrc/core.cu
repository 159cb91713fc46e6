// core.cu — error reporting, device check, launch counter.
#include <stdarg.h>
#include <string.h>
#include "tma.cuh"

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

// cuTensorMapEncodeTiled is fetched through the runtime so that the library has no link-time
// dependency on libcuda (it must load, and fail loudly, on machines without a driver).
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int make_tensor_map(CUtensorMap* map, CUtensorMapDataType dt, int rank, const void* ptr,
                    const uint64_t* dims, const uint64_t* strides_bytes, const uint32_t* box,
                    CUtensorMapSwizzle swz) {
    static EncodeTiledFn enc = nullptr;
    if (!enc) {
        void* fp = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &qres) != cudaSuccess ||
            qres != cudaDriverEntryPointSuccess || !fp) {
            cudaGetLastError();
            set_error("cuTensorMapEncodeTiled is not available from the driver");
            return TB200_ECUDA;
        }
        enc = (EncodeTiledFn)fp;
    }
    cuuint64_t d[5] = {1, 1, 1, 1, 1}, st[4] = {0, 0, 0, 0};
    cuuint32_t bx[5] = {1, 1, 1, 1, 1}, es[5] = {1, 1, 1, 1, 1};
    for (int i = 0; i < rank; ++i) { d[i] = dims[i]; bx[i] = box[i]; }
    for (int i = 0; i + 1 < rank; ++i) st[i] = strides_bytes[i];
    CUresult r = enc(map, dt, (cuuint32_t)rank, const_cast<void*>(ptr), d, st, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     swz, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed with CUresult %d (rank %d dims %llu,%llu,%llu)", (int)r, rank,
                  (unsigned long long)d[0], (unsigned long long)d[1], (unsigned long long)d[2]);
        return TB200_ECUDA;
    }
    return TB200_OK;
}

}  // namespace tb200

extern "C" {
int tb200_abi_version(void) { return TB200_ABI_VERSION; }
const char* tb200_last_error(void) { return tb200::t_err; }
uint64_t tb200_launch_count(void) { return tb200::g_launches.load(); }
}
