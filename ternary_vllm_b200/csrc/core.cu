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

// per-device: 1 = unknown, 0 = ok, <0 = error
static std::atomic<int> g_dev_state[TB_MAX_DEVICES];
static std::atomic<int> g_dev_sms[TB_MAX_DEVICES];
static std::atomic<bool> g_dev_init{false};
static std::atomic<int> g_no_device{0};          // cudaGetDevice itself failed (no driver / no GPU): cached

int current_device() {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return -1; }
    return dev;
}

int check_device() {
    if (g_no_device.load(std::memory_order_relaxed)) {
        set_error("no CUDA device (libternary_b200 has no CPU fallback)");
        return TB200_ENODEV;
    }
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) {
        cudaGetLastError();
        g_no_device.store(1);
        set_error("no CUDA device: %s (libternary_b200 has no CPU fallback)", cudaGetErrorString(e));
        return TB200_ENODEV;
    }
    if (dev < 0 || dev >= TB_MAX_DEVICES) { set_error("device ordinal %d out of range", dev); return TB200_ENODEV; }
    // states: 0 = unknown (zero-initialised), 1 = ok, 2 = not sm_100
    const int st = g_dev_state[dev].load(std::memory_order_acquire);
    if (st == 1) return TB200_OK;
    if (st == 2) {
        set_error("device %d is not sm_100; libternary_b200 is built for sm_100a only (no CPU fallback)", dev);
        return TB200_ENODEV;
    }
    int major = 0, sms = 0;
    e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) {
        cudaGetLastError();
        set_error("cannot query device %d: %s (libternary_b200 has no CPU fallback)", dev, cudaGetErrorString(e));
        return TB200_ENODEV;
    }
    if (major != 10) {
        g_dev_state[dev].store(2, std::memory_order_release);
        set_error("device %d is sm_%dx; libternary_b200 is built for sm_100a only (no CPU fallback)", dev, major);
        return TB200_ENODEV;
    }
    g_dev_sms[dev].store(sms, std::memory_order_relaxed);
    g_dev_state[dev].store(1, std::memory_order_release);
    return TB200_OK;
}

int sm_count() {
    const int dev = current_device();
    if (dev >= 0 && dev < TB_MAX_DEVICES) {
        int sms = g_dev_sms[dev].load(std::memory_order_relaxed);
        if (sms > 0) return sms;
        if (check_device() == TB200_OK) return g_dev_sms[dev].load(std::memory_order_relaxed);
    }
    return 148;
}

// cuTensorMapEncodeTiled is fetched through the runtime so that the library has no link-time
// dependency on libcuda (it must load, and fail loudly, on machines without a driver).
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int make_tensor_map(CUtensorMap* map, CUtensorMapDataType dt, int rank, const void* ptr,
                    const uint64_t* dims, const uint64_t* strides_bytes, const uint32_t* box,
                    CUtensorMapSwizzle swz) {
    static std::atomic<EncodeTiledFn> enc_slot{nullptr};
    EncodeTiledFn enc = enc_slot.load(std::memory_order_acquire);
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
        enc_slot.store(enc, std::memory_order_release);
    }
    cuuint64_t d[5] = {1, 1, 1, 1, 1}, st[4] = {0, 0, 0, 0};
    cuuint32_t bx[5] = {1, 1, 1, 1, 1}, es[5] = {1, 1, 1, 1, 1};
    for (int i = 0; i < rank; ++i) { d[i] = dims[i]; bx[i] = box[i]; }
    for (int i = 0; i + 1 < rank; ++i) st[i] = strides_bytes[i];
    // Per-thread cache: an eager per-layer call re-encoded 3-4 maps (~1-2 us of driver time each).  A map is a pure function of
    // the key below (address, shape, pitches, box, type, swizzle), so a hit can never be stale; 64 direct-mapped slots.
    struct MapKey { uint64_t v[16]; };
    struct MapSlot { MapKey key; CUtensorMap map; bool used; };
    static thread_local MapSlot cache[64];
    MapKey key = {};
    key.v[0] = (uint64_t)(uintptr_t)ptr; key.v[1] = ((uint64_t)dt << 32) | ((uint64_t)rank << 8) | (uint64_t)swz;
    for (int i = 0; i < 5; ++i) { key.v[2 + i] = d[i]; key.v[7 + i] = bx[i]; }
    for (int i = 0; i < 4; ++i) key.v[12 + i] = st[i];
    uint64_t h = 1469598103934665603ull;
    for (int i = 0; i < 16; ++i) { h ^= key.v[i]; h *= 1099511628211ull; }
    MapSlot& slot = cache[(h >> 20) & 63];
    if (slot.used && memcmp(&slot.key, &key, sizeof(key)) == 0) { *map = slot.map; return TB200_OK; }
    CUresult r = enc(map, dt, (cuuint32_t)rank, const_cast<void*>(ptr), d, st, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     swz, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed with CUresult %d (rank %d dims %llu,%llu,%llu)", (int)r, rank,
                  (unsigned long long)d[0], (unsigned long long)d[1], (unsigned long long)d[2]);
        return TB200_ECUDA;
    }
    slot.key = key; slot.map = *map; slot.used = true;
    return TB200_OK;
}

}  // namespace tb200

extern "C" {
int tb200_abi_version(void) { return TB200_ABI_VERSION; }
const char* tb200_last_error(void) { return tb200::t_err; }
uint64_t tb200_launch_count(void) { return tb200::g_launches.load(); }
}
