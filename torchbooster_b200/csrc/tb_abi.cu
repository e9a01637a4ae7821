// Version, error reporting, launch accounting and cached device attributes of the C ABI.
#include <stdarg.h>
#include <atomic>

#include "tb_common.cuh"

static thread_local char g_tb_error[512] = "";
static std::atomic<int64_t> g_tb_launches{0};

int tb_fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_tb_error, sizeof(g_tb_error), fmt, ap);
    va_end(ap);
    return code;
}

int tb_cuda_fail(cudaError_t err, const char* what) {
    snprintf(g_tb_error, sizeof(g_tb_error), "CUDA error %d (%s) in %s", (int)err, cudaGetErrorString(err), what);
    return (int)err;
}

void tb_count_launch(int n) { g_tb_launches.fetch_add(n, std::memory_order_relaxed); }

int tb_device(TbDevice* out) {
    // Attributes never change for a device; a benign race on first use writes identical values.
    static TbDevice cache[64];
    static std::atomic<bool> have[64];
    int dev = 0;
    TB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return tb_fail(TB_EINVAL, "device ordinal %d out of range", dev);
    if (!have[dev].load(std::memory_order_acquire)) {
        TbDevice d;
        d.device = dev;
        TB_CUDA(cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, dev));
        TB_CUDA(cudaDeviceGetAttribute(&d.max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
        cache[dev] = d;
        have[dev].store(true, std::memory_order_release);
    }
    *out = cache[dev];
    return TB_OK;
}

extern "C" int tb_version(void) { return TB_ABI_VERSION; }
extern "C" const char* tb_last_error(void) { return g_tb_error; }
extern "C" int64_t tb_launch_count(void) { return g_tb_launches.load(std::memory_order_relaxed); }
