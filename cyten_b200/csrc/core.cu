// Library plumbing: error string, device info, stream-ordered scratch.
#include "common.cuh"
#include <mutex>

namespace cyb {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int sm_count()
{
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
            n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

static void configure_pool_once()
{
    static std::once_flag flags[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return;
    std::call_once(flags[dev], [dev]() {
        cudaMemPool_t pool;
        if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
            uint64_t thr = UINT64_MAX;  // keep freed scratch cached in the pool
            cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
        }
    });
}

Scratch::Scratch(size_t bytes, cudaStream_t s) : stream(s)
{
    configure_pool_once();
    if (bytes == 0) bytes = 16;
    err = cudaMallocAsync(&ptr, bytes, s);
    if (err != cudaSuccess) {
        ptr = nullptr;
        set_error("cudaMallocAsync(%zu) failed: %s", bytes, cudaGetErrorString(err));
    }
}

Scratch::~Scratch()
{
    if (ptr) cudaFreeAsync(ptr, stream);
}

}  // namespace cyb

extern "C" {

int cyb_version(void) { return CYB_VERSION; }

const char* cyb_last_error(void) { return cyb::g_err; }

int cyb_device_info(int* sm_count, int* cc_major, int* cc_minor)
{
    int dev = 0;
    CYB_CHECK_CUDA(cudaGetDevice(&dev));
    if (sm_count) CYB_CHECK_CUDA(cudaDeviceGetAttribute(sm_count, cudaDevAttrMultiProcessorCount, dev));
    if (cc_major) CYB_CHECK_CUDA(cudaDeviceGetAttribute(cc_major, cudaDevAttrComputeCapabilityMajor, dev));
    if (cc_minor) CYB_CHECK_CUDA(cudaDeviceGetAttribute(cc_minor, cudaDevAttrComputeCapabilityMinor, dev));
    return CYB_OK;
}

}  // extern "C"
