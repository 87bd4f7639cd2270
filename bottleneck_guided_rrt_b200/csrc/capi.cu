// Error plumbing and device queries of the C ABI (include/bgrrt.h).
#include <cstdarg>
#include <cstdio>
#include <atomic>
#include "common.cuh"

namespace bgr {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
const char* get_error() { return g_err; }

static std::atomic<unsigned long long> g_launches{0};
void note_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
unsigned long long launches() { return g_launches.load(std::memory_order_relaxed); }

int num_sms() {
    static int sms = 0;
    if (sms == 0) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess ||
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0)
            sms = kNumSMsB200;
    }
    return sms;
}

}  // namespace bgr

extern "C" {

const char* bgr_last_error(void) { return bgr::get_error(); }
int bgr_version(void) { return BGR_VERSION; }
uint64_t bgr_launch_count(void) { return bgr::launches(); }

int bgr_device_count(int* count) {
    BGR_ARG(count != nullptr, "count is null");
    *count = 0;
    BGR_CUDA(cudaGetDeviceCount(count));
    return BGR_OK;
}

int bgr_device_info(int* sms, int* sm_clock_khz, int64_t* l2_bytes) {
    int dev = 0, v = 0;
    BGR_CUDA(cudaGetDevice(&dev));
    if (sms) {
        BGR_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev));
        *sms = v;
    }
    if (sm_clock_khz) {
        BGR_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrClockRate, dev));
        *sm_clock_khz = v;
    }
    if (l2_bytes) {
        BGR_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrL2CacheSize, dev));
        *l2_bytes = v;
    }
    return BGR_OK;
}

}  // extern "C"
