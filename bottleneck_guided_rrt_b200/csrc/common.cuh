// Shared host/device helpers for libbgrrt (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <cstdio>
#include <cstdarg>
#include <cstdlib>
#include <string>
#include "../../include/bgrrt.h"

namespace bgr {

void set_error(const char* fmt, ...);
const char* get_error();
void note_launch();

#define BGR_CUDA(call)                                                                   \
    do {                                                                                 \
        cudaError_t e__ = (call);                                                        \
        if (e__ != cudaSuccess) {                                                        \
            bgr::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call,                 \
                           cudaGetErrorString(e__));                                     \
            return BGR_ERR_CUDA;                                                         \
        }                                                                                \
    } while (0)

// after every kernel launch: count it (bench.py's gpu_launches) and surface launch errors
#define BGR_LAUNCH_CHECK()                                                               \
    do {                                                                                 \
        bgr::note_launch();                                                              \
        BGR_CUDA(cudaGetLastError());                                                    \
    } while (0)

#define BGR_CHECK(rc_expr)                                                               \
    do {                                                                                 \
        int rc__ = (rc_expr);                                                            \
        if (rc__ != BGR_OK) return rc__;                                                 \
    } while (0)

#define BGR_ARG(cond, ...)                                                               \
    do {                                                                                 \
        if (!(cond)) {                                                                   \
            bgr::set_error(__VA_ARGS__);                                                 \
            return BGR_ERR_ARG;                                                          \
        }                                                                                \
    } while (0)

constexpr int kNumSMsB200 = 148;
constexpr unsigned kFull = 0xffffffffu;

inline int64_t round_up(int64_t a, int64_t b) { return (a + b - 1) / b * b; }
inline int64_t div_up(int64_t a, int64_t b) { return (a + b - 1) / b; }

int num_sms();  // of the current device (cached)

// a handle's memory and kernels live on the device that was current at its creation: calls made with
// another device current are refused (BGR_ERR_STATE) instead of failing inside a launch
inline int check_device(int handle_device) {
    int cur = -1;
    BGR_CUDA(cudaGetDevice(&cur));
    if (cur != handle_device) {
        set_error("handle belongs to CUDA device %d but device %d is current (cudaSetDevice first)", handle_device, cur);
        return BGR_ERR_STATE;
    }
    return BGR_OK;
}

// Filter image of a node for the brute-force kNN scan (knn_brute.cuh): {-2x,-2y,-2z,|p|^2''},
// |p|^2'' = f32 sum of squares shrunk by 16 ulp and rounded down — it absorbs the rounding of
// this sum and of the 3-FMA filter chain so the filter can never reject a true neighbour.
// Every writer of the node store (append, planner commit) must write it with the node.
#ifdef __CUDACC__
__device__ __forceinline__ float4 filter_image(float px, float py, float pz) {
    const float pn = __fadd_rn(__fadd_rn(__fmul_rn(px, px), __fmul_rn(py, py)), __fmul_rn(pz, pz));
    return make_float4(-2.f * px, -2.f * py, -2.f * pz, __fmul_rd(pn, 0.99999904632568359375f));
}
#endif

// growable device buffer (append-only stores double their capacity)
template <typename T>
struct DevBuf {
    T* p = nullptr;
    int64_t cap = 0;
    int reserve(int64_t n, cudaStream_t st, bool keep, int64_t keep_n = 0) {
        if (n <= cap) return BGR_OK;
        int64_t nc = cap ? cap : 1024;
        while (nc < n) nc *= 2;
        T* np = nullptr;
        BGR_CUDA(cudaMalloc(&np, (size_t)nc * sizeof(T)));
        if (keep && p && keep_n > 0)
            BGR_CUDA(cudaMemcpyAsync(np, p, (size_t)keep_n * sizeof(T), cudaMemcpyDeviceToDevice, st));
        if (p) {
            BGR_CUDA(cudaStreamSynchronize(st));
            BGR_CUDA(cudaFree(p));
        }
        p = np;
        cap = nc;
        return BGR_OK;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

// Side streams + events of the host-pointer entry points: a batch is cut into chunks so that
// the H2D copy of chunk c+1 and the D2H copy of chunk c-1 overlap the kernels of chunk c
// (the caller's buffers must be pinned for the copies to be asynchronous; pageable memory is
// still correct, just serialised by the driver).
struct CopyPipe {
    static constexpr int kMaxChunks = 8;
    cudaStream_t s_in = nullptr, s_out = nullptr;
    cudaEvent_t ev_in[kMaxChunks] = {}, ev_out[kMaxChunks] = {}, ev_start = nullptr;
    int init() {
        if (s_in) return BGR_OK;
        BGR_CUDA(cudaStreamCreateWithFlags(&s_in, cudaStreamNonBlocking));
        BGR_CUDA(cudaStreamCreateWithFlags(&s_out, cudaStreamNonBlocking));
        BGR_CUDA(cudaEventCreateWithFlags(&ev_start, cudaEventDisableTiming));
        for (int i = 0; i < kMaxChunks; ++i) {
            BGR_CUDA(cudaEventCreateWithFlags(&ev_in[i], cudaEventDisableTiming));
            BGR_CUDA(cudaEventCreateWithFlags(&ev_out[i], cudaEventDisableTiming));
        }
        return BGR_OK;
    }
    void release() {
        if (!s_in) return;
        cudaStreamDestroy(s_in);
        cudaStreamDestroy(s_out);
        cudaEventDestroy(ev_start);
        for (int i = 0; i < kMaxChunks; ++i) {
            cudaEventDestroy(ev_in[i]);
            cudaEventDestroy(ev_out[i]);
        }
        s_in = s_out = nullptr;
    }
};

// measurement aid: BGR_NO_PIPE=1 in the environment turns the chunked overlap off (read per call)
inline bool pipe_disabled() {
    const char* e = getenv("BGR_NO_PIPE");
    return e && e[0] == '1';
}

// pinned host staging buffer
struct PinBuf {
    void* p = nullptr;
    size_t cap = 0;
    int reserve(size_t bytes) {
        if (bytes <= cap) return BGR_OK;
        if (p) cudaFreeHost(p);
        p = nullptr;
        cap = 0;
        size_t nc = 1 << 16;
        while (nc < bytes) nc *= 2;
        BGR_CUDA(cudaMallocHost(&p, nc));
        cap = nc;
        return BGR_OK;
    }
    void release() {
        if (p) cudaFreeHost(p);
        p = nullptr;
        cap = 0;
    }
};

}  // namespace bgr
