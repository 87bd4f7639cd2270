// bgr_peer_* — exchange step of the index-range-sharded kNN over NVLink peer
// memory, one process per GPU.
//
// Every rank owns one device buffer ("window") that its peers map with CUDA IPC:
//     [ flags | slab_idx (k x Q) | slab_d2 (k x Q) | out_idx (k x Q) | out_d2 (k x Q) ]
// A sharded search is then ONE kernel after the local search instead of all-gather + merge:
//   1. the local search writes this rank's top-k slab (global indices) into its own window;
//   2. peer_merge_kernel: signal "slab ready" to all peers, wait for theirs, then merge — rank r
//      owns the query slice [r*Q/G, (r+1)*Q/G): it LOADS that slice of every peer's slab over
//      NVLink (G-1 remote k-lists per query), keeps the k best by (d2, index) — the same rule as
//      knn_merge_kernel, so the result is independent of G — and STORES the merged list into
//      every peer's out region; the last CTA to finish signals "merged" to all peers;
//   3. peer_wait_kernel: wait until every peer's "merged" flag reached this epoch (all slices
//      of the local out region are written, and no peer still reads the local slab).
// Traffic per rank: (G-1)/G * 8kQ bytes in + the same out, against (G-1) * 8kQ bytes in for the
// all-gather form (8x less at G=8).  The NCCL all-gather + bgr_knn_merge_dev path stays as the
// reference point (dist.py, exchange="nccl").
//
// Flags are 32-bit epochs written with system-scope release stores and polled with acquire
// loads; spins are bounded (kSpinLimitNs) and report through an error word instead of hanging.
#include <cstring>
#include <vector>
#include "warp_topk.cuh"
#include "internal.cuh"

using namespace bgr;

namespace {

constexpr int kMaxPeers = 16;
constexpr long long kSpinLimitNs = 4000000000LL;   // 4 s
constexpr size_t kFlagBytes = 1024;

struct PeerFlags {                 // lives at the start of every window
    unsigned ready[kMaxPeers];     // ready[p]  : rank p's slab of epoch e is complete
    unsigned merged[kMaxPeers];    // merged[p] : rank p wrote its slice of epoch e into this window's out region
    unsigned error;                // != 0: a bounded spin expired
    unsigned blocks_done;          // last-CTA counter of the local merge kernel
};

struct PeerView {
    unsigned char* win[kMaxPeers];   // window base of every rank (own one included), this process' mapping
    int rank, world;
    int64_t cap;                     // capacity in entries (k*Q) of each region
};

__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ long long now_ns() {
    long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// epochs only grow; compare with wrap-around in mind
__device__ __forceinline__ bool reached(unsigned have, unsigned want) { return (int)(have - want) >= 0; }

__device__ __forceinline__ int32_t* region_i32(unsigned char* win, int64_t cap, int which) {
    return reinterpret_cast<int32_t*>(win + kFlagBytes) + (int64_t)which * cap;
}

// wait (one thread) until flag[p] of the LOCAL window reached epoch for every p; false on time-out
__device__ bool wait_all(const unsigned* flags, int world, unsigned epoch, unsigned* error) {
    const long long t0 = now_ns();
    for (int p = 0; p < world; ++p) {
        while (!reached(ld_acquire_sys(flags + p), epoch)) {
            if (now_ns() - t0 > kSpinLimitNs) {
                atomicExch(error, 1u + (unsigned)p);
                return false;
            }
            __nanosleep(200);
        }
    }
    return true;
}


// flat copy of n 32-bit words, 16 bytes per lane when both sides allow it (all threads of the CTA)
__device__ __forceinline__ void cta_copy_in(int32_t* __restrict__ dst, const int32_t* __restrict__ src, int n, bool vec) {
    if (vec) {
        const int4* s4 = reinterpret_cast<const int4*>(src);
        int4* d4 = reinterpret_cast<int4*>(dst);
        for (int e = threadIdx.x; e < (n >> 2); e += blockDim.x) d4[e] = __ldcg(s4 + e);
        for (int e = (n & ~3) + threadIdx.x; e < n; e += blockDim.x) dst[e] = __ldcg(src + e);
    } else {
        for (int e = threadIdx.x; e < n; e += blockDim.x) dst[e] = __ldcg(src + e);
    }
}
__device__ __forceinline__ void cta_copy_out(int32_t* __restrict__ dst, const int32_t* __restrict__ src, int n, bool vec) {
    if (vec) {
        const int4* s4 = reinterpret_cast<const int4*>(src);
        int4* d4 = reinterpret_cast<int4*>(dst);
        for (int e = threadIdx.x; e < (n >> 2); e += blockDim.x) d4[e] = s4[e];
        for (int e = (n & ~3) + threadIdx.x; e < n; e += blockDim.x) dst[e] = src[e];
    } else {
        for (int e = threadIdx.x; e < n; e += blockDim.x) dst[e] = src[e];
    }
}

// Persistent grid (all CTAs resident).  Per chunk of `chunk` queries of this rank's slice:
// stage the chunk of every slab in shared memory with wide loads — all of a thread's loads, to
// all peers, are in flight together, which is what hides the NVLink round trip — merge from
// shared memory (warp per query), then store the merged chunk to every peer with wide stores.
__global__ void __launch_bounds__(256) peer_merge_kernel(PeerView v, int64_t nq, int k, unsigned epoch, int chunk) {
    extern __shared__ __align__(16) int32_t stage[];   // [world][2][chunk*k]; chunk = queries staged per CTA iteration
    __shared__ int ok_s;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warps = blockDim.x >> 5;
    PeerFlags* mine = reinterpret_cast<PeerFlags*>(v.win[v.rank]);
    if (blockIdx.x == 0 && threadIdx.x < v.world) {
        // the local search finished before this kernel started (stream order): publish the slab
        __threadfence_system();
        st_release_sys(&reinterpret_cast<PeerFlags*>(v.win[threadIdx.x])->ready[v.rank], epoch);
    }
    if (threadIdx.x == 0) ok_s = wait_all(mine->ready, v.world, epoch, &mine->error) ? 1 : 0;
    __syncthreads();
    if (ok_s) {
        const int64_t q_lo = nq * v.rank / v.world, q_hi = nq * (v.rank + 1) / v.world;
        const int per = chunk * k;
        for (int64_t q0 = q_lo + (int64_t)blockIdx.x * chunk; q0 < q_hi; q0 += (int64_t)gridDim.x * chunk) {
            const int nqc = (int)min((int64_t)chunk, q_hi - q0), n = nqc * k;
            const int64_t o = q0 * k;
            const bool vec = (o & 3) == 0;
            for (int s = 0; s < v.world; ++s) {
                const int p = (v.rank + s) % v.world;               // slot 0 = home, remote reads staggered
                cta_copy_in(stage + (2 * s) * per, region_i32(v.win[p], v.cap, 0) + o, n, vec);
                cta_copy_in(stage + (2 * s + 1) * per, region_i32(v.win[p], v.cap, 1) + o, n, vec);
            }
            __syncthreads();
            for (int ql = warp; ql < nqc; ql += warps) {
                WarpList L;
                L.init(kKeyMax);
                L.i = kKeyMax;
                for (int s = 0; s < v.world; ++s) {
                    uint32_t bd = kKeyMax, bi = kKeyMax;
                    if (lane < k) {
                        const int32_t ii = stage[(2 * s) * per + ql * k + lane];
                        if (ii >= 0) {
                            bd = (uint32_t)stage[(2 * s + 1) * per + ql * k + lane];
                            bi = (uint32_t)ii;
                        }
                    }
                    L.merge_sorted(bd, bi, lane);
                }
                __syncwarp();
                if (lane < k) {                                     // the warp's own entries of slot 0
                    const bool empty = (L.d == kKeyMax);
                    stage[ql * k + lane] = empty ? -1 : (int32_t)L.i;
                    stage[per + ql * k + lane] = (int32_t)(empty ? kInfBits : L.d);
                }
            }
            __syncthreads();
            for (int s = 0; s < v.world; ++s) {
                const int p = (v.rank + s) % v.world;
                cta_copy_out(region_i32(v.win[p], v.cap, 2) + o, stage, n, vec);
                cta_copy_out(region_i32(v.win[p], v.cap, 3) + o, stage + per, n, vec);
            }
            __syncthreads();
        }
    }
    // last CTA: all slices this rank owes are written -> tell every peer
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned done = atomicAdd(&mine->blocks_done, 1u) + 1u;
        ok_s = (done == gridDim.x) ? 1 : 0;
    }
    __syncthreads();
    if (ok_s) {
        if (threadIdx.x == 0) mine->blocks_done = 0;
        if (threadIdx.x < v.world) {
            __threadfence_system();
            st_release_sys(&reinterpret_cast<PeerFlags*>(v.win[threadIdx.x])->merged[v.rank], epoch);
        }
    }
}

__global__ void peer_wait_kernel(PeerView v, unsigned epoch) {
    PeerFlags* mine = reinterpret_cast<PeerFlags*>(v.win[v.rank]);
    if (threadIdx.x == 0) wait_all(mine->merged, v.world, epoch, &mine->error);
}

}  // namespace

struct bgr_peer {
    int rank = 0, world = 1;
    int64_t cap = 0;               // entries per region
    size_t bytes = 0;
    unsigned char* local = nullptr;
    unsigned char* win[kMaxPeers] = {nullptr};
    bool opened[kMaxPeers] = {false};
    unsigned epoch = 0;
};

extern "C" {

int bgr_peer_handle_bytes(void) { return (int)sizeof(cudaIpcMemHandle_t); }

int bgr_peer_create(int rank, int world, int64_t max_entries, bgr_peer** out) {
    BGR_ARG(out != nullptr, "out is null");
    *out = nullptr;
    BGR_ARG(world >= 1 && world <= kMaxPeers && rank >= 0 && rank < world, "bad rank/world (%d/%d, max %d)", rank, world, kMaxPeers);
    BGR_ARG(max_entries > 0, "max_entries must be > 0");
    static_assert(sizeof(PeerFlags) <= kFlagBytes, "flag block too large");
    bgr_peer* h = new bgr_peer();
    h->rank = rank;
    h->world = world;
    h->cap = round_up(max_entries, 64);
    h->bytes = kFlagBytes + (size_t)4 * h->cap * sizeof(int32_t);
    cudaError_t e = cudaMalloc(&h->local, h->bytes);
    if (e != cudaSuccess) {
        set_error("cudaMalloc(%zu) -> %s", h->bytes, cudaGetErrorString(e));
        delete h;
        return BGR_ERR_ALLOC;
    }
    e = cudaMemset(h->local, 0, kFlagBytes);
    if (e != cudaSuccess) {
        set_error("cudaMemset -> %s", cudaGetErrorString(e));
        cudaFree(h->local);
        delete h;
        return BGR_ERR_CUDA;
    }
    h->win[rank] = h->local;
    *out = h;
    return BGR_OK;
}

int bgr_peer_export(bgr_peer* h, void* handle_out) {
    BGR_ARG(h != nullptr && handle_out != nullptr, "bad arguments");
    cudaIpcMemHandle_t mh;
    BGR_CUDA(cudaIpcGetMemHandle(&mh, h->local));
    std::memcpy(handle_out, &mh, sizeof(mh));
    return BGR_OK;
}

int bgr_peer_open(bgr_peer* h, int peer, const void* handle) {
    BGR_ARG(h != nullptr && handle != nullptr && peer >= 0 && peer < h->world, "bad arguments");
    if (peer == h->rank) return BGR_OK;
    BGR_ARG(!h->opened[peer], "peer %d already opened", peer);
    cudaIpcMemHandle_t mh;
    std::memcpy(&mh, handle, sizeof(mh));
    void* p = nullptr;
    BGR_CUDA(cudaIpcOpenMemHandle(&p, mh, cudaIpcMemLazyEnablePeerAccess));
    h->win[peer] = static_cast<unsigned char*>(p);
    h->opened[peer] = true;
    return BGR_OK;
}

int bgr_peer_slab(bgr_peer* h, int32_t** d_idx, float** d_d2) {
    BGR_ARG(h != nullptr, "null peer context");
    if (d_idx) *d_idx = reinterpret_cast<int32_t*>(h->local + kFlagBytes);
    if (d_d2) *d_d2 = reinterpret_cast<float*>(h->local + kFlagBytes) + h->cap;
    return BGR_OK;
}

int bgr_peer_result(bgr_peer* h, int32_t** d_idx, float** d_d2) {
    BGR_ARG(h != nullptr, "null peer context");
    if (d_idx) *d_idx = reinterpret_cast<int32_t*>(h->local + kFlagBytes) + 2 * h->cap;
    if (d_d2) *d_d2 = reinterpret_cast<float*>(h->local + kFlagBytes) + 3 * h->cap;
    return BGR_OK;
}

int bgr_peer_merge_dev(bgr_peer* h, int64_t nq, int k, void* stream) {
    BGR_ARG(h != nullptr && nq >= 0 && k >= 1 && k <= 32, "bad merge arguments (k<=32)");
    BGR_ARG(nq * k <= h->cap, "window too small: %lld entries > %lld", (long long)(nq * k), (long long)h->cap);
    for (int p = 0; p < h->world; ++p) BGR_ARG(h->win[p] != nullptr, "peer %d not opened", p);
    PeerView v;
    for (int p = 0; p < kMaxPeers; ++p) v.win[p] = h->win[p];
    v.rank = h->rank;
    v.world = h->world;
    v.cap = h->cap;
    const unsigned epoch = ++h->epoch;
    const int64_t slice = div_up(nq > 0 ? nq : 1, h->world);
    int chunk = 64;                                  // more bytes in flight per CTA hide the NVLink round trip
    while (chunk > 8 && (size_t)h->world * 2 * chunk * k * sizeof(int32_t) > 64 * 1024) chunk >>= 1;
    const size_t smem = (size_t)h->world * 2 * chunk * k * sizeof(int32_t);
    static size_t smem_set = 0;
    if (smem > 48 * 1024 && smem > smem_set) {
        BGR_CUDA(cudaFuncSetAttribute(peer_merge_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        smem_set = smem;
    }
    // persistent grid: resident CTAs only (every CTA polls the ready flags before it works)
    const int per_sm = (int)std::max<size_t>(1, std::min<size_t>(4, (200 * 1024) / std::max<size_t>(smem, 1)));
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((int64_t)per_sm * num_sms(), div_up(slice, chunk)));
    cudaStream_t st = (cudaStream_t)stream;
    peer_merge_kernel<<<grid, 256, smem, st>>>(v, nq, k, epoch, chunk);
    BGR_LAUNCH_CHECK();
    peer_wait_kernel<<<1, 32, 0, st>>>(v, epoch);
    BGR_LAUNCH_CHECK();
    return BGR_OK;
}

int bgr_peer_status(bgr_peer* h, void* stream, int* error) {
    BGR_ARG(h != nullptr && error != nullptr, "bad arguments");
    unsigned e = 0;
    BGR_CUDA(cudaMemcpyAsync(&e, h->local + offsetof(PeerFlags, error), sizeof(e), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    BGR_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    *error = (int)e;
    return BGR_OK;
}

int bgr_peer_destroy(bgr_peer* h) {
    if (!h) return BGR_OK;
    cudaDeviceSynchronize();
    for (int p = 0; p < h->world; ++p)
        if (h->opened[p]) cudaIpcCloseMemHandle(h->win[p]);
    if (h->local) cudaFree(h->local);
    delete h;
    return BGR_OK;
}

}  // extern "C"
