// Exact brute-force kNN scan (reference surface: nns->knn, MS:1309-1312; libnabo leaf
// arithmetic).
//
// Layout: node coordinates are SoA float arrays x[],y[],z[] in HBM, padded with NaN to a
// multiple of kTile.  One CTA = 8 consumer warps + 1 producer warp.  The producer streams
// node tiles HBM/L2 -> shared memory with 1-D bulk copies (TMA engine, cp.async.bulk ->
// SASS UBLKCP) through a 3-stage mbarrier ring; every consumer warp owns R queries whose
// coordinates and thresholds live in registers, so one conflict-free LDS of a node serves
// R distance evaluations.  d2 = ((dx*dx)+dy*dy)+dz*dz with __fmul_rn/__fadd_rn (never
// contracted to FMA): 8 FP32 instructions per (query,node) pair — the FP32-pipe roofline.
// The top-k of each query is a lane-distributed sorted list (warp_topk.cuh); the common
// case per 32 nodes is R compares OR-ed into one predicate and one warp vote.
#pragma once
#include <cfloat>
#include "warp_topk.cuh"

namespace bgr {

constexpr int kTile = 1024;          // nodes per stage: 3 arrays * 4 KB
constexpr int kStages = 3;
constexpr int kConsumerWarps = 8;
constexpr int kBruteThreads = (kConsumerWarps + 1) * 32;

struct BruteArgs {
    const float* x;
    const float* y;
    const float* z;
    int64_t n;            // valid nodes (arrays are NaN padded to a multiple of kTile)
    const float* q;       // 3 x Q interleaved
    int64_t nq;
    int k;
    unsigned flags;
    float max_radius;
    const float* max_radii;   // nullable
    int64_t chunk_nodes;  // nodes per chunk (multiple of kTile)
    int32_t* idx;         // [chunks][Q][k]
    float* d2;
    int64_t index_base;
};

__device__ __forceinline__ uint32_t radius_sentinel(float mr) {
    // admission d2 <= mr*mr (libnabo maxRadius2) and d2 < +inf  =>  d2 <= min(mr2, FLT_MAX);
    // sentinel key (bits(lim)+1, 0) is greater than every admissible key and no other.
    float mr2 = __fmul_rn(mr, mr);
    if (!(mr2 <= FLT_MAX)) mr2 = FLT_MAX;   // also maps NaN to "no limit"
    return __float_as_uint(mr2) + 1u;
}

template <int R>
__global__ void __launch_bounds__(kBruteThreads) knn_brute_kernel(BruteArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float* tiles = reinterpret_cast<float*>(smem_raw);                       // [stage][3][kTile]
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(tiles + kStages * 3 * kTile);
    uint64_t* empty_bar = full_bar + kStages;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t chunk_begin = (int64_t)blockIdx.y * a.chunk_nodes;
    const int64_t n_pad = (a.n + kTile - 1) / kTile * kTile;
    const int64_t chunk_end = min(chunk_begin + a.chunk_nodes, n_pad);
    const int ntiles = (int)((chunk_end - chunk_begin) / kTile);

    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], kConsumerWarps);
        }
        mbar_fence_init();
    }
    __syncthreads();

    if (warp == kConsumerWarps) {
        // ------------------------------------------------ producer warp (one elected lane)
        if (lane == 0) {
            for (int t = 0; t < ntiles; ++t) {
                const int s = t % kStages;
                if (t >= kStages) mbar_wait(&empty_bar[s], ((t / kStages) - 1) & 1);
                float* dst = tiles + s * 3 * kTile;
                const int64_t off = chunk_begin + (int64_t)t * kTile;
                mbar_expect_tx(&full_bar[s], 3 * kTile * sizeof(float));
                bulk_g2s(dst, a.x + off, kTile * sizeof(float), &full_bar[s]);
                bulk_g2s(dst + kTile, a.y + off, kTile * sizeof(float), &full_bar[s]);
                bulk_g2s(dst + 2 * kTile, a.z + off, kTile * sizeof(float), &full_bar[s]);
            }
        }
        return;
    }

    // ---------------------------------------------------- consumer warps
    const int64_t q0 = ((int64_t)blockIdx.x * kConsumerWarps + warp) * R;
    const bool self = (a.flags & BGR_ALLOW_SELF_MATCH) != 0;
    float qx[R], qy[R], qz[R], tau_f[R];
    uint32_t tau_d[R], tau_i[R], sent[R];
    WarpList L[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int64_t qi = q0 + r;
        if (qi < a.nq) {
            qx[r] = __ldg(a.q + 3 * qi);
            qy[r] = __ldg(a.q + 3 * qi + 1);
            qz[r] = __ldg(a.q + 3 * qi + 2);
            sent[r] = radius_sentinel(a.max_radii ? __ldg(a.max_radii + qi) : a.max_radius);
        } else {
            qx[r] = qy[r] = qz[r] = 0.f;
            sent[r] = 0u;   // nothing is admissible
        }
        L[r].init(sent[r]);
        tau_d[r] = sent[r];
        tau_i[r] = 0u;
        tau_f[r] = __uint_as_float(sent[r]);
    }

    for (int t = 0; t < ntiles; ++t) {
        const int s = t % kStages;
        mbar_wait(&full_bar[s], (t / kStages) & 1);
        const float* sx = tiles + s * 3 * kTile;
        const float* sy = sx + kTile;
        const float* sz = sy + kTile;
        const int64_t tile_base = chunk_begin + (int64_t)t * kTile;
#pragma unroll 2
        for (int j = lane; j < kTile; j += 32) {
            const float px = sx[j], py = sy[j], pz = sz[j];
            float d[R];
            bool any = false;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const float dx = __fsub_rn(qx[r], px), dy = __fsub_rn(qy[r], py),
                            dz = __fsub_rn(qz[r], pz);
                d[r] = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
                any |= (d[r] <= tau_f[r]);
            }
            if (__any_sync(kFull, any)) {
                const uint32_t ci = (uint32_t)(tile_base + j);
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const bool pass = (d[r] <= tau_f[r]) && (self || d[r] > FLT_EPSILON);
                    warp_offer(L[r], pass, __float_as_uint(d[r]), ci, tau_d[r], tau_i[r], a.k, lane);
                    tau_f[r] = __uint_as_float(tau_d[r]);
                }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[s]);
    }

    // ---------------------------------------------------- write this chunk's slab
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int64_t qi = q0 + r;
        if (qi < a.nq && lane < a.k) {
            const bool empty = (L[r].d == sent[r]);
            const int64_t o = ((int64_t)blockIdx.y * a.nq + qi) * a.k + lane;
            a.idx[o] = empty ? -1 : (int32_t)((int64_t)L[r].i + a.index_base);
            a.d2[o] = empty ? __uint_as_float(kInfBits) : __uint_as_float(L[r].d);
        }
    }
}

// Merge G slabs (each k x Q ascending, padded with idx=-1,d2=+inf) into one: warp per query.
__global__ void __launch_bounds__(256) knn_merge_kernel(const int32_t* __restrict__ idx_in,
                                                        const float* __restrict__ d2_in, int g,
                                                        int64_t nq, int k,
                                                        int32_t* __restrict__ idx_out,
                                                        float* __restrict__ d2_out) {
    const int lane = threadIdx.x & 31;
    const int64_t qi = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (qi >= nq) return;
    WarpList L;
    L.init(kKeyMax);
    L.i = kKeyMax;
    for (int s = 0; s < g; ++s) {
        uint32_t bd = kKeyMax, bi = kKeyMax;
        if (lane < k) {
            const int64_t o = ((int64_t)s * nq + qi) * k + lane;
            const int32_t ii = idx_in[o];
            if (ii >= 0) {
                bd = __float_as_uint(d2_in[o]);
                bi = (uint32_t)ii;
            }
        }
        L.merge_sorted(bd, bi, lane);
    }
    if (lane < k) {
        const bool empty = (L.d == kKeyMax);
        idx_out[qi * k + lane] = empty ? -1 : (int32_t)L.i;
        d2_out[qi * k + lane] = empty ? __uint_as_float(kInfBits) : __uint_as_float(L.d);
    }
}

}  // namespace bgr
