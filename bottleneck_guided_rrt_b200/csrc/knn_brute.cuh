// Exact brute-force kNN scan (reference surface: nns->knn, MS:1309-1312; libnabo leaf
// arithmetic).
//
// Layout: node coordinates are SoA float arrays x[],y[],z[] in HBM, padded with NaN to a
// multiple of kTile, plus a float4 "filter image" {-2x,-2y,-2z,|p|^2''} of the same store.
// One CTA = 8 consumer warps + 1 producer warp.  The producer streams filter tiles
// HBM/L2 -> shared memory with 1-D bulk copies (TMA engine, cp.async.bulk -> SASS UBLKCP)
// through a 3-stage mbarrier ring; every consumer warp owns R queries, held as R/2 packed
// f32x2 register pairs, so one conflict-free LDS.128 of a node serves R evaluations.
// Two-level evaluation keeps the result bit-exact while using the FMA pipe at full rate:
//   filter : v = fl(|p|^2'' - 2 q.p)  — 3 FFMA2 per query PAIR (sm_100a packed FP32: two
//            element-ops per issue slot), compared against a conservative threshold T(tau);
//            a node with libnabo d2 <= tau can never be filtered out (filter_threshold()).
//   exact  : survivors only (a few per thousand pairs): d2 = ((dx*dx)+dy*dy)+dz*dz with
//            __fsub_rn/__fmul_rn/__fadd_rn on the original coordinates (never contracted),
//            admission and (d2,index) order exactly as before.
// The top-k of each query is a lane-distributed sorted list (warp_topk.cuh).
#pragma once
#include <cfloat>
#include "warp_topk.cuh"

namespace bgr {

constexpr int kTile = 1024;          // nodes per stage
constexpr int kStages = 3;
constexpr int kConsumerWarps = 8;
constexpr int kBruteThreads = (kConsumerWarps + 1) * 32;

struct BruteArgs {
    const float* x;
    const float* y;
    const float* z;
    const float4* f4;     // filter image of the store: {-2x, -2y, -2z, |p|^2 shrunk by 16 ulp}
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
    const long long* n_dev;   // optional: the store's size lives on the device (CUDA-graph replays); n = its upper bound
};

__device__ __forceinline__ uint32_t radius_sentinel(float mr) {
    // admission d2 <= mr*mr (libnabo maxRadius2) and d2 < +inf  =>  d2 <= min(mr2, FLT_MAX);
    // sentinel key (bits(lim)+1, 0) is greater than every admissible key and no other.
    float mr2 = __fmul_rn(mr, mr);
    if (!(mr2 <= FLT_MAX)) mr2 = FLT_MAX;   // also maps NaN to "no limit"
    return __float_as_uint(mr2) + 1u;
}

// ---- packed FP32 (sm_100a FFMA2: two element-ops per issue slot)
__device__ __forceinline__ unsigned long long pack2(float lo, float hi) {
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(unsigned long long v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
    unsigned long long r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}

// Conservative filter threshold for a query with squared norm qn (double) and current
// admission bound tau (f32): every node whose libnabo-arithmetic d2 is <= tau satisfies
//     fl(|p|^2'' - 2 q.p)  <=  T = up( tau (1 + 8u) - qn + 4u qn ),   u = 2^-24
// (3-FMA chain error <= 3.01u (qn + 2|p|^2), absorbed by the 16-ulp shrink of |p|^2'' and the
// 4u qn term; f32 d2 vs real d2 <= 5u relative; derivation in DESIGN.md §3.1).
__device__ __forceinline__ float filter_threshold(float tau, double qn) {
    const double u = 5.9604644775390625e-8;
    if (!(tau <= FLT_MAX)) return tau;                       // +inf sentinel (or NaN): pass / reject everything
    return __double2float_ru((double)tau * (1.0 + 8.0 * u) - qn + 4.0 * u * qn);
}

// Tiles are visited in a scrambled order (t * stride mod ntiles, stride coprime to ntiles, near
// the golden ratio).  An append-only tree stores its nodes in growth order: for a query ahead of
// the growth front nearly every later node is closer than every earlier one, and an in-order
// scan would update its running bound (and run the exact path) on almost every node — measured
// on a 150k-node planner tree: a third of all lanes took the slow path.  Results do not depend
// on the order (keys are (d2, index)).
__device__ __forceinline__ int tile_stride(int ntiles) {
    if (ntiles < 3) return 1;
    for (int s = max(1, (int)(0.6180339887f * (float)ntiles));; ++s) {
        int a = s, b = ntiles;
        while (b) { const int c = a % b; a = b; b = c; }
        if (a == 1) return s;
    }
}
__device__ __forceinline__ int tile_perm(int t, int stride, int ntiles) {
    return (int)(((long long)t * stride) % ntiles);
}

template <int R>
__global__ void __launch_bounds__(kBruteThreads, 2) knn_brute_kernel(BruteArgs a) {
    static_assert(R % 2 == 0 || R == 1, "queries are processed in packed pairs");
    constexpr int P = (R + 1) / 2;                            // packed pairs per warp
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float4* tiles = reinterpret_cast<float4*>(smem_raw);                     // [stage][kTile]
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(tiles + kStages * kTile);
    uint64_t* empty_bar = full_bar + kStages;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t chunk_begin = (int64_t)blockIdx.y * a.chunk_nodes;
    const int64_t n_now = a.n_dev ? (int64_t)*a.n_dev : a.n;
    const int64_t n_pad = (n_now + kTile - 1) / kTile * kTile;
    const int64_t chunk_end = min(chunk_begin + a.chunk_nodes, n_pad);
    const int ntiles = (int)((chunk_end - chunk_begin) / kTile);
    const int tstride = tile_stride(ntiles);

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
                const int64_t off = chunk_begin + (int64_t)tile_perm(t, tstride, ntiles) * kTile;
                mbar_expect_tx(&full_bar[s], kTile * sizeof(float4));
                bulk_g2s(tiles + s * kTile, a.f4 + off, kTile * sizeof(float4), &full_bar[s]);
            }
        }
        return;
    }

    // ---------------------------------------------------- consumer warps
    const int64_t q0 = ((int64_t)blockIdx.x * kConsumerWarps + warp) * R;
    const bool self = (a.flags & BGR_ALLOW_SELF_MATCH) != 0;
    float thr[2 * P];
    unsigned long long qx2[P], qy2[P], qz2[P];
    WarpList L[2 * P];
    auto sentinel_of = [&](int r) -> uint32_t {
        const int64_t qi = q0 + r;
        if (r < R && qi < a.nq) return radius_sentinel(a.max_radii ? __ldg(a.max_radii + qi) : a.max_radius);
        return 0u;   // nothing is admissible
    };
#pragma unroll
    for (int p = 0; p < P; ++p) {
        float c[2][3];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int r = 2 * p + h;
            const int64_t qi = q0 + r;
            const bool live = r < R && qi < a.nq;
            c[h][0] = live ? __ldg(a.q + 3 * qi) : 0.f;
            c[h][1] = live ? __ldg(a.q + 3 * qi + 1) : 0.f;
            c[h][2] = live ? __ldg(a.q + 3 * qi + 2) : 0.f;
            const uint32_t sent = sentinel_of(r);
            L[r].init(sent);
            const double qn = (double)c[h][0] * (double)c[h][0] + (double)c[h][1] * (double)c[h][1] + (double)c[h][2] * (double)c[h][2];
            thr[r] = sent ? filter_threshold(__uint_as_float(sent), qn) : -__uint_as_float(kInfBits);
        }
        qx2[p] = pack2(c[0][0], c[1][0]);
        qy2[p] = pack2(c[0][1], c[1][1]);
        qz2[p] = pack2(c[0][2], c[1][2]);
    }

    for (int t = 0; t < ntiles; ++t) {
        const int s = t % kStages;
        mbar_wait(&full_bar[s], (t / kStages) & 1);
        const float4* tile = tiles + s * kTile;
        const int64_t tile_base = chunk_begin + (int64_t)tile_perm(t, tstride, ntiles) * kTile;
#pragma unroll 2
        for (int j = lane; j < kTile; j += 32) {
            const float4 f = tile[j];                         // {-2x, -2y, -2z, |p|^2''}
            const unsigned long long ax = pack2(f.x, f.x), ay = pack2(f.y, f.y), az = pack2(f.z, f.z),
                                     pn = pack2(f.w, f.w);
            float v[2 * P];
            bool any = false;
#pragma unroll
            for (int p = 0; p < P; ++p) {
                unsigned long long acc = fma2(qx2[p], ax, pn);
                acc = fma2(qy2[p], ay, acc);
                acc = fma2(qz2[p], az, acc);
                unpack2(acc, v[2 * p], v[2 * p + 1]);
                any |= (v[2 * p] <= thr[2 * p]);
                any |= (v[2 * p + 1] <= thr[2 * p + 1]);
            }
            if (__any_sync(kFull, any)) {
                // ---- exact pass for the survivors: libnabo's arithmetic on the original
                // coordinates, recovered exactly from the filter image (x = -0.5 * (-2x))
                const float px = -0.5f * f.x, py = -0.5f * f.y, pz = -0.5f * f.z;
                const uint32_t ci = (uint32_t)(tile_base + j);
                unsigned mine = 0;
#pragma unroll
                for (int r = 0; r < 2 * P; ++r) mine |= (v[r] <= thr[r]) ? (1u << r) : 0u;
                const unsigned active = __reduce_or_sync(kFull, mine);   // queries with a survivor in this warp
#pragma unroll
                for (int r = 0; r < 2 * P; ++r) {
                    if ((active >> r) & 1u) {
                        float qa[2], qb[2], qc[2];
                        unpack2(qx2[r >> 1], qa[0], qa[1]);
                        unpack2(qy2[r >> 1], qb[0], qb[1]);
                        unpack2(qz2[r >> 1], qc[0], qc[1]);
                        const float dx = __fsub_rn(qa[r & 1], px), dy = __fsub_rn(qb[r & 1], py), dz = __fsub_rn(qc[r & 1], pz);
                        const float d = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
                        // key of entry k-1 = the admission bound (kept in the list, not in registers)
                        uint32_t tau_d = __shfl_sync(kFull, L[r].d, a.k - 1), tau_i = __shfl_sync(kFull, L[r].i, a.k - 1);
                        const bool pass = ((mine >> r) & 1u) && (d <= __uint_as_float(tau_d)) && (self || d > FLT_EPSILON);
                        const uint32_t before = tau_d;
                        warp_offer(L[r], pass, __float_as_uint(d), ci, tau_d, tau_i, a.k, lane);
                        if (tau_d != before) {
                            const double qn = (double)qa[r & 1] * (double)qa[r & 1] + (double)qb[r & 1] * (double)qb[r & 1] +
                                              (double)qc[r & 1] * (double)qc[r & 1];
                            thr[r] = filter_threshold(__uint_as_float(tau_d), qn);
                        }
                    }
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
            const bool empty = (L[r].d == sentinel_of(r));
            const int64_t o = ((int64_t)blockIdx.y * a.nq + qi) * a.k + lane;
            a.idx[o] = empty ? -1 : (int32_t)((int64_t)L[r].i + a.index_base);
            a.d2[o] = empty ? __uint_as_float(kInfBits) : __uint_as_float(L[r].d);
        }
    }
}

// Merge G slabs (each k x Q ascending, padded with idx=-1,d2=+inf; slab s starts at entry
// s*slab_stride of idx_in / d2_in) into one: warp per query.
__global__ void __launch_bounds__(256) knn_merge_kernel(const int32_t* __restrict__ idx_in,
                                                        const float* __restrict__ d2_in, int g,
                                                        int64_t slab_stride, int64_t nq, int k,
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
            const int64_t o = (int64_t)s * slab_stride + qi * k + lane;
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
