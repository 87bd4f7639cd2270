// Batched Nearest() — reference MS:138-154 with Norm() MS:130-134:
//   Norm = float( pow( pow(dx,2)+pow(dy,2)+pow(dz,2), 0.5 ) ),  dx = x2-x1 in float,
//   squares / sum / root in double;  `if (temp <= min)` => LAST index among equal floats.
// Same tile pipeline as knn_brute.cuh.  Per (query,node): an f32 FILTER first — 3 FSUB, FMUL,
// 2 FFMA, one FSETP against Tf — and only for its survivors the reference's arithmetic:
// 3 F2D, 3 DMUL, 2 DADD, one DSETP against T = nextup(best)^2 (a superset test:
// float(sqrt(s)) <= best implies s < nextup(best)^2) and the DSQRT for candidates that pass.
// The filter cannot drop a candidate: the f32 sum of squares of the same float differences is
// within 2e-7 relative of the double one, and Tf = T * (1 + 1e-6) rounded up.
#pragma once
#include "knn_brute.cuh"

namespace bgr {

struct NearestArgs {
    const float* x;
    const float* y;
    const float* z;
    int64_t n;
    const float* q;
    int64_t nq;
    int64_t chunk_nodes;
    int32_t* idx;    // [chunks][Q]
    float* dist;     // [chunks][Q]
    int64_t index_base;
    const long long* nq_dev;   // optional: only the first nq_base + *nq_dev queries are valid (<= nq)
    int64_t nq_base;
    const long long* n_dev;    // optional: the store's size lives on the device (CUDA-graph replays); n = its upper bound
};

__device__ __forceinline__ int64_t valid_queries(const long long* nq_dev, int64_t nq_base, int64_t nq) {
    return nq_dev ? min(nq, nq_base + (int64_t)__ldg(nq_dev)) : nq;
}

__device__ __forceinline__ double nextup_sq(float f) {
    const float nu = __uint_as_float(__float_as_uint(f) + 1u);  // f >= 0, finite or inf
    const double d = (double)nu;
    return (f < __uint_as_float(kInfBits)) ? d * d : __longlong_as_double(0x7ff0000000000000LL);
}

template <int R>
__global__ void __launch_bounds__(kBruteThreads) nearest_kernel(NearestArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    // query count known only on the device (planner: distinct sample points of the batch)
    if ((int64_t)blockIdx.x * kConsumerWarps * R >= valid_queries(a.nq_dev, a.nq_base, a.nq)) return;
    float* tiles = reinterpret_cast<float*>(smem_raw);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(tiles + kStages * 3 * kTile);
    uint64_t* empty_bar = full_bar + kStages;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t chunk_begin = (int64_t)blockIdx.y * a.chunk_nodes;
    const int64_t n_now = a.n_dev ? (int64_t)*a.n_dev : a.n;
    const int64_t n_pad = (n_now + kTile - 1) / kTile * kTile;
    const int64_t chunk_end = min(chunk_begin + a.chunk_nodes, n_pad);
    const int ntiles = (int)((chunk_end - chunk_begin) / kTile);
    const int tstride = tile_stride(ntiles);   // scrambled tile order: see knn_brute.cuh

    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], kConsumerWarps);
        }
        mbar_fence_init();
    }
    __syncthreads();
    if (warp == kConsumerWarps) {
        if (lane == 0) {
            for (int t = 0; t < ntiles; ++t) {
                const int s = t % kStages;
                if (t >= kStages) mbar_wait(&empty_bar[s], ((t / kStages) - 1) & 1);
                float* dst = tiles + s * 3 * kTile;
                const int64_t off = chunk_begin + (int64_t)tile_perm(t, tstride, ntiles) * kTile;
                mbar_expect_tx(&full_bar[s], 3 * kTile * sizeof(float));
                bulk_g2s(dst, a.x + off, kTile * sizeof(float), &full_bar[s]);
                bulk_g2s(dst + kTile, a.y + off, kTile * sizeof(float), &full_bar[s]);
                bulk_g2s(dst + 2 * kTile, a.z + off, kTile * sizeof(float), &full_bar[s]);
            }
        }
        return;
    }

    const int64_t q0 = ((int64_t)blockIdx.x * kConsumerWarps + warp) * R;
    float qx[R], qy[R], qz[R], best_f[R], Tf[R];
    int32_t best_i[R];
    double T[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int64_t qi = min(q0 + r, a.nq - 1);
        qx[r] = __ldg(a.q + 3 * qi);
        qy[r] = __ldg(a.q + 3 * qi + 1);
        qz[r] = __ldg(a.q + 3 * qi + 2);
        best_f[r] = __uint_as_float(kInfBits);
        best_i[r] = -1;
        T[r] = __longlong_as_double(0x7ff0000000000000LL);
        Tf[r] = __uint_as_float(kInfBits);
    }
    for (int t = 0; t < ntiles; ++t) {
        const int s = t % kStages;
        mbar_wait(&full_bar[s], (t / kStages) & 1);
        const float* sx = tiles + s * 3 * kTile;
        const float* sy = sx + kTile;
        const float* sz = sy + kTile;
        const int64_t tile_base = chunk_begin + (int64_t)tile_perm(t, tstride, ntiles) * kTile;
#pragma unroll 2
        for (int j = lane; j < kTile; j += 32) {
            const float px = sx[j], py = sy[j], pz = sz[j];
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const float fx = __fsub_rn(qx[r], px), fy = __fsub_rn(qy[r], py), fz = __fsub_rn(qz[r], pz);
                if (fmaf(fz, fz, fmaf(fy, fy, fx * fx)) <= Tf[r]) {   // f32 filter; NaN padding never passes
                    const double dx = (double)fx, dy = (double)fy, dz = (double)fz;
                    const double sq = __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
                    if (sq <= T[r]) {
                        const float f = __double2float_rn(__dsqrt_rn(sq));
                        const int32_t ci = (int32_t)(tile_base + j);
                        if (f < best_f[r] || (f == best_f[r] && ci > best_i[r])) {   // equal Norm: the last index wins
                            best_f[r] = f;
                            best_i[r] = ci;
                            T[r] = nextup_sq(f);
                            Tf[r] = __double2float_ru(T[r] * (1.0 + 1e-6));
                        }
                    }
                }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[s]);
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
        float f = best_f[r];
        int32_t i = best_i[r];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float of = __shfl_xor_sync(kFull, f, o);
            const int32_t oi = __shfl_xor_sync(kFull, i, o);
            if (of < f || (of == f && oi > i)) {
                f = of;
                i = oi;
            }
        }
        const int64_t qi = q0 + r;
        if (lane == 0 && qi < a.nq) {
            const int64_t o = (int64_t)blockIdx.y * a.nq + qi;
            a.idx[o] = i < 0 ? -1 : (int32_t)(i + a.index_base);
            a.dist[o] = f;
        }
    }
}

// reduce chunk partials: min dist, then max index
__global__ void nearest_reduce_kernel(const int32_t* __restrict__ idx_in,
                                      const float* __restrict__ dist_in, int chunks, int64_t nq,
                                      int32_t* __restrict__ idx_out, float* __restrict__ dist_out) {
    const int64_t qi = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (qi >= nq) return;
    float f = dist_in[qi];
    int32_t i = idx_in[qi];
    for (int c = 1; c < chunks; ++c) {
        const float of = dist_in[(int64_t)c * nq + qi];
        const int32_t oi = idx_in[(int64_t)c * nq + qi];
        if (oi >= 0 && (i < 0 || of < f || (of == f && oi > i))) {
            f = of;
            i = oi;
        }
    }
    idx_out[qi] = i;
    if (dist_out) dist_out[qi] = f;
}

}  // namespace bgr
