// Exact grid-bucketed kNN (reference surface nns->knn, MS:1309-1312).
//
// Index: the indexed prefix [0, n_indexed) of the append-only store is bucketed into a
// uniform grid (about kPtsPerCell nodes per cell, <= 1024 cells per axis) and stored as
// one float4 {x,y,z,index} per node, sorted by cell with x fastest — so the cells of one
// x-row of a query's neighbourhood are ONE contiguous, 16-byte-aligned range that a warp
// reads with coalesced float4 loads.  Nodes appended after the last build form an unsorted
// tail that every query scans linearly; the index is rebuilt when the tail exceeds a
// quarter of the indexed part (nns.cu).
//
// Query: one warp per query.  Ring 1 (3x3x3 cells) is scanned first, then cubic shells;
// after each ring the search stops once the k-th key is closer than the nearest unscanned
// cell face (minus a slack that covers the float rounding of the cell computation), which
// makes the result identical to the brute-force scan: same f32 d2, same (d2,index) order.
#pragma once
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <vector>
#include "warp_topk.cuh"

namespace bgr {

constexpr int kPtsPerCell = 8;
constexpr int kMaxCellsPerAxis = 1024;

struct GridGeom {
    float mn[3];
    float cs;      // cell size
    float inv_cs;
    int g[3];
};

struct GridIndex {
    int64_t n_indexed = 0;
    GridGeom geom{};
    DevBuf<float4> pts;
    DevBuf<int32_t> cell_start;   // ncells + 1
    DevBuf<int32_t> cell_id;      // per node (build scratch)
    DevBuf<int32_t> cursor;       // per cell (build scratch)
    DevBuf<int32_t> block_sums;
    DevBuf<float> partial;        // bbox partials
    DevBuf<double> sumsq;         // build scratch
    DevBuf<long long> tile_prof;
    DevBuf<int32_t> tile_count, tile_list, tile_fb;   // tile kNN (tile_knn.cuh): per-tile query lists, fallback list
    int64_t tile_zeroed = 0;      // tile counters known to be zero (every call leaves them zeroed)
    int tile_parity = 0;          // which of the two fallback counts the next call uses
    uint64_t n_tile = 0;          // tile_knn_kernel launches
    int tile_shape[3] = {-1, 0, 0};  // tile kNN: -1 = off (the default, see tile_knn.cuh), 0 = tile size chosen from the query density, else forced
    int tile_warps = 0;           // forced warps per tile CTA (0 = default)
    uint64_t tile_last_shape = 0, tile_last_count = 0;   // of the last tile search (bgr_nns_tile_stats)
    DevBuf<int32_t> heavy;        // queries handed to grid_knn_heavy_kernel and its scratch (layout: kHeavyInts)
    double crowd = 0.0;           // sum(cell count^2) / n: the occupancy of the cell an average NODE sits in
                                  // (~kPtsPerCell+1 for a uniform cloud, thousands for a piled-up tree)
    // which code paths ran (bgr_nns_debug_stats): index builds, grid_knn_kernel<false> / <true> launches,
    // grid_nearest_kernel launches
    uint64_t n_builds = 0, n_lean = 0, n_piled = 0, n_grid_nearest = 0;
};

__device__ __forceinline__ int cell_coord(float v, float mn, float inv_cs, int g) {
    const float t = __fmul_rn(__fsub_rn(v, mn), inv_cs);
    int c = (int)floorf(t);
    // NaN / out-of-range queries clamp into the grid
    if (!(t >= 0.f)) c = 0;
    return min(c, g - 1);
}

// ---------------------------------------------------------------- build kernels
__global__ void bbox_kernel(const float* __restrict__ x, const float* __restrict__ y,
                            const float* __restrict__ z, int64_t n, float* __restrict__ partial) {
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float v[3] = {x[i], y[i], z[i]};
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            mn[a] = fminf(mn[a], v[a]);
            mx[a] = fmaxf(mx[a], v[a]);
        }
    }
    __shared__ float sm[6][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        for (int o = 16; o > 0; o >>= 1) {
            mn[a] = fminf(mn[a], __shfl_xor_sync(kFull, mn[a], o));
            mx[a] = fmaxf(mx[a], __shfl_xor_sync(kFull, mx[a], o));
        }
        if (lane == 0) {
            sm[a][warp] = mn[a];
            sm[3 + a][warp] = mx[a];
        }
    }
    __syncthreads();
    if (warp == 0) {
        const int nw = blockDim.x >> 5;
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            float m0 = lane < nw ? sm[a][lane] : FLT_MAX, m1 = lane < nw ? sm[3 + a][lane] : -FLT_MAX;
            for (int o = 16; o > 0; o >>= 1) {
                m0 = fminf(m0, __shfl_xor_sync(kFull, m0, o));
                m1 = fmaxf(m1, __shfl_xor_sync(kFull, m1, o));
            }
            if (lane == 0) {
                partial[blockIdx.x * 6 + a] = m0;
                partial[blockIdx.x * 6 + 3 + a] = m1;
            }
        }
    }
}

__global__ void cell_count_kernel(const float* __restrict__ x, const float* __restrict__ y,
                                  const float* __restrict__ z, int64_t n, GridGeom gm,
                                  int32_t* __restrict__ cell_id, int32_t* __restrict__ counts) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int cx = cell_coord(x[i], gm.mn[0], gm.inv_cs, gm.g[0]);
    const int cy = cell_coord(y[i], gm.mn[1], gm.inv_cs, gm.g[1]);
    const int cz = cell_coord(z[i], gm.mn[2], gm.inv_cs, gm.g[2]);
    const int32_t c = (cz * gm.g[1] + cy) * gm.g[0] + cx;
    cell_id[i] = c;
    atomicAdd(&counts[c], 1);
}

// exclusive scan, 3 phases over 1024-element blocks
__global__ void scan_block_kernel(int32_t* __restrict__ data, int64_t n, int32_t* __restrict__ sums) {
    __shared__ int32_t warp_tot[32];
    const int64_t i = (int64_t)blockIdx.x * 1024 + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int32_t v = i < n ? data[i] : 0;
    int32_t inc = v;
    for (int o = 1; o < 32; o <<= 1) {
        const int32_t t = __shfl_up_sync(kFull, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        int32_t w = warp_tot[lane];
        for (int o = 1; o < 32; o <<= 1) {
            const int32_t t = __shfl_up_sync(kFull, w, o);
            if (lane >= o) w += t;
        }
        warp_tot[lane] = w;
    }
    __syncthreads();
    const int32_t off = warp ? warp_tot[warp - 1] : 0;
    if (i < n) data[i] = off + inc - v;
    if (threadIdx.x == 1023) sums[blockIdx.x] = off + inc;
}
__global__ void scan_add_kernel(int32_t* __restrict__ data, int64_t n, const int32_t* __restrict__ sums) {
    const int64_t i = (int64_t)blockIdx.x * 1024 + threadIdx.x;
    if (i < n) data[i] += sums[blockIdx.x];
}

__global__ void scatter_kernel(const float* __restrict__ x, const float* __restrict__ y,
                               const float* __restrict__ z, int64_t n,
                               const int32_t* __restrict__ cell_id, const int32_t* __restrict__ cell_start,
                               int32_t* __restrict__ cursor, float4* __restrict__ pts) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int32_t c = cell_id[i];
    const int32_t pos = cell_start[c] + atomicAdd(&cursor[c], 1);
    pts[pos] = make_float4(x[i], y[i], z[i], __int_as_float((int)i));
}

__global__ void cell_sumsq_kernel(const int32_t* __restrict__ cell_start, int64_t ncells, double* __restrict__ out) {
    double acc = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < ncells; i += (int64_t)gridDim.x * blockDim.x) {
        const double c = (double)(cell_start[i + 1] - cell_start[i]);
        acc += c * c;
    }
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(kFull, acc, o);
    if ((threadIdx.x & 31) == 0 && acc != 0.0) atomicAdd(out, acc);
}

inline int exclusive_scan(int32_t* d, int64_t n, DevBuf<int32_t>& sums, cudaStream_t st) {
    // recursive over levels; the sums of level L live at offset in `sums`
    std::vector<int64_t> sizes;
    int64_t m = n, total = 0;
    while (m > 1024) {
        m = div_up(m, 1024);
        sizes.push_back(m);
        total += round_up(m, 4);
    }
    sizes.push_back(1);
    total += 4;
    BGR_CHECK(sums.reserve(total, st, false));
    std::vector<int32_t*> lv;
    int32_t* p = sums.p;
    for (int64_t s : sizes) {
        lv.push_back(p);
        p += round_up(s, 4);
    }
    std::vector<int32_t*> data{d};
    std::vector<int64_t> ns{n};
    for (size_t l = 0; l + 1 < sizes.size() + 1 && ns.back() > 0; ++l) {
        const int64_t cur = ns.back();
        scan_block_kernel<<<(unsigned)div_up(cur, 1024), 1024, 0, st>>>(data.back(), cur, lv[l]);
        BGR_LAUNCH_CHECK();
        if (cur <= 1024) break;
        data.push_back(lv[l]);
        ns.push_back(div_up(cur, 1024));
    }
    for (int l = (int)data.size() - 2; l >= 0; --l) {
        scan_add_kernel<<<(unsigned)div_up(ns[l], 1024), 1024, 0, st>>>(data[l], ns[l], data[l + 1]);
        BGR_LAUNCH_CHECK();
    }
    return BGR_OK;
}

inline int grid_build(GridIndex& G, const float* x, const float* y, const float* z, int64_t n,
                      cudaStream_t st) {
    if (n <= 0) {
        G.n_indexed = 0;
        return BGR_OK;
    }
    // 1. bounding box
    const int nb = 512;
    BGR_CHECK(G.partial.reserve(nb * 6, st, false));
    bbox_kernel<<<nb, 256, 0, st>>>(x, y, z, n, G.partial.p);
    BGR_LAUNCH_CHECK();
    std::vector<float> hp((size_t)nb * 6);
    BGR_CUDA(cudaMemcpyAsync(hp.data(), G.partial.p, hp.size() * sizeof(float), cudaMemcpyDeviceToHost, st));
    BGR_CUDA(cudaStreamSynchronize(st));
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int b = 0; b < nb; ++b)
        for (int a = 0; a < 3; ++a) {
            mn[a] = std::min(mn[a], hp[(size_t)b * 6 + a]);
            mx[a] = std::max(mx[a], hp[(size_t)b * 6 + 3 + a]);
        }
    for (int a = 0; a < 3; ++a)
        BGR_ARG(std::isfinite(mn[a]) && std::isfinite(mx[a]), "non-finite node coordinates");
    // 2. cell size: smallest cs with prod(ceil(ext/cs)) <= n / kPtsPerCell (bisection)
    double ext[3], emax = 0;
    for (int a = 0; a < 3; ++a) {
        ext[a] = (double)mx[a] - (double)mn[a];
        emax = std::max(emax, ext[a]);
    }
    const double target = std::max<double>(1.0, (double)n / kPtsPerCell);
    auto cells = [&](double cs) {
        double c = 1;
        for (int a = 0; a < 3; ++a) c *= std::max(1.0, std::ceil(ext[a] / cs));
        return c;
    };
    double lo = std::max(emax / kMaxCellsPerAxis, 1e-30), hi = std::max(emax, 1e-30) * 1.0001;
    if (cells(lo) <= target) hi = lo;
    else
        for (int it = 0; it < 60; ++it) {
            const double mid = 0.5 * (lo + hi);
            if (cells(mid) <= target) hi = mid; else lo = mid;
        }
    GridGeom gm;
    gm.cs = (float)(hi * 1.000001);
    gm.inv_cs = 1.0f / gm.cs;
    int64_t ncells = 1;
    for (int a = 0; a < 3; ++a) {
        gm.mn[a] = mn[a];
        gm.g[a] = (int)std::min<double>(kMaxCellsPerAxis, std::max(1.0, std::floor(ext[a] / (double)gm.cs) + 1.0));
        ncells *= gm.g[a];
    }
    // 3. count, scan, scatter
    BGR_CHECK(G.cell_id.reserve(n, st, false));
    BGR_CHECK(G.cell_start.reserve(ncells + 1, st, false));
    BGR_CHECK(G.cursor.reserve(ncells, st, false));
    BGR_CHECK(G.pts.reserve(n, st, false));
    BGR_CUDA(cudaMemsetAsync(G.cell_start.p, 0, (size_t)(ncells + 1) * sizeof(int32_t), st));
    BGR_CUDA(cudaMemsetAsync(G.cursor.p, 0, (size_t)ncells * sizeof(int32_t), st));
    cell_count_kernel<<<(unsigned)div_up(n, 256), 256, 0, st>>>(x, y, z, n, gm, G.cell_id.p, G.cell_start.p);
    BGR_LAUNCH_CHECK();
    BGR_CHECK(exclusive_scan(G.cell_start.p, ncells + 1, G.block_sums, st));
    scatter_kernel<<<(unsigned)div_up(n, 256), 256, 0, st>>>(x, y, z, n, G.cell_id.p, G.cell_start.p, G.cursor.p, G.pts.p);
    BGR_LAUNCH_CHECK();
    // crowding statistic for the AUTO method choice (nns.cu)
    BGR_CHECK(G.sumsq.reserve(1, st, false));
    BGR_CUDA(cudaMemsetAsync(G.sumsq.p, 0, sizeof(double), st));
    cell_sumsq_kernel<<<256, 256, 0, st>>>(G.cell_start.p, ncells, G.sumsq.p);
    BGR_LAUNCH_CHECK();
    double ss = 0.0;
    BGR_CUDA(cudaMemcpyAsync(&ss, G.sumsq.p, sizeof(double), cudaMemcpyDeviceToHost, st));
    BGR_CUDA(cudaStreamSynchronize(st));
    G.crowd = ss / (double)n;
    G.geom = gm;
    G.n_indexed = n;
    G.n_builds++;
    return BGR_OK;
}

inline void grid_release(GridIndex& G) {
    G.pts.release(); G.cell_start.release(); G.cell_id.release(); G.cursor.release();
    G.block_sums.release(); G.partial.release(); G.sumsq.release(); G.heavy.release();
    G.tile_count.release(); G.tile_list.release(); G.tile_fb.release(); G.tile_prof.release(); G.tile_zeroed = 0;
    G.n_indexed = 0;
}

// ---------------------------------------------------------------- query kernel
struct GridKnnArgs {
    GridGeom gm;
    const float4* pts;
    const int32_t* cell_start;
    int64_t n_indexed;
    const float* x;  // raw store (tail scan)
    const float* y;
    const float* z;
    int64_t n;
    const float* q;
    int64_t nq;
    int k;
    unsigned flags;
    float max_radius;
    const float* max_radii;
    int32_t* idx;
    float* d2;
    int64_t index_base;
    unsigned long long* counter;
    int fast;            // enable the one-pass ring-1 path
    const uint8_t* skip; // optional per-query flag: non-zero = leave this query's rows untouched
    float fast_target;   // expected survivors the a-priori threshold aims at (k + ~2.5 sqrt k)
    int32_t* heavy;      // optional: [0] count, [1..heavy_cap] query numbers (see grid_knn_heavy_kernel)
    int heavy_cap;
    int64_t heavy_min;   // ring-1 block + tail at least this long => heavy
};

__device__ __forceinline__ uint32_t radius_sentinel_g(float mr) {
    float mr2 = __fmul_rn(mr, mr);
    if (!(mr2 <= FLT_MAX)) mr2 = FLT_MAX;
    return __float_as_uint(mr2) + 1u;
}

constexpr int kFastCap = 64;          // per-warp candidate buffer of the fast path
constexpr int kGridWarps = 8;

// Fast path of a query whose k neighbours lie inside ring 1 (the common case by construction:
// ~8 nodes per cell).  ONE pass over the 3x3x3 block with an a-priori threshold tau (from the
// local node density), survivors compacted into a per-warp shared-memory buffer, then a single
// 64-bit bitonic sort — instead of maintaining the sorted list through ~200 candidates.
// Exactness: every node of the block (and of the unsorted tail) with d2 <= tau is in the
// buffer; tau < b2 (nearest unscanned face), so if the buffer holds >= k admissible nodes its
// k smallest keys are the global answer; if tau already equals the radius limit the (shorter)
// answer is complete too.  Otherwise tau is rescaled from the observed count and the pass is
// repeated; after kFastTries the caller falls back to the general ring search.
constexpr int kFastTries = 3;

// PILED = the index is crowded (see grid_knn): pile queries are set aside for grid_knn_heavy_kernel
// and the long scans (unsorted tail, rows longer than a warp) keep 3 x 32 nodes in flight; the lean
// instantiation serves uniform clouds, where neither occurs and the extra registers would spill.
template <bool PILED>
__device__ __forceinline__ void grid_knn_query(const GridKnnArgs& a, const int64_t qi, unsigned long long* cand_w, const int lane,
                                               const bool allow_fast = true) {
    const float qx = __ldg(a.q + 3 * qi), qy = __ldg(a.q + 3 * qi + 1), qz = __ldg(a.q + 3 * qi + 2);
    const bool self = (a.flags & BGR_ALLOW_SELF_MATCH) != 0;
    const uint32_t sent = radius_sentinel_g(a.max_radii ? __ldg(a.max_radii + qi) : a.max_radius);
    const float lim = __uint_as_float(sent - 1u);
    unsigned long long touched = 0;
    const GridGeom& gm = a.gm;
    int cx = 0, cy = 0, cz = 0;
    if (a.n_indexed > 0) {
        cx = cell_coord(qx, gm.mn[0], gm.inv_cs, gm.g[0]);
        cy = cell_coord(qy, gm.mn[1], gm.inv_cs, gm.g[1]);
        cz = cell_coord(qz, gm.mn[2], gm.inv_cs, gm.g[2]);
    }
    auto dist2 = [&](float px, float py, float pz) {
        const float dx = __fsub_rn(qx, px), dy = __fsub_rn(qy, py), dz = __fsub_rn(qz, pz);
        return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
    };
    // nearest unscanned face after `ring` rings, squared and shrunk by the rounding slack of the
    // cell computation (+inf when the block reaches every border)
    auto face2 = [&](int ring) {
        float b = __uint_as_float(kInfBits);
        const float qv[3] = {qx, qy, qz};
        const int cc[3] = {cx, cy, cz};
#pragma unroll
        for (int ax = 0; ax < 3; ++ax) {
            if (cc[ax] - ring > 0) b = fminf(b, qv[ax] - (gm.mn[ax] + (float)(cc[ax] - ring) * gm.cs));
            if (cc[ax] + ring + 1 < gm.g[ax]) b = fminf(b, (gm.mn[ax] + (float)(cc[ax] + ring + 1) * gm.cs) - qv[ax]);
        }
        if (b == __uint_as_float(kInfBits)) return b;
        const float bs = fmaxf(0.f, b - 2e-3f * gm.cs);
        return bs * bs * (1.f - 2e-6f);
    };

    // ------------------------------------------------------------------ fast path
    if (a.n_indexed > 0 && a.fast && allow_fast) {
        // the 9 x-rows of ring 1: lane t < 9 owns row (dy,dz)
        int rs = 0, re = 0;
        if (lane < 9) {
            const int yy = cy + lane % 3 - 1, zz = cz + lane / 3 - 1;
            if (yy >= 0 && yy < gm.g[1] && zz >= 0 && zz < gm.g[2]) {
                const int x0 = max(cx - 1, 0), x1 = min(cx + 1, gm.g[0] - 1);
                const int64_t rb = ((int64_t)zz * gm.g[1] + yy) * gm.g[0];
                rs = __ldg(a.cell_start + rb + x0);
                re = __ldg(a.cell_start + rb + x1 + 1);
            }
        }
        int n27 = re - rs;
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) n27 += __shfl_xor_sync(kFull, n27, o);   // lanes 0..15 hold rows
        n27 = __shfl_sync(kFull, n27, 0);
        if (PILED && a.heavy && (int64_t)n27 + (a.n - a.n_indexed) >= a.heavy_min) {
            // a piled-up block (thousands of coincident nodes): one warp would walk it alone, several times
            int slot = 0;
            if (lane == 0) slot = atomicAdd(a.heavy, 1);
            slot = __shfl_sync(kFull, slot, 0);
            if (slot < a.heavy_cap) {
                if (lane == 0) a.heavy[1 + slot] = (int32_t)qi;
                return;
            }
        }
        const float b2 = face2(1);
        // largest usable threshold: strictly below the unscanned region, within the radius limit
        const float tmax = fminf(lim, b2 == __uint_as_float(kInfBits) ? lim : __uint_as_float(max(__float_as_uint(b2), 1u) - 1u));
        if (n27 > 0) {
            const int nxr = min(cx + 1, gm.g[0] - 1) - max(cx - 1, 0) + 1, nyr = min(cy + 1, gm.g[1] - 1) - max(cy - 1, 0) + 1,
                      nzr = min(cz + 1, gm.g[2] - 1) - max(cz - 1, 0) + 1;
            const float vol = (float)(nxr * nyr * nzr) * gm.cs * gm.cs * gm.cs;
            const float r3 = a.fast_target * vol / (4.18879f * (float)n27);
            const float r0 = exp2f(__log2f(r3) * 0.33333334f);      // the threshold only steers the work: any value is exact
            float tau = fminf(r0 * r0, tmax);
            const uint32_t cand_s = smem_u32(cand_w);
            const unsigned lt_mask = (1u << lane) - 1u;
            const float eps_lo = self ? -1.f : FLT_EPSILON;         // the self-match rule d2 > FLT_EPSILON as one comparison (d2 >= 0 > -1)
            const int last = (int)a.n_indexed - 1;
            for (int attempt = 0; attempt < kFastTries; ++attempt) {
                int c = 0;
                auto consider = [&](float px, float py, float pz, uint32_t pidx, bool valid) {
                    const float d = dist2(px, py, pz);
                    const bool pass = valid && (d <= tau) && (d > eps_lo);
                    const unsigned m = __ballot_sync(kFull, pass);
                    const int slot = c + __popc(m & lt_mask);
                    // key = (d2 bits << 32) | index: low word first
                    if (pass && slot < kFastCap)
                        asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(cand_s + 8u * (unsigned)slot), "r"(pidx), "r"(__float_as_uint(d)) : "memory");
                    c += __popc(m);
                };
                if (PILED) {
                    for (int64_t base = a.n_indexed; base < a.n; base += 96) {          // unsorted tail, 3 x 32 nodes in flight
                        float tx[3], ty[3], tz[3];
    #pragma unroll
                        for (int u = 0; u < 3; ++u) {
                            const int64_t j = base + 32 * u + lane;
                            const bool valid = j < a.n;
                            tx[u] = valid ? a.x[j] : 0.f; ty[u] = valid ? a.y[j] : 0.f; tz[u] = valid ? a.z[j] : 0.f;
                        }
    #pragma unroll
                        for (int u = 0; u < 3; ++u) {
                            const int64_t j = base + 32 * u + lane;
                            if (base + 32 * u < a.n) consider(tx[u], ty[u], tz[u], (uint32_t)j, j < a.n);
                        }
                    }
                } else {
                    for (int64_t base = a.n_indexed; base < a.n; base += 32) {          // unsorted tail
                        const int64_t j = base + lane;
                        const bool valid = j < a.n;
                        consider(valid ? a.x[j] : 0.f, valid ? a.y[j] : 0.f, valid ? a.z[j] : 0.f, (uint32_t)j, valid);
                    }
                }
                // rows in groups of three: the three leading LDG.128 are in flight together
#pragma unroll 1
                for (int g = 0; g < 9; g += 3) {
                    int s0[3], e0[3];
                    float4 p[3];
#pragma unroll
                    for (int i = 0; i < 3; ++i) {
                        s0[i] = __shfl_sync(kFull, rs, g + i);
                        e0[i] = __shfl_sync(kFull, re, g + i);
                        p[i] = __ldg(a.pts + min(s0[i] + lane, last));      // lanes past the row's end load a node they ignore
                    }
#pragma unroll
                    for (int i = 0; i < 3; ++i) {
                        if (s0[i] < e0[i]) consider(p[i].x, p[i].y, p[i].z, (uint32_t)__float_as_int(p[i].w), s0[i] + lane < e0[i]);
                        if (PILED) {
                            for (int b = s0[i] + 32; b < e0[i]; b += 96) {      // rows longer than a warp: 3 x 32 nodes in flight
                                float4 q4[3];
#pragma unroll
                                for (int u = 0; u < 3; ++u) {
                                    const int j = b + 32 * u + lane;
                                    q4[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                                    if (j < e0[i]) q4[u] = __ldg(a.pts + j);
                                }
#pragma unroll
                                for (int u = 0; u < 3; ++u)
                                    if (b + 32 * u < e0[i])
                                        consider(q4[u].x, q4[u].y, q4[u].z, (uint32_t)__float_as_int(q4[u].w), b + 32 * u + lane < e0[i]);
                            }
                        } else {
                            for (int b = s0[i] + 32; b < e0[i]; b += 32) {      // rows longer than a warp (rare)
                                const int j = b + lane;
                                const bool valid = j < e0[i];
                                float4 q4 = make_float4(0.f, 0.f, 0.f, 0.f);
                                if (valid) q4 = __ldg(a.pts + j);
                                consider(q4.x, q4.y, q4.z, (uint32_t)__float_as_int(q4.w), valid);
                            }
                        }
                    }
                }
                touched += (unsigned long long)(n27 + (a.n - a.n_indexed));
                if (c <= kFastCap && (c >= a.k || (tau >= tmax && tmax >= lim))) {
                    __syncwarp();
                    unsigned long long key = lane < c ? cand_w[lane] : kKey64Max;
                    if (c <= 32) {
                        // rank by counting (d2 alone orders the survivors unless two of them share a d2: then
                        // the bitonic sort below breaks the tie by index): lane -> row position directly
                        const uint32_t kd = (uint32_t)(key >> 32);
                        const unsigned peers = __match_any_sync(kFull, kd);
                        if (!__any_sync(kFull, lane < c && (peers & (peers - 1u)) != 0u)) {
                            const uint32_t* cd = reinterpret_cast<const uint32_t*>(cand_w) + 1;
                            int rank = 0;
#pragma unroll 4
                            for (int j = 0; j < c; ++j) rank += cd[2 * j] < kd ? 1 : 0;
                            if (lane < c) {
                                if (rank < a.k) {
                                    a.idx[qi * a.k + rank] = (int32_t)((int64_t)(uint32_t)key + a.index_base);
                                    a.d2[qi * a.k + rank] = __uint_as_float(kd);
                                }
                            } else if (lane < a.k) {                 // fewer than k inside the radius limit
                                a.idx[qi * a.k + lane] = -1;
                                a.d2[qi * a.k + lane] = __uint_as_float(kInfBits);
                            }
                            if (a.counter && lane == 0) atomicAdd(a.counter, touched);
                            return;
                        }
                    }
                    warp_sort32_u64(key, lane);
                    if (c > 32) {
                        unsigned long long k2 = 32 + lane < c ? cand_w[32 + lane] : kKey64Max;
                        warp_sort32_u64(k2, lane);
                        warp_merge32_u64(key, k2, lane);
                    }
                    if (lane < a.k) {
                        const bool empty = key == kKey64Max;
                        a.idx[qi * a.k + lane] = empty ? -1 : (int32_t)((int64_t)(uint32_t)key + a.index_base);
                        a.d2[qi * a.k + lane] = empty ? __uint_as_float(kInfBits) : __uint_as_float((uint32_t)(key >> 32));
                    }
                    if (a.counter && lane == 0) atomicAdd(a.counter, touched);
                    return;
                }
                __syncwarp();
                if (c > kFastCap) {
                    const float f = exp2f(__log2f(a.fast_target / (float)c) * 0.33333334f);
                    tau = tau * f * f;
                } else {
                    if (tau >= tmax) break;                       // needs ring 2: general search
                    const float f = c > 0 ? exp2f(__log2f((a.fast_target + 6.f) / (float)c) * 0.33333334f) : 2.f;
                    tau = fminf(tmax, tau * f * f * 1.15f);
                }
            }
        }
    }

    // ------------------------------------------------------------------ general ring search
    WarpList L;
    L.init(sent);
    uint32_t tau_d = sent, tau_i = 0u;
    float tau_f = __uint_as_float(sent);
    auto offer = [&](float px, float py, float pz, uint32_t pidx, bool valid) {
        const float d = dist2(px, py, pz);
        const bool pass = valid && (d <= tau_f) && (self || d > FLT_EPSILON);
        if (__any_sync(kFull, pass)) {
            warp_offer(L, pass, __float_as_uint(d), pidx, tau_d, tau_i, a.k, lane);
            tau_f = __uint_as_float(tau_d);
        }
    };

    // ---- unsorted tail [n_indexed, n)
    if (PILED) {
        for (int64_t base = a.n_indexed; base < a.n; base += 96) {
            float tx[3], ty[3], tz[3];
    #pragma unroll
            for (int u = 0; u < 3; ++u) {
                const int64_t j = base + 32 * u + lane;
                const bool valid = j < a.n;
                tx[u] = valid ? a.x[j] : 0.f; ty[u] = valid ? a.y[j] : 0.f; tz[u] = valid ? a.z[j] : 0.f;
            }
    #pragma unroll
            for (int u = 0; u < 3; ++u) {
                const int64_t j = base + 32 * u + lane;
                if (base + 32 * u < a.n) offer(tx[u], ty[u], tz[u], (uint32_t)j, j < a.n);
            }
        }
    } else {
        for (int64_t base = a.n_indexed; base < a.n; base += 32) {
            const int64_t j = base + lane;
            const bool valid = j < a.n;
            const float px = valid ? a.x[j] : 0.f, py = valid ? a.y[j] : 0.f, pz = valid ? a.z[j] : 0.f;
            offer(px, py, pz, (uint32_t)j, valid);
        }
    }
    touched += (unsigned long long)(a.n - a.n_indexed);

    if (a.n_indexed > 0) {
        const int gmax = max(gm.g[0], max(gm.g[1], gm.g[2]));
        for (int ring = 1; ring <= gmax; ++ring) {
            // items of this ring: every (y,z) row of the block contributes two items;
            // ring 1 scans the full 3x3x3 block, later rings only the cubic shell.
            const int side = 2 * ring + 1;
            const int nitems = 2 * side * side;
            for (int t0 = 0; t0 < nitems; t0 += 32) {
                const int t = t0 + lane;
                int s = 0, e = 0;
                if (t < nitems) {
                    const int row = t >> 1, which = t & 1;
                    const int dy = row % side - ring, dz = row / side - ring;
                    const int yy = cy + dy, zz = cz + dz;
                    if (yy >= 0 && yy < gm.g[1] && zz >= 0 && zz < gm.g[2]) {
                        const bool shell_row = (ring == 1) || abs(dy) == ring || abs(dz) == ring;
                        int x0, x1;   // inclusive cell range
                        if (shell_row) {
                            x0 = which == 0 ? cx - ring : 1;
                            x1 = which == 0 ? cx + ring : 0;
                        } else {
                            x0 = x1 = which == 0 ? cx - ring : cx + ring;
                            if (x0 < 0 || x0 >= gm.g[0]) { x0 = 1; x1 = 0; }
                        }
                        x0 = max(x0, 0);
                        x1 = min(x1, gm.g[0] - 1);
                        if (x0 <= x1) {
                            const int64_t rb = ((int64_t)zz * gm.g[1] + yy) * gm.g[0];
                            s = __ldg(a.cell_start + rb + x0);
                            e = __ldg(a.cell_start + rb + x1 + 1);
                        }
                    }
                }
                unsigned nonempty = __ballot_sync(kFull, e > s);
                while (nonempty) {
                    const int src = __ffs(nonempty) - 1;
                    nonempty &= nonempty - 1;
                    const int rs = __shfl_sync(kFull, s, src), re = __shfl_sync(kFull, e, src);
                    touched += (unsigned long long)(re - rs);
                    if (PILED) {
                        for (int b = rs; b < re; b += 96) {
                            float4 p[3];
    #pragma unroll
                            for (int u = 0; u < 3; ++u) {
                                const int j = b + 32 * u + lane;
                                p[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                                if (j < re) p[u] = __ldg(a.pts + j);
                            }
    #pragma unroll
                            for (int u = 0; u < 3; ++u)
                                if (b + 32 * u < re) offer(p[u].x, p[u].y, p[u].z, (uint32_t)__float_as_int(p[u].w), b + 32 * u + lane < re);
                        }
                    } else {
                        for (int b = rs; b < re; b += 32) {
                            const int j = b + lane;
                            const bool valid = j < re;
                            float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
                            if (valid) p = __ldg(a.pts + j);
                            offer(p.x, p.y, p.z, (uint32_t)__float_as_int(p.w), valid);
                        }
                    }
                }
            }
            // ---- termination: nearest unscanned face (inf when the block reaches the border)
            const float b2 = face2(ring);
            if (b2 == __uint_as_float(kInfBits)) break;               // whole grid scanned
            if (b2 > lim) break;                                      // nothing farther is admissible
            if (tau_d != sent && tau_f < b2) break;                   // k found, all closer than the rest
        }
    }

    if (lane < a.k) {
        const bool empty = (L.d == sent);
        a.idx[qi * a.k + lane] = empty ? -1 : (int32_t)((int64_t)L.i + a.index_base);
        a.d2[qi * a.k + lane] = empty ? __uint_as_float(kInfBits) : __uint_as_float(L.d);
    }
    if (a.counter && lane == 0) atomicAdd(a.counter, touched);
}

template <bool PILED>
__global__ void __launch_bounds__(kGridWarps * 32, PILED ? 4 : 5) grid_knn_kernel(GridKnnArgs a) {
    __shared__ unsigned long long cand[kGridWarps][kFastCap];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t qi = (int64_t)blockIdx.x * (blockDim.x >> 5) + wib;
    if (qi >= a.nq) return;
    if (a.skip && __ldg(a.skip + qi)) return;
    grid_knn_query<PILED>(a, qi, cand[wib], lane);
}

// The same search for a device-side list of query numbers (the queries tile_knn_kernel hands back:
// a handful per thousand): persistent warps stride over the list.
__global__ void __launch_bounds__(kGridWarps * 32, 5) grid_knn_list_kernel(GridKnnArgs a, const int32_t* __restrict__ list,
                                                                           const int32_t* __restrict__ count) {
    __shared__ unsigned long long cand[kGridWarps][kFastCap];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int n = *count;
    for (int w = blockIdx.x * kGridWarps + wib; w < n; w += gridDim.x * kGridWarps) {
        grid_knn_query<false>(a, (int64_t)list[w], cand[wib], lane);
        __syncwarp();
    }
}

// Queries the main kernel set aside because their ring-1 block is piled up (a planner tree holds
// thousands of coincident copies of a reached milestone; no grid separates them, and one warp
// walking the pile alone, several times, was 0.19 ms of a 0.52 ms planner step).  Each such query
// is scanned over the whole raw store by kHeavySlices CTAs on different SMs: CTA (b, s) takes the
// 128-node tiles s, s + S, s + 2S, ... in a scrambled order (growth-ordered store: an in-order scan
// tightens its bound on almost every tile), its 4 warp lists are merged through shared memory, and
// the CTA that finishes last merges the S partial lists and writes the row.  Same arithmetic, same
// admission rules and the same (d2, index) order as the ring search => the same rows.
constexpr int kHeavyCap = 256;
constexpr int kHeavySlices = 32;
constexpr int kHeavyWarps = 4;
// layout of GridIndex::heavy (int32): [0] count | [1, 1+cap) queries | [1+cap, 1+2cap) finished slices |
// then cap * S * 32 * 2 words of partial lists
constexpr int kHeavyInts = 1 + 2 * kHeavyCap + kHeavyCap * kHeavySlices * 64;

__global__ void __launch_bounds__(kHeavyWarps * 32) grid_knn_heavy_kernel(GridKnnArgs a) {
    __shared__ uint32_t sd[kHeavyWarps][32], si[kHeavyWarps][32];
    const int b = blockIdx.x, sl = blockIdx.y;
    const int cnt = min(a.heavy[0], a.heavy_cap);
    if (b >= cnt) return;
    const int64_t qi = a.heavy[1 + b];
    int* done = a.heavy + 1 + a.heavy_cap;
    uint32_t* part = reinterpret_cast<uint32_t*>(a.heavy + 1 + 2 * a.heavy_cap);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float qx = __ldg(a.q + 3 * qi), qy = __ldg(a.q + 3 * qi + 1), qz = __ldg(a.q + 3 * qi + 2);
    const bool self = (a.flags & BGR_ALLOW_SELF_MATCH) != 0;
    const uint32_t sent = radius_sentinel_g(a.max_radii ? __ldg(a.max_radii + qi) : a.max_radius);
    WarpList L;
    L.init(sent);
    uint32_t tau_d = sent, tau_i = 0u;
    float tau_f = __uint_as_float(sent);
    constexpr int kTileN = kHeavyWarps * 32;
    const int64_t ntiles = (a.n + kTileN - 1) / kTileN;
    const int M = ntiles > sl ? (int)((ntiles - sl + kHeavySlices - 1) / kHeavySlices) : 0;
    int stride = 1;
    if (M >= 3) {
        for (stride = max(1, (int)(0.6180339887f * (float)M));; ++stride) {
            int u = stride, v = M;
            while (v) { const int w = u % v; u = v; v = w; }
            if (u == 1) break;
        }
    }
    for (int m = 0; m < M; ++m) {
        const int mm = (int)(((long long)m * stride) % M);
        const int64_t j = ((int64_t)sl + (int64_t)kHeavySlices * mm) * kTileN + threadIdx.x;
        const bool valid = j < a.n;
        const float px = valid ? a.x[j] : 0.f, py = valid ? a.y[j] : 0.f, pz = valid ? a.z[j] : 0.f;
        const float dx = __fsub_rn(qx, px), dy = __fsub_rn(qy, py), dz = __fsub_rn(qz, pz);
        const float d = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
        const bool pass = valid && (d <= tau_f) && (self || d > FLT_EPSILON);
        if (__any_sync(kFull, pass)) {
            warp_offer(L, pass, __float_as_uint(d), (uint32_t)j, tau_d, tau_i, a.k, lane);
            tau_f = __uint_as_float(tau_d);
        }
    }
    sd[warp][lane] = L.d;
    si[warp][lane] = L.i;
    __syncthreads();
    if (warp != 0) return;
    for (int w = 1; w < kHeavyWarps; ++w) {
        const uint32_t cd = sd[w][lane], ci = si[w][lane];
        const bool pass = lane < a.k && cd != sent;
        if (__any_sync(kFull, pass)) warp_offer(L, pass, cd, ci, tau_d, tau_i, a.k, lane);
    }
    uint32_t* mine = part + (((size_t)b * kHeavySlices + sl) * 32 + lane) * 2;
    __stcg(mine, L.d);
    __stcg(mine + 1, L.i);
    __threadfence();
    __syncwarp();
    int prev = 0;
    if (lane == 0) prev = atomicAdd(done + b, 1);
    prev = __shfl_sync(kFull, prev, 0);
    if (prev != kHeavySlices - 1) return;
    __threadfence();
    for (int s0 = 0; s0 < kHeavySlices; s0 += 8) {          // 8 partial lists in flight
        uint32_t cd[8], ci[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const uint32_t* other = part + (((size_t)b * kHeavySlices + s0 + u) * 32 + lane) * 2;
            cd[u] = __ldcg(other);
            ci[u] = __ldcg(other + 1);
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const bool pass = s0 + u != sl && lane < a.k && cd[u] != sent;
            if (__any_sync(kFull, pass)) warp_offer(L, pass, cd[u], ci[u], tau_d, tau_i, a.k, lane);
        }
    }
    if (lane < a.k) {
        const bool empty = (L.d == sent);
        a.idx[qi * a.k + lane] = empty ? -1 : (int32_t)((int64_t)L.i + a.index_base);
        a.d2[qi * a.k + lane] = empty ? __uint_as_float(kInfBits) : __uint_as_float(L.d);
    }
    if (a.counter && lane == 0) atomicAdd(a.counter, (unsigned long long)a.n);
}

// ---------------------------------------------------------------- grid Nearest (MS:138-154)
// Same index, the reference's 1-NN: Norm = float(sqrt(double(dx^2+dy^2+dz^2))) with dx = x2-x1
// in float, `<=` scan => the LAST index among nodes with equal float Norm.  One warp per
// query, ring by ring; a ring ends the search when the best Norm is strictly below the
// nearest unscanned cell face (shrunk by the cell-rounding slack and 1e-6 relative), so no
// unscanned node can beat or tie it: the result equals the brute-force scan's.
struct GridNearestArgs {
    GridGeom gm;
    const float4* pts;
    const int32_t* cell_start;
    int64_t n_indexed;
    const float* x;
    const float* y;
    const float* z;
    int64_t n;
    const float* q;
    int64_t nq;
    int32_t* idx;
    float* dist;
    int64_t index_base;
    const long long* nq_dev;   // optional device-side query count (see NearestArgs)
    int64_t nq_base;
};

__global__ void __launch_bounds__(kGridWarps * 32) grid_nearest_kernel(GridNearestArgs a) {
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t qi = (int64_t)blockIdx.x * (blockDim.x >> 5) + wib;
    if (qi >= a.nq) return;
    if (a.nq_dev && qi >= a.nq_base + (int64_t)__ldg(a.nq_dev)) return;
    const float qx = __ldg(a.q + 3 * qi), qy = __ldg(a.q + 3 * qi + 1), qz = __ldg(a.q + 3 * qi + 2);
    const GridGeom& gm = a.gm;
    const double kInfD = __longlong_as_double(0x7ff0000000000000LL);
    float best_f = __uint_as_float(kInfBits);
    int32_t best_i = -1;
    double T = kInfD;
    auto consider = [&](float px, float py, float pz, int32_t pidx, bool valid) {
        const double dx = (double)__fsub_rn(qx, px), dy = (double)__fsub_rn(qy, py), dz = (double)__fsub_rn(qz, pz);
        const double sq = __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
        if (valid && sq <= T) {   // NaN never passes
            const float f = __double2float_rn(__dsqrt_rn(sq));
            if (f < best_f || (f == best_f && pidx > best_i)) {
                best_f = f;
                best_i = pidx;
                T = f < __uint_as_float(kInfBits) ? (double)__uint_as_float(__float_as_uint(f) + 1u) * (double)__uint_as_float(__float_as_uint(f) + 1u) : kInfD;
            }
        }
    };
    // ---- unsorted tail [n_indexed, n)
    for (int64_t base = a.n_indexed; base < a.n; base += 32) {
        const int64_t j = base + lane;
        const bool valid = j < a.n;
        consider(valid ? a.x[j] : 0.f, valid ? a.y[j] : 0.f, valid ? a.z[j] : 0.f, (int32_t)j, valid);
    }
    const bool finite_q = (qx - qx == 0.f) && (qy - qy == 0.f) && (qz - qz == 0.f);
    if (a.n_indexed > 0 && !finite_q) {
        // infinite / NaN query: no geometry to prune with; scan the whole index like the plain loop
        for (int64_t base = 0; base < a.n_indexed; base += 32) {
            const int64_t j = base + lane;
            const bool valid = j < a.n_indexed;
            float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
            if (valid) p = __ldg(a.pts + j);
            consider(p.x, p.y, p.z, __float_as_int(p.w), valid);
        }
    }
    if (a.n_indexed > 0 && finite_q) {
        const int cx = cell_coord(qx, gm.mn[0], gm.inv_cs, gm.g[0]);
        const int cy = cell_coord(qy, gm.mn[1], gm.inv_cs, gm.g[1]);
        const int cz = cell_coord(qz, gm.mn[2], gm.inv_cs, gm.g[2]);
        const int gmax = max(gm.g[0], max(gm.g[1], gm.g[2]));
        for (int ring = 1; ring <= gmax; ++ring) {
            const int side = 2 * ring + 1;
            const int nitems = 2 * side * side;
            for (int t0 = 0; t0 < nitems; t0 += 32) {
                const int t = t0 + lane;
                int s = 0, e = 0;
                if (t < nitems) {
                    const int row = t >> 1, which = t & 1;
                    const int dy = row % side - ring, dz = row / side - ring;
                    const int yy = cy + dy, zz = cz + dz;
                    if (yy >= 0 && yy < gm.g[1] && zz >= 0 && zz < gm.g[2]) {
                        const bool shell_row = (ring == 1) || abs(dy) == ring || abs(dz) == ring;
                        int x0, x1;   // inclusive cell range
                        if (shell_row) {
                            x0 = which == 0 ? cx - ring : 1;
                            x1 = which == 0 ? cx + ring : 0;
                        } else {
                            x0 = x1 = which == 0 ? cx - ring : cx + ring;
                            if (x0 < 0 || x0 >= gm.g[0]) { x0 = 1; x1 = 0; }
                        }
                        x0 = max(x0, 0);
                        x1 = min(x1, gm.g[0] - 1);
                        if (x0 <= x1) {
                            const int64_t rb = ((int64_t)zz * gm.g[1] + yy) * gm.g[0];
                            s = __ldg(a.cell_start + rb + x0);
                            e = __ldg(a.cell_start + rb + x1 + 1);
                        }
                    }
                }
                unsigned nonempty = __ballot_sync(kFull, e > s);
                while (nonempty) {
                    const int src = __ffs(nonempty) - 1;
                    nonempty &= nonempty - 1;
                    const int rs = __shfl_sync(kFull, s, src), re = __shfl_sync(kFull, e, src);
                    for (int b = rs; b < re; b += 32) {
                        const int j = b + lane;
                        const bool valid = j < re;
                        float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (valid) p = __ldg(a.pts + j);
                        consider(p.x, p.y, p.z, __float_as_int(p.w), valid);
                    }
                }
            }
            // ---- termination: nearest unscanned face
            float bw = best_f;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) bw = fminf(bw, __shfl_xor_sync(kFull, bw, o));
            if (bw < __uint_as_float(kInfBits)) {   // lanes only need candidates that can still win or tie
                const float nu = __uint_as_float(__float_as_uint(bw) + 1u);
                T = fmin(T, (double)nu * (double)nu);
            }
            float b = __uint_as_float(kInfBits);
            const float qv[3] = {qx, qy, qz};
            const int cc[3] = {cx, cy, cz};
#pragma unroll
            for (int ax = 0; ax < 3; ++ax) {
                if (cc[ax] - ring > 0) b = fminf(b, qv[ax] - (gm.mn[ax] + (float)(cc[ax] - ring) * gm.cs));
                if (cc[ax] + ring + 1 < gm.g[ax]) b = fminf(b, (gm.mn[ax] + (float)(cc[ax] + ring + 1) * gm.cs) - qv[ax]);
            }
            if (b == __uint_as_float(kInfBits)) break;                       // whole grid scanned
            const float bs = fmaxf(0.f, b - 2e-3f * gm.cs) * (1.f - 1e-6f);
            if (bw < bs) break;
        }
    }
    // ---- warp reduce: min Norm, then max index
    float f = best_f;
    int32_t i = best_i;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float of = __shfl_xor_sync(kFull, f, o);
        const int32_t oi = __shfl_xor_sync(kFull, i, o);
        if (oi >= 0 && (i < 0 || of < f || (of == f && oi > i))) {
            f = of;
            i = oi;
        }
    }
    if (lane == 0) {
        a.idx[qi] = i < 0 ? -1 : (int32_t)(i + a.index_base);
        if (a.dist) a.dist[qi] = f;
    }
}

inline int grid_nearest(GridIndex& G, const float* x, const float* y, const float* z, int64_t n, const float* d_q,
                        int64_t nq, int32_t* d_idx, float* d_dist, int64_t index_base, cudaStream_t st,
                        const long long* nq_dev = nullptr, int64_t nq_base = 0) {
    GridNearestArgs a;
    a.gm = G.geom;
    a.pts = G.pts.p;
    a.cell_start = G.cell_start.p;
    a.n_indexed = G.n_indexed;
    a.x = x; a.y = y; a.z = z; a.n = n;
    a.q = d_q; a.nq = nq; a.idx = d_idx; a.dist = d_dist; a.index_base = index_base;
    a.nq_dev = nq_dev; a.nq_base = nq_base;
    grid_nearest_kernel<<<(unsigned)div_up(nq, kGridWarps), kGridWarps * 32, 0, st>>>(a);
    BGR_LAUNCH_CHECK();
    G.n_grid_nearest++;
    return BGR_OK;
}

}  // namespace bgr
