// Tile kNN: the exact grid search for LARGE query batches, one LANE per query
// (reference surface nns->knn, MS:1309-1312; same results as grid_knn_kernel / the brute-force scan).
//
// grid_knn_kernel spends a whole warp on one query: ~1,100 warp instructions per query, most of
// them per-query bookkeeping (cell rows, threshold, a 64-bit bitonic sort) repeated by 32 lanes.
// Here the queries are first binned into tiles of Tx x Ty x Tz grid cells (tile_bin_kernel: one
// atomicAdd per query).  One CTA owns a tile: it stages the nodes of the tile's halo
// ((Tx+2)(Ty+2)(Tz+2) cells: every ring-1 block of every query of the tile) into shared memory —
// each halo x-row is ONE contiguous range of the cell-sorted index, moved by one 1-D bulk copy
// (cp.async.bulk, SASS UBLKCP) — and then every lane runs its own query against its own 3x3x3
// block out of shared memory:
//   pass 1   : d2 = ((dx*dx)+dy*dy)+dz*dz (never contracted) for the ~216 nodes of the block; nodes with
//              d2 <= tau (a-priori threshold from the local density, strictly below the nearest
//              unscanned cell face) are remembered as 16-bit staging positions in a per-lane column
//              of shared memory;
//   select   : the <= 64 survivors' distances are recomputed into REGISTERS as 32-bit keys
//              (d2 bits with the low 6 bits replaced by the slot number) and the 16 smallest are
//              found with a fully unrolled compare-exchange network (min/max only: no shuffles, no
//              64-bit compares), every lane sorting its own query at once;
//   exactness: truncating 6 bits can only misorder two candidates whose d2 agree in the upper 26
//              bits; such a tie among the first k+1 keys, fewer than k survivors, or more than 64
//              sends THAT query to the fallback list, which grid_knn_list_kernel (the warp-per-query
//              search) answers.  Everything written here is therefore the exact (d2, index) order.
//
// Status: exact (tests/test_gpu_tile_knn.py) but NOT the default.  Measured on configs[1] (1 M nodes, 65,536
// queries, k = 16; profiles/r2_tile_knn_tune.txt, r2_ncu_full_g_tile_knn.txt): half the warp instructions of
// grid_knn_kernel (34 M against 67 M) yet 0.143 ms against 0.100 ms, because a tile needs ~37 KB of shared
// memory: 6 single-warp CTAs per SM, 8 % of the SM's warp slots, every warp a bare latency chain (pass 1:
// 170 cycles per 2-node trip; issue slots 32 % busy), plus 6 us of binning and a serial tail of handed-back
// queries.  Turned on per handle with bgr_nns_set_tile; what it would take to win is in DESIGN.md 3.2b.
#pragma once
#include <atomic>
#include <cstdio>
#include "grid.cuh"
#include "grid_group.cuh"

namespace bgr {

constexpr int kTileSlots = 64;        // survivors a lane can remember (6 slot bits in the key)
constexpr int kTileMaxRows = 64;      // (Ty+2)*(Tz+2)
constexpr int kTileMaxT0 = 8;
constexpr int kTileMaxWarps = 4;

struct TileKnnArgs {
    GridKnnArgs g;
    int T[3];             // tile size in cells
    int nt[3];            // tiles per axis
    int node_cap;         // staged nodes per tile
    int list_cap;         // queries per tile list
    int32_t* tile_count;  // [ntiles], all zero between calls
    int32_t* tile_list;   // [ntiles][list_cap]
    int32_t* fb;          // [0], [1]: fallback counts of even / odd calls; [2 ...): query numbers
    int parity;           // which count this call uses
    int vec_out;          // rows of idx / d2 are 16-byte aligned and k % 4 == 0
    long long* prof;      // measurement aid (BGR_TILE_PROF=1): [8] cycle sums of the kernel's phases, warp 0 of every tile
};

__global__ void tile_bin_kernel(TileKnnArgs a) {
    const int64_t qi = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (qi >= a.g.nq) return;
    if (a.g.skip && __ldg(a.g.skip + qi)) return;
    const GridGeom& gm = a.g.gm;
    const int cx = cell_coord(__ldg(a.g.q + 3 * qi), gm.mn[0], gm.inv_cs, gm.g[0]);
    const int cy = cell_coord(__ldg(a.g.q + 3 * qi + 1), gm.mn[1], gm.inv_cs, gm.g[1]);
    const int cz = cell_coord(__ldg(a.g.q + 3 * qi + 2), gm.mn[2], gm.inv_cs, gm.g[2]);
    const int t = ((cz / a.T[2]) * a.nt[1] + cy / a.T[1]) * a.nt[0] + cx / a.T[0];
    const int slot = atomicAdd(a.tile_count + t, 1);
    if (slot < a.list_cap) {
        a.tile_list[(int64_t)t * a.list_cap + slot] = (int32_t)qi;
    } else {
        const int s = atomicAdd(a.fb + a.parity, 1);
        a.fb[2 + s] = (int32_t)qi;
    }
}

// ---- register-resident compare-exchange networks on 32-bit keys (fully unrolled: all indices static)
__device__ __forceinline__ void cex(uint32_t& lo, uint32_t& hi) {
    const uint32_t a = min(lo, hi), b = max(lo, hi);
    lo = a;
    hi = b;
}
// ascending bitonic sort of 16 registers
__device__ __forceinline__ void sort16(uint32_t (&v)[16]) {
#pragma unroll
    for (int size = 2; size <= 16; size <<= 1) {
#pragma unroll
        for (int s = size >> 1; s > 0; s >>= 1) {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const int j = i ^ s;
                if (j > i) {
                    if ((i & size) == 0) cex(v[i], v[j]);
                    else cex(v[j], v[i]);
                }
            }
        }
    }
}
// a, b ascending: a <- the 16 smallest of the union, ascending; rest <- min(rest, smallest key dropped)
__device__ __forceinline__ void merge16(uint32_t (&a)[16], const uint32_t (&b)[16], uint32_t& rest) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const uint32_t lo = min(a[i], b[15 - i]), hi = max(a[i], b[15 - i]);
        a[i] = lo;
        rest = min(rest, hi);
    }
#pragma unroll
    for (int s = 8; s > 0; s >>= 1) {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const int j = i ^ s;
            if (j > i) cex(a[i], a[j]);
        }
    }
}

// dynamic shared memory of tile_knn_kernel
__host__ __device__ inline size_t tile_knn_smem(int node_cap, int warps) {
    return 16 + (size_t)node_cap * sizeof(float4) + (size_t)kTileMaxRows * (kTileMaxT0 + 3) * sizeof(int) +
           (size_t)(kTileMaxRows + 1) * sizeof(int) + (size_t)warps * kTileSlots * 32 * sizeof(uint16_t);
}

__global__ void __launch_bounds__(kTileMaxWarps * 32) tile_knn_kernel(TileKnnArgs a) {
    extern __shared__ __align__(128) unsigned char tsm[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(tsm);
    float4* nodes = reinterpret_cast<float4*>(tsm + 16);
    int* cs = reinterpret_cast<int*>(nodes + a.node_cap);                 // [row][T0+3]: cell boundaries (staging positions)
    int* rowoff = cs + kTileMaxRows * (kTileMaxT0 + 3);                   // [rows+1]
    uint16_t* bufs = reinterpret_cast<uint16_t*>(rowoff + kTileMaxRows + 1);

    const long long pt0 = clock64();
    const int t = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    if (t == 0 && tid == 0) a.fb[a.parity ^ 1] = 0;    // the next call's fallback count (its last reader has finished)
    int count = a.tile_count[t];
    if (count == 0) return;
    count = min(count, a.list_cap);
    const GridKnnArgs& g = a.g;
    const GridGeom& gm = g.gm;
    const int T0 = a.T[0], T1 = a.T[1], T2 = a.T[2];
    const int tx = t % a.nt[0], ty = (t / a.nt[0]) % a.nt[1], tz = t / (a.nt[0] * a.nt[1]);
    const int hx0 = tx * T0 - 1, hy0 = ty * T1 - 1, hz0 = tz * T2 - 1;   // halo origin (may be -1)
    const int nb = T0 + 3, nrows = (T1 + 2) * (T2 + 2);

    // ---- cell boundaries of the halo rows (global index positions first)
    for (int e = tid; e < nrows * nb; e += blockDim.x) {
        const int r = e / nb, i = e - r * nb;
        const int yy = hy0 + r % (T1 + 2), zz = hz0 + r / (T1 + 2);
        int v = 0;
        if (yy >= 0 && yy < gm.g[1] && zz >= 0 && zz < gm.g[2]) {
            const int x = min(max(hx0 + i, 0), gm.g[0]);
            v = __ldg(g.cell_start + ((int64_t)zz * gm.g[1] + yy) * gm.g[0] + x);
        }
        cs[r * nb + i] = v;
    }
    if (tid == 0) {
        mbar_init(bar, 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (warp == 0) {
        int run = 0;
        for (int r0 = 0; r0 < nrows; r0 += 32) {
            const int r = r0 + lane;
            const int len = r < nrows ? cs[r * nb + nb - 1] - cs[r * nb] : 0;
            int inc = len;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(kFull, inc, o);
                if (lane >= o) inc += v;
            }
            if (r < nrows) rowoff[r] = run + inc - len;
            run += __shfl_sync(kFull, inc, 31);
        }
        if (lane == 0) rowoff[nrows] = run;
    }
    __syncthreads();
    const int total = rowoff[nrows];
    if (total > a.node_cap || total == 0) {
        // denser than the staging buffer (or an empty neighbourhood): the warp-per-query search takes the tile
        for (int li = tid; li < count; li += blockDim.x) {
            const int s = atomicAdd(a.fb + a.parity, 1);
            a.fb[2 + s] = a.tile_list[(int64_t)t * a.list_cap + li];
        }
        if (tid == 0) a.tile_count[t] = 0;
        return;
    }
    const long long pt1 = clock64();
    if (tid == 0) mbar_expect_tx(bar, (uint32_t)total * 16u);
    __syncthreads();
    for (int r = tid; r < nrows; r += blockDim.x) {
        const int len = cs[r * nb + nb - 1] - cs[r * nb];
        if (len > 0) bulk_g2s(nodes + rowoff[r], g.pts + cs[r * nb], (uint32_t)len * 16u, bar);
    }
    __syncthreads();                       // every row's global start has been consumed by its copy
    for (int e = tid; e < nrows * nb; e += blockDim.x) {
        const int r = e / nb;
        // (reads the row's first boundary before anyone rewrites it: rewritten last, below)
        const int first = cs[r * nb];
        if (e - r * nb != 0) cs[e] = rowoff[r] + (cs[e] - first);
    }
    __syncthreads();
    for (int r = tid; r < nrows; r += blockDim.x) cs[r * nb] = rowoff[r];
    if (tid == 0) a.tile_count[t] = 0;     // leave the counters zeroed for the next call
    __syncthreads();
    const long long pt2 = clock64();
    mbar_wait(bar, 0);
    const long long pt3 = clock64();
    long long pq = 0, pp1 = 0, psel = 0, pout = 0;

    // self-match rule d2 > FLT_EPSILON as one comparison (d2 >= 0 always passes -1)
    const float eps_lo = (g.flags & BGR_ALLOW_SELF_MATCH) ? -1.f : FLT_EPSILON;
    uint16_t* buf = bufs + (size_t)warp * kTileSlots * 32;
    // queries of the tile, spread evenly over the warps (at most 32 per warp and pass)
    const int passes = (count + 32 * nwarps - 1) / (32 * nwarps);
    const int per = (count + passes * nwarps - 1) / (passes * nwarps);
    unsigned long long touched = 0;
    for (int pass = 0; pass < passes; ++pass) {
        const long long qa = clock64();
        const int chunk = pass * nwarps + warp;
        const int li = chunk * per + lane;
        const bool active = lane < per && li < count;
        int64_t qi = 0;
        float qx = 0.f, qy = 0.f, qz = 0.f;
        if (active) {
            qi = __ldg(a.tile_list + (int64_t)t * a.list_cap + li);
            qx = __ldg(g.q + 3 * qi); qy = __ldg(g.q + 3 * qi + 1); qz = __ldg(g.q + 3 * qi + 2);
        }
        const int cx = cell_coord(qx, gm.mn[0], gm.inv_cs, gm.g[0]);
        const int cy = cell_coord(qy, gm.mn[1], gm.inv_cs, gm.g[1]);
        const int cz = cell_coord(qz, gm.mn[2], gm.inv_cs, gm.g[2]);
        // inactive lanes sit on the tile's first cell with empty ranges
        const int lx = active ? cx - hx0 : 1, ly = active ? cy - hy0 : 1, lz = active ? cz - hz0 : 1;
        auto dist2 = [&](const float4& p) {
            const float dx = __fsub_rn(qx, p.x), dy = __fsub_rn(qy, p.y), dz = __fsub_rn(qz, p.z);
            return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
        };
        const int rbase = (lz - 1) * (T1 + 2) + (ly - 1);
        int n27 = 0;
        if (active) {
#pragma unroll
            for (int r9 = 0; r9 < 9; ++r9) {
                const int r = rbase + (r9 / 3) * (T1 + 2) + r9 % 3;
                n27 += cs[r * nb + lx + 2] - cs[r * nb + lx - 1];
            }
        }
        // threshold: as grid_knn_kernel's fast path (any tau < the nearest unscanned face is exact)
        const uint32_t sent = radius_sentinel_g(g.max_radii && active ? __ldg(g.max_radii + qi) : g.max_radius);
        const float lim = __uint_as_float(sent - 1u);
        float b = __uint_as_float(kInfBits);
        {
            const float qv[3] = {qx, qy, qz};
            const int cc[3] = {cx, cy, cz};
#pragma unroll
            for (int ax = 0; ax < 3; ++ax) {
                if (cc[ax] - 1 > 0) b = fminf(b, qv[ax] - (gm.mn[ax] + (float)(cc[ax] - 1) * gm.cs));
                if (cc[ax] + 2 < gm.g[ax]) b = fminf(b, (gm.mn[ax] + (float)(cc[ax] + 2) * gm.cs) - qv[ax]);
            }
        }
        float tmax = lim;
        if (b != __uint_as_float(kInfBits)) {
            const float bs = fmaxf(0.f, b - 2e-3f * gm.cs);
            const float b2 = bs * bs * (1.f - 2e-6f);
            tmax = fminf(lim, __uint_as_float(max(__float_as_uint(b2), 1u) - 1u));
        }
        float tau = 0.f;
        if (n27 > 0) {
            const int nxr = min(cx + 1, gm.g[0] - 1) - max(cx - 1, 0) + 1, nyr = min(cy + 1, gm.g[1] - 1) - max(cy - 1, 0) + 1,
                      nzr = min(cz + 1, gm.g[2] - 1) - max(cz - 1, 0) + 1;
            const float vol = (float)(nxr * nyr * nzr) * gm.cs * gm.cs * gm.cs;
            const float r0 = cbrtf(g.fast_target * vol / (4.18879f * (float)n27));
            tau = fminf(r0 * r0, tmax);
        }
        const bool scan = active && n27 > 0;
        touched += scan ? (unsigned long long)n27 : 0ull;

        const long long qb = clock64();
        // ---- pass 1: survivors of the 3x3x3 block
        int c = 0;
#pragma unroll 1
        for (int r9 = 0; r9 < 9; ++r9) {
            const int r = rbase + (r9 / 3) * (T1 + 2) + r9 % 3;
            int j = 0, e = 0;
            if (scan) {
                j = cs[r * nb + lx - 1];
                e = cs[r * nb + lx + 2];
            }
            // two nodes per trip, both loads first; branch-free (predicated) body
            while (__any_sync(kFull, j < e)) {
                const bool v0 = j < e, v1 = j + 1 < e;
                const float4 p0 = nodes[v0 ? j : 0], p1 = nodes[v1 ? j + 1 : 0];
                const float d0 = dist2(p0), d1 = dist2(p1);
                const bool s0 = v0 && d0 <= tau && d0 > eps_lo;
                if (s0 && c < kTileSlots) buf[c * 32 + lane] = (uint16_t)j;
                c += s0 ? 1 : 0;
                const bool s1 = v1 && d1 <= tau && d1 > eps_lo;
                if (s1 && c < kTileSlots) buf[c * 32 + lane] = (uint16_t)(j + 1);
                c += s1 ? 1 : 0;
                j += 2;
            }
        }
        // complete answer: at least k survivors, or everything inside the radius limit has been seen
        bool ok = scan && c <= kTileSlots && (c >= g.k || (tau >= tmax && tmax >= lim));
        __syncwarp();
        const long long qc = clock64();

        // ---- select: 32-bit keys in registers, the 16 smallest by a compare-exchange network
        const int cmax = min(kTileSlots, __reduce_max_sync(kFull, ok ? c : 0));
        uint32_t top[16];
        uint32_t rest = 0xFFFFFFFFu;
        auto load16 = [&](uint32_t (&v)[16], int base) {
#pragma unroll
            for (int s = 0; s < 16; ++s) {
                v[s] = 0xFFFFFFFFu;
                if (ok && base + s < c) {
                    const float4 p = nodes[buf[(base + s) * 32 + lane]];
                    v[s] = (__float_as_uint(dist2(p)) & ~63u) | (uint32_t)(base + s);
                }
            }
        };
        load16(top, 0);
        sort16(top);
#pragma unroll 1
        for (int base = 16; base < cmax; base += 16) {
            uint32_t grp[16];
            load16(grp, base);
            sort16(grp);
            merge16(top, grp, rest);
        }
        // a tie of the upper 26 bits among the first k+1 keys could hide a misordering
#pragma unroll
        for (int r = 0; r < 16; ++r) {
            const uint32_t nxt = r < 15 ? top[(r + 1) & 15] : rest;
            if (r < g.k && top[r] != 0xFFFFFFFFu && (top[r] >> 6) == (nxt >> 6)) ok = false;
        }
        const long long qd = clock64();
        if (active && !ok) {
            const int s = atomicAdd(a.fb + a.parity, 1);
            a.fb[2 + s] = (int32_t)qi;
        }
        if (ok) {
            int32_t oi[16];
            float od[16];
#pragma unroll
            for (int r = 0; r < 16; ++r) {
                oi[r] = -1;
                od[r] = __uint_as_float(kInfBits);
                if (r < g.k && top[r] != 0xFFFFFFFFu) {
                    const float4 p = nodes[buf[(top[r] & 63u) * 32 + lane]];
                    oi[r] = (int32_t)((int64_t)__float_as_int(p.w) + g.index_base);
                    od[r] = dist2(p);
                }
            }
            int32_t* ri = g.idx + qi * g.k;
            float* rd = g.d2 ? g.d2 + qi * g.k : nullptr;
            if (a.vec_out) {
#pragma unroll
                for (int r = 0; r < 16; r += 4) {
                    if (r < g.k) {
                        *reinterpret_cast<int4*>(ri + r) = make_int4(oi[r], oi[r + 1], oi[r + 2], oi[r + 3]);
                        if (rd) *reinterpret_cast<float4*>(rd + r) = make_float4(od[r], od[r + 1], od[r + 2], od[r + 3]);
                    }
                }
            } else {
#pragma unroll
                for (int r = 0; r < 16; ++r) {
                    if (r < g.k) {
                        ri[r] = oi[r];
                        if (rd) rd[r] = od[r];
                    }
                }
            }
        }
        __syncwarp();
        const long long qe = clock64();
        pq += qb - qa; pp1 += qc - qb; psel += qd - qc; pout += qe - qd;
    }
    if (a.prof && tid == 0) {
        atomicAdd((unsigned long long*)a.prof + 0, (unsigned long long)(pt1 - pt0));
        atomicAdd((unsigned long long*)a.prof + 1, (unsigned long long)(pt2 - pt1));
        atomicAdd((unsigned long long*)a.prof + 2, (unsigned long long)(pt3 - pt2));
        atomicAdd((unsigned long long*)a.prof + 3, (unsigned long long)pq);
        atomicAdd((unsigned long long*)a.prof + 4, (unsigned long long)pp1);
        atomicAdd((unsigned long long*)a.prof + 5, (unsigned long long)psel);
        atomicAdd((unsigned long long*)a.prof + 6, (unsigned long long)pout);
        atomicAdd((unsigned long long*)a.prof + 7, 1ull);
    }
    if (g.counter) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) touched += __shfl_xor_sync(kFull, touched, o);
        if (lane == 0 && touched) atomicAdd(g.counter, touched);
    }
}

// ---------------------------------------------------------------- host side
struct TilePlan {
    bool ok = false;
    int T[3] = {0, 0, 0}, nt[3] = {0, 0, 0};
    int node_cap = 0, list_cap = 0, warps = 1;
    int64_t ntiles = 0;
    size_t smem = 0;
};

// Tile size from the query density (queries are assumed to be spread like the nodes; a batch that is
// not overflows its tile lists into the fallback list and is still answered exactly): the largest tile
// whose expected query count still fits one warp pass, smallest halo overhead first.
inline TilePlan tile_plan(const GridIndex& G, int64_t n, int64_t nq) {
    TilePlan P;
    const GridGeom& gm = G.geom;
    const double ncells = (double)gm.g[0] * gm.g[1] * gm.g[2];
    const double rho = (double)nq / ncells, npc = (double)n / ncells;
    int forced[4] = {G.tile_shape[0], G.tile_shape[1], G.tile_shape[2], G.tile_warps};
    if (const char* e = getenv("BGR_TILE")) sscanf(e, "%d,%d,%d,%d", &forced[0], &forced[1], &forced[2], &forced[3]);
    double best = 1e300;
    for (int t0 = 1; t0 <= kTileMaxT0; ++t0)
        for (int t1 = 1; t1 <= 6; ++t1)
            for (int t2 = 1; t2 <= 6; ++t2) {
                if (forced[0] > 0 && (t0 != forced[0] || t1 != forced[1] || t2 != forced[2])) continue;
                if ((t1 + 2) * (t2 + 2) > kTileMaxRows) continue;
                const double vol = (double)t0 * t1 * t2, qexp = rho * vol;
                const double halo = (double)(t0 + 2) * (t1 + 2) * (t2 + 2), nodes = halo * npc;
                if (nodes * 1.05 + 6.0 * std::sqrt(nodes) + 64.0 > 6144.0) continue;
                if (forced[0] <= 0) {
                    if (qexp + 1.3 * std::sqrt(qexp) > 32.0 && vol > 1.0) continue;   // more than one warp pass: smaller tile
                    if (qexp < 6.0) continue;                                          // lanes mostly idle: not worth staging
                }
                const double cost = halo / vol + 1e-3 * (t1 + 2) * (t2 + 2);
                if (cost < best) {
                    best = cost;
                    P.T[0] = t0; P.T[1] = t1; P.T[2] = t2;
                }
            }
    if (best == 1e300) return P;
    const double vol = (double)P.T[0] * P.T[1] * P.T[2], qexp = rho * vol;
    const double nodes = (double)(P.T[0] + 2) * (P.T[1] + 2) * (P.T[2] + 2) * npc;
    P.node_cap = (int)round_up((int64_t)(nodes * 1.05 + 6.0 * std::sqrt(nodes) + 64.0), 64);
    P.list_cap = (int)round_up((int64_t)(2.0 * qexp + 6.0 * std::sqrt(qexp) + 32.0), 32);
    P.warps = forced[3] > 0 ? std::min(forced[3], kTileMaxWarps) : std::max(1, std::min(kTileMaxWarps, (int)std::ceil(qexp / 26.0)));
    for (int a = 0; a < 3; ++a) P.nt[a] = (gm.g[a] + P.T[a] - 1) / P.T[a];
    P.ntiles = (int64_t)P.nt[0] * P.nt[1] * P.nt[2];
    P.smem = tile_knn_smem(P.node_cap, P.warps);
    P.ok = P.ntiles < (1 << 30) && P.smem <= 200 * 1024;
    return P;
}

inline int tile_knn_launch(GridIndex& G, const GridKnnArgs& a, const TilePlan& P, int device, cudaStream_t st) {
    static std::atomic<bool> attr_set[64];
    if (!attr_set[device & 63].load(std::memory_order_acquire)) {
        BGR_CUDA(cudaFuncSetAttribute(tile_knn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr_set[device & 63].store(true, std::memory_order_release);
    }
    const int32_t* old_count = G.tile_count.p;
    BGR_CHECK(G.tile_count.reserve(P.ntiles, st, false));
    if (G.tile_count.p != old_count || G.tile_zeroed < P.ntiles) {
        BGR_CUDA(cudaMemsetAsync(G.tile_count.p, 0, (size_t)G.tile_count.cap * sizeof(int32_t), st));
        G.tile_zeroed = G.tile_count.cap;
    }
    BGR_CHECK(G.tile_list.reserve(P.ntiles * P.list_cap, st, false));
    const int32_t* old_fb = G.tile_fb.p;
    BGR_CHECK(G.tile_fb.reserve(2 + a.nq, st, false));
    if (G.tile_fb.p != old_fb) {
        BGR_CUDA(cudaMemsetAsync(G.tile_fb.p, 0, 2 * sizeof(int32_t), st));
        G.tile_parity = 0;
    }
    TileKnnArgs t;
    t.g = a;
    t.g.fast = 1;
    t.g.heavy = nullptr; t.g.heavy_cap = 0; t.g.heavy_min = 0;
    t.g.fast_target = (float)a.k + 3.5f * sqrtf((float)a.k) + 6.f;
    for (int i = 0; i < 3; ++i) { t.T[i] = P.T[i]; t.nt[i] = P.nt[i]; }
    t.node_cap = P.node_cap; t.list_cap = P.list_cap;
    t.tile_count = G.tile_count.p; t.tile_list = G.tile_list.p; t.fb = G.tile_fb.p;
    t.parity = G.tile_parity;
    t.prof = nullptr;
    static const bool want_prof = getenv("BGR_TILE_PROF") != nullptr;
    if (want_prof) {
        BGR_CHECK(G.tile_prof.reserve(8, st, false));
        BGR_CUDA(cudaMemsetAsync(G.tile_prof.p, 0, 8 * sizeof(long long), st));
        t.prof = G.tile_prof.p;
    }
    t.vec_out = (a.k % 4 == 0) && ((uintptr_t)a.idx % 16 == 0) && ((uintptr_t)a.d2 % 16 == 0);
    tile_bin_kernel<<<(unsigned)div_up(a.nq, 256), 256, 0, st>>>(t);
    BGR_LAUNCH_CHECK();
    tile_knn_kernel<<<(unsigned)P.ntiles, P.warps * 32, P.smem, st>>>(t);
    BGR_LAUNCH_CHECK();
    GridKnnArgs fa = a;                      // the queries handed back: warp-per-query search, its own threshold
    fa.fast = 1; fa.skip = nullptr;
    fa.heavy = nullptr; fa.heavy_cap = 0; fa.heavy_min = 0;
    const unsigned fgrid = (unsigned)std::min<int64_t>(5LL * num_sms(), div_up(a.nq, kGridWarps));
    grid_knn_list_kernel<<<fgrid, kGridWarps * 32, 0, st>>>(fa, G.tile_fb.p + 2, G.tile_fb.p + G.tile_parity);
    BGR_LAUNCH_CHECK();
    if (want_prof) {
        long long hp[8];
        BGR_CUDA(cudaMemcpyAsync(hp, G.tile_prof.p, sizeof(hp), cudaMemcpyDeviceToHost, st));
        BGR_CUDA(cudaStreamSynchronize(st));
        const double nt = (double)std::max<long long>(1, hp[7]);
        fprintf(stderr, "[tile prof] tiles %lld: setup %.0f issue %.0f wait %.0f | query %.0f pass1 %.0f select %.0f out %.0f cycles per tile (warp 0)\n", hp[7],
                hp[0] / nt, hp[1] / nt, hp[2] / nt, hp[3] / nt, hp[4] / nt, hp[5] / nt, hp[6] / nt);
    }
    G.tile_parity ^= 1;
    G.n_tile++;
    G.tile_last_shape = (uint64_t)P.T[0] | ((uint64_t)P.T[1] << 8) | ((uint64_t)P.T[2] << 16) | ((uint64_t)P.warps << 24);
    G.tile_last_count = (uint64_t)P.ntiles;
    return BGR_OK;
}

inline int grid_knn(GridIndex& G, const float* x, const float* y, const float* z, int64_t n,
                    const float* d_q, int64_t nq, int k, unsigned flags, float max_radius,
                    const float* d_radii, int32_t* d_idx, float* d_d2, int64_t index_base,
                    unsigned long long* counter, int fast, int device, cudaStream_t st, const uint8_t* d_skip = nullptr) {
    GridKnnArgs a;
    a.gm = G.geom;
    a.pts = G.pts.p;
    a.cell_start = G.cell_start.p;
    a.n_indexed = G.n_indexed;
    a.x = x; a.y = y; a.z = z; a.n = n;
    a.q = d_q; a.nq = nq; a.k = k; a.flags = flags;
    a.max_radius = max_radius; a.max_radii = d_radii;
    a.idx = d_idx; a.d2 = d_d2; a.index_base = index_base; a.counter = counter;
    a.fast = fast;
    a.skip = d_skip;
    a.fast_target = (float)k + 2.5f * sqrtf((float)k) + 1.f;
    if (const char* e = getenv("BGR_FAST_TARGET")) { const float f = (float)atof(e); if (f > 0.f) a.fast_target = (float)k + f * sqrtf((float)k) + 1.f; }   // measurement aid
    a.heavy = nullptr; a.heavy_cap = 0; a.heavy_min = 0;
    const bool crowded = G.crowd > 64.0;
    // large batches over a fully indexed, uncrowded cloud: one lane per query out of staged tiles
    if (fast == 1 && G.tile_shape[0] >= 0 && !crowded && G.n_indexed > 0 && G.n_indexed == n && k <= 16 && nq >= 2048) {
        const TilePlan P = tile_plan(G, n, nq);
        if (P.ok) return tile_knn_launch(G, a, P, device, st);
    }
    // piled-up index (crowd: occupancy of the cell an average node sits in; ~9 for a uniform cloud):
    // queries whose ring-1 block is a pile go to the CTA-per-query scan
    if (fast && G.n_indexed > 0 && crowded) {
        BGR_CHECK(G.heavy.reserve(kHeavyInts, st, false));
        BGR_CUDA(cudaMemsetAsync(G.heavy.p, 0, (1 + 2 * kHeavyCap) * sizeof(int32_t), st));
        a.heavy = G.heavy.p; a.heavy_cap = kHeavyCap;
        a.heavy_min = std::max<int64_t>(2048, n / 128);
    }
    // fully indexed, uncrowded cloud (no unsorted tail, no piles): several queries per warp (grid_group.cuh);
    // BGR_KNN_GROUP=0 keeps the warp-per-query kernel, 16 / 8 picks the lanes per query (measurement / test aid, read per call)
    int group = (fast == 1 && !a.heavy && G.n_indexed > 0 && G.n_indexed == n && nq >= 2) ? kGroupLanesDefault : 0;
    if (group) { const char* e = getenv("BGR_KNN_GROUP"); if (e) { const int v = atoi(e); if (v == 0 || v == 8 || v == 16) group = v; } }
    if (a.heavy) grid_knn_kernel<true><<<(unsigned)div_up(nq, 8), 256, 0, st>>>(a);
    else if (group == 16) grid_knn_group_kernel<16><<<(unsigned)div_up(nq, 2 * kGroupWarps), kGroupWarps * 32, 0, st>>>(a);
    else if (group == 8) grid_knn_group_kernel<8><<<(unsigned)div_up(nq, 4 * kGroupWarps), kGroupWarps * 32, 0, st>>>(a);
    else grid_knn_kernel<false><<<(unsigned)div_up(nq, 8), 256, 0, st>>>(a);
    BGR_LAUNCH_CHECK();
    if (a.heavy) G.n_piled++; else G.n_lean++;
    if (a.heavy) {
        grid_knn_heavy_kernel<<<dim3(kHeavyCap, kHeavySlices), kHeavyWarps * 32, 0, st>>>(a);
        BGR_LAUNCH_CHECK();
    }
    return BGR_OK;
}

}  // namespace bgr
