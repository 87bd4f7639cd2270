// Sphere state check (isStateValid MS:310-351) as a warp-level device function: shared by
// collide_spheres_kernel (map.cu) and the planner's persistent small-batch kernel (planner.cu).
#pragma once
#include "mapview.cuh"

namespace bgr {

// Exact closed form: gap = dist(centre, leaf box) - r in double; hit = gap < 0.
__device__ __forceinline__ int nth_set_bit64(unsigned long long v, int r) {   // r-th (0-based) set bit
    // popcount bisection (the __fns intrinsic is a long software loop)
    unsigned w = (unsigned)v;
    int base = 0, c = __popc(w);
    if (r >= c) { r -= c; w = (unsigned)(v >> 32); base = 32; }
    c = __popc(w & 0xFFFFu); if (r >= c) { r -= c; w >>= 16; base += 16; }
    c = __popc(w & 0xFFu);   if (r >= c) { r -= c; w >>= 8;  base += 8; }
    c = __popc(w & 0xFu);    if (r >= c) { r -= c; w >>= 4;  base += 4; }
    c = __popc(w & 0x3u);    if (r >= c) { r -= c; w >>= 2;  base += 2; }
    return base + ((r >= (int)(w & 1u)) ? 1 : 0);
}

// The cells of the sphere's box (radius + margin) are read as whole 4x4x4 brick words, one brick per
// lane (a 0.1 m sphere at 0.05 m cells touches <= 3x3x3 bricks: 27 loads instead of 125+ bit probes),
// masked to the box; the set bits are spread over the lanes: exact point-to-cell distance, min over
// the warp.  G = spheres per warp: with G = 32 (large batches) each lane first runs the gate of its
// own sphere and the warp then walks only the undecided ones, one after the other; G = 1 (small
// batches, e.g. a planner step) keeps one warp per sphere so the walks run side by side.
template <int G>
// (no __restrict__ here: the planner's persistent kernel rewrites the centres between calls inside one launch,
// so they must not be read through the non-coherent path; the kernel's own parameters keep theirs)
__device__ __forceinline__ void collide_spheres_warp(const MapView& m, const float* c3, int64_t n, double radius,
                                                     uint8_t* hit, uint8_t* band, const LevelGrid& dil,
                                                     const int64_t first, const int lane) {
    const int64_t my = first + (G == 1 ? 0 : lane);
    double mc[3] = {0.0, 0.0, 0.0};
    bool need = false;
    if (my < n) {
        mc[0] = (double)c3[3 * my]; mc[1] = (double)c3[3 * my + 1]; mc[2] = (double)c3[3 * my + 2];
        need = true;
        const int gx = (int)floor((mc[0] - m.origin[0]) * m.inv_res), gy = (int)floor((mc[1] - m.origin[1]) * m.inv_res),
                  gz = (int)floor((mc[2] - m.origin[2]) * m.inv_res);
        if (gx >= 0 && gx < m.dims[0] && gy >= 0 && gy < m.dims[1] && gz >= 0 && gz < m.dims[2]) {
            // gate: level 0 dilated by ceil((radius + 1e-4) / res) cells.  The centre's cell clear in it =>
            // no leaf within radius + 1e-4 of the centre: free, and not a tolerance-band contact.
            // The centre's own cell occupied => distance 0, gap = -radius: decided without the box walk.
            const int64_t w = ((int64_t)(gz >> 2) * m.lv[0].nb[1] + (gy >> 2)) * m.lv[0].nb[0] + (gx >> 2);   // same geometry in both maps
            const int b = ((gz & 3) << 4) | ((gy & 3) << 2) | (gx & 3);
            const bool clear = dil.bits && !((__ldg(dil.bits + w) >> b) & 1ull);
            if (clear || ((__ldg(m.lv[0].bits + w) >> b) & 1ull)) {
                if (G != 1 || lane == 0) {
                    hit[my] = (!clear && radius > 0.0) ? 1 : 0;
                    if (band) band[my] = (!clear && radius <= kBandTol) ? 1 : 0;
                }
                need = false;
            }
        }
    }
    unsigned pending = __ballot_sync(kFull, need);
    if (G == 1) pending &= 1u;
    while (pending) {
    const int src = __ffs(pending) - 1;
    pending &= pending - 1u;
    const int64_t i = first + src;
    const double c[3] = {__shfl_sync(kFull, mc[0], src), __shfl_sync(kFull, mc[1], src), __shfl_sync(kFull, mc[2], src)};
    // a cell closer than radius - tolerance decides the query (hit, not in the band): stop looking
    const double thr2 = radius > kBandTol ? (radius - kBandTol) * (radius - kBandTol) * (1.0 - 1e-12) : -1.0;
    bool decided = false;
    int lo[3], hi[3];
    bool empty = false;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const double r = radius + kSweepMargin;
        // cells outside [c - r, c + r] are farther than radius + 1e-4: they decide neither hit nor band
        // (the 1e-4 m margin is ~1e9 times the rounding of the cell computation)
        const double g = (c[a] - m.origin[a]) * m.inv_res, rc = r * m.inv_res;
        lo[a] = max(0, (int)floor(g - rc));
        hi[a] = min(m.dims[a] - 1, (int)floor(g + rc));
        empty |= lo[a] > hi[a];
    }
    double d2min = 1e300;
    if (!empty) {
        const int bl[3] = {lo[0] >> 2, lo[1] >> 2, lo[2] >> 2};
        const int nbr[3] = {(hi[0] >> 2) - bl[0] + 1, (hi[1] >> 2) - bl[1] + 1, (hi[2] >> 2) - bl[2] + 1};
        const int total = nbr[0] * nbr[1] * nbr[2];
        const int n01 = nbr[0] * nbr[1];
        const float inv0 = __frcp_rn((float)nbr[0]), inv01 = __frcp_rn((float)n01);   // exact quotients below 2^20 bricks
        const double half = 0.5 * m.res;
        for (int t0 = 0; t0 < total && !decided; t0 += 32) {
            // ---- one brick per lane: its word masked to the box
            const int t = t0 + lane;
            unsigned long long W = 0ull;
            int bxyz = 0;
            if (t < total) {
                int oz, oy, ox;
                if (total < (1 << 20)) {
                    oz = (int)(((float)t + 0.5f) * inv01);
                    const int rem = t - oz * n01;
                    oy = (int)(((float)rem + 0.5f) * inv0);
                    ox = rem - oy * nbr[0];
                } else {
                    oz = t / n01; oy = (t / nbr[0]) % nbr[1]; ox = t % nbr[0];
                }
                const int bx = bl[0] + ox, by = bl[1] + oy, bz = bl[2] + oz;
                W = __ldg(m.lv[0].bits + ((int64_t)bz * m.lv[0].nb[1] + by) * m.lv[0].nb[0] + bx);
                // cells of this brick inside the box, per axis (x: nibble pattern, y: 4-bit groups, z: 16-bit groups)
                const int x0 = max(lo[0] - 4 * bx, 0), x1 = min(hi[0] - 4 * bx, 3);
                const int y0 = max(lo[1] - 4 * by, 0), y1 = min(hi[1] - 4 * by, 3);
                const int z0 = max(lo[2] - 4 * bz, 0), z1 = min(hi[2] - 4 * bz, 3);
                const unsigned long long mx = (unsigned long long)(((2u << x1) - (1u << x0)) & 0xFu) * 0x1111111111111111ull;
                const unsigned long long my = (unsigned long long)(((1u << (4 * y1 + 4)) - (1u << (4 * y0))) & 0xFFFFu) * 0x0001000100010001ull;
                const unsigned long long mz = (z1 == 3 ? ~0ull : ((1ull << (16 * z1 + 16)) - 1ull)) & ~((1ull << (16 * z0)) - 1ull);
                W &= mx & my & mz;
                bxyz = bx | (by << 10) | (bz << 20);
            }
            // ---- the set bits of the 32 words, one per lane (dense: the words are mostly empty or sparse)
            const int cnt = __popcll(W);
            int incl = cnt;
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(kFull, incl, o);
                if (lane >= o) incl += v;
            }
            const int T = __shfl_sync(kFull, incl, 31);
            for (int j0 = 0; j0 < T; j0 += 32) {
                const int j = j0 + lane;
                int src = 0;
#pragma unroll
                for (int step = 16; step > 0; step >>= 1) {
                    const int v = __shfl_sync(kFull, incl, src + step - 1);
                    if (v <= j) src += step;
                }
                src = min(src, 31);
                const int s_incl = __shfl_sync(kFull, incl, src), s_cnt = __shfl_sync(kFull, cnt, src);
                const unsigned long long s_W = __shfl_sync(kFull, W, src);
                const int s_b = __shfl_sync(kFull, bxyz, src);
                if (j < T) {
                    // The leaves are unions of their finest cells (level 0 holds every cell of a pruned leaf), so
                    // the distance to the nearest leaf is the distance to the nearest occupied CELL: no leaf lookup.
                    const int bit = nth_set_bit64(s_W, j - (s_incl - s_cnt));
                    const int x = 4 * (s_b & 1023) + (bit & 3), y = 4 * ((s_b >> 10) & 1023) + ((bit >> 2) & 3), z = 4 * (s_b >> 20) + (bit >> 4);
                    const double ccx = m.origin[0] + ((double)x + 0.5) * m.res, ccy = m.origin[1] + ((double)y + 0.5) * m.res,
                                 ccz = m.origin[2] + ((double)z + 0.5) * m.res;
                    const double tx = fmax(0.0, fabs(c[0] - ccx) - half), ty = fmax(0.0, fabs(c[1] - ccy) - half),
                                 tz = fmax(0.0, fabs(c[2] - ccz) - half);
                    double d2 = tx * tx;
                    d2 += ty * ty;
                    d2 += tz * tz;
                    d2min = fmin(d2min, d2);
                }
                if (__any_sync(kFull, d2min < thr2)) { decided = true; break; }
            }
        }
    }
    if (decided) {
        if (lane == 0) {
            hit[i] = 1;
            if (band) band[i] = 0;
        }
        continue;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) d2min = fmin(d2min, __shfl_xor_sync(kFull, d2min, o));
    const double gmin = d2min < 1e299 ? sqrt(d2min) - radius : 1e300;   // sqrt is monotone: min of the distances
    if (lane == 0) {
        hit[i] = gmin < 0.0 ? 1 : 0;
        if (band) band[i] = fabs(gmin) <= kBandTol ? 1 : 0;
    }
    }   // while (pending)
}

// ---------------------------------------------------------------------------------------------------------
// Lane-per-sphere form for LARGE batches of small spheres (box of radius + margin spans <= kLaneSpan cells per
// axis, i.e. <= 3x3x3 bricks).  collide_spheres_warp<32> walks every undecided sphere with the whole warp: ~250
// warp instructions per sphere, most of them the warp-wide bookkeeping of spreading ~13 occupied cells over 32
// lanes (prefix scan, binary search, n-th set bit), at 80 % of the issue slots (profiles/r2_ncu_full_s_edges_spheres.txt).
// With a dilation of 3 cells the gate decides only the ~10 % of spheres whose own cell is occupied, so nearly every
// sphere is walked.  Here a lane keeps its sphere: it reads the <= 27 brick words itself, masks them to the box and
// walks their set bits; the squared per-axis distances of the <= 8 cell coordinates per axis are tabulated once
// per sphere in shared memory (the same double operations, in the same order, as the warp form: identical bits),
// so an occupied cell costs three LDS.64 and two DADD.  No shuffles, no ballots.
constexpr int kLaneSpan = 8;
constexpr int kLaneThreads = 128;

// largest radius the lane form takes on a map of resolution res (the box spans floor(2 rc) + 2 cells at most)
__host__ __device__ inline bool spheres_lane_ok(double radius, double inv_res) {
    return 2.0 * (radius + kSweepMargin) * inv_res + 2.0 <= (double)kLaneSpan;
}

static __global__ void __launch_bounds__(kLaneThreads) collide_spheres_lane_kernel(MapView m, const float* __restrict__ c3, int64_t n,
                                                                            double radius, uint8_t* __restrict__ hit,
                                                                            uint8_t* __restrict__ band, LevelGrid dil) {
    __shared__ double t2[3 * kLaneSpan][kLaneThreads];     // [axis * kLaneSpan + cell - lo][thread]
    const int tid = threadIdx.x;
    const int64_t i = (int64_t)blockIdx.x * kLaneThreads + tid;
    if (i >= n) return;
    const double c[3] = {(double)c3[3 * i], (double)c3[3 * i + 1], (double)c3[3 * i + 2]};
    {
        const int gx = (int)floor((c[0] - m.origin[0]) * m.inv_res), gy = (int)floor((c[1] - m.origin[1]) * m.inv_res),
                  gz = (int)floor((c[2] - m.origin[2]) * m.inv_res);
        if (gx >= 0 && gx < m.dims[0] && gy >= 0 && gy < m.dims[1] && gz >= 0 && gz < m.dims[2]) {
            // the gate of collide_spheres_warp: centre cell clear in the dilated map => free; occupied => distance 0
            const int64_t w = ((int64_t)(gz >> 2) * m.lv[0].nb[1] + (gy >> 2)) * m.lv[0].nb[0] + (gx >> 2);
            const int b = ((gz & 3) << 4) | ((gy & 3) << 2) | (gx & 3);
            const bool clear = dil.bits && !((__ldg(dil.bits + w) >> b) & 1ull);
            if (clear || ((__ldg(m.lv[0].bits + w) >> b) & 1ull)) {
                hit[i] = (!clear && radius > 0.0) ? 1 : 0;
                if (band) band[i] = (!clear && radius <= kBandTol) ? 1 : 0;
                return;
            }
        }
    }
    const double thr2 = radius > kBandTol ? (radius - kBandTol) * (radius - kBandTol) * (1.0 - 1e-12) : -1.0;
    int lo[3], hi[3];
    bool empty = false;
    const double half = 0.5 * m.res;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const double r = radius + kSweepMargin;
        const double g = (c[a] - m.origin[a]) * m.inv_res, rc = r * m.inv_res;
        lo[a] = max(0, (int)floor(g - rc));
        hi[a] = min(m.dims[a] - 1, (int)floor(g + rc));
        empty |= lo[a] > hi[a];
        // squared distance along this axis from the centre to each cell of the box
#pragma unroll
        for (int k = 0; k < kLaneSpan; ++k) {
            if (lo[a] + k <= hi[a]) {
                const double cc = m.origin[a] + ((double)(lo[a] + k) + 0.5) * m.res;
                const double t = fmax(0.0, fabs(c[a] - cc) - half);
                t2[a * kLaneSpan + k][tid] = t * t;
            }
        }
    }
    double d2min = 1e300;
    if (!empty) {
        const int bl[3] = {lo[0] >> 2, lo[1] >> 2, lo[2] >> 2};
        const int bh[3] = {hi[0] >> 2, hi[1] >> 2, hi[2] >> 2};
        for (int bz = bl[2]; bz <= bh[2]; ++bz) {
            const int z0 = max(lo[2] - 4 * bz, 0), z1 = min(hi[2] - 4 * bz, 3);
            const unsigned long long mz = (z1 == 3 ? ~0ull : ((1ull << (16 * z1 + 16)) - 1ull)) & ~((1ull << (16 * z0)) - 1ull);
            const int kz = 2 * kLaneSpan + 4 * bz - lo[2];
            for (int by = bl[1]; by <= bh[1]; ++by) {
                const int y0 = max(lo[1] - 4 * by, 0), y1 = min(hi[1] - 4 * by, 3);
                const unsigned long long myz =
                    ((unsigned long long)(((1u << (4 * y1 + 4)) - (1u << (4 * y0))) & 0xFFFFu) * 0x0001000100010001ull) & mz;
                const int ky = kLaneSpan + 4 * by - lo[1];
                const uint64_t* row = m.lv[0].bits + ((int64_t)bz * m.lv[0].nb[1] + by) * m.lv[0].nb[0];
                // the (up to three) words of this brick row are loaded together, then walked one after the other
                unsigned long long Wr[3];
#pragma unroll
                for (int j = 0; j < 3; ++j) Wr[j] = bl[0] + j <= bh[0] ? __ldg(row + bl[0] + j) : 0ull;
#pragma unroll
                for (int j = 0; j < 3; ++j) {
                    const int bx = bl[0] + j;
                    const int x0 = max(lo[0] - 4 * bx, 0), x1 = min(hi[0] - 4 * bx, 3);
                    const unsigned long long mx = x1 >= x0 ? (unsigned long long)(((2u << x1) - (1u << x0)) & 0xFu) * 0x1111111111111111ull : 0ull;
                    const unsigned long long W = Wr[j] & mx & myz;
                    const int kx = 4 * bx - lo[0];
                    // the set bits, 32 at a time (z planes 0-1, then 2-3).  A cell closer than radius - tolerance
                    // decides the sphere (hit, not in the band): stop at once — the lanes next to a clump of
                    // occupied cells would otherwise walk dozens of bits while the rest of the warp waits
#pragma unroll
                    for (int hw = 0; hw < 2; ++hw) {
                        unsigned w = (unsigned)(W >> (32 * hw));
                        while (w) {
                            const int bit = __ffs((int)w) - 1;
                            w &= w - 1u;
                            double d2 = t2[kx + (bit & 3)][tid];
                            d2 += t2[ky + ((bit >> 2) & 3)][tid];
                            d2 += t2[kz + 2 * hw + (bit >> 4)][tid];
                            if (d2 < thr2) {
                                hit[i] = 1;
                                if (band) band[i] = 0;
                                return;
                            }
                            d2min = fmin(d2min, d2);
                        }
                    }
                }
            }
        }
    }
    const double gmin = d2min < 1e299 ? sqrt(d2min) - radius : 1e300;
    hit[i] = gmin < 0.0 ? 1 : 0;
    if (band) band[i] = fabs(gmin) <= kBandTol ? 1 : 0;
}

}  // namespace bgr
