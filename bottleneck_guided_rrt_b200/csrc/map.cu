// bgr_map_* — Octomap occupied leaves flattened to bit grids in HBM + batched collision
// queries.  Reference surface: octomap_Callback/generateBoxesFromOctomap MS:273-306,
// isStateValid MS:310-351, isLineValid MS:385-471.
//
// Data layout in HBM
//   level 0 : one bit per finest cell, 4x4x4 cells per 64-bit brick word
//             word = (bz*nby + by)*nbx + bx,  bit = (z&3)<<4 | (y&3)<<2 | (x&3)
//             (512^3 cells -> 128^3 words = 16 MiB, L2 resident on B200)
//   level L : same brick format over the grid of aligned 2^L-cell blocks; a set bit says
//             "one pruned Octomap leaf of size res*2^L covers exactly this block".
//             Together they are the sparse leaf-box list in implicit form: an occupied cell
//             finds its leaf with at most kMaxLevel extra bit probes, and the narrow phase
//             tests the SAME boxes the reference gets from fcl::OcTree::toBoxes().
//
// Edge query = major-axis DDA sweep (supercover): an edge is cut into slabs one cell thick
// along its dominant axis; each (edge, slab) work item covers a <= few-cell rectangle in the
// other two axes (conservative interval bounds around the capsule hull of the shape), so the
// items are independent and are spread over the lanes of a CTA by a prefix sum instead of
// being walked serially per edge (Amanatides-Woo's tMax stepping is inherently sequential).
// Occupied candidate cells are classified by two closed-form ball tests (deep hit / clear
// miss, margin 1e-5 m >> MPR tolerance); only ambiguous ones run the exact MPR narrow phase.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <iterator>
#include <vector>
#include "spheres.cuh"

using namespace bgr;

namespace {

// ------------------------------------------------------------------ rasterisation
struct LeafRec {
    int32_t x, y, z, level;   // min corner (finest cells), level
};

__global__ void raster_kernel(MapView m, const LeafRec* __restrict__ recs, int64_t n, int set) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const LeafRec r = recs[i];
    const int L = r.level;
    if (L >= 1) {
        const LevelGrid& g = m.lv[L];
        const int X = r.x >> L, Y = r.y >> L, Z = r.z >> L;
        const int64_t w = ((int64_t)(Z >> 2) * g.nb[1] + (Y >> 2)) * g.nb[0] + (X >> 2);
        const unsigned long long bit = 1ull << (((Z & 3) << 4) | ((Y & 3) << 2) | (X & 3));
        if (set) atomicOr((unsigned long long*)g.bits + w, bit);
        else atomicAnd((unsigned long long*)g.bits + w, ~bit);
    }
    const LevelGrid& g0 = m.lv[0];
    if (L >= 2) {
        const int nbk = 1 << (L - 2);
        const int bx0 = r.x >> 2, by0 = r.y >> 2, bz0 = r.z >> 2;
        for (int bz = 0; bz < nbk; ++bz)
            for (int by = 0; by < nbk; ++by)
                for (int bx = 0; bx < nbk; ++bx) {
                    const int64_t w = ((int64_t)(bz0 + bz) * g0.nb[1] + (by0 + by)) * g0.nb[0] + (bx0 + bx);
                    g0.bits[w] = set ? ~0ull : 0ull;
                }
    } else {
        unsigned long long mask = 0;
        const int s = 1 << L;
        for (int dz = 0; dz < s; ++dz)
            for (int dy = 0; dy < s; ++dy)
                for (int dx = 0; dx < s; ++dx)
                    mask |= 1ull << ((((r.z + dz) & 3) << 4) | (((r.y + dy) & 3) << 2) | ((r.x + dx) & 3));
        const int64_t w = ((int64_t)(r.z >> 2) * g0.nb[1] + (r.y >> 2)) * g0.nb[0] + (r.x >> 2);
        if (set) atomicOr((unsigned long long*)g0.bits + w, mask);
        else atomicAnd((unsigned long long*)g0.bits + w, ~mask);
    }
}

__global__ void popcount_kernel(const uint64_t* __restrict__ bits, int64_t n, unsigned long long* out) {
    unsigned long long c = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        c += __popcll(bits[i]);
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(kFull, c, o);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, c);
}

// ------------------------------------------------------------------ spheres (isStateValid): spheres.cuh
template <int G>
__global__ void __launch_bounds__(128) collide_spheres_kernel(MapView m, const float* __restrict__ c3,
                                                              int64_t n, double radius,
                                                              uint8_t* __restrict__ hit,
                                                              uint8_t* __restrict__ band, LevelGrid dil) {
    const int64_t first = (((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5) * G;
    if (first >= n) return;
    collide_spheres_warp<G>(m, c3, n, radius, hit, band, dil, first, (int)(threadIdx.x & 31));
}

// ------------------------------------------------------------------ edges (isLineValid)
constexpr int kEdgeThreads = 256;
constexpr int kEdgesPerCta = 256;
constexpr int kQueueCap = 2048;

struct EdgeSetup {       // per-edge broadphase frame, shared memory
    double c[3], u[3], a0[3];   // centre, unit axis, endpoint A = c - h*u
    double h;
    int ax, i_lo, nslab;
};

enum { F_HIT = 1, F_DEEP = 2, F_INFL = 4, F_DEFL = 8, F_CAP = 16, F_OVERFLOW = 32 };

struct EdgeArgs {
    MapView m;
    const float* p1;
    const float* p2;
    int64_t n;
    double radius;
    int shape;
    int want_band;
    int only_overflow;           // process only edges whose narrow-phase queue overflowed
    const unsigned* n_overflow;  // with only_overflow: number of such edges (0: every CTA returns at once)
    unsigned* gflags;            // per-edge result flags (F_*), consumed by edges_finalize_kernel
    unsigned long long* stats;   // [4]
};

__device__ __forceinline__ double gap_center_ball(const EdgeSetup& s, double radius, int shape, const D3& bc,
                                                  double dr) {
    // signed "uniform inflation" distance from the shape inflated by dr to the point bc
    const double wx = bc.x - s.c[0], wy = bc.y - s.c[1], wz = bc.z - s.c[2];
    const double a = wx * s.u[0] + wy * s.u[1] + wz * s.u[2];
    if (shape == BGR_SHAPE_CAPSULE) {
        const double t = fmax(-s.h, fmin(s.h, a));
        const double ex = wx - t * s.u[0], ey = wy - t * s.u[1], ez = wz - t * s.u[2];
        return sqrt(ex * ex + ey * ey + ez * ez) - (radius + dr);
    }
    const double ex = wx - a * s.u[0], ey = wy - a * s.u[1], ez = wz - a * s.u[2];
    const double rho = sqrt(ex * ex + ey * ey + ez * ez);
    const double da = fmax(0.0, fabs(a) - (s.h + dr)), dq = fmax(0.0, rho - (radius + dr));
    return sqrt(da * da + dq * dq);   // Euclidean distance to the (inflated) cylinder, 0 inside
}

__global__ void __launch_bounds__(kEdgeThreads) collide_edges_kernel(EdgeArgs a) {
    __shared__ EdgeSetup es[kEdgesPerCta];
    __shared__ int prefix[kEdgesPerCta + 1];
    __shared__ unsigned flags[kEdgesPerCta];
    __shared__ unsigned long long queue[kQueueCap];
    __shared__ int qcount;
    __shared__ int warp_sums[kEdgeThreads / 32];
    __shared__ unsigned long long st_items, st_cand, st_mpr, st_cap;

    const MapView& m = a.m;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (a.only_overflow && a.n_overflow && __ldg(a.n_overflow) == 0u) return;
    // blocks of kEdgesPerCta edges, strided over the grid (the overflow pass runs on a small grid: it
    // almost always returns on the line above, and 3,907 empty CTAs cost 12 us)
    const int64_t nblk = (a.n + kEdgesPerCta - 1) / kEdgesPerCta;
    for (int64_t blk = blockIdx.x; blk < nblk; blk += gridDim.x) {
    __syncthreads();                       // the previous block's shared state has been consumed
    const int64_t e0 = blk * kEdgesPerCta;
    const int ne = (int)min((int64_t)kEdgesPerCta, a.n - e0);
    const bool mine = tid < ne && (!a.only_overflow || (a.gflags[e0 + tid] & F_OVERFLOW));
    if (!__syncthreads_or(mine)) continue;
    if (tid == 0) { qcount = 0; st_items = st_cand = st_mpr = st_cap = 0; }

    // ---- phase 0: per-edge setup (thread per edge)
    int myslabs = 0;
    flags[tid] = 0;
    if (mine) {
        const int64_t e = e0 + tid;
        const float p1[3] = {a.p1[3 * e], a.p1[3 * e + 1], a.p1[3 * e + 2]};
        const float p2[3] = {a.p2[3 * e], a.p2[3 * e + 1], a.p2[3 * e + 2]};
        const ShapeObj o = make_edge_obj(p1, p2, a.radius, a.shape);
        D3 u = {0, 0, 1};
        quat_rot(u, o.rot);
        d3_scale(u, 1.0 / sqrt(d3_dot(u, u)));
        EdgeSetup s;
        s.c[0] = o.pos.x; s.c[1] = o.pos.y; s.c[2] = o.pos.z;
        s.u[0] = u.x; s.u[1] = u.y; s.u[2] = u.z;
        s.h = o.height;
        for (int k = 0; k < 3; ++k) s.a0[k] = s.c[k] - s.h * s.u[k];
        s.ax = 0;
        if (fabs(s.u[1]) > fabs(s.u[s.ax])) s.ax = 1;
        if (fabs(s.u[2]) > fabs(s.u[s.ax])) s.ax = 2;
        const double rm = a.radius + kSweepMargin;
        const double pa = s.a0[s.ax], pb = s.c[s.ax] + s.h * s.u[s.ax];
        int lo = (int)floor((fmin(pa, pb) - rm - m.origin[s.ax]) * m.inv_res);
        int hi = (int)floor((fmax(pa, pb) + rm - m.origin[s.ax]) * m.inv_res);
        lo = max(lo, 0);
        hi = min(hi, m.dims[s.ax] - 1);
        s.i_lo = lo;
        s.nslab = max(0, hi - lo + 1);
        const bool finite = isfinite(s.c[0] + s.c[1] + s.c[2] + s.h);
        if (!finite) s.nslab = 0;
        es[tid] = s;
        myslabs = s.nslab;
    }
    // block exclusive prefix sum of slab counts
    {
        int inc = myslabs;
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(kFull, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) warp_sums[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            int w = lane < kEdgeThreads / 32 ? warp_sums[lane] : 0;
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(kFull, w, o);
                if (lane >= o) w += t;
            }
            if (lane < kEdgeThreads / 32) warp_sums[lane] = w;
        }
        __syncthreads();
        const int off = warp ? warp_sums[warp - 1] : 0;
        prefix[tid + 1] = off + inc;
        if (tid == 0) prefix[0] = 0;
    }
    __syncthreads();
    const int total = prefix[kEdgesPerCta];

    // ---- phase 1: sweep.  work item = (edge, slab)
    unsigned long long n_cand = 0, n_mpr = 0, n_cap = 0;
    const double rm = a.radius + kSweepMargin;
    const double sqrt3 = 1.7320508075688772;
    for (int w = tid; w < total; w += kEdgeThreads) {
        int lo = 0, hi = kEdgesPerCta;   // largest el with prefix[el] <= w
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (prefix[mid] <= w) lo = mid; else hi = mid;
        }
        const int el = lo;
        const EdgeSetup& s = es[el];
        const int ax = s.ax, bx = (ax + 1) % 3, cx = (ax + 2) % 3;
        const int i = s.i_lo + (w - prefix[el]);
        const double x0 = m.origin[ax] + (double)i * m.res - rm, x1 = m.origin[ax] + (double)(i + 1) * m.res + rm;
        // parameter range of the axis segment A + t*(2h u), t in [0,1], inside [x0,x1] on ax
        const double dax = 2.0 * s.h * s.u[ax];
        double t0 = 0.0, t1 = 1.0;
        if (fabs(dax) > 1e-12) {
            const double ta = (x0 - s.a0[ax]) / dax, tb = (x1 - s.a0[ax]) / dax;
            t0 = fmax(0.0, fmin(ta, tb));
            t1 = fmin(1.0, fmax(ta, tb));
        }
        if (t0 > t1) continue;   // slab only touched by the hull's end caps: covered below by margins
        int r0[3], r1[3];
        r0[ax] = r1[ax] = i;
        bool none = false;
        for (int k = 0; k < 2; ++k) {
            const int b = k ? cx : bx;
            const double pa = s.a0[b] + t0 * 2.0 * s.h * s.u[b], pb = s.a0[b] + t1 * 2.0 * s.h * s.u[b];
            int lo_b = (int)floor((fmin(pa, pb) - rm - m.origin[b]) * m.inv_res);
            int hi_b = (int)floor((fmax(pa, pb) + rm - m.origin[b]) * m.inv_res);
            lo_b = max(lo_b, 0);
            hi_b = min(hi_b, m.dims[b] - 1);
            r0[b] = lo_b;
            r1[b] = hi_b;
            none |= lo_b > hi_b;
        }
        if (none) continue;
        for (int jc = r0[cx]; jc <= r1[cx]; ++jc)
            for (int jb = r0[bx]; jb <= r1[bx]; ++jb) {
                int v[3];
                v[ax] = i; v[bx] = jb; v[cx] = jc;
                if (!test_bit(m.lv[0], v[0], v[1], v[2])) continue;
                int l[3];
                const int L = find_leaf(m, v[0], v[1], v[2], l[0], l[1], l[2]);
                // one representative cell per (item, leaf)
                if (jb != max(r0[bx], l[bx]) || jc != max(r0[cx], l[cx])) continue;
                ++n_cand;
                D3 bc;
                double e;
                leaf_box(m, L, l[0], l[1], l[2], bc, e);
                // certificates from the leaf's circumscribed / inscribed balls
                const double g_out = gap_center_ball(s, a.radius, a.shape, bc, kCertMargin);
                if (g_out > e * sqrt3 * (1.0 + 1e-9)) continue;                 // clear miss
                bool deep = false;
                if (a.shape == BGR_SHAPE_CAPSULE || s.h > kCertMargin) {
                    const double g_in = gap_center_ball(s, a.radius, a.shape, bc, -kCertMargin);
                    deep = a.shape == BGR_SHAPE_CAPSULE ? (g_in < e) : (g_in < e && a.radius > kCertMargin);
                }
                if (deep) {
                    atomicOr(&flags[el], F_HIT | F_DEEP);
                    continue;
                }
                const unsigned long long rec = ((unsigned long long)el << 40) | ((unsigned long long)L << 36) |
                                               ((unsigned long long)l[0] << 24) | ((unsigned long long)l[1] << 12) |
                                               (unsigned long long)l[2];
                const int slot = atomicAdd(&qcount, 1);
                if (slot < kQueueCap) queue[slot] = rec;
                else {
                    // queue overflow: run the narrow phase inline (rare)
                    const int64_t eg = e0 + el;
                    const float p1[3] = {a.p1[3 * eg], a.p1[3 * eg + 1], a.p1[3 * eg + 2]};
                    const float p2[3] = {a.p2[3 * eg], a.p2[3 * eg + 1], a.p2[3 * eg + 2]};
                    ShapeObj o = make_edge_obj(p1, p2, a.radius, a.shape);
                    int capped = 0;
                    unsigned f = 0;
                    if (mpr_shape_box(o, bc, e, kBandTol, &capped)) f |= F_HIT;
                    ++n_mpr;
                    if (a.want_band) {
                        ShapeObj oi = o, od = o;
                        oi.radius += kBandFlag; od.radius = fmax(0.0, od.radius - kBandFlag);
                        if (o.type == kShapeCyl) { oi.height += kBandFlag; od.height = fmax(0.0, od.height - kBandFlag); }
                        if (mpr_shape_box(oi, bc, e, kBandTol, &capped)) f |= F_INFL;
                        if (mpr_shape_box(od, bc, e, kBandTol, &capped)) f |= F_DEFL;
                        n_mpr += 2;
                    }
                    if (capped) { f |= F_CAP; ++n_cap; }
                    if (f) atomicOr(&flags[el], f);
                }
            }
    }
    __syncthreads();

    // ---- phase 2: exact narrow phase on the queued (edge, leaf) pairs
    const int nq = min(qcount, kQueueCap);
    for (int ci = tid; ci < nq; ci += kEdgeThreads) {
        const unsigned long long rec = queue[ci];
        const int el = (int)(rec >> 40);
        if (flags[el] & F_DEEP) continue;
        const int L = (int)((rec >> 36) & 15), lx = (int)((rec >> 24) & 4095), ly = (int)((rec >> 12) & 4095),
                  lz = (int)(rec & 4095);
        D3 bc;
        double e;
        leaf_box(m, L, lx, ly, lz, bc, e);
        const int64_t eg = e0 + el;
        const float p1[3] = {a.p1[3 * eg], a.p1[3 * eg + 1], a.p1[3 * eg + 2]};
        const float p2[3] = {a.p2[3 * eg], a.p2[3 * eg + 1], a.p2[3 * eg + 2]};
        ShapeObj o = make_edge_obj(p1, p2, a.radius, a.shape);
        int capped = 0;
        unsigned f = 0;
        const bool h = mpr_shape_box(o, bc, e, kBandTol, &capped) != 0;
        ++n_mpr;
        if (h) f |= F_HIT;
        if (a.want_band) {
            // flag results that flip under a +-2.5e-6 m inflation (covers MPR's 1e-6 band)
            ShapeObj o2 = o;
            if (h) {
                o2.radius = fmax(0.0, o2.radius - kBandFlag);
                if (o.type == kShapeCyl) o2.height = fmax(0.0, o2.height - kBandFlag);
                if (mpr_shape_box(o2, bc, e, kBandTol, &capped)) f |= F_DEFL;
            } else {
                o2.radius += kBandFlag;
                if (o.type == kShapeCyl) o2.height += kBandFlag;
                if (mpr_shape_box(o2, bc, e, kBandTol, &capped)) f |= F_INFL;
            }
            ++n_mpr;
        }
        if (capped) { f |= F_CAP; ++n_cap; }
        if (f) atomicOr(&flags[el], f);
    }
    __syncthreads();

    if (mine) a.gflags[e0 + tid] = flags[tid];
    if (a.stats) {
        atomicAdd(&st_cand, n_cand);
        atomicAdd(&st_mpr, n_mpr);
        atomicAdd(&st_cap, n_cap);
        __syncthreads();
        if (tid == 0) {
            atomicAdd(a.stats + 0, (unsigned long long)total);
            atomicAdd(a.stats + 1, st_cand);
            atomicAdd(a.stats + 2, st_mpr);
            atomicAdd(a.stats + 3, st_cap);
        }
    }
    }   // for blk
}

// ------------------------------------------------------------------ edges, fast path
// sweep kernel  : items (edge, slab) -> 64-bit occupancy mask of the item's <=8x8-cell
//                 rectangle; the set bits of the 32 items of a warp are then processed
//                 DENSELY (one candidate per lane, prefix sum + n-th-set-bit) so the
//                 double-precision certificate math runs with full warps.  Ambiguous
//                 (edge, leaf) pairs go to a global queue (one atomicAdd per warp).
// narrow kernel : grid-stride over the queue, exact MPR (+ band flag), atomicOr on flags.
// finalize      : flags -> hit / in_band bytes.
constexpr int kSweepThreads = 256;
constexpr int kSweepEdges = 256;
constexpr int kRectMax = 8;
constexpr int kSqCap = 1024;      // shared-memory queue of ambiguous (edge, leaf) pairs per CTA

struct SweepEdge {
    double c[3], u[3], h;
    double sb, sc;      // lateral slopes d(b)/d(ax), d(c)/d(ax) of the axis line (|.| <= 1: ax is dominant)
    double xlo, xhi;    // extent of the axis segment on ax
    int ax, i_lo;
};
struct NarrowRec {
    uint32_t edge, a, b, pad;   // a = L<<12 | lx, b = ly<<12 | lz
};
struct Edge2Args {
    MapView m;
    const float* p1;
    const float* p2;
    int64_t n;
    double radius;
    int shape;
    int want_band;
    unsigned* gflags;
    NarrowRec* queue;
    unsigned queue_cap;
    unsigned* qctl;              // [0] queue fill, [1] edges that overflowed the queue
    unsigned long long* stats;   // [6]: [4] = items that reached the rectangle phase
    uint8_t* hit;
    uint8_t* band;
    LevelGrid dil;               // level 0 dilated by gate_d cells (Chebyshev), same brick layout
    int gate_d;                  // 0 = gate off (every slab runs the rectangle phase)
    int gate_seg, gate_rounds;   // slabs per gate item (divides 64, <= 32); segments per edge walked in early-out rounds
    int64_t e_begin;             // first edge of this launch (n = one past its last edge): chunked host calls
};

// ------------------------------------------------------------------ dilated occupancy (sweep gate)
// One separable pass: out = in OR-ed with its copies shifted by -d..+d cells along `axis`
// (d <= 3), computed on whole 4x4x4 brick words: the cells that leave a brick on one side come
// in from the neighbouring brick on the other.  x, y, z passes in turn give the box dilation.
__device__ __forceinline__ uint64_t axis_mask_eq(int axis, int v) {
    return axis == 0 ? (0x1111111111111111ull << v) : axis == 1 ? (0x000F000F000F000Full << (4 * v)) : (0xFFFFull << (16 * v));
}
// `W` OR-ed with its copies shifted by -d..+d cells along `axis`; Wm / Wp = the neighbouring brick words
// on the low / high side (0 outside the map)
__device__ __forceinline__ uint64_t dilate_word(int axis, int d, uint64_t W, uint64_t Wm, uint64_t Wp) {
    const int unit = axis == 0 ? 1 : (axis == 1 ? 4 : 16);
    uint64_t r = W;
    for (int s = 1; s <= d; ++s) {
        uint64_t lt = 0, ge = 0;   // cells with coordinate < 4-s / >= s along the axis
        for (int v = 0; v < 4; ++v) {
            if (v < 4 - s) lt |= axis_mask_eq(axis, v);
            if (v >= s) ge |= axis_mask_eq(axis, v);
        }
        r |= (W & lt) << (unit * s);
        r |= (Wm & ~lt) >> (unit * (4 - s));
        r |= (W & ge) >> (unit * s);
        r |= (Wp & ~ge) << (unit * (4 - s));
    }
    return r;
}
__global__ void dilate_axis_kernel(const uint64_t* __restrict__ in, uint64_t* __restrict__ out, int nbx, int nby,
                                   int nbz, int axis, int d) {
    const int64_t total = (int64_t)nbx * nby * nbz;
    const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= total) return;
    const int bx = (int)(w % nbx), by = (int)((w / nbx) % nby), bz = (int)(w / ((int64_t)nbx * nby));
    const int pos = axis == 0 ? bx : (axis == 1 ? by : bz), n = axis == 0 ? nbx : (axis == 1 ? nby : nbz);
    const int64_t stride = axis == 0 ? 1 : (axis == 1 ? nbx : (int64_t)nbx * nby);
    out[w] = dilate_word(axis, d, in[w], pos > 0 ? in[w - stride] : 0ull, pos + 1 < n ? in[w + stride] : 0ull);
}

// Incremental form (map updates): the dilated words of the bricks within d cells of each changed leaf are
// recomputed from level 0 — the same three separable steps, on the brick's 3x3x3 neighbourhood in registers —
// so a map update touches a few bricks per changed leaf instead of re-dilating all of level 0.
// One thread per (leaf, brick slot); leaves larger than the slot count loop.  Two leaves may recompute the
// same brick: both write the same word.
constexpr int kRedilateSlots = 27;
__global__ void redilate_kernel(const uint64_t* __restrict__ lv0, uint64_t* __restrict__ dil, int nbx, int nby, int nbz,
                                int d, const LeafRec* __restrict__ recs, int64_t n) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n * kRedilateSlots) return;
    const LeafRec r = recs[i / kRedilateSlots];
    const int s = 1 << r.level;
    const int b0[3] = {max(0, (r.x - d) >> 2), max(0, (r.y - d) >> 2), max(0, (r.z - d) >> 2)};
    const int b1[3] = {min(nbx - 1, (r.x + s - 1 + d) >> 2), min(nby - 1, (r.y + s - 1 + d) >> 2), min(nbz - 1, (r.z + s - 1 + d) >> 2)};
    const int c0 = b1[0] - b0[0] + 1, c1 = b1[1] - b0[1] + 1, c2 = b1[2] - b0[2] + 1;
    if (c0 <= 0 || c1 <= 0 || c2 <= 0) return;
    auto load = [&](int bx, int by, int bz) -> uint64_t {
        if (bx < 0 || bx >= nbx || by < 0 || by >= nby || bz < 0 || bz >= nbz) return 0ull;
        return lv0[((int64_t)bz * nby + by) * nbx + bx];
    };
    for (int t = (int)(i % kRedilateSlots); t < c0 * c1 * c2; t += kRedilateSlots) {
        const int bx = b0[0] + t % c0, by = b0[1] + (t / c0) % c1, bz = b0[2] + t / (c0 * c1);
        uint64_t yw[3];
#pragma unroll
        for (int dz = -1; dz <= 1; ++dz) {
            uint64_t xw[3];
#pragma unroll
            for (int dy = -1; dy <= 1; ++dy)
                xw[dy + 1] = dilate_word(0, d, load(bx, by + dy, bz + dz), load(bx - 1, by + dy, bz + dz), load(bx + 1, by + dy, bz + dz));
            yw[dz + 1] = dilate_word(1, d, xw[1], xw[0], xw[2]);
        }
        dil[((int64_t)bz * nby + by) * nbx + bx] = dilate_word(2, d, yw[1], yw[0], yw[2]);
    }
}

// CTA-wide: move the shared-memory queue to the global queue, dropping the pairs of edges
// that already hold a deep-hit certificate (all pairs of an edge come from its own CTA, so the
// filter is exact at the final flush).  Must be called by all threads of the CTA.
__device__ __forceinline__ void sweep_flush(const Edge2Args& a, const unsigned long long* sq, int* sq_count,
                                            unsigned* eflags, int64_t e0) {
    __syncthreads();
    const int n = min(*sq_count, kSqCap);
    const int lane = threadIdx.x & 31;
    for (int base = 0; base < n; base += kSweepThreads) {
        const int j = base + (int)threadIdx.x;
        unsigned long long rec = 0;
        bool keep = false;
        if (j < n) {
            rec = sq[j];
            keep = !(eflags[(int)(rec >> 40)] & F_DEEP);
        }
        const unsigned km = __ballot_sync(kFull, keep);
        if (km) {
            unsigned gbase = 0;
            if (lane == __ffs(km) - 1) gbase = atomicAdd(a.qctl, (unsigned)__popc(km));
            gbase = __shfl_sync(kFull, gbase, __ffs(km) - 1);
            if (keep) {
                const unsigned slot = gbase + __popc(km & ((1u << lane) - 1u));
                const int el = (int)(rec >> 40);
                if (slot < a.queue_cap) {
                    NarrowRec r;
                    r.edge = (uint32_t)(e0 + el);
                    r.a = (uint32_t)((rec >> 24) & 0xFFFFu);          // L<<12 | lx
                    r.b = (uint32_t)(rec & 0xFFFFFFu);                // ly<<12 | lz
                    r.pad = 0;
                    a.queue[slot] = r;
                } else {
                    atomicOr(&eflags[el], F_OVERFLOW);
                }
            }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) *sq_count = 0;
    __syncthreads();
}

// Sweep-plane axes of a dominant axis ax: b = the lower, c = the higher of the two others
// (ax=0: b=y,c=z; ax=1: b=x,c=z; ax=2: b=x,c=y), so that a brick word yields its 4x4 slice at
// a fixed ax coordinate with shifts only.  (i, jb, jc) -> (x, y, z); selects, no local arrays.
__device__ __forceinline__ void unperm(int ax, int i, int jb, int jc, int& x, int& y, int& z) {
    x = ax == 0 ? i : jb;
    y = ax == 0 ? jb : (ax == 1 ? i : jc);
    z = ax == 2 ? i : jc;
}

// Occupancy of the rectangle [b0,b1] x [c0,c1] (<= 8x8 cells) of the plane ax = i, as a 64-bit
// mask with bit (jc-c0)*8 + (jb-b0): one 64-bit load per 4x4x4 brick the rectangle touches
// (instead of one per cell); empty bricks cost one load + one test.
__device__ __forceinline__ unsigned long long rect_occupancy(const LevelGrid& g, int ax, int i, int b0, int b1,
                                                             int c0, int c1) {
    const int64_t sy = g.nb[0], sz = (int64_t)g.nb[0] * g.nb[1];
    const int64_t si = ax == 0 ? 1 : (ax == 1 ? sy : sz), sb = ax == 0 ? sy : 1, sc = ax == 2 ? sy : sz;
    const int il = i & 3;
    const unsigned colmask = (1u << (b1 - b0 + 1)) - 1u;
    // bit offset of row r (c-local) / of the slice inside the word, and the bit stride along b
    const int row_mul = ax == 2 ? 4 : 16, base_sh = ax == 2 ? 16 * il : (ax == 1 ? 4 * il : il);
    const uint64_t* base = g.bits + (int64_t)(i >> 2) * si;
    unsigned long long occ = 0;
    for (int cb = c0 >> 2; cb <= (c1 >> 2); ++cb)
        for (int bb = b0 >> 2; bb <= (b1 >> 2); ++bb) {
            const unsigned long long W = __ldg(base + bb * sb + cb * sc);
            if (W == 0ull) continue;
            unsigned E = 0;   // 4 rows (c-local) in 4 bytes, 4 columns (b-local) in the low nibble of each
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const unsigned v = (unsigned)(W >> (base_sh + row_mul * r));
                // ax=0: the row's cells are 4 bits apart (bit = z<<4 | y<<2 | x): gather them with a multiply
                const unsigned nib = ax == 0 ? (((v & 0x1111u) * 0x1248u) >> 12) & 0xFu : (v & 0xFu);
                E |= nib << (8 * r);
            }
            const int db = bb * 4 - b0, dc = cb * 4 - c0;
            const unsigned pre = (db >= 0 ? (colmask >> db) : (colmask << (-db))) & 0xFu;
            E &= pre * 0x01010101u;
            E = db >= 0 ? (E << db) : (E >> (-db));
            occ |= dc >= 0 ? ((unsigned long long)E << (8 * dc)) : ((unsigned long long)E >> (-8 * dc));
        }
    const int rows = c1 - c0 + 1;
    if (rows < 8) occ &= (1ull << (8 * rows)) - 1ull;
    return occ;
}

// Closed-form certificates of one (edge, leaf) pair: 0 = clear miss, 1 = deep hit, 2 = ambiguous
// (needs the exact narrow phase).  Both certificates hold a margin kCertMargin >> MPR tolerance:
//   clear miss : a unit direction n separates the shapes by more than the margin (support
//                functions of the cylinder / capsule and of the box along n); tried for the box
//                face normals, the cylinder axis and the direction between the closest points
//                found by a few alternating projections (axis segment <-> box);
//   deep hit   : a point of the box lies inside the shape shrunk by the margin, so the shapes
//                overlap by at least the margin.
// Any n and any box point give a valid certificate; the projections only pick good candidates.
__device__ __forceinline__ int classify_leaf(const SweepEdge& s, double radius, int shape, const D3& bc, double e) {
    const double mg = kCertMargin;
    const bool caps = shape == BGR_SHAPE_CAPSULE;
    const double w[3] = {bc.x - s.c[0], bc.y - s.c[1], bc.z - s.c[2]};
    const double u[3] = {s.u[0], s.u[1], s.u[2]};
    // box face normals
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const double supp = s.h * fabs(u[k]) + (caps ? radius : radius * sqrt(fmax(0.0, 1.0 - u[k] * u[k])));
        if (fabs(w[k]) - supp - e > mg) return 0;
    }
    // cylinder axis
    const double al = w[0] * u[0] + w[1] * u[1] + w[2] * u[2];
    const double box_u = e * (fabs(u[0]) + fabs(u[1]) + fabs(u[2]));
    if (fabs(al) - s.h - (caps ? radius : 0.0) - box_u > mg) return 0;
    // alternating projections between the axis segment (pulled 2 margins inside the caps, so an
    // axis point inside the box certifies a flat-capped cylinder too) and the box
    const double hs = fmax(0.0, s.h - 2.0 * mg);
    double t = fmax(-hs, fmin(hs, al));
    double q[3];
#pragma unroll
    for (int it = 0; it < 3; ++it) {
#pragma unroll
        for (int k = 0; k < 3; ++k) q[k] = fmax(w[k] - e, fmin(w[k] + e, t * u[k]));
        t = fmax(-hs, fmin(hs, q[0] * u[0] + q[1] * u[1] + q[2] * u[2]));
    }
    const double d[3] = {q[0] - t * u[0], q[1] - t * u[1], q[2] - t * u[2]};
    const double dn2 = d[0] * d[0] + d[1] * d[1] + d[2] * d[2];
    const double rin = radius - mg;
    if (radius > mg && s.h > 2.0 * mg) {
        if (caps) {
            if (dn2 < rin * rin) return 1;
        } else {
            const double aq = q[0] * u[0] + q[1] * u[1] + q[2] * u[2];
            const double ex = q[0] - aq * u[0], ey = q[1] - aq * u[1], ez = q[2] - aq * u[2];
            if (fabs(aq) <= s.h - mg && ex * ex + ey * ey + ez * ez < rin * rin) return 1;
        }
    }
    if (dn2 > 1e-24) {
        const double inv = 1.0 / sqrt(dn2);
        const double n[3] = {d[0] * inv, d[1] * inv, d[2] * inv};
        const double un = u[0] * n[0] + u[1] * n[1] + u[2] * n[2];
        const double supp_c = s.h * fabs(un) + (caps ? radius : radius * sqrt(fmax(0.0, 1.0 - un * un)));
        const double supp_b = e * (fabs(n[0]) + fabs(n[1]) + fabs(n[2]));
        if (w[0] * n[0] + w[1] * n[1] + w[2] * n[2] - supp_c - supp_b > mg) return 0;
    }
    return 2;
}

// CTA-wide exclusive prefix of one int per thread: prefix[t] = sum of val[0..t), prefix[nthreads] = total
__device__ __forceinline__ void cta_prefix(int val, int* prefix, int* warp_sums) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int inc = val;
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(kFull, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        int w = lane < kSweepThreads / 32 ? warp_sums[lane] : 0;
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(kFull, w, o);
            if (lane >= o) w += t;
        }
        if (lane < kSweepThreads / 32) warp_sums[lane] = w;
    }
    __syncthreads();
    const int off = warp ? warp_sums[warp - 1] : 0;
    prefix[tid + 1] = off + inc;
    if (tid == 0) prefix[0] = 0;
    __syncthreads();
}

constexpr int kGateSlabs = 128;   // slabs per edge the gate tracks (longer edges run every slab)
constexpr float kGateSlop = 2e-3f;   // cells: covers the f32 rounding of the gate walk
constexpr int kGateSeg = 8;           // slabs one thread walks (64 % kGateSeg == 0)
constexpr int kGateRounds = 2;        // segments of an edge walked in early-out rounds before the flat list (measured:
                                      // profiles/r2_gate_rounds_tune.txt; more rounds cost more in barriers than they skip)
struct GateEdge {                     // the axis line in cell units relative to the edge's first slab
    float tlo, thi;                   // extent of the segment along ax
    float sb, sc, beta_b, beta_c;     // lateral cell coordinate = beta + slope * t
};

// The kernel has two item passes over the (edge, slab) work items of the CTA's 256 edges.
//   gate pass      : an item looks only at the cells the AXIS LINE crosses inside its slab
//                    (<= 2x2, slope <= 1) in the dilated map (level 0 grown by d = ceil(rm/res)
//                    cells).  No dilated cell on the whole axis  =>  every leaf is farther than
//                    rm from every axis point: clear miss, the edge is finished.  An occupied
//                    level-0 cell on the axis whose leaf holds a deep-hit certificate finishes
//                    the edge as a hit.  Otherwise the item sets its bit in the edge's slab mask.
//   rectangle pass : only slabs within d of a set bit (any leaf within rm of the axis has a cell
//                    in such a slab), of edges that are not finished: the rm-inflated <= 8x8
//                    rectangle, candidates processed densely as before.
__global__ void __launch_bounds__(kSweepThreads, 4) edges_sweep_kernel(Edge2Args a) {
    __shared__ SweepEdge es[kSweepEdges];
    __shared__ unsigned long long sq[kSqCap];
    __shared__ unsigned long long emask[kSweepEdges][2];
    __shared__ int prefix[kSweepEdges + 1];    // over all slabs
    __shared__ int prefix2[kSweepEdges + 1];   // over the slabs the rectangle pass runs
    __shared__ unsigned eflags[kSweepEdges];
    __shared__ int warp_sums[kSweepThreads / 32];
    __shared__ GateEdge ge[kSweepEdges];
    __shared__ uint8_t alist[2][kSweepEdges];   // undecided edges of the current / next gate round
    __shared__ int sq_count;
    __shared__ unsigned long long st_cand;

    const MapView& m = a.m;
    const int tid = threadIdx.x, lane = tid & 31;
    const int64_t e0 = a.e_begin + (int64_t)blockIdx.x * kSweepEdges;
    const int ne = (int)min((int64_t)kSweepEdges, a.n - e0);
    const double rm = a.radius + kSweepMargin;
    const bool gate = a.gate_d > 0;
    if (tid == 0) { st_cand = 0; sq_count = 0; }

    // ---- phase 0: per-edge frame (thread per edge)
    int myslabs = 0;
    eflags[tid] = 0;
    emask[tid][0] = emask[tid][1] = 0ull;
    if (tid < ne) {
        const int64_t e = e0 + tid;
        const float p1[3] = {a.p1[3 * e], a.p1[3 * e + 1], a.p1[3 * e + 2]};
        const float p2[3] = {a.p2[3 * e], a.p2[3 * e + 1], a.p2[3 * e + 2]};
        const ShapeObj o = make_edge_obj(p1, p2, a.radius, a.shape);
        D3 u = {0, 0, 1};
        quat_rot(u, o.rot);
        d3_scale(u, 1.0 / sqrt(d3_dot(u, u)));
        SweepEdge s;
        s.c[0] = o.pos.x; s.c[1] = o.pos.y; s.c[2] = o.pos.z;
        s.u[0] = u.x; s.u[1] = u.y; s.u[2] = u.z;
        s.h = o.height;
        s.ax = 0;
        if (fabs(s.u[1]) > fabs(s.u[s.ax])) s.ax = 1;
        if (fabs(s.u[2]) > fabs(s.u[s.ax])) s.ax = 2;
        const int bx = s.ax == 0 ? 1 : 0, cx = s.ax == 2 ? 1 : 2;
        const double inv_uax = 1.0 / s.u[s.ax];            // |u[ax]| >= 1/sqrt(3): u is a unit vector
        s.sb = s.u[bx] * inv_uax;
        s.sc = s.u[cx] * inv_uax;
        const double pa = s.c[s.ax] - s.h * s.u[s.ax], pb = s.c[s.ax] + s.h * s.u[s.ax];
        s.xlo = fmin(pa, pb);
        s.xhi = fmax(pa, pb);
        int lo = (int)floor((s.xlo - rm - m.origin[s.ax]) * m.inv_res);
        int hi = (int)floor((s.xhi + rm - m.origin[s.ax]) * m.inv_res);
        lo = max(lo, 0);
        hi = min(hi, m.dims[s.ax] - 1);
        s.i_lo = lo;
        myslabs = max(0, hi - lo + 1);
        if (!isfinite(s.c[0] + s.c[1] + s.c[2] + s.h)) myslabs = 0;
        es[tid] = s;
        if (gate) {
            GateEdge g;
            const double t_ref = (s.c[s.ax] - m.origin[s.ax]) * m.inv_res - (double)lo;   // centre, slab units from i_lo
            g.tlo = (float)((s.xlo - m.origin[s.ax]) * m.inv_res - (double)lo);
            g.thi = (float)((s.xhi - m.origin[s.ax]) * m.inv_res - (double)lo);
            g.sb = (float)s.sb;
            g.sc = (float)s.sc;
            g.beta_b = (float)((s.c[bx] - m.origin[bx]) * m.inv_res - t_ref * s.sb);
            g.beta_c = (float)((s.c[cx] - m.origin[cx]) * m.inv_res - t_ref * s.sc);
            ge[tid] = g;
        }
    }
    cta_prefix(myslabs, prefix, warp_sums);
    const int total = prefix[kSweepEdges];
    unsigned long long n_cand = 0;

    // ---- gate pass: an item = (edge, segment of gate_seg consecutive slabs), walked by one thread
    // serially in f32 cell units relative to the edge's first slab (|values| <= 4096, error
    // < 5e-4 cell, covered by kGateSlop).  The first gate_rounds segments of every edge run in
    // ROUNDS (round r = segment r of every edge that is still undecided, compacted so that whole
    // warps stay busy): a deep hit on the axis ends the edge, and half of all edges are such hits
    // found within their first ~10 slabs — the later segments of those edges are never walked
    // (17.6 M -> ~9 M slabs on configs[2]).  The segments left after the rounds run as one flat
    // item list.  The order only decides how much is skipped: an edge ends with F_DEEP (hit, not in
    // the band) whichever deep cell is found first, and undecided edges have walked every segment.
    if (gate) {
        const int kSeg = a.gate_seg, kRounds = a.gate_rounds;
        const unsigned nbx = (unsigned)a.dil.nb[0], nbxy = (unsigned)a.dil.nb[0] * (unsigned)a.dil.nb[1];
        // deep-hit certificate on an axis cell needs a ball of radius >= 1e-3 m around the axis
        // point inside the shape: radius >= 1e-3 and the point >= 0.05 cell (res >= 0.02) from both ends
        const bool allow_deep = a.radius >= 1e-3 && m.res >= 0.02;
        const float rm_c = (float)(rm * m.inv_res) + kGateSlop;
        // walks slabs [k0, k1) of edge el; returns true on a deep hit, else the fired slabs in fired_mask
        auto gate_segment = [&](const int el, const int k0, const int k1, unsigned& fired_mask) -> bool {
            const GateEdge g = ge[el];
            const int ax = es[el].ax, bx = ax == 0 ? 1 : 0, cx = ax == 2 ? 1 : 2;
            const int i_lo = es[el].i_lo, dim_a = m.dims[ax], dim_b = m.dims[bx], dim_c = m.dims[cx];
            // brick strides / log2 bit units of the permuted axes
            const unsigned Si = ax == 0 ? 1u : (ax == 1 ? nbx : nbxy), Sb = bx == 0 ? 1u : nbx, Sc = cx == 1 ? nbx : nbxy;
            const int Ui = ax == 0 ? 0 : (ax == 1 ? 2 : 4), Ub = bx == 0 ? 0 : 2, Uc = cx == 1 ? 2 : 4;
            fired_mask = 0;
            for (int k = k0; k < k1; ++k) {
                const int i = i_lo + k;
                const float t0 = i == 0 ? (float)k - rm_c : (float)k;
                const float t1 = i == dim_a - 1 ? (float)(k + 1) + rm_c : (float)(k + 1);
                const float ta = fminf(fmaxf(t0, g.tlo), g.thi), tb = fminf(fmaxf(t1, g.tlo), g.thi);
                const float ba = fmaf(g.sb, ta, g.beta_b), bb = fmaf(g.sb, tb, g.beta_b);
                const float ca = fmaf(g.sc, ta, g.beta_c), cb = fmaf(g.sc, tb, g.beta_c);
                int b0 = (int)floorf(fminf(ba, bb) - kGateSlop), b1 = (int)floorf(fmaxf(ba, bb) + kGateSlop);
                int c0 = (int)floorf(fminf(ca, cb) - kGateSlop), c1 = (int)floorf(fmaxf(ca, cb) + kGateSlop);
                // axis points up to d cells outside the map still see the border cells' dilation
                if (b1 < -a.gate_d || b0 > dim_b - 1 + a.gate_d || c1 < -a.gate_d || c0 > dim_c - 1 + a.gate_d) continue;
                b0 = min(max(b0, 0), dim_b - 1); b1 = min(max(b1, 0), dim_b - 1);
                c0 = min(max(c0, 0), dim_c - 1); c1 = min(max(c1, 0), dim_c - 1);
                bool fired = (b1 - b0 > 1) || (c1 - c0 > 1);   // wider than the 2x2 corner probes: keep the slab
                const unsigned wi = (unsigned)(i >> 2) * Si;
                const int bi = (i & 3) << Ui;
                if (!fired) {
                    const unsigned wb0 = (unsigned)(b0 >> 2) * Sb, wb1 = (unsigned)(b1 >> 2) * Sb;
                    const unsigned wc0 = (unsigned)(c0 >> 2) * Sc, wc1 = (unsigned)(c1 >> 2) * Sc;
                    const int bb0 = (b0 & 3) << Ub, bb1 = (b1 & 3) << Ub, bc0 = (c0 & 3) << Uc, bc1 = (c1 & 3) << Uc;
                    const uint64_t W00 = __ldg(a.dil.bits + (wi + wb0 + wc0)), W10 = __ldg(a.dil.bits + (wi + wb1 + wc0));
                    const uint64_t W01 = __ldg(a.dil.bits + (wi + wb0 + wc1)), W11 = __ldg(a.dil.bits + (wi + wb1 + wc1));
                    const uint64_t any = (W00 >> (bi + bb0 + bc0)) | (W10 >> (bi + bb1 + bc0)) | (W01 >> (bi + bb0 + bc1)) |
                                         (W11 >> (bi + bb1 + bc1));
                    fired = (any & 1ull) != 0ull;
                }
                if (fired) fired_mask |= 1u << (k - k0);
            }
            // deep-hit probe of the fired slabs (after the walk, so that the lanes that fired run it
            // together): the cell holding the axis at the middle of the slab portion (kept 0.05 cell
            // inside the segment and inside the slab) is occupied => the shapes overlap by ~radius
            bool deep = false;
            if (allow_deep) {
                unsigned fm = fired_mask;
                while (fm && !deep) {
                    const int k = k0 + __ffs(fm) - 1;
                    fm &= fm - 1;
                    const int i = i_lo + k;
                    const float ua = fmaxf((float)k, g.tlo + 0.05f), ub = fminf((float)(k + 1), g.thi - 0.05f);
                    if (ua <= ub) {
                        const float um = 0.5f * (ua + ub);
                        const int jb = (int)floorf(fmaf(g.sb, um, g.beta_b)), jc = (int)floorf(fmaf(g.sc, um, g.beta_c));
                        if (jb >= 0 && jb < dim_b && jc >= 0 && jc < dim_c) {
                            const uint64_t W = __ldg(m.lv[0].bits + ((unsigned)(i >> 2) * Si + (unsigned)(jb >> 2) * Sb + (unsigned)(jc >> 2) * Sc));
                            deep = ((W >> (((i & 3) << Ui) + ((jb & 3) << Ub) + ((jc & 3) << Uc))) & 1ull) != 0ull;
                        }
                    }
                }
            }
            return deep;
        };
        // ---- rounds: alist = the undecided edges that still have a segment to walk
        const int gslabs_mine = min(myslabs, kGateSlabs);
        cta_prefix(gslabs_mine > 0 ? 1 : 0, prefix2, warp_sums);
        if (gslabs_mine > 0) alist[0][prefix2[tid]] = (uint8_t)tid;
        int n_alive = prefix2[kSweepEdges];
        __syncthreads();
        int cur = 0;
        for (int r = 0; r < kRounds && n_alive > 0; ++r) {
            int el = 0;
            bool go_on = false;
            if (tid < n_alive) {
                el = alist[cur][tid];
                const int gs = min(prefix[el + 1] - prefix[el], kGateSlabs);
                const int k0 = r * kSeg, k1 = min(k0 + kSeg, gs);
                unsigned fired_mask;
                if (gate_segment(el, k0, k1, fired_mask)) eflags[el] = F_HIT | F_DEEP;    // one thread per edge and round
                else {
                    if (fired_mask) emask[el][k0 >> 6] |= (unsigned long long)fired_mask << (k0 & 63);
                    go_on = k1 < gs;
                }
            }
            cta_prefix(go_on ? 1 : 0, prefix2, warp_sums);
            if (go_on) alist[cur ^ 1][prefix2[tid]] = (uint8_t)el;
            n_alive = prefix2[kSweepEdges];
            cur ^= 1;
            __syncthreads();
        }
        // ---- the segments left: one flat item list over the edges still undecided
        if (n_alive > 0) {
            int nseg_left = 0;
            if (tid < n_alive) {
                const int el = alist[cur][tid];
                const int gs = min(prefix[el + 1] - prefix[el], kGateSlabs);
                nseg_left = (gs + kSeg - 1) / kSeg - kRounds;
            }
            cta_prefix(nseg_left, prefix2, warp_sums);
            const int total_seg = prefix2[kSweepEdges];
            for (int w0 = 0; w0 < total_seg; w0 += kSweepThreads) {
                const int w = w0 + tid;
                if (w >= total_seg) break;
                int lo = 0, hi = n_alive;
                while (hi - lo > 1) {
                    const int mid = (lo + hi) >> 1;
                    if (prefix2[mid] <= w) lo = mid; else hi = mid;
                }
                const int el = alist[cur][lo];
                if (eflags[el] & F_DEEP) continue;
                const int k0 = (kRounds + w - prefix2[lo]) * kSeg;
                const int k1 = min(k0 + kSeg, min(prefix[el + 1] - prefix[el], kGateSlabs));
                unsigned fired_mask;
                if (gate_segment(el, k0, k1, fired_mask)) atomicOr(&eflags[el], F_HIT | F_DEEP);
                else if (fired_mask) atomicOr(&emask[el][k0 >> 6], (unsigned long long)fired_mask << (k0 & 63));
            }
        }
        __syncthreads();
        // slabs of the rectangle pass: within d of a fired slab; none for finished edges
        int cnt = myslabs;
        if (eflags[tid] & F_DEEP) {
            cnt = 0;
        } else if (myslabs <= kGateSlabs) {
            unsigned long long lo = emask[tid][0], hi = emask[tid][1];
            unsigned long long rl = lo, rh = hi;
            for (int sft = 1; sft <= a.gate_d; ++sft) {
                rl |= (lo << sft) | (lo >> sft) | (hi << (64 - sft));
                rh |= (hi << sft) | (hi >> sft) | (lo >> (64 - sft));
            }
            if (myslabs < 64) { rl &= (1ull << myslabs) - 1ull; rh = 0ull; }
            else if (myslabs < 128) rh &= (1ull << (myslabs - 64)) - 1ull;
            emask[tid][0] = rl;
            emask[tid][1] = rh;
            cnt = __popcll(rl) + __popcll(rh);
        }
        cta_prefix(cnt, prefix2, warp_sums);
    } else {
        prefix2[tid + 1] = prefix[tid + 1];
        if (tid == 0) prefix2[0] = 0;
        __syncthreads();
    }
    const int total2 = prefix2[kSweepEdges];

    // ---- rectangle pass
    for (int w0 = 0; w0 < total2; w0 += kSweepThreads) {
        // ---- one item per thread: occupancy mask of its rectangle
        const int w = w0 + tid;
        unsigned long long occ = 0;
        int info_a = 0, info_b = 0;   // el | slab<<8 ; b0 | c0<<12
        if (w < total2) {
            int lo = 0, hi = kSweepEdges;
            while (hi - lo > 1) {
                const int mid = (lo + hi) >> 1;
                if (prefix2[mid] <= w) lo = mid; else hi = mid;
            }
            const int el = lo;
            const SweepEdge& s = es[el];
            const int ax = s.ax, bx = ax == 0 ? 1 : 0, cx = ax == 2 ? 1 : 2;
            int k = w - prefix2[el];
            if (gate && prefix[el + 1] - prefix[el] <= kGateSlabs) {
                const unsigned long long ml = emask[el][0];
                const int cl = __popcll(ml);
                k = k < cl ? nth_set_bit64(ml, k) : 64 + nth_set_bit64(emask[el][1], k - cl);
            }
            const int i = s.i_lo + k;
            // the part of the axis segment whose ax coordinate lies in the slab inflated by rm
            // (clamped to the segment: slabs past an end see the end point + the rm margin)
            const double x0 = m.origin[ax] + (double)i * m.res - rm, x1 = m.origin[ax] + (double)(i + 1) * m.res + rm;
            const double xa = fmin(fmax(x0, s.xlo), s.xhi) - s.c[ax], xb = fmax(fmin(x1, s.xhi), s.xlo) - s.c[ax];
            int r0[2], r1[2];
            bool none = false;
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const int b = q ? cx : bx;
                const double sl = q ? s.sc : s.sb;
                const double pa = s.c[b] + xa * sl, pb = s.c[b] + xb * sl;
                int lo_b = (int)floor((fmin(pa, pb) - rm - m.origin[b]) * m.inv_res);
                int hi_b = (int)floor((fmax(pa, pb) + rm - m.origin[b]) * m.inv_res);
                lo_b = max(lo_b, 0);
                hi_b = min(hi_b, m.dims[b] - 1);
                hi_b = min(hi_b, lo_b + kRectMax - 1);   // host guarantees the hull fits kRectMax cells
                r0[q] = lo_b;
                r1[q] = hi_b;
                none |= lo_b > hi_b;
            }
            if (!none) {
                occ = rect_occupancy(m.lv[0], ax, i, r0[0], r1[0], r0[1], r1[1]);
                info_a = el | (i << 8);
                info_b = r0[0] | (r0[1] << 12);
            }
        }
        // ---- dense processing of the warp's candidates
        const int cnt = __popcll(occ);
        int incl = cnt;
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(kFull, incl, o);
            if (lane >= o) incl += t;
        }
        const int T = __shfl_sync(kFull, incl, 31);
        for (int j0 = 0; j0 < T; j0 += 32) {
            const int j = j0 + lane;
            const bool active = j < T;
            int src = 0;
#pragma unroll
            for (int step = 16; step > 0; step >>= 1) {
                const int v = __shfl_sync(kFull, incl, src + step - 1);
                if (v <= j) src += step;
            }
            src = min(src, 31);
            const int s_incl = __shfl_sync(kFull, incl, src), s_cnt = __shfl_sync(kFull, cnt, src);
            const unsigned long long s_occ = __shfl_sync(kFull, occ, src);
            const int s_a = __shfl_sync(kFull, info_a, src), s_b = __shfl_sync(kFull, info_b, src);
            bool ambiguous = false;
            unsigned long long rec = 0;
            int el = 0;
            if (active) {
                const int bit = nth_set_bit64(s_occ, j - (s_incl - s_cnt));
                el = s_a & 255;
                const int i = s_a >> 8, b0 = s_b & 4095, c0 = s_b >> 12;
                const SweepEdge& s = es[el];
                const int ax = s.ax;
                const int vb = b0 + (bit & (kRectMax - 1)), vc = c0 + (bit >> 3);
                int vx, vy, vz, l[3];
                unperm(ax, i, vb, vc, vx, vy, vz);
                const int L = find_leaf(m, vx, vy, vz, l[0], l[1], l[2]);
                const int lb = ax == 0 ? l[1] : l[0], lc = ax == 2 ? l[1] : l[2];
                // one representative cell per (item, leaf)
                if (vb == max(b0, lb) && vc == max(c0, lc)) {
                    ++n_cand;
                    D3 bc;
                    double e;
                    leaf_box(m, L, l[0], l[1], l[2], bc, e);
                    const int cls = classify_leaf(s, a.radius, a.shape, bc, e);
                    if (cls == 1) {
                        atomicOr(&eflags[el], F_HIT | F_DEEP);
                    } else if (cls == 2) {
                        ambiguous = !(eflags[el] & F_DEEP);
                        rec = ((unsigned long long)el << 40) | ((unsigned long long)L << 36) |
                              ((unsigned long long)l[0] << 24) | ((unsigned long long)l[1] << 12) | (unsigned long long)l[2];
                    }
                }
            }
            const unsigned am = __ballot_sync(kFull, ambiguous);
            if (am) {
                int base = 0;
                if (lane == __ffs(am) - 1) base = atomicAdd(&sq_count, __popc(am));
                base = __shfl_sync(kFull, base, __ffs(am) - 1);
                if (ambiguous) {
                    const int slot = base + __popc(am & ((1u << lane) - 1u));
                    if (slot < kSqCap) {
                        sq[slot] = rec;
                    } else {
                        // shared queue full: straight to the global queue (not filtered by later deep hits)
                        const unsigned g = atomicAdd(a.qctl, 1u);
                        if (g < a.queue_cap) {
                            NarrowRec r;
                            r.edge = (uint32_t)(e0 + el);
                            r.a = (uint32_t)((rec >> 24) & 0xFFFFu);
                            r.b = (uint32_t)(rec & 0xFFFFFFu);
                            r.pad = 0;
                            a.queue[g] = r;
                        } else {
                            atomicOr(&eflags[el], F_OVERFLOW);
                        }
                    }
                }
            }
        }
    }
    sweep_flush(a, sq, &sq_count, eflags, e0);
    if (a.stats) {
        for (int o = 16; o > 0; o >>= 1) n_cand += __shfl_xor_sync(kFull, n_cand, o);
        if (lane == 0 && n_cand) atomicAdd(&st_cand, n_cand);
    }
    __syncthreads();
    if (tid < ne) {
        a.gflags[e0 + tid] = eflags[tid];
        if (eflags[tid] & F_OVERFLOW) atomicAdd(a.qctl + 1, 1u);
    }
    if (a.stats && tid == 0) {
        atomicAdd(a.stats + 0, (unsigned long long)total);
        atomicAdd(a.stats + 1, st_cand);
        atomicAdd(a.stats + 4, (unsigned long long)total2);
    }
}

__global__ void __launch_bounds__(128) edges_narrow_kernel(Edge2Args a) {
    const MapView& m = a.m;
    const unsigned count = min(a.qctl[0], a.queue_cap);
    unsigned long long n_mpr = 0, n_cap = 0;
    if (a.want_band) {
        // With the band test a pair needs the exact test of the shape AND of the shape deflated
        // (after a hit) or inflated (after a miss) by kBandFlag.  The narrow phase is a latency
        // chain (one MPR run after the other per thread, few pairs), so a pair takes 4 lanes: lane 0
        // runs the shape, lane 1 the deflated, lane 2 the inflated one, all at once.
        const int lane = threadIdx.x & 31, sub = lane & 3, grp = lane & ~3;
        const unsigned stride = gridDim.x * blockDim.x / 4;
        for (unsigned base = (blockIdx.x * blockDim.x + (threadIdx.x & ~31u)) / 4; base < count; base += stride) {
            const unsigned i = base + (unsigned)(lane >> 2);
            bool active = i < count;
            NarrowRec rec = {0, 0, 0, 0};
            if (active) rec = a.queue[i];
            unsigned cur = 0;
            if (active && sub == 0) cur = __ldcg(a.gflags + rec.edge);
            cur = __shfl_sync(kFull, cur, grp);            // one read per pair: the 4 lanes must agree
            if (cur & (F_DEEP | F_OVERFLOW)) active = false;
            if ((cur & F_HIT) && (cur & F_DEFL)) active = false;      // already decided
            int hitv = 0, capped = 0;
            if (active && sub < 3) {
                const int L = (int)(rec.a >> 12), lx = (int)(rec.a & 4095), ly = (int)(rec.b >> 12), lz = (int)(rec.b & 4095);
                D3 bc;
                double e;
                leaf_box(m, L, lx, ly, lz, bc, e);
                const int64_t eg = rec.edge;
                const float p1[3] = {a.p1[3 * eg], a.p1[3 * eg + 1], a.p1[3 * eg + 2]};
                const float p2[3] = {a.p2[3 * eg], a.p2[3 * eg + 1], a.p2[3 * eg + 2]};
                ShapeObj o = make_edge_obj(p1, p2, a.radius, a.shape);
                if (sub == 1) {
                    o.radius = fmax(0.0, o.radius - kBandFlag);
                    if (o.type == kShapeCyl) o.height = fmax(0.0, o.height - kBandFlag);
                } else if (sub == 2) {
                    o.radius += kBandFlag;
                    if (o.type == kShapeCyl) o.height += kBandFlag;
                }
                hitv = mpr_shape_box(o, bc, e, kBandTol, &capped) != 0;
            }
            const int h0 = __shfl_sync(kFull, hitv, grp), h1 = __shfl_sync(kFull, hitv, grp + 1), h2 = __shfl_sync(kFull, hitv, grp + 2);
            const int c0 = __shfl_sync(kFull, capped, grp), c1 = __shfl_sync(kFull, capped, grp + 1), c2 = __shfl_sync(kFull, capped, grp + 2);
            if (active && sub == 0) {
                unsigned f = 0;
                if (h0) f |= F_HIT;
                if (h0 && h1) f |= F_DEFL;
                if (!h0 && h2) f |= F_INFL;
                n_mpr += 2;                                 // the two runs the result depends on
                if (c0 || (h0 ? c1 : c2)) { f |= F_CAP; ++n_cap; }
                if (f) atomicOr(a.gflags + rec.edge, f);
            }
        }
    } else {
        for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
            const NarrowRec rec = a.queue[i];
            const unsigned cur = __ldcg(a.gflags + rec.edge);
            if (cur & (F_DEEP | F_OVERFLOW | F_HIT)) continue;   // decided
            const int L = (int)(rec.a >> 12), lx = (int)(rec.a & 4095), ly = (int)(rec.b >> 12), lz = (int)(rec.b & 4095);
            D3 bc;
            double e;
            leaf_box(m, L, lx, ly, lz, bc, e);
            const int64_t eg = rec.edge;
            const float p1[3] = {a.p1[3 * eg], a.p1[3 * eg + 1], a.p1[3 * eg + 2]};
            const float p2[3] = {a.p2[3 * eg], a.p2[3 * eg + 1], a.p2[3 * eg + 2]};
            ShapeObj o = make_edge_obj(p1, p2, a.radius, a.shape);
            int capped = 0;
            unsigned f = 0;
            if (mpr_shape_box(o, bc, e, kBandTol, &capped)) f |= F_HIT;
            ++n_mpr;
            if (capped) { f |= F_CAP; ++n_cap; }
            if (f) atomicOr(a.gflags + rec.edge, f);
        }
    }
    if (a.stats) {
        for (int o = 16; o > 0; o >>= 1) {
            n_mpr += __shfl_xor_sync(kFull, n_mpr, o);
            n_cap += __shfl_xor_sync(kFull, n_cap, o);
        }
        if ((threadIdx.x & 31) == 0) {
            if (n_mpr) atomicAdd(a.stats + 2, n_mpr);
            if (n_cap) atomicAdd(a.stats + 3, n_cap);
        }
    }
}

__global__ void edges_finalize_kernel(const unsigned* __restrict__ gflags, int64_t n, uint8_t* __restrict__ hit,
                                      uint8_t* __restrict__ band) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned f = gflags[i];
    const bool h = (f & F_HIT) != 0;
    hit[i] = h ? 1 : 0;
    if (band) band[i] = (!(f & F_DEEP) && ((h && !(f & F_DEFL)) || (!h && (f & F_INFL)) || (f & F_CAP))) ? 1 : 0;
}

}  // namespace

// ====================================================================== host side
struct bgr_map {
    int device = 0;                    // device current at creation: every call must run with it current
    cudaStream_t stream = nullptr;
    void* aux = nullptr;               // scratch of the path post-processing entry points (path.cu), owned by the map
    void (*aux_free)(void*) = nullptr;
    double res = 0;
    double origin[3] = {0, 0, 0};
    int dims[3] = {0, 0, 0};
    DevBuf<uint64_t> bits[kMaxLevel + 1];
    int64_t words[kMaxLevel + 1] = {0};
    int nb[kMaxLevel + 1][3];
    unsigned level_mask = 0;
    int64_t level_count[kMaxLevel + 1] = {0};
    DevBuf<LeafRec> recs;
    DevBuf<float> p1_dev, p2_dev;
    DevBuf<uint8_t> hit_dev, band_dev;
    DevBuf<unsigned long long> stats;
    DevBuf<unsigned> eflags, qctl;
    DevBuf<NarrowRec> queue;
    DevBuf<uint64_t> dil[2], dil_tmp;  // level 0 dilated by dil_cells[s] cells (0 = stale): the gate maps of the
    int dil_cells[2] = {0, 0};         // edge sweep and of the sphere check, rebuilt lazily after a map change
    int dil_next = 0;
    uint64_t n_dilate_full = 0, n_dilate_incr = 0;   // full three-pass dilations / incremental updates of a cached one
    int last_gate_d = 0;
    bool gate_on = true;
    int pipe_chunks = 0;               // > 1: the sweep of chunk c waits for pipe.ev_in[c] (host-pointer call)
    int64_t pipe_begin[CopyPipe::kMaxChunks + 1] = {0};   // chunk c = edges [pipe_begin[c], pipe_begin[c+1])
    CopyPipe pipe;                     // side streams of the host-pointer collide_edges
    uint64_t last_stats[4] = {0, 0, 0, 0};
    cudaEvent_t tev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};   // kernel boundaries of the last collide_edges (timing on)
    bool timing = false;
    std::vector<uint64_t> leaf_keys;   // sorted keys (level, z, y, x) of the leaves set through set/refresh_boxes
    bool keys_valid = false;           // false after update_boxes (the caller manages the diff itself)

    MapView view() const {
        MapView v;
        v.res = res;
        v.inv_res = 1.0 / res;
        for (int a = 0; a < 3; ++a) { v.origin[a] = origin[a]; v.dims[a] = dims[a]; }
        for (int L = 0; L <= kMaxLevel; ++L) {
            v.lv[L].bits = bits[L].p;
            for (int a = 0; a < 3; ++a) v.lv[L].nb[a] = nb[L][a];
        }
        v.level_mask = level_mask;
        return v;
    }
};

namespace bgr {
MapView map_view(const bgr_map* h) { return h->view(); }
cudaStream_t map_stream(const bgr_map* h) { return h->stream; }
int map_device(const bgr_map* h) { return h->device; }
void** map_aux(bgr_map* h, void (*free_fn)(void*)) {
    h->aux_free = free_fn;
    return &h->aux;
}
}  // namespace bgr

namespace {

// toBoxes() rows -> validated integer leaf records (leaves larger than 2^kMaxLevel cells are
// split into 2^kMaxLevel blocks: the union is identical, only MPR's sub-tolerance behaviour
// on the interior faces can differ)
int boxes_to_recs(const bgr_map* h, const double* boxes, int64_t nb, std::vector<LeafRec>& out, int64_t counts[kMaxLevel + 1]) {
    out.clear();
    out.reserve((size_t)nb);
    for (int64_t i = 0; i < nb; ++i) {
        const double* b = boxes + 6 * i;
        if (!(b[4] >= b[5])) continue;   // toBoxes keeps occupancy >= threshold; be safe on raw lists
        const double cells = b[3] / h->res;
        const int L = (int)std::lround(std::log2(cells));
        BGR_ARG(L >= 0 && L <= 16 && std::fabs(cells - (double)(1 << L)) < 1e-6 * cells,
                "box %lld: size %g is not res*2^L", (long long)i, b[3]);
        int32_t c[3];
        for (int a = 0; a < 3; ++a) {
            const double v = (b[a] - h->origin[a]) / h->res - 0.5 * (double)(1 << L);
            const double r = std::round(v);
            BGR_ARG(std::fabs(v - r) < 1e-4, "box %lld: centre not aligned to the grid on axis %d", (long long)i, a);
            c[a] = (int32_t)r;
            BGR_ARG(c[a] >= 0 && (int64_t)c[a] + (1 << L) <= h->dims[a] && (c[a] & ((1 << L) - 1)) == 0,
                    "box %lld: outside the map or not octree-aligned on axis %d", (long long)i, a);
        }
        if (L <= kMaxLevel) {
            out.push_back({c[0], c[1], c[2], L});
            counts[L]++;
        } else {
            const int s = 1 << kMaxLevel, n = 1 << (L - kMaxLevel);
            for (int z = 0; z < n; ++z)
                for (int y = 0; y < n; ++y)
                    for (int x = 0; x < n; ++x) {
                        out.push_back({c[0] + x * s, c[1] + y * s, c[2] + z * s, kMaxLevel});
                        counts[kMaxLevel]++;
                    }
        }
    }
    return BGR_OK;
}

inline uint64_t leaf_key(const LeafRec& r) {
    return ((uint64_t)r.level << 48) | ((uint64_t)r.z << 32) | ((uint64_t)r.y << 16) | (uint64_t)r.x;
}
inline LeafRec key_leaf(uint64_t k) {
    return {(int32_t)(k & 0xFFFF), (int32_t)((k >> 16) & 0xFFFF), (int32_t)((k >> 32) & 0xFFFF), (int32_t)(k >> 48)};
}
void keys_of(const std::vector<LeafRec>& recs, std::vector<uint64_t>& keys) {
    keys.resize(recs.size());
    for (size_t i = 0; i < recs.size(); ++i) keys[i] = leaf_key(recs[i]);
    std::sort(keys.begin(), keys.end());
    keys.erase(std::unique(keys.begin(), keys.end()), keys.end());
}

int raster(bgr_map* h, const std::vector<LeafRec>& recs, int set) {
    if (recs.empty()) return BGR_OK;
    // the cached dilations (gate maps) follow the change: a small update re-dilates only the bricks within d
    // cells of the changed leaves (redilate_kernel); a large one marks them stale and the next query rebuilds
    // them in three passes over level 0.  Stale until the update below has gone through.
    const int keep[2] = {h->dil_cells[0], h->dil_cells[1]};
    h->dil_cells[0] = h->dil_cells[1] = 0;
    BGR_CHECK(check_device(h->device));
    BGR_CHECK(h->recs.reserve((int64_t)recs.size(), h->stream, false));
    BGR_CUDA(cudaMemcpyAsync(h->recs.p, recs.data(), recs.size() * sizeof(LeafRec), cudaMemcpyHostToDevice, h->stream));
    raster_kernel<<<(unsigned)div_up((int64_t)recs.size(), 256), 256, 0, h->stream>>>(h->view(), h->recs.p, (int64_t)recs.size(), set);
    BGR_LAUNCH_CHECK();
    int64_t big = 0;                                   // leaves whose region exceeds the slot count: they loop
    for (const LeafRec& r : recs) big += r.level >= 2;
    const bool incremental = (int64_t)recs.size() * kRedilateSlots + big * 512 < h->words[0] / 4;
    for (int k = 0; k < 2; ++k) {
        if (keep[k] <= 0 || !incremental) continue;
        redilate_kernel<<<(unsigned)div_up((int64_t)recs.size() * kRedilateSlots, 256), 256, 0, h->stream>>>(
            h->bits[0].p, h->dil[k].p, h->nb[0][0], h->nb[0][1], h->nb[0][2], keep[k], h->recs.p, (int64_t)recs.size());
        BGR_LAUNCH_CHECK();
        h->dil_cells[k] = keep[k];
        h->n_dilate_incr++;
    }
    BGR_CUDA(cudaStreamSynchronize(h->stream));   // recs is a host temporary
    return BGR_OK;
}

// level 0 grown by d cells in every direction (three separable brick-word passes); lazily
// rebuilt after a map change or when the query radius asks for another d
int ensure_dilated(bgr_map* h, int d, const uint64_t** out) {
    for (int k = 0; k < 2; ++k)
        if (h->dil_cells[k] == d) { *out = h->dil[k].p; return BGR_OK; }
    const int k = h->dil_cells[0] == 0 ? 0 : (h->dil_cells[1] == 0 ? 1 : h->dil_next);
    h->dil_next = k ^ 1;
    h->dil_cells[k] = 0;
    const int64_t w = h->words[0];
    BGR_CHECK(h->dil[k].reserve(w, h->stream, false));
    BGR_CHECK(h->dil_tmp.reserve(w, h->stream, false));
    const unsigned nblk = (unsigned)div_up(w, 256);
    dilate_axis_kernel<<<nblk, 256, 0, h->stream>>>(h->bits[0].p, h->dil[k].p, h->nb[0][0], h->nb[0][1], h->nb[0][2], 0, d);
    BGR_LAUNCH_CHECK();
    dilate_axis_kernel<<<nblk, 256, 0, h->stream>>>(h->dil[k].p, h->dil_tmp.p, h->nb[0][0], h->nb[0][1], h->nb[0][2], 1, d);
    BGR_LAUNCH_CHECK();
    dilate_axis_kernel<<<nblk, 256, 0, h->stream>>>(h->dil_tmp.p, h->dil[k].p, h->nb[0][0], h->nb[0][1], h->nb[0][2], 2, d);
    BGR_LAUNCH_CHECK();
    h->dil_cells[k] = d;
    h->n_dilate_full++;
    *out = h->dil[k].p;
    return BGR_OK;
}

void refresh_mask(bgr_map* h) {
    h->level_mask = 0;
    for (int L = 1; L <= kMaxLevel; ++L)
        if (h->level_count[L] > 0) h->level_mask |= 1u << L;
}

}  // namespace

namespace bgr {
// everything a captured planner step depends on inside the map: the scratch pointers of the collision calls, the
// gate maps and their state, the gate / timing switches.  A different value = re-capture.
uint64_t map_signature(const bgr_map* h) {
    uint64_t v = 1469598103934665603ull;
    auto mix = [&](uint64_t x) { v = (v ^ x) * 1099511628211ull; };
    mix((uint64_t)(uintptr_t)h->eflags.p); mix((uint64_t)(uintptr_t)h->qctl.p); mix((uint64_t)(uintptr_t)h->queue.p);
    mix((uint64_t)h->queue.cap); mix((uint64_t)(uintptr_t)h->stats.p);
    mix((uint64_t)(uintptr_t)h->dil[0].p); mix((uint64_t)(uintptr_t)h->dil[1].p);
    mix((uint64_t)h->dil_cells[0]); mix((uint64_t)h->dil_cells[1]); mix((uint64_t)h->n_dilate_full);
    mix((uint64_t)h->gate_on); mix((uint64_t)h->timing); mix((uint64_t)h->level_mask);
    for (int L = 0; L <= kMaxLevel; ++L) mix((uint64_t)(uintptr_t)h->bits[L].p);
    mix((uint64_t)(uintptr_t)h->stream);
    return v;
}
int map_sphere_gate(bgr_map* h, double radius, LevelGrid* dil) {
    dil->bits = nullptr;
    for (int a = 0; a < 3; ++a) dil->nb[a] = h->nb[0][a];
    const int gate_d = (int)std::ceil((radius + kSweepMargin) / h->res + 1e-9);
    if (h->gate_on && gate_d >= 1 && gate_d <= 3) {
        const uint64_t* bits = nullptr;
        BGR_CHECK(ensure_dilated(h, gate_d, &bits));
        dil->bits = const_cast<uint64_t*>(bits);
    }
    return BGR_OK;
}
}  // namespace bgr

extern "C" {

int bgr_map_create(double res, const double origin[3], const int dims[3], bgr_map** out) {
    BGR_ARG(out != nullptr, "out is null");
    *out = nullptr;
    BGR_ARG(res > 0 && origin && dims, "bad map geometry");
    for (int a = 0; a < 3; ++a) BGR_ARG(dims[a] >= 1 && dims[a] <= 4096, "dims[%d]=%d out of range [1,4096]", a, dims[a]);
    int dev = 0;
    BGR_CUDA(cudaGetDevice(&dev));
    bgr_map* h = new bgr_map;
    h->device = dev;
    h->res = res;
    for (int a = 0; a < 3; ++a) { h->origin[a] = origin[a]; h->dims[a] = dims[a]; }
    for (int L = 0; L <= kMaxLevel; ++L) {
        int64_t w = 1;
        for (int a = 0; a < 3; ++a) {
            const int cells = (dims[a] + (1 << L) - 1) >> L;
            h->nb[L][a] = (cells + 3) >> 2;
            w *= h->nb[L][a];
        }
        h->words[L] = w;
        int rc = h->bits[L].reserve(w, h->stream, false);
        if (rc == BGR_OK && cudaMemsetAsync(h->bits[L].p, 0, (size_t)w * 8, h->stream) != cudaSuccess) {
            set_error("cudaMemset failed");
            rc = BGR_ERR_CUDA;
        }
        if (rc != BGR_OK) { bgr_map_destroy(h); return rc; }
    }
    if (cudaStreamSynchronize(h->stream) != cudaSuccess) { set_error("sync failed"); bgr_map_destroy(h); return BGR_ERR_CUDA; }
    *out = h;
    return BGR_OK;
}

int bgr_map_set_stream(bgr_map* h, void* stream) {
    BGR_ARG(h != nullptr, "null handle");
    BGR_CUDA(cudaStreamSynchronize(h->stream));
    h->stream = (cudaStream_t)stream;
    return BGR_OK;
}

int bgr_map_set_boxes(bgr_map* h, const double* boxes, int64_t nb) {
    BGR_ARG(h != nullptr && nb >= 0 && (nb == 0 || boxes), "bad set_boxes arguments");
    std::vector<LeafRec> recs;
    int64_t counts[kMaxLevel + 1] = {0};
    BGR_CHECK(boxes_to_recs(h, boxes, nb, recs, counts));
    std::vector<uint64_t> keys;
    keys_of(recs, keys);                // de-duplicated: refresh_boxes counts leaves per unique key too
    for (int L = 0; L <= kMaxLevel; ++L) {
        BGR_CUDA(cudaMemsetAsync(h->bits[L].p, 0, (size_t)h->words[L] * 8, h->stream));
        h->level_count[L] = 0;
    }
    h->dil_cells[0] = h->dil_cells[1] = 0;   // level 0 was cleared: the cached dilations are rebuilt in full
    for (uint64_t k : keys) h->level_count[(int)(k >> 48)] += 1;
    refresh_mask(h);
    // the leaf list is trusted by refresh_boxes only once the bitmap really holds it
    h->keys_valid = false;
    BGR_CHECK(raster(h, recs, 1));
    h->leaf_keys.swap(keys);
    h->keys_valid = true;
    return BGR_OK;
}

// Map refresh (the octomap server republishes the whole tree; abb_sense.cpp:447-452 resets it
// every 4 frames): diff the new leaf list against the one currently rasterised and touch only
// the leaves that changed.  Falls back to a full set_boxes when the current list is unknown.
int bgr_map_refresh_boxes(bgr_map* h, const double* boxes, int64_t nb, int64_t* n_added, int64_t* n_removed) {
    BGR_ARG(h != nullptr && nb >= 0 && (nb == 0 || boxes), "bad refresh_boxes arguments");
    if (n_added) *n_added = 0;
    if (n_removed) *n_removed = 0;
    if (!h->keys_valid) {
        BGR_CHECK(bgr_map_set_boxes(h, boxes, nb));
        if (n_added) *n_added = (int64_t)h->leaf_keys.size();
        return BGR_OK;
    }
    std::vector<LeafRec> recs;
    int64_t counts[kMaxLevel + 1] = {0};
    BGR_CHECK(boxes_to_recs(h, boxes, nb, recs, counts));
    std::vector<uint64_t> keys, add_k, rem_k;
    keys_of(recs, keys);
    std::set_difference(keys.begin(), keys.end(), h->leaf_keys.begin(), h->leaf_keys.end(), std::back_inserter(add_k));
    std::set_difference(h->leaf_keys.begin(), h->leaf_keys.end(), keys.begin(), keys.end(), std::back_inserter(rem_k));
    std::vector<LeafRec> add_r(add_k.size()), rem_r(rem_k.size());
    for (size_t i = 0; i < add_k.size(); ++i) add_r[i] = key_leaf(add_k[i]);
    for (size_t i = 0; i < rem_k.size(); ++i) rem_r[i] = key_leaf(rem_k[i]);
    // a failure between the two rasters leaves the bitmap between the two lists: the stored list is
    // invalid until both succeed, and the next refresh then falls back to a full set_boxes
    h->keys_valid = false;
    for (const LeafRec& r : add_r) h->level_count[r.level] += 1;
    refresh_mask(h);                    // a superset of the levels in use while the rasters run
    BGR_CHECK(raster(h, rem_r, 0));     // removals first: a leaf split into children clears, then its children set
    BGR_CHECK(raster(h, add_r, 1));
    for (const LeafRec& r : rem_r) h->level_count[r.level] = std::max<int64_t>(0, h->level_count[r.level] - 1);
    refresh_mask(h);
    h->leaf_keys.swap(keys);
    h->keys_valid = true;
    if (n_added) *n_added = (int64_t)add_k.size();
    if (n_removed) *n_removed = (int64_t)rem_k.size();
    return BGR_OK;
}

int bgr_map_update_boxes(bgr_map* h, const double* add, int64_t n_add, const double* rem, int64_t n_rem) {
    BGR_ARG(h != nullptr && n_add >= 0 && n_rem >= 0, "bad update arguments");
    std::vector<LeafRec> recs;
    int64_t counts[kMaxLevel + 1] = {0};
    if (n_rem > 0) {
        BGR_CHECK(boxes_to_recs(h, rem, n_rem, recs, counts));
        for (int L = 0; L <= kMaxLevel; ++L) h->level_count[L] = std::max<int64_t>(0, h->level_count[L] - counts[L]);
        BGR_CHECK(raster(h, recs, 0));
    }
    if (n_add > 0) {
        std::memset(counts, 0, sizeof(counts));
        BGR_CHECK(boxes_to_recs(h, add, n_add, recs, counts));
        for (int L = 0; L <= kMaxLevel; ++L) h->level_count[L] += counts[L];
        BGR_CHECK(raster(h, recs, 1));
    }
    refresh_mask(h);
    h->keys_valid = false;
    return BGR_OK;
}

int bgr_map_occupied_cells(bgr_map* h, int64_t* cells) {
    BGR_ARG(h != nullptr && cells, "bad arguments");
    BGR_CHECK(h->stats.reserve(4, h->stream, false));
    BGR_CUDA(cudaMemsetAsync(h->stats.p, 0, 8, h->stream));
    popcount_kernel<<<256, 256, 0, h->stream>>>(h->bits[0].p, h->words[0], h->stats.p);
    BGR_LAUNCH_CHECK();
    unsigned long long c = 0;
    BGR_CUDA(cudaMemcpyAsync(&c, h->stats.p, 8, cudaMemcpyDeviceToHost, h->stream));
    BGR_CUDA(cudaStreamSynchronize(h->stream));
    *cells = (int64_t)c;
    return BGR_OK;
}

int bgr_map_download_bitmap(bgr_map* h, uint64_t* out, int64_t* n_words) {
    BGR_ARG(h != nullptr && n_words, "bad arguments");
    *n_words = h->words[0];
    if (!out) return BGR_OK;
    BGR_CUDA(cudaMemcpyAsync(out, h->bits[0].p, (size_t)h->words[0] * 8, cudaMemcpyDeviceToHost, h->stream));
    BGR_CUDA(cudaStreamSynchronize(h->stream));
    return BGR_OK;
}

int bgr_map_collide_spheres_dev(bgr_map* h, const float* d_c, int64_t n, double radius, uint8_t* d_hit, uint8_t* d_band) {
    BGR_ARG(h != nullptr && n >= 0 && radius >= 0, "bad collide_spheres arguments");
    if (n == 0) return BGR_OK;
    BGR_CHECK(check_device(h->device));
    LevelGrid dil;
    BGR_CHECK(map_sphere_gate(h, radius, &dil));
    // large batches of small spheres: a lane per sphere (spheres.cuh); large batches of larger spheres: 32 spheres per
    // warp (lane-per-sphere gate, then the warp walks the undecided ones); small batches: a warp per sphere so the
    // walks overlap
    // BGR_SPHERES=lane|warp32|warp1 forces one form (test / measurement aid, read per call)
    // (measured crossover of the lane form against a warp per sphere: ~20 k spheres, profiles/r2_spheres_lane.txt)
    int form = (n >= 24576 && spheres_lane_ok(radius, h->view().inv_res)) ? 2 : (n >= (int64_t)num_sms() * 1024 ? 1 : 0);
    if (const char* e = getenv("BGR_SPHERES")) {
        if (!strcmp(e, "lane") && spheres_lane_ok(radius, h->view().inv_res)) form = 2;
        else if (!strcmp(e, "warp32")) form = 1;
        else if (!strcmp(e, "warp1")) form = 0;
    }
    if (form == 2)
        collide_spheres_lane_kernel<<<(unsigned)div_up(n, kLaneThreads), kLaneThreads, 0, h->stream>>>(h->view(), d_c, n, radius, d_hit, d_band, dil);
    else if (form == 1)
        collide_spheres_kernel<32><<<(unsigned)div_up(n, 128), 128, 0, h->stream>>>(h->view(), d_c, n, radius, d_hit, d_band, dil);
    else
        collide_spheres_kernel<1><<<(unsigned)div_up(n, 4), 128, 0, h->stream>>>(h->view(), d_c, n, radius, d_hit, d_band, dil);
    BGR_LAUNCH_CHECK();
    return BGR_OK;
}

int bgr_map_collide_spheres(bgr_map* h, const float* c, int64_t n, double radius, uint8_t* hit, uint8_t* band) {
    BGR_ARG(h != nullptr && n >= 0 && (n == 0 || (c && hit)), "bad collide_spheres arguments");
    if (n == 0) return BGR_OK;
    BGR_CHECK(h->p1_dev.reserve(3 * n, h->stream, false));
    BGR_CHECK(h->hit_dev.reserve(n, h->stream, false));
    BGR_CHECK(h->band_dev.reserve(n, h->stream, false));
    // large batches: cut into chunks so the upload of chunk c+1 and the download of chunk c-1 overlap the
    // kernel of chunk c (results are per sphere: chunking cannot change them)
    const int chunks = pipe_disabled() ? 1 : (int)std::min<int64_t>(4, n / 131072);
    if (chunks <= 1) {
        BGR_CUDA(cudaMemcpyAsync(h->p1_dev.p, c, (size_t)3 * n * sizeof(float), cudaMemcpyHostToDevice, h->stream));
        BGR_CHECK(bgr_map_collide_spheres_dev(h, h->p1_dev.p, n, radius, h->hit_dev.p, band ? h->band_dev.p : nullptr));
        BGR_CUDA(cudaMemcpyAsync(hit, h->hit_dev.p, (size_t)n, cudaMemcpyDeviceToHost, h->stream));
        if (band) BGR_CUDA(cudaMemcpyAsync(band, h->band_dev.p, (size_t)n, cudaMemcpyDeviceToHost, h->stream));
        BGR_CUDA(cudaStreamSynchronize(h->stream));
        return BGR_OK;
    }
    BGR_CHECK(h->pipe.init());
    CopyPipe& P = h->pipe;
    BGR_CUDA(cudaEventRecord(P.ev_start, h->stream));
    BGR_CUDA(cudaStreamWaitEvent(P.s_in, P.ev_start, 0));
    BGR_CUDA(cudaStreamWaitEvent(P.s_out, P.ev_start, 0));
    const int64_t per = round_up(div_up(n, chunks), 128);
    for (int k = 0; k < chunks; ++k) {
        const int64_t s0 = (int64_t)k * per, m = std::min<int64_t>(per, n - s0);
        if (m <= 0) break;
        BGR_CUDA(cudaMemcpyAsync(h->p1_dev.p + 3 * s0, c + 3 * s0, (size_t)3 * m * sizeof(float), cudaMemcpyHostToDevice, P.s_in));
        BGR_CUDA(cudaEventRecord(P.ev_in[k], P.s_in));
    }
    // every CUDA return is checked; on a failure the three streams are still drained before returning,
    // so no copy into the caller's buffers is left in flight
    auto chunk = [&](int k) -> int {
        const int64_t s0 = (int64_t)k * per, m = std::min<int64_t>(per, n - s0);
        if (m <= 0) return BGR_OK;
        BGR_CUDA(cudaStreamWaitEvent(h->stream, P.ev_in[k], 0));
        BGR_CHECK(bgr_map_collide_spheres_dev(h, h->p1_dev.p + 3 * s0, m, radius, h->hit_dev.p + s0, band ? h->band_dev.p + s0 : nullptr));
        BGR_CUDA(cudaEventRecord(P.ev_out[k], h->stream));
        BGR_CUDA(cudaStreamWaitEvent(P.s_out, P.ev_out[k], 0));
        BGR_CUDA(cudaMemcpyAsync(hit + s0, h->hit_dev.p + s0, (size_t)m, cudaMemcpyDeviceToHost, P.s_out));
        if (band) BGR_CUDA(cudaMemcpyAsync(band + s0, h->band_dev.p + s0, (size_t)m, cudaMemcpyDeviceToHost, P.s_out));
        return BGR_OK;
    };
    int rc = BGR_OK;
    for (int k = 0; k < chunks && rc == BGR_OK; ++k) rc = chunk(k);
    const cudaError_t e1 = cudaStreamSynchronize(P.s_in), e2 = cudaStreamSynchronize(P.s_out), e3 = cudaStreamSynchronize(h->stream);
    if (rc == BGR_OK && (e1 != cudaSuccess || e2 != cudaSuccess || e3 != cudaSuccess)) {
        set_error("collide_spheres: stream synchronisation failed: %s", cudaGetErrorString(e1 != cudaSuccess ? e1 : (e2 != cudaSuccess ? e2 : e3)));
        rc = BGR_ERR_CUDA;
    }
    return rc;
}

// fast path needs the swept hull of one slab to fit a kRectMax x kRectMax cell rectangle:
// extent <= res (slope <= 1 on the dominant axis) + 2*(radius+margin) + 2*(radius+margin)
static bool fast_ok(const bgr_map* h, double radius) {
    return (h->res + 4.0 * (radius + kSweepMargin)) / h->res + 1.0 <= (double)kRectMax;
}

int bgr_map_collide_edges_dev(bgr_map* h, const float* d_p1, const float* d_p2, int64_t n, double radius,
                              int shape, uint8_t* d_hit, uint8_t* d_band) {
    BGR_ARG(h != nullptr && n >= 0 && radius >= 0, "bad collide_edges arguments");
    BGR_ARG(shape == BGR_SHAPE_CYLINDER || shape == BGR_SHAPE_CAPSULE, "unknown shape %d", shape);
    BGR_ARG(radius + kSweepMargin < 64 * h->res, "radius too large for the sweep (>= 64 cells)");
    BGR_ARG(n < ((int64_t)1 << 31), "too many edges in one call");
    if (n == 0) return BGR_OK;
    BGR_CHECK(check_device(h->device));
    BGR_CHECK(h->stats.reserve(6, h->stream, false));
    BGR_CHECK(h->eflags.reserve(n, h->stream, false));
    BGR_CHECK(h->qctl.reserve(2, h->stream, false));
    BGR_CUDA(cudaMemsetAsync(h->stats.p, 0, 6 * sizeof(unsigned long long), h->stream));
    BGR_CUDA(cudaMemsetAsync(h->qctl.p, 0, 2 * sizeof(unsigned), h->stream));
    EdgeArgs ma;
    ma.m = h->view();
    ma.p1 = d_p1; ma.p2 = d_p2; ma.n = n; ma.radius = radius; ma.shape = shape;
    ma.want_band = d_band != nullptr; ma.gflags = h->eflags.p; ma.stats = h->stats.p;
    const bool fast = fast_ok(h, radius);
    // sweep gate: level 0 dilated by d = ceil(rm / res) cells (built outside the timed kernels:
    // it belongs to the map, not to the query batch)
    const int gate_d = (int)std::ceil((radius + kSweepMargin) / h->res + 1e-9);
    const bool gate = fast_ok(h, radius) && h->gate_on && gate_d >= 1 && gate_d <= 3;
    const uint64_t* dil_bits = nullptr;
    if (gate) BGR_CHECK(ensure_dilated(h, gate_d, &dil_bits));
    h->last_gate_d = gate ? gate_d : 0;
    if (h->timing) BGR_CUDA(cudaEventRecord(h->tev[0], h->stream));
    if (fast) {
        const int64_t cap = std::max<int64_t>(1 << 16, 4 * n);
        BGR_CHECK(h->queue.reserve(cap, h->stream, false));
        Edge2Args a;
        a.m = ma.m;
        a.p1 = d_p1; a.p2 = d_p2; a.n = n; a.radius = radius; a.shape = shape;
        a.want_band = ma.want_band; a.gflags = h->eflags.p;
        a.queue = h->queue.p; a.queue_cap = (unsigned)std::min<int64_t>(cap, 0x7fffffff);
        a.qctl = h->qctl.p; a.stats = h->stats.p; a.hit = d_hit; a.band = d_band;
        a.gate_d = gate ? gate_d : 0;
        a.gate_seg = kGateSeg; a.gate_rounds = kGateRounds;
        if (const char* e = getenv("BGR_GATE")) {          // measurement aid: "seg,rounds"
            int sg = 0, rd = 0;
            if (sscanf(e, "%d,%d", &sg, &rd) == 2 && sg >= 1 && sg <= 32 && 64 % sg == 0 && rd >= 0) { a.gate_seg = sg; a.gate_rounds = rd; }
        }
        a.dil.bits = const_cast<uint64_t*>(dil_bits);
        for (int ax = 0; ax < 3; ++ax) a.dil.nb[ax] = h->nb[0][ax];
        // one sweep launch per chunk of a chunked host call (each waits for its H2D copy), one
        // narrow / finalize pass over the whole batch
        const int nch = h->pipe_chunks > 1 ? h->pipe_chunks : 1;
        for (int c = 0; c < nch; ++c) {
            a.e_begin = nch > 1 ? h->pipe_begin[c] : 0;
            a.n = nch > 1 ? h->pipe_begin[c + 1] : n;
            if (a.n <= a.e_begin) continue;
            if (nch > 1) BGR_CUDA(cudaStreamWaitEvent(h->stream, h->pipe.ev_in[c], 0));
            edges_sweep_kernel<<<(unsigned)div_up(a.n - a.e_begin, kSweepEdges), kSweepThreads, 0, h->stream>>>(a);
            BGR_LAUNCH_CHECK();
        }
        a.e_begin = 0;
        a.n = n;
        if (h->timing) BGR_CUDA(cudaEventRecord(h->tev[1], h->stream));
        const int nblk = (int)std::min<int64_t>((ma.want_band ? 16LL : 4LL) * num_sms(), std::max<int64_t>(1, div_up(cap, 32)));
        edges_narrow_kernel<<<nblk, 128, 0, h->stream>>>(a);
        BGR_LAUNCH_CHECK();
        ma.only_overflow = 1;   // no-op unless the queue overflowed
        ma.n_overflow = h->qctl.p + 1;
    } else {
        for (int c = 0; c < h->pipe_chunks; ++c) BGR_CUDA(cudaStreamWaitEvent(h->stream, h->pipe.ev_in[c], 0));
        ma.only_overflow = 0;
        ma.n_overflow = nullptr;
        if (h->timing) BGR_CUDA(cudaEventRecord(h->tev[1], h->stream));
    }
    if (h->timing) BGR_CUDA(cudaEventRecord(h->tev[2], h->stream));
    collide_edges_kernel<<<(unsigned)std::min<int64_t>(div_up(n, kEdgesPerCta), ma.only_overflow ? 2LL * num_sms() : (1LL << 30)), kEdgeThreads, 0, h->stream>>>(ma);
    BGR_LAUNCH_CHECK();
    if (h->timing) BGR_CUDA(cudaEventRecord(h->tev[3], h->stream));
    edges_finalize_kernel<<<(unsigned)div_up(n, 256), 256, 0, h->stream>>>(h->eflags.p, n, d_hit, d_band);
    BGR_LAUNCH_CHECK();
    if (h->timing) BGR_CUDA(cudaEventRecord(h->tev[4], h->stream));
    return BGR_OK;
}

int bgr_map_collide_edges(bgr_map* h, const float* p1, const float* p2, int64_t n, double radius, int shape,
                          uint8_t* hit, uint8_t* band) {
    BGR_ARG(h != nullptr && n >= 0 && (n == 0 || (p1 && p2 && hit)), "bad collide_edges arguments");
    if (n == 0) return BGR_OK;
    // experiment (BGR_EDGES_ZEROCOPY=1): page-locked caller buffers are read and written by the kernels themselves over
    // the host link (UVA device pointers), instead of staged copies
    static const bool zero_copy = getenv("BGR_EDGES_ZEROCOPY") && getenv("BGR_EDGES_ZEROCOPY")[0] == '1';
    if (zero_copy) {
        cudaPointerAttributes a1{}, a2{}, ah{}, ab{};
        const bool ok = cudaPointerGetAttributes(&a1, p1) == cudaSuccess && a1.type == cudaMemoryTypeHost && a1.devicePointer &&
                        cudaPointerGetAttributes(&a2, p2) == cudaSuccess && a2.type == cudaMemoryTypeHost && a2.devicePointer &&
                        cudaPointerGetAttributes(&ah, hit) == cudaSuccess && ah.type == cudaMemoryTypeHost && ah.devicePointer &&
                        (!band || (cudaPointerGetAttributes(&ab, band) == cudaSuccess && ab.type == cudaMemoryTypeHost && ab.devicePointer));
        (void)cudaGetLastError();
        if (ok) {
            BGR_CHECK(bgr_map_collide_edges_dev(h, (const float*)a1.devicePointer, (const float*)a2.devicePointer, n, radius, shape,
                                                (uint8_t*)ah.devicePointer, band ? (uint8_t*)ab.devicePointer : nullptr));
            BGR_CUDA(cudaStreamSynchronize(h->stream));
            return BGR_OK;
        }
    }
    BGR_CHECK(h->p1_dev.reserve(3 * n, h->stream, false));
    BGR_CHECK(h->p2_dev.reserve(3 * n, h->stream, false));
    BGR_CHECK(h->hit_dev.reserve(n, h->stream, false));
    BGR_CHECK(h->band_dev.reserve(n, h->stream, false));
    // large batches are cut into chunks: the H2D copy of chunk c+1 and the D2H copy of chunk c-1
    // overlap the kernels of chunk c (edges are independent, so chunking cannot change results)
    const int chunks = pipe_disabled() ? 1 : (int)std::min<int64_t>(4, n / 65536);
    if (chunks <= 1) {
        BGR_CUDA(cudaMemcpyAsync(h->p1_dev.p, p1, (size_t)3 * n * sizeof(float), cudaMemcpyHostToDevice, h->stream));
        BGR_CUDA(cudaMemcpyAsync(h->p2_dev.p, p2, (size_t)3 * n * sizeof(float), cudaMemcpyHostToDevice, h->stream));
        BGR_CHECK(bgr_map_collide_edges_dev(h, h->p1_dev.p, h->p2_dev.p, n, radius, shape, h->hit_dev.p,
                                            band ? h->band_dev.p : nullptr));
        BGR_CUDA(cudaMemcpyAsync(hit, h->hit_dev.p, (size_t)n, cudaMemcpyDeviceToHost, h->stream));
        if (band) BGR_CUDA(cudaMemcpyAsync(band, h->band_dev.p, (size_t)n, cudaMemcpyDeviceToHost, h->stream));
        BGR_CUDA(cudaStreamSynchronize(h->stream));
        return BGR_OK;
    }
    BGR_CHECK(h->pipe.init());
    CopyPipe& P = h->pipe;
    // the staging buffers may still be read by earlier work on the map's stream
    BGR_CUDA(cudaEventRecord(P.ev_start, h->stream));
    BGR_CUDA(cudaStreamWaitEvent(P.s_in, P.ev_start, 0));
    // equal chunks (measured: tapering them towards the end, or more than 4, is slower — the first
    // sweep then starts later / the launches get too small)
    h->pipe_begin[0] = 0;
    for (int c = 0; c < chunks; ++c)
        h->pipe_begin[c + 1] = c + 1 == chunks ? n : std::min<int64_t>(n, round_up(n * (int64_t)(c + 1) / chunks, 256));
    for (int c = 0; c < chunks; ++c) {
        const int64_t s0 = h->pipe_begin[c], m = h->pipe_begin[c + 1] - s0;
        if (m > 0) {
            BGR_CUDA(cudaMemcpyAsync(h->p1_dev.p + 3 * s0, p1 + 3 * s0, (size_t)3 * m * sizeof(float), cudaMemcpyHostToDevice, P.s_in));
            BGR_CUDA(cudaMemcpyAsync(h->p2_dev.p + 3 * s0, p2 + 3 * s0, (size_t)3 * m * sizeof(float), cudaMemcpyHostToDevice, P.s_in));
        }
        BGR_CUDA(cudaEventRecord(P.ev_in[c], P.s_in));
    }
    h->pipe_chunks = chunks;
    const int rc = bgr_map_collide_edges_dev(h, h->p1_dev.p, h->p2_dev.p, n, radius, shape, h->hit_dev.p,
                                             band ? h->band_dev.p : nullptr);
    h->pipe_chunks = 0;
    cudaError_t ec = cudaSuccess;
    if (rc == BGR_OK) {
        ec = cudaMemcpyAsync(hit, h->hit_dev.p, (size_t)n, cudaMemcpyDeviceToHost, h->stream);
        if (ec == cudaSuccess && band) ec = cudaMemcpyAsync(band, h->band_dev.p, (size_t)n, cudaMemcpyDeviceToHost, h->stream);
    }
    const cudaError_t e1 = cudaStreamSynchronize(P.s_in), e2 = cudaStreamSynchronize(h->stream);
    if (rc == BGR_OK && (ec != cudaSuccess || e1 != cudaSuccess || e2 != cudaSuccess)) {
        set_error("collide_edges: copy / synchronisation failed: %s", cudaGetErrorString(ec != cudaSuccess ? ec : (e1 != cudaSuccess ? e1 : e2)));
        return BGR_ERR_CUDA;
    }
    return rc;
}

int bgr_map_set_timing(bgr_map* h, int on) {
    BGR_ARG(h != nullptr, "null map");
    if (on && !h->tev[0])
        for (int i = 0; i < 5; ++i) BGR_CUDA(cudaEventCreate(&h->tev[i]));
    h->timing = on != 0;
    return BGR_OK;
}

int bgr_map_last_kernel_ms(bgr_map* h, float out[4]) {
    BGR_ARG(h != nullptr && out != nullptr && h->tev[0] != nullptr, "timing was never enabled on this map");
    BGR_CUDA(cudaEventSynchronize(h->tev[4]));
    for (int i = 0; i < 4; ++i) BGR_CUDA(cudaEventElapsedTime(&out[i], h->tev[i], h->tev[i + 1]));
    return BGR_OK;
}

int bgr_map_last_stats(bgr_map* h, uint64_t out[4]) {
    BGR_ARG(h != nullptr && out, "bad arguments");
    if (!h->stats.p) { out[0] = out[1] = out[2] = out[3] = 0; return BGR_OK; }
    unsigned long long s[4];
    BGR_CUDA(cudaMemcpyAsync(s, h->stats.p, sizeof(s), cudaMemcpyDeviceToHost, h->stream));
    BGR_CUDA(cudaStreamSynchronize(h->stream));
    for (int i = 0; i < 4; ++i) out[i] = s[i];
    return BGR_OK;
}

int bgr_map_set_gate(bgr_map* h, int on) {
    BGR_ARG(h != nullptr, "null map");
    h->gate_on = on != 0;
    return BGR_OK;
}

int bgr_map_last_gate_stats(bgr_map* h, uint64_t out[2]) {
    BGR_ARG(h != nullptr && out, "bad arguments");
    out[0] = 0;
    out[1] = (uint64_t)h->last_gate_d;
    if (!h->stats.p || h->stats.cap < 6) return BGR_OK;
    unsigned long long s = 0;
    BGR_CUDA(cudaMemcpyAsync(&s, h->stats.p + 4, sizeof(s), cudaMemcpyDeviceToHost, h->stream));
    BGR_CUDA(cudaStreamSynchronize(h->stream));
    out[0] = s;
    return BGR_OK;
}

int bgr_map_download_gate(bgr_map* h, int d, uint64_t* out, int64_t* n_words) {
    BGR_ARG(h != nullptr && n_words && d >= 1 && d <= 3, "bad arguments (d in 1..3)");
    *n_words = h->words[0];
    if (!out) return BGR_OK;
    BGR_CHECK(check_device(h->device));
    const uint64_t* bits = nullptr;
    BGR_CHECK(ensure_dilated(h, d, &bits));
    BGR_CUDA(cudaMemcpyAsync(out, bits, (size_t)h->words[0] * 8, cudaMemcpyDeviceToHost, h->stream));
    BGR_CUDA(cudaStreamSynchronize(h->stream));
    return BGR_OK;
}

int bgr_map_dilation_stats(bgr_map* h, uint64_t out[2]) {
    BGR_ARG(h != nullptr && out, "bad arguments");
    out[0] = h->n_dilate_full;
    out[1] = h->n_dilate_incr;
    return BGR_OK;
}

int bgr_map_destroy(bgr_map* h) {
    if (!h) return BGR_OK;
    cudaStreamSynchronize(h->stream);
    for (int L = 0; L <= kMaxLevel; ++L) h->bits[L].release();
    h->recs.release(); h->p1_dev.release(); h->p2_dev.release(); h->hit_dev.release(); h->band_dev.release();
    h->stats.release(); h->eflags.release(); h->qctl.release(); h->queue.release();
    h->dil[0].release(); h->dil[1].release(); h->dil_tmp.release();
    h->pipe.release();
    if (h->aux && h->aux_free) h->aux_free(h->aux);
    for (int i = 0; i < 5; ++i)
        if (h->tev[i]) cudaEventDestroy(h->tev[i]);
    delete h;
    return BGR_OK;
}

}  // extern "C"
