// bgr_path_* / bgr_map_prune_path / bgr_map_clearance — the planner's path post-processing
//  around the edge and sphere queries of map.cu.
//   PrunePlan     MS:562-609  greedy shortcutting: ALL P(P-1)/2 anchor-waypoint edges are checked
//                             in one collide_edges launch, the greedy walk reads the hit matrix
//   SmoothenPlan  MS:645-799  Bezier resampling (11 samples; host arithmetic, the reference's types)
//   find_labels   MS:943-996  per-waypoint clearance = min over all leaves of
//                             fcl::distance(Sphere(r), Box); 3 smallest -> milestone labels
#include <algorithm>
#include <cmath>
#include <cstring>
#include <numeric>
#include <vector>
#include "mapview.cuh"

using namespace bgr;

namespace {

// edge list of all pairs i < j of a P-waypoint path; pair (i, j) sits at
// e = i*P - i*(i+1)/2 + (j - i - 1)
__global__ void pair_edges_kernel(const float* __restrict__ path, int p, float* __restrict__ p1, float* __restrict__ p2) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
    if (j <= i || j >= p) return;
    const int64_t e = (int64_t)i * p - (int64_t)i * (i + 1) / 2 + (j - i - 1);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        p1[3 * e + c] = path[3 * i + c];
        p2[3 * e + c] = path[3 * j + c];
    }
}

// non-negative doubles order like their bit patterns
__device__ __forceinline__ void atomic_min_pos_double(double* addr, double v) {
    atomicMin(reinterpret_cast<unsigned long long*>(addr), (unsigned long long)__double_as_longlong(v));
}

constexpr int kClrPts = 128;   // waypoints staged per CTA pass

// One thread per level-0 brick word (grid-stride); every occupied cell resolves to its Octomap
// leaf and the point-box distance is taken to that leaf (the boxes the reference iterates).
// best[] holds min over leaves of max(dist(point, box), 0) as a double; the caller turns it into
// FCL's convention (distance - r, or -1 when the shapes intersect).
__global__ void __launch_bounds__(256) clearance_kernel(MapView m, const double* __restrict__ pts, int npts,
                                                        double* __restrict__ best) {
    __shared__ double sp[kClrPts][3];
    __shared__ double sbest[kClrPts];
    const LevelGrid& g = m.lv[0];
    const int64_t words = (int64_t)g.nb[0] * g.nb[1] * g.nb[2];
    for (int p0 = 0; p0 < npts; p0 += kClrPts) {
        const int np = min(kClrPts, npts - p0);
        __syncthreads();
        if (threadIdx.x < np) {
            sp[threadIdx.x][0] = pts[3 * (p0 + threadIdx.x)];
            sp[threadIdx.x][1] = pts[3 * (p0 + threadIdx.x) + 1];
            sp[threadIdx.x][2] = pts[3 * (p0 + threadIdx.x) + 2];
            sbest[threadIdx.x] = best[p0 + threadIdx.x];
        }
        __syncthreads();
        for (int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w < words; w += (int64_t)gridDim.x * blockDim.x) {
            unsigned long long bits = __ldg(g.bits + w);
            if (!bits) continue;
            const int bx = (int)(w % g.nb[0]), by = (int)((w / g.nb[0]) % g.nb[1]), bz = (int)(w / ((int64_t)g.nb[0] * g.nb[1]));
            // brick box (4 cells per axis)
            const double h = 2.0 * m.res;
            const double cx = m.origin[0] + (4.0 * bx + 2.0) * m.res, cy = m.origin[1] + (4.0 * by + 2.0) * m.res,
                         cz = m.origin[2] + (4.0 * bz + 2.0) * m.res;
            for (int p = 0; p < np; ++p) {
                const double tx = fmax(0.0, fabs(sp[p][0] - cx) - h), ty = fmax(0.0, fabs(sp[p][1] - cy) - h),
                             tz = fmax(0.0, fabs(sp[p][2] - cz) - h);
                // lower bound of every cell of the brick (and of every leaf through it: a larger leaf
                // reaches beyond the brick only away from the cells tested here, never closer than them)
                const double lb = sqrt(tx * tx + ty * ty + tz * tz) * (1.0 - 1e-12);
                if (lb > sbest[p]) continue;
                double loc = sbest[p];
                unsigned long long b = bits;
                while (b) {
                    const int bit = __ffsll((long long)b) - 1;
                    b &= b - 1;
                    const int x = 4 * bx + (bit & 3), y = 4 * by + ((bit >> 2) & 3), z = 4 * bz + (bit >> 4);
                    int lx, ly, lz;
                    const int L = find_leaf(m, x, y, z, lx, ly, lz);
                    D3 bc;
                    double e;
                    leaf_box(m, L, lx, ly, lz, bc, e);
                    const double ux = fmax(0.0, fabs(sp[p][0] - bc.x) - e), uy = fmax(0.0, fabs(sp[p][1] - bc.y) - e),
                                 uz = fmax(0.0, fabs(sp[p][2] - bc.z) - e);
                    double d2 = ux * ux;
                    d2 += uy * uy;
                    d2 += uz * uz;
                    loc = fmin(loc, sqrt(d2));
                }
                if (loc < sbest[p]) atomic_min_pos_double(&sbest[p], loc);
            }
        }
        __syncthreads();
        if (threadIdx.x < np) atomic_min_pos_double(&best[p0 + threadIdx.x], sbest[threadIdx.x]);
    }
}

struct Scratch {
    DevBuf<float> path, p1, p2;
    DevBuf<uint8_t> hit, band;
    DevBuf<double> pts, best;
};
// the scratch belongs to the map handle (its stream and device), not to the process or the thread
void free_scratch(void* p) {
    Scratch* s = static_cast<Scratch*>(p);
    s->path.release(); s->p1.release(); s->p2.release(); s->hit.release(); s->band.release();
    s->pts.release(); s->best.release();
    delete s;
}
Scratch& scratch_of(bgr_map* map) {
    void** slot = map_aux(map, free_scratch);
    if (!*slot) *slot = new Scratch;
    return *static_cast<Scratch*>(*slot);
}

}  // namespace

extern "C" {

int bgr_map_prune_path(bgr_map* map, const float* path_3xP, int64_t p, double radius, int shape,
                       int32_t* keep, int64_t* n_keep, int64_t* n_band) {
    BGR_ARG(map != nullptr && path_3xP != nullptr && keep != nullptr && n_keep != nullptr, "bad prune_path arguments");
    BGR_ARG(p >= 1, "empty path");
    if (p > 4096) {
        set_error("prune_path: %lld waypoints > 4096 (the all-pairs edge batch would exceed 8.4M edges)", (long long)p);
        return BGR_ERR_UNSUPPORTED;
    }
    BGR_ARG(*n_keep >= p, "keep buffer too small (%lld < %lld)", (long long)*n_keep, (long long)p);
    BGR_CHECK(check_device(map_device(map)));
    cudaStream_t st = map_stream(map);
    Scratch& s = scratch_of(map);
    const int64_t m = p * (p - 1) / 2;
    std::vector<uint8_t> hit((size_t)m), band((size_t)m);
    if (m > 0) {
        BGR_CHECK(s.path.reserve(3 * p, st, false));
        BGR_CHECK(s.p1.reserve(3 * m, st, false));
        BGR_CHECK(s.p2.reserve(3 * m, st, false));
        BGR_CHECK(s.hit.reserve(m, st, false));
        BGR_CHECK(s.band.reserve(m, st, false));
        BGR_CUDA(cudaMemcpyAsync(s.path.p, path_3xP, (size_t)3 * p * sizeof(float), cudaMemcpyHostToDevice, st));
        dim3 grid((unsigned)div_up(p, 128), (unsigned)p);
        pair_edges_kernel<<<grid, 128, 0, st>>>(s.path.p, (int)p, s.p1.p, s.p2.p);
        BGR_LAUNCH_CHECK();
        BGR_CHECK(bgr_map_collide_edges_dev(map, s.p1.p, s.p2.p, m, radius, shape, s.hit.p, s.band.p));
        BGR_CUDA(cudaMemcpyAsync(hit.data(), s.hit.p, (size_t)m, cudaMemcpyDeviceToHost, st));
        BGR_CUDA(cudaMemcpyAsync(band.data(), s.band.p, (size_t)m, cudaMemcpyDeviceToHost, st));
        BGR_CUDA(cudaStreamSynchronize(st));
    }
    // the reference's greedy walk (MS:579-606) over the precomputed answers
    auto pair = [&](int64_t i, int64_t j) { return i * p - i * (i + 1) / 2 + (j - i - 1); };
    int64_t nk = 0, nb = 0, cur = 0;
    keep[nk++] = 0;
    while (path_3xP[3 * keep[nk - 1]] != path_3xP[3 * (p - 1)]) {     // MS:579: the x coordinate alone ends the loop
        bool found = false;
        for (int64_t it = p - 1; it > cur; --it) {                   // MS:584
            const int64_t e = pair(cur, it);
            nb += band[(size_t)e];
            if (!hit[(size_t)e]) {                                    // MS:591: line is safe
                keep[nk++] = (int32_t)it;
                cur = it;
                found = true;
                break;
            }
        }
        if (!found) {
            set_error("PrunePlan: no collision-free successor from waypoint %lld (the reference loops forever here)", (long long)cur);
            *n_keep = nk;
            return BGR_ERR_STATE;
        }
    }
    *n_keep = nk;
    if (n_band) *n_band = nb;
    return BGR_OK;
}

int bgr_path_smoothen(const double* path_3xP, int64_t p, double* out_3xS, int64_t* n_out) {
    BGR_ARG(path_3xP != nullptr && n_out != nullptr, "bad smoothen arguments");
    BGR_ARG(p >= 2, "SmoothenPlan reads rrt_path[1]: needs >= 2 waypoints");
    std::vector<double> out;
    if (p == 3) {                                                     // MS:653-711
        float a[3], b[3], c[3], ctl[3];
        for (int j = 0; j < 3; ++j) {
            a[j] = (float)path_3xP[j];
            b[j] = (float)path_3xP[3 + j];
            c[j] = (float)path_3xP[6 + j];
            ctl[j] = b[j] * 2 - ((a[j] + c[j]) / 2);                  // MS:671-673
        }
        for (float t = 0; t < 1.1; t = t + 0.1)                       // MS:676
            for (int j = 0; j < 3; ++j) out.push_back(std::pow((double)(1 - t), 2.0) * a[j] + (1 - t) * 2 * t * ctl[j] + t * t * c[j]);   // pow(float,int) -> double
    } else if (p > 3) {                                               // MS:713-752: the first four waypoints only
        for (double t = 0.0; t < 1.0; t += 0.1)
            for (int j = 0; j < 3; ++j) {
                const float x0 = (float)path_3xP[j], x1 = (float)path_3xP[3 + j], x2 = (float)path_3xP[6 + j],
                            x3 = (float)path_3xP[9 + j];
                out.push_back(std::pow(1 - t, 3.0) * x0 + 3 * t * std::pow(1 - t, 2.0) * x1 + 3 * std::pow(t, 2.0) * (1 - t) * x2 + std::pow(t, 3.0) * x3);
            }
    } else {                                                          // MS:754-790: straight line, resampled
        float a[3], u[3];
        for (int j = 0; j < 3; ++j) {
            a[j] = (float)path_3xP[j];
            u[j] = (float)path_3xP[3 + j] - a[j];
        }
        for (float t = 0; t < 1.1; t = t + 0.1)
            for (int j = 0; j < 3; ++j) out.push_back((double)(a[j] + (u[j] * t)));
    }
    const int64_t ns = (int64_t)out.size() / 3;
    if (out_3xS) {
        BGR_ARG(*n_out >= ns, "output buffer too small (%lld < %lld)", (long long)*n_out, (long long)ns);
        std::memcpy(out_3xS, out.data(), out.size() * sizeof(double));
    }
    *n_out = ns;
    return BGR_OK;
}

int bgr_map_clearance(bgr_map* map, const double* pts_3xM, int64_t m, double radius, double* dist) {
    BGR_ARG(map != nullptr && m >= 0 && (m == 0 || (pts_3xM && dist)) && radius >= 0, "bad clearance arguments");
    if (m == 0) return BGR_OK;
    BGR_ARG(m <= (1 << 20), "too many points in one call");
    BGR_CHECK(check_device(map_device(map)));
    cudaStream_t st = map_stream(map);
    Scratch& s = scratch_of(map);
    BGR_CHECK(s.pts.reserve(3 * m, st, false));
    BGR_CHECK(s.best.reserve(m, st, false));
    std::vector<double> init((size_t)m, 1000.0 + radius);            // MS:962 min_dist = 1000
    BGR_CUDA(cudaMemcpyAsync(s.pts.p, pts_3xM, (size_t)3 * m * sizeof(double), cudaMemcpyHostToDevice, st));
    BGR_CUDA(cudaMemcpyAsync(s.best.p, init.data(), (size_t)m * sizeof(double), cudaMemcpyHostToDevice, st));
    const MapView v = map_view(map);
    const int64_t words = (int64_t)v.lv[0].nb[0] * v.lv[0].nb[1] * v.lv[0].nb[2];
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>(8LL * num_sms(), div_up(words, 256)));
    clearance_kernel<<<grid, 256, 0, st>>>(v, s.pts.p, (int)m, s.best.p);
    BGR_LAUNCH_CHECK();
    BGR_CUDA(cudaMemcpyAsync(dist, s.best.p, (size_t)m * sizeof(double), cudaMemcpyDeviceToHost, st));
    BGR_CUDA(cudaStreamSynchronize(st));
    for (int64_t i = 0; i < m; ++i) {
        const double d = dist[i] - radius;                           // sphere surface to box
        dist[i] = d <= 0.0 ? -1.0 : std::min(d, 1000.0);            // fcl::distance: -1 when the shapes intersect
    }
    return BGR_OK;
}

int bgr_map_find_labels(bgr_map* map, const double* path_3xP, int64_t p, double radius, int n_labels,
                        int32_t* idx, float* labels_3xL, float* clearance, int* n_ties) {
    BGR_ARG(map != nullptr && path_3xP != nullptr && idx != nullptr && labels_3xL != nullptr, "bad find_labels arguments");
    const int64_t n = p - 2;                                          // MS:950-952: start and goal removed
    BGR_ARG(n_labels >= 1 && n >= n_labels, "find_labels needs >= %d interior waypoints (have %lld)", n_labels, (long long)n);
    std::vector<double> d((size_t)n);
    BGR_CHECK(bgr_map_clearance(map, path_3xP + 3, n, radius, d.data()));
    std::vector<float> df((size_t)n);
    for (int64_t i = 0; i < n; ++i) df[(size_t)i] = (float)d[(size_t)i];   // MS:945 vector<float> dist_array
    std::vector<int32_t> order((size_t)n);
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int32_t a, int32_t b) { return df[(size_t)a] < df[(size_t)b]; });
    int ties = 0;
    for (int q = 0; q < n_labels; ++q) {
        idx[q] = order[(size_t)q];
        for (int c = 0; c < 3; ++c) labels_3xL[3 * q + c] = (float)path_3xP[3 * (1 + order[(size_t)q]) + c];   // MS:989-991
        if ((size_t)q + 1 < order.size() && df[(size_t)order[(size_t)q]] == df[(size_t)order[(size_t)q + 1]]) ++ties;
    }
    if (clearance) std::memcpy(clearance, df.data(), (size_t)n * sizeof(float));
    if (n_ties) *n_ties = ties;   // float-equal clearances among the selected: std::sort's order is unspecified there
    return BGR_OK;
}

}  // extern "C"
