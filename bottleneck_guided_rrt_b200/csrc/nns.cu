// bgr_nns_* — libnabo-shaped nearest-neighbour object over an append-only SoA node store
// in HBM.  Host side of the C ABI declared in include/bgrrt.h (reference surface:
// MS:138-154 Nearest, MS:1277-1315 libnabo create/knn/delete).
#include <atomic>
#include <cmath>
#include <cstring>
#include <vector>
#include "tile_knn.cuh"
#include "knn_brute.cuh"
#include "nearest.cuh"
#include "internal.cuh"

using namespace bgr;

struct bgr_nns {
    int device = 0;
    cudaStream_t stream = nullptr;   // borrowed (set_stream) or the legacy default stream
    unsigned create_flags = 0;
    int method = BGR_METHOD_AUTO;
    int64_t index_base = 0;
    // append-only SoA store, NaN padded to a multiple of kTile
    DevBuf<float> x, y, z;
    DevBuf<float4> f4;           // filter image {-2x,-2y,-2z,|p|^2''} (knn_brute.cuh)
    int64_t n = 0;
    // scratch
    DevBuf<float> stage_f;       // interleaved staging for host API / appends
    DevBuf<float> q_dev, radii_dev, d2_dev, part_d2;
    DevBuf<int32_t> idx_dev, part_idx;
    DevBuf<unsigned long long> counter;
    GridIndex grid;
    CopyPipe pipe;               // side streams of the host-pointer kNN (chunked D2H overlap)
    double tail_work = 0.0;      // pair evaluations spent on the unsorted tail since the last index build
    uint64_t n_brute = 0, n_scan_nearest = 0;   // knn_brute_kernel / nearest_kernel launches (bgr_nns_debug_stats)
};

namespace {

__global__ void deinterleave_kernel(const float* __restrict__ src, int64_t m, float* __restrict__ x,
                                    float* __restrict__ y, float* __restrict__ z, float4* __restrict__ f4) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < m) {
        const float px = src[3 * i], py = src[3 * i + 1], pz = src[3 * i + 2];
        x[i] = px;
        y[i] = py;
        z[i] = pz;
        f4[i] = filter_image(px, py, pz);
    }
}

int grow_store(bgr_nns* h, int64_t need) {
    need = round_up(need > 0 ? need : 1, kTile);
    if (need <= h->x.cap) return BGR_OK;
    const int64_t old_cap = h->x.cap;
    BGR_CHECK(h->x.reserve(need, h->stream, true, h->n));
    BGR_CHECK(h->y.reserve(need, h->stream, true, h->n));
    BGR_CHECK(h->z.reserve(need, h->stream, true, h->n));
    BGR_CHECK(h->f4.reserve(need, h->stream, true, h->n));
    (void)old_cap;
    // 0xFF bytes = NaN: padded nodes can never be admitted by any comparison
    const int64_t cap = h->x.cap;
    BGR_CUDA(cudaMemsetAsync(h->x.p + h->n, 0xFF, (size_t)(cap - h->n) * sizeof(float), h->stream));
    BGR_CUDA(cudaMemsetAsync(h->y.p + h->n, 0xFF, (size_t)(cap - h->n) * sizeof(float), h->stream));
    BGR_CUDA(cudaMemsetAsync(h->z.p + h->n, 0xFF, (size_t)(cap - h->n) * sizeof(float), h->stream));
    BGR_CUDA(cudaMemsetAsync(h->f4.p + h->n, 0xFF, (size_t)(cap - h->n) * sizeof(float4), h->stream));
    return BGR_OK;
}

int append_dev(bgr_nns* h, const float* d_pts, int64_t m) {
    if (m == 0) return BGR_OK;
    BGR_CHECK(check_device(h->device));
    BGR_ARG(h->n + m < (int64_t)1 << 31, "node count exceeds int32 index range");
    BGR_CHECK(grow_store(h, h->n + m));
    deinterleave_kernel<<<(unsigned)div_up(m, 256), 256, 0, h->stream>>>(d_pts, m, h->x.p + h->n,
                                                                          h->y.p + h->n, h->z.p + h->n, h->f4.p + h->n);
    BGR_LAUNCH_CHECK();
    h->n += m;
    return BGR_OK;
}

constexpr int kMaxDevices = 64;   // power of two
constexpr size_t kBruteSmem = (size_t)kStages * kTile * sizeof(float4) + 2 * kStages * sizeof(uint64_t);
constexpr size_t kNearestSmem = (size_t)kStages * 3 * kTile * sizeof(float) + 2 * kStages * sizeof(uint64_t);

template <int R>
int launch_brute(bgr_nns* h, BruteArgs a, int chunks) {
    // the opt-in above 48 KB is per device (and run_many's worker threads race here): one flag per device
    static std::atomic<bool> attr_set[kMaxDevices];
    if (!attr_set[h->device & (kMaxDevices - 1)].load(std::memory_order_acquire)) {
        BGR_CUDA(cudaFuncSetAttribute(knn_brute_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kBruteSmem));
        attr_set[h->device & (kMaxDevices - 1)].store(true, std::memory_order_release);
    }
    dim3 grid((unsigned)div_up(a.nq, (int64_t)kConsumerWarps * R), (unsigned)chunks);
    knn_brute_kernel<R><<<grid, kBruteThreads, kBruteSmem, h->stream>>>(a);
    BGR_LAUNCH_CHECK();
    return BGR_OK;
}

template <int R>
int launch_nearest(bgr_nns* h, NearestArgs a, int chunks) {
    static std::atomic<bool> attr_set[kMaxDevices];
    if (!attr_set[h->device & (kMaxDevices - 1)].load(std::memory_order_acquire)) {
        BGR_CUDA(cudaFuncSetAttribute(nearest_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kNearestSmem));
        attr_set[h->device & (kMaxDevices - 1)].store(true, std::memory_order_release);
    }
    dim3 grid((unsigned)div_up(a.nq, (int64_t)kConsumerWarps * R), (unsigned)chunks);
    nearest_kernel<R><<<grid, kBruteThreads, kNearestSmem, h->stream>>>(a);
    BGR_LAUNCH_CHECK();
    return BGR_OK;
}

// queries per warp and node chunks so that the grid fills the machine for >= ~2 waves
void plan_scan(int64_t nq, int64_t n, int* R, int* chunks, int64_t* chunk_nodes) {
    const int sms = num_sms();
    // (measured and dropped: preferring R = 4 with many node chunks for the planner's ~800-query Nearest — every
    // chunk restarts its running bound, the f32 filter passes more nodes to the fp64 path: cfg 3 0.40 -> 0.48 s)
    int r = 8;
    while (r > 1 && div_up(nq, (int64_t)kConsumerWarps * r) < 2 * sms) r >>= 1;
    if (const char* e = getenv("BGR_SCAN_R")) { const int f = atoi(e); if (f == 1 || f == 2 || f == 4 || f == 8) r = f; }   // measurement aid
    const int64_t qblocks = div_up(nq, (int64_t)kConsumerWarps * r);
    const int64_t tiles = div_up(n > 0 ? n : 1, kTile);
    const int64_t resident = 3LL * sms;
    int64_t c = 1;
    if (qblocks < 2 * resident) c = div_up(2 * resident, qblocks);
    static const int min_tiles = getenv("BGR_SCAN_MIN_TILES") ? std::max(1, atoi(getenv("BGR_SCAN_MIN_TILES"))) : 4;   // measurement aid
    c = std::min<int64_t>(c, std::max<int64_t>(1, tiles / min_tiles));
    c = std::min<int64_t>(c, 64);
    int64_t per = div_up(tiles, c) * kTile;
    c = div_up(tiles * kTile, per);
    *R = r;
    *chunks = (int)c;
    *chunk_nodes = per;
}

// n_dev != nullptr: the kernels read the store's size from the device and the launch is laid out for n_plan nodes
// (CUDA-graph replays of a planner step: the same launch parameters serve a growing tree)
int knn_brute_dev(bgr_nns* h, const float* d_q, int64_t nq, int k, unsigned flags, float max_radius,
                  const float* d_radii, int32_t* d_idx, float* d_d2, int64_t n_plan = -1, const long long* n_dev = nullptr) {
    int R, chunks;
    int64_t per;
    if (n_plan < 0) n_plan = h->n;
    plan_scan(nq, n_plan, &R, &chunks, &per);
    BruteArgs a;
    a.x = h->x.p; a.y = h->y.p; a.z = h->z.p; a.f4 = h->f4.p; a.n = n_plan; a.n_dev = n_dev;
    a.q = d_q; a.nq = nq; a.k = k; a.flags = flags;
    a.max_radius = max_radius; a.max_radii = d_radii;
    a.chunk_nodes = per;
    a.index_base = h->index_base;
    if (chunks == 1) {
        a.idx = d_idx; a.d2 = d_d2;
    } else {
        BGR_CHECK(h->part_idx.reserve((int64_t)chunks * nq * k, h->stream, false));
        BGR_CHECK(h->part_d2.reserve((int64_t)chunks * nq * k, h->stream, false));
        a.idx = h->part_idx.p; a.d2 = h->part_d2.p;
    }
    switch (R) {
        case 8: BGR_CHECK(launch_brute<8>(h, a, chunks)); break;
        case 4: BGR_CHECK(launch_brute<4>(h, a, chunks)); break;
        case 2: BGR_CHECK(launch_brute<2>(h, a, chunks)); break;
        default: BGR_CHECK(launch_brute<1>(h, a, chunks)); break;
    }
    h->n_brute++;
    if (chunks > 1) {
        knn_merge_kernel<<<(unsigned)div_up(nq, 8), 256, 0, h->stream>>>(h->part_idx.p, h->part_d2.p, chunks,
                                                                         nq * k, nq, k, d_idx, d_d2);
        BGR_LAUNCH_CHECK();
    }
    return BGR_OK;
}

int check_knn_args(const bgr_nns* h, int64_t nq, int k, float eps, unsigned flags, float max_radius) {
    BGR_ARG(h != nullptr, "null nns handle");
    BGR_ARG(nq >= 0, "negative query count");
    BGR_ARG((flags & ~(BGR_ALLOW_SELF_MATCH | BGR_SORT_RESULTS)) == 0,
            "OR-ed value of option flags (%u) is larger than maximal valid value (3)", flags);
    BGR_ARG(k >= 1, "k must be >= 1");
    if (flags & BGR_ALLOW_SELF_MATCH)
        BGR_ARG(k <= h->n, "Requesting more points (%d) than available in cloud (%lld)", k, (long long)h->n);
    else
        BGR_ARG(k <= h->n - 1,
                "Requesting more points (%d) than available in cloud minus 1 (%lld) (as self match is forbidden)",
                k, (long long)(h->n - 1));
    if (eps != 0.f) {
        set_error("approximate search (epsilon=%g) is not offered: exact search only", (double)eps);
        return BGR_ERR_UNSUPPORTED;
    }
    BGR_ARG(!(max_radius < 0.f) && max_radius == max_radius, "maxRadius must be >= 0");
    if (k > 32) {
        set_error("k=%d > 32 is not supported by the warp top-k in this build", k);
        return BGR_ERR_UNSUPPORTED;
    }
    return BGR_OK;
}

}  // namespace

namespace bgr {
int nns_reserve(bgr_nns* h, int64_t n) { return grow_store(h, n); }
void nns_soa(bgr_nns* h, float** x, float** y, float** z, float4** f4) {
    *x = h->x.p; *y = h->y.p; *z = h->z.p;
    if (f4) *f4 = h->f4.p;
}
void nns_set_size(bgr_nns* h, int64_t n) { h->n = n; }
int nns_reset(bgr_nns* h) {
    // back to an empty store: the scans read whole 1,024-node tiles and rely on NaN padding past the last node
    const int64_t used = std::min<int64_t>(h->x.cap, round_up(h->n > 0 ? h->n : 1, kTile));
    if (h->x.p && used > 0) {
        BGR_CUDA(cudaMemsetAsync(h->x.p, 0xFF, (size_t)used * sizeof(float), h->stream));
        BGR_CUDA(cudaMemsetAsync(h->y.p, 0xFF, (size_t)used * sizeof(float), h->stream));
        BGR_CUDA(cudaMemsetAsync(h->z.p, 0xFF, (size_t)used * sizeof(float), h->stream));
        BGR_CUDA(cudaMemsetAsync(h->f4.p, 0xFF, (size_t)used * sizeof(float4), h->stream));
    }
    h->n = 0;
    h->grid.n_indexed = 0;
    h->grid.crowd = 0.0;
    h->tail_work = 0.0;
    return BGR_OK;
}
cudaStream_t nns_stream(bgr_nns* h) { return h->stream; }
}  // namespace bgr

extern "C" {

int bgr_nns_create_dev(const float* d_cloud, int64_t n, int dim, unsigned create_flags, void* stream,
                       bgr_nns** out) {
    BGR_ARG(out != nullptr, "out is null");
    *out = nullptr;
    BGR_ARG(n >= 0, "negative point count");
    if (dim != 3) {
        set_error("dim=%d: only 3-D clouds are supported (the planner's state space)", dim);
        return BGR_ERR_UNSUPPORTED;
    }
    BGR_ARG((create_flags & ~BGR_TOUCH_STATISTICS) == 0, "unknown creation flags %u", create_flags);
    int dev = 0;
    BGR_CUDA(cudaGetDevice(&dev));
    bgr_nns* h = new bgr_nns;
    h->device = dev;
    h->stream = (cudaStream_t)stream;
    h->create_flags = create_flags;
    int rc = grow_store(h, n);
    if (rc == BGR_OK && n > 0) rc = append_dev(h, d_cloud, n);
    if (rc != BGR_OK) {
        bgr_nns_destroy(h);
        return rc;
    }
    *out = h;
    return BGR_OK;
}

int bgr_nns_create(const float* cloud, int64_t n, int dim, unsigned create_flags, bgr_nns** out) {
    BGR_ARG(out != nullptr, "out is null");
    *out = nullptr;
    BGR_ARG(n == 0 || cloud != nullptr, "cloud is null");
    bgr_nns* h = nullptr;
    BGR_CHECK(bgr_nns_create_dev(nullptr, 0, dim, create_flags, nullptr, &h));
    int rc = bgr_nns_append(h, cloud, n);
    if (rc != BGR_OK) {
        bgr_nns_destroy(h);
        return rc;
    }
    *out = h;
    return BGR_OK;
}

int bgr_nns_append_dev(bgr_nns* h, const float* d_pts, int64_t m) {
    BGR_ARG(h != nullptr && m >= 0, "bad append arguments");
    return append_dev(h, d_pts, m);
}

int bgr_nns_append(bgr_nns* h, const float* pts, int64_t m) {
    BGR_ARG(h != nullptr && m >= 0 && (m == 0 || pts != nullptr), "bad append arguments");
    if (m == 0) return BGR_OK;
    BGR_CHECK(h->stage_f.reserve(3 * m, h->stream, false));
    BGR_CUDA(cudaMemcpyAsync(h->stage_f.p, pts, (size_t)3 * m * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    BGR_CHECK(append_dev(h, h->stage_f.p, m));
    BGR_CUDA(cudaStreamSynchronize(h->stream));
    return BGR_OK;
}

int64_t bgr_nns_size(const bgr_nns* h) { return h ? h->n : -1; }

int bgr_nns_set_stream(bgr_nns* h, void* stream) {
    BGR_ARG(h != nullptr, "null handle");
    BGR_CUDA(cudaStreamSynchronize(h->stream));
    h->stream = (cudaStream_t)stream;
    return BGR_OK;
}

int bgr_nns_set_method(bgr_nns* h, int method) {
    BGR_ARG(h != nullptr && method >= 0 && method <= 3, "bad method");
    h->method = method;
    return BGR_OK;
}

int bgr_nns_set_index_base(bgr_nns* h, int64_t base) {
    BGR_ARG(h != nullptr && base >= 0 && base + h->n < ((int64_t)1 << 31), "bad index base");
    h->index_base = base;
    return BGR_OK;
}

int bgr_nns_build_index(bgr_nns* h) {
    BGR_ARG(h != nullptr, "null handle");
    if (h->n == 0) return BGR_OK;
    h->tail_work = 0.0;
    return grid_build(h->grid, h->x.p, h->y.p, h->z.p, h->n, h->stream);
}

// Re-bucket policy of the grid index.  Every query scans the unsorted tail (nodes appended since
// the last build) linearly, so a batch costs nq * tail pair evaluations on top of the grid
// search; the index is rebuilt once that accumulated work would exceed the cost of a rebuild
// (~64 pair evaluations per indexed node, measured), or the tail passes a quarter of the index.
static int grid_refresh(bgr_nns* h, int64_t nq) {
    const int64_t tail = h->n - h->grid.n_indexed;
    const double work = (double)nq * (double)tail;
    if (h->grid.n_indexed == 0 || tail > std::max<int64_t>(4096, h->grid.n_indexed / 4) ||
        (tail > 0 && h->tail_work + work > 64.0 * (double)h->n)) {
        BGR_CHECK(grid_build(h->grid, h->x.p, h->y.p, h->z.p, h->n, h->stream));
        h->tail_work = 0.0;
    } else {
        h->tail_work += work;
    }
    return BGR_OK;
}

// Search method of this call; keeps the grid index current when it is (or may be) used.
// AUTO: small trees scan, large trees use the grid.  `for_nearest`: the 1-NN of queries spread
// over the whole map (the planner's samples) has to scan every crowded cell that faces the query,
// so when the nodes are piled up in a few cells (crowd = occupancy of the cell an average node
// sits in: ~9 for a uniform cloud, thousands for a tree stuck behind a narrow passage) the plain
// scan wins (measured at 150k nodes: 1.1 ms against 2.3 ms per 4096 queries); kNN queries lie
// next to the tree and stay on the grid (0.4 ms against 2 ms).
static int pick_method(bgr_nns* h, int64_t nq, bool for_nearest, int* method) {
    int m = h->method;
    if (m == BGR_METHOD_AUTO && h->n < 16384) m = BGR_METHOD_BRUTE;
    if (m != BGR_METHOD_BRUTE) BGR_CHECK(grid_refresh(h, nq));
    if (m == BGR_METHOD_AUTO) {
        m = (!for_nearest || 27.0 * h->grid.crowd < (double)h->n / 16.0) ? BGR_METHOD_GRID : BGR_METHOD_BRUTE;
        // a Nearest batch whose plain scan is short anyway (< 2e8 pairs, ~0.25 ms) stays on the scan: the ring
        // search is at its worst for the planner's map-wide samples far outside a tree that fills a corner of the
        // map (measured on a 55 k-node query tree, ~820 distinct samples: scan 0.05 ms, grid 0.9 ms per step)
        if (for_nearest && (double)nq * (double)h->n < 2.0e8) m = BGR_METHOD_BRUTE;
    }
    *method = m;
    return BGR_OK;
}

static int knn_dev_impl(bgr_nns* h, const float* d_q, int64_t nq, int k, float eps, unsigned flags,
                        float max_radius, const float* d_radii, int32_t* d_idx, float* d_d2,
                        uint64_t* touched, const uint8_t* d_skip) {
    BGR_CHECK(check_knn_args(h, nq, k, eps, flags, max_radius));
    BGR_CHECK(check_device(h->device));
    if (touched) *touched = 0;
    if (nq == 0) return BGR_OK;
    int method;
    BGR_CHECK(pick_method(h, nq, false, &method));
    const int grid_fast = method != BGR_METHOD_GRID_GENERAL;
    if (method == BGR_METHOD_GRID_GENERAL) method = BGR_METHOD_GRID;
    if (method == BGR_METHOD_GRID) {
        const bool stats = (h->create_flags & BGR_TOUCH_STATISTICS) && touched;
        if (stats) {
            BGR_CHECK(h->counter.reserve(1, h->stream, false));
            BGR_CUDA(cudaMemsetAsync(h->counter.p, 0, sizeof(unsigned long long), h->stream));
        }
        BGR_CHECK(grid_knn(h->grid, h->x.p, h->y.p, h->z.p, h->n, d_q, nq, k, flags, max_radius, d_radii,
                           d_idx, d_d2, h->index_base, stats ? h->counter.p : nullptr, grid_fast, h->device, h->stream, d_skip));
        if (stats) {
            unsigned long long c = 0;
            BGR_CUDA(cudaMemcpyAsync(&c, h->counter.p, sizeof(c), cudaMemcpyDeviceToHost, h->stream));
            BGR_CUDA(cudaStreamSynchronize(h->stream));
            *touched = c;
        }
        return BGR_OK;
    }
    BGR_CHECK(knn_brute_dev(h, d_q, nq, k, flags, max_radius, d_radii, d_idx, d_d2));
    if (touched) *touched = (uint64_t)nq * (uint64_t)h->n;
    return BGR_OK;
}

int bgr_nns_knn_dev(bgr_nns* h, const float* d_q, int64_t nq, int k, float eps, unsigned flags,
                    float max_radius, const float* d_radii, int32_t* d_idx, float* d_d2,
                    uint64_t* touched) {
    return knn_dev_impl(h, d_q, nq, k, eps, flags, max_radius, d_radii, d_idx, d_d2, touched, nullptr);
}

int bgr_nns_knn_skip_dev(bgr_nns* h, const float* d_q, int64_t nq, int k, unsigned flags, const uint8_t* d_skip,
                         int32_t* d_idx, float* d_d2) {
    return knn_dev_impl(h, d_q, nq, k, 0.f, flags, INFINITY, nullptr, d_idx, d_d2, nullptr, d_skip);
}

int bgr_nns_knn(bgr_nns* h, const float* q, int64_t nq, int k, float eps, unsigned flags,
                float max_radius, const float* max_radii, int32_t* idx, float* d2, uint64_t* touched) {
    BGR_CHECK(check_knn_args(h, nq, k, eps, flags, max_radius));
    BGR_ARG(nq == 0 || (q && idx), "null query/result buffer");
    if (touched) *touched = 0;
    if (nq == 0) return BGR_OK;
    BGR_CHECK(h->q_dev.reserve(3 * nq, h->stream, false));
    BGR_CHECK(h->idx_dev.reserve(nq * k, h->stream, false));
    BGR_CHECK(h->d2_dev.reserve(nq * k, h->stream, false));
    BGR_CUDA(cudaMemcpyAsync(h->q_dev.p, q, (size_t)3 * nq * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    const float* d_radii = nullptr;
    if (max_radii) {
        BGR_CHECK(h->radii_dev.reserve(nq, h->stream, false));
        BGR_CUDA(cudaMemcpyAsync(h->radii_dev.p, max_radii, (size_t)nq * sizeof(float), cudaMemcpyHostToDevice, h->stream));
        d_radii = h->radii_dev.p;
    }
    // grid search of a large batch: cut it into chunks so the D2H copy of chunk c overlaps the
    // kernel of chunk c+1 (results are per query, so chunking cannot change them)
    int method = h->method;
    if (method == BGR_METHOD_AUTO) method = (h->n >= 16384) ? BGR_METHOD_GRID : BGR_METHOD_BRUTE;
    const bool want_stats = (h->create_flags & BGR_TOUCH_STATISTICS) && touched;
    int chunks = 1;
    if ((method == BGR_METHOD_GRID || method == BGR_METHOD_GRID_GENERAL) && !want_stats && !pipe_disabled())
        chunks = (int)std::min<int64_t>(d2 ? 4 : 2, nq / 8192);   // measured (profiles/r2_e2e_chunks.txt): 4 with both result arrays, 2 with indices only
    if (const char* e = getenv("BGR_KNN_CHUNKS")) chunks = std::max(1, std::min(atoi(e), (int)CopyPipe::kMaxChunks));   // measurement aid
    // Page-locked result buffers are written by the search kernel itself, straight over the host link (UVA: a
    // pinned host allocation has a device pointer), instead of being staged in HBM and copied back in chunks:
    // measured on configs[1] 0.256 -> 0.234 ms (profiles/r2_e2e_chunks.txt).  BGR_KNN_ZEROCOPY=0 turns it off;
    // pageable buffers take the staged path below.
    static const bool zero_copy = !(getenv("BGR_KNN_ZEROCOPY") && getenv("BGR_KNN_ZEROCOPY")[0] == '0');
    if (zero_copy && !want_stats) {
        cudaPointerAttributes ai{}, ad{};
        const bool ok_i = cudaPointerGetAttributes(&ai, idx) == cudaSuccess && ai.type == cudaMemoryTypeHost && ai.devicePointer;
        const bool ok_d = !d2 || (cudaPointerGetAttributes(&ad, d2) == cudaSuccess && ad.type == cudaMemoryTypeHost && ad.devicePointer);
        (void)cudaGetLastError();
        if (ok_i && ok_d) {
            BGR_CHECK(bgr_nns_knn_dev(h, h->q_dev.p, nq, k, eps, flags, max_radius, d_radii, (int32_t*)ai.devicePointer,
                                      d2 ? (float*)ad.devicePointer : h->d2_dev.p, nullptr));
            BGR_CUDA(cudaStreamSynchronize(h->stream));
            return BGR_OK;
        }
    }
    if (chunks <= 1) {
        BGR_CHECK(bgr_nns_knn_dev(h, h->q_dev.p, nq, k, eps, flags, max_radius, d_radii, h->idx_dev.p,
                                  h->d2_dev.p, touched));
        BGR_CUDA(cudaMemcpyAsync(idx, h->idx_dev.p, (size_t)nq * k * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
        if (d2) BGR_CUDA(cudaMemcpyAsync(d2, h->d2_dev.p, (size_t)nq * k * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
        BGR_CUDA(cudaStreamSynchronize(h->stream));
        return BGR_OK;
    }
    BGR_CHECK(h->pipe.init());
    const int64_t per = round_up(div_up(nq, chunks), 8);
    for (int c = 0; c < chunks; ++c) {
        const int64_t s0 = (int64_t)c * per, m = std::min<int64_t>(per, nq - s0);
        if (m <= 0) break;
        BGR_CHECK(bgr_nns_knn_dev(h, h->q_dev.p + 3 * s0, m, k, eps, flags, max_radius, d_radii ? d_radii + s0 : nullptr,
                                  h->idx_dev.p + s0 * k, h->d2_dev.p + s0 * k, nullptr));
        BGR_CUDA(cudaEventRecord(h->pipe.ev_out[c], h->stream));
        BGR_CUDA(cudaStreamWaitEvent(h->pipe.s_out, h->pipe.ev_out[c], 0));
        BGR_CUDA(cudaMemcpyAsync(idx + s0 * k, h->idx_dev.p + s0 * k, (size_t)m * k * sizeof(int32_t), cudaMemcpyDeviceToHost, h->pipe.s_out));
        if (d2) BGR_CUDA(cudaMemcpyAsync(d2 + s0 * k, h->d2_dev.p + s0 * k, (size_t)m * k * sizeof(float), cudaMemcpyDeviceToHost, h->pipe.s_out));
    }
    BGR_CUDA(cudaStreamSynchronize(h->pipe.s_out));
    BGR_CUDA(cudaStreamSynchronize(h->stream));
    return BGR_OK;
}

// nq = capacity of the query / result arrays; with d_count only the first count_base + *d_count
// queries are valid (the rest of the rows is left undefined) and nq_hint (host guess of that
// number, e.g. the previous batch's) sizes the scan.
// n_dev != nullptr: plain scan laid out for n_plan nodes, the kernels read the store's size from the device
// (CUDA-graph replays of a planner step)
int nearest_dev_impl(bgr_nns* h, const float* d_q, int64_t nq, int32_t* d_idx, float* d_dist,
                            const long long* d_count, int64_t count_base, int64_t nq_hint, int64_t n_plan = -1,
                            const long long* n_dev = nullptr) {
    BGR_ARG(h != nullptr && nq >= 0, "bad nearest arguments");
    BGR_ARG(h->n >= 1, "Nearest on an empty tree");
    if (nq == 0) return BGR_OK;
    BGR_CHECK(check_device(h->device));
    if (n_plan < 0) n_plan = h->n;
    if (!n_dev) {
        // large trees: the grid index answers Nearest exactly (grid_nearest_kernel); the plain scan
        // below stays for small trees and for METHOD_BRUTE
        int method;
        BGR_CHECK(pick_method(h, d_count ? std::max<int64_t>(1, std::min(nq, nq_hint)) : nq, true, &method));
        if (method == BGR_METHOD_GRID || method == BGR_METHOD_GRID_GENERAL) {
            return grid_nearest(h->grid, h->x.p, h->y.p, h->z.p, h->n, d_q, nq, d_idx, d_dist, h->index_base, h->stream,
                                d_count, count_base);
        }
    }
    int R, chunks;
    int64_t per;
    plan_scan(d_count ? std::max<int64_t>(1, std::min(nq, nq_hint)) : nq, n_plan, &R, &chunks, &per);
    if (R > 4) R = 4;   // double-precision state: keep registers in check
    NearestArgs a;
    a.x = h->x.p; a.y = h->y.p; a.z = h->z.p; a.n = n_plan; a.n_dev = n_dev;
    a.q = d_q; a.nq = nq; a.chunk_nodes = per; a.index_base = h->index_base;
    a.nq_dev = d_count; a.nq_base = count_base;
    BGR_CHECK(h->part_idx.reserve((int64_t)chunks * nq, h->stream, false));
    BGR_CHECK(h->part_d2.reserve((int64_t)chunks * nq, h->stream, false));
    a.idx = h->part_idx.p; a.dist = h->part_d2.p;
    switch (R) {
        case 4: BGR_CHECK(launch_nearest<4>(h, a, chunks)); break;
        case 2: BGR_CHECK(launch_nearest<2>(h, a, chunks)); break;
        default: BGR_CHECK(launch_nearest<1>(h, a, chunks)); break;
    }
    nearest_reduce_kernel<<<(unsigned)div_up(nq, 256), 256, 0, h->stream>>>(h->part_idx.p, h->part_d2.p, chunks,
                                                                            nq, d_idx, d_dist);
    BGR_LAUNCH_CHECK();
    h->n_scan_nearest++;
    return BGR_OK;
}

int bgr_nns_nearest_dev(bgr_nns* h, const float* d_q, int64_t nq, int32_t* d_idx, float* d_dist) {
    return nearest_dev_impl(h, d_q, nq, d_idx, d_dist, nullptr, 0, nq);
}


int bgr_nns_nearest(bgr_nns* h, const float* q, int64_t nq, int32_t* idx, float* dist) {
    BGR_ARG(h != nullptr && nq >= 0 && (nq == 0 || (q && idx)), "bad nearest arguments");
    BGR_ARG(h->n >= 1, "Nearest on an empty tree");
    if (nq == 0) return BGR_OK;
    BGR_CHECK(h->q_dev.reserve(3 * nq, h->stream, false));
    BGR_CHECK(h->idx_dev.reserve(nq, h->stream, false));
    BGR_CHECK(h->d2_dev.reserve(nq, h->stream, false));
    BGR_CUDA(cudaMemcpyAsync(h->q_dev.p, q, (size_t)3 * nq * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    BGR_CHECK(bgr_nns_nearest_dev(h, h->q_dev.p, nq, h->idx_dev.p, h->d2_dev.p));
    BGR_CUDA(cudaMemcpyAsync(idx, h->idx_dev.p, (size_t)nq * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
    if (dist)
        BGR_CUDA(cudaMemcpyAsync(dist, h->d2_dev.p, (size_t)nq * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    BGR_CUDA(cudaStreamSynchronize(h->stream));
    return BGR_OK;
}

int bgr_nns_debug_stats(bgr_nns* h, uint64_t out[8]) {
    BGR_ARG(h != nullptr && out != nullptr, "bad arguments");
    out[0] = h->grid.n_builds; out[1] = h->grid.n_lean; out[2] = h->grid.n_piled; out[3] = 0;
    out[4] = h->grid.n_grid_nearest; out[5] = h->n_scan_nearest; out[6] = h->n_brute; out[7] = (uint64_t)h->grid.n_indexed;
    if (h->grid.n_piled && h->grid.heavy.p) {
        int32_t cnt = 0;
        BGR_CUDA(cudaMemcpyAsync(&cnt, h->grid.heavy.p, sizeof(cnt), cudaMemcpyDeviceToHost, h->stream));
        BGR_CUDA(cudaStreamSynchronize(h->stream));
        out[3] = (uint64_t)std::max(0, cnt);
    }
    return BGR_OK;
}

int bgr_nns_set_tile(bgr_nns* h, int tx, int ty, int tz, int warps) {
    BGR_ARG(h != nullptr, "null handle");
    BGR_ARG(tx < 0 || (tx == 0 && ty == 0 && tz == 0) || (tx >= 1 && tx <= kTileMaxT0 && ty >= 1 && ty <= 6 && tz >= 1 && tz <= 6),
            "tile size out of range (1..%d, 1..6, 1..6)", kTileMaxT0);
    BGR_ARG(warps >= 0 && warps <= kTileMaxWarps, "warps per tile CTA out of range (0..%d)", kTileMaxWarps);
    h->grid.tile_shape[0] = tx; h->grid.tile_shape[1] = ty; h->grid.tile_shape[2] = tz;
    h->grid.tile_warps = warps;
    return BGR_OK;
}

int bgr_nns_tile_stats(bgr_nns* h, uint64_t out[4]) {
    BGR_ARG(h != nullptr && out != nullptr, "bad arguments");
    out[0] = h->grid.n_tile; out[1] = 0; out[2] = h->grid.tile_last_shape; out[3] = h->grid.tile_last_count;
    if (h->grid.n_tile && h->grid.tile_fb.p) {
        int32_t cnt = 0;
        BGR_CUDA(cudaMemcpyAsync(&cnt, h->grid.tile_fb.p + (h->grid.tile_parity ^ 1), sizeof(cnt), cudaMemcpyDeviceToHost, h->stream));
        BGR_CUDA(cudaStreamSynchronize(h->stream));
        out[1] = (uint64_t)std::max(0, cnt);
    }
    return BGR_OK;
}

int bgr_nns_destroy(bgr_nns* h) {
    if (!h) return BGR_OK;
    cudaStreamSynchronize(h->stream);
    h->x.release(); h->y.release(); h->z.release(); h->f4.release();
    h->stage_f.release(); h->q_dev.release(); h->radii_dev.release(); h->d2_dev.release();
    h->part_d2.release(); h->idx_dev.release(); h->part_idx.release(); h->counter.release();
    grid_release(h->grid);
    h->pipe.release();
    delete h;
    return BGR_OK;
}

int bgr_knn_merge_strided_dev(const int32_t* d_idx_in, const float* d_d2_in, int g, int64_t slab_stride,
                              int64_t nq, int k, int32_t* d_idx, float* d_d2, void* stream) {
    BGR_ARG(g >= 1 && nq >= 0 && k >= 1 && k <= 32, "bad merge arguments (k<=32)");
    BGR_ARG(slab_stride >= nq * k, "slab stride %lld < k*Q", (long long)slab_stride);
    if (nq == 0) return BGR_OK;
    knn_merge_kernel<<<(unsigned)div_up(nq, 8), 256, 0, (cudaStream_t)stream>>>(d_idx_in, d_d2_in, g, slab_stride, nq, k, d_idx, d_d2);
    BGR_LAUNCH_CHECK();
    return BGR_OK;
}

int bgr_knn_merge_dev(const int32_t* d_idx_in, const float* d_d2_in, int g, int64_t nq, int k,
                      int32_t* d_idx, float* d_d2, void* stream) {
    return bgr_knn_merge_strided_dev(d_idx_in, d_d2_in, g, nq * (int64_t)k, nq, k, d_idx, d_d2, stream);
}

}  // extern "C"

namespace bgr {
int nns_nearest_counted(bgr_nns* h, const float* d_q, int64_t nq_cap, int64_t nq_hint, const long long* d_count,
                        int64_t count_base, int32_t* d_idx, float* d_dist) {
    return nearest_dev_impl(h, d_q, nq_cap, d_idx, d_dist, d_count, count_base, nq_hint);
}
int nns_nearest_scan_cap(bgr_nns* h, const float* d_q, int64_t nq_cap, const long long* d_count, int64_t count_base,
                         int32_t* d_idx, float* d_dist, int64_t n_cap, const long long* n_dev) {
    return nearest_dev_impl(h, d_q, nq_cap, d_idx, d_dist, d_count, count_base, nq_cap, n_cap, n_dev);
}
int nns_knn_brute_cap(bgr_nns* h, const float* d_q, int64_t nq, int k, int32_t* d_idx, float* d_d2, int64_t n_cap,
                      const long long* n_dev) {
    BGR_CHECK(check_device(h->device));
    return knn_brute_dev(h, d_q, nq, k, 0u, INFINITY, nullptr, d_idx, d_d2, n_cap, n_dev);
}
void nns_scan_scratch(bgr_nns* h, const void** a, const void** b) { *a = h->part_idx.p; *b = h->part_d2.p; }
int nns_method(const bgr_nns* h) { return h->method; }
}  // namespace bgr

// ---------------------------------------------------------------------------------------------------------
// bgr_nns_knn_scan_static_dev — the brute-force kNN scan for callers that may not allocate (no counterpart in
// the reference).  knn_brute_dev above keeps the per-chunk
// partial lists of a chunked scan in the handle's part_idx / part_d2 buffers, which grow with cudaMalloc.  Here
// the same kernels (knn_brute_kernel, knn_merge_kernel) write / read those partial lists in ONE file-scope
// device array of fixed size, so the call allocates nothing: no cudaMalloc / cudaFree, no allocating library call.
//   * limit  : nq * k <= BGR_KNN_STATIC_MAX_RESULTS (32,768 results, e.g. 2,048 queries x k = 16) and at most
//              kStaticMaxChunks (16) node chunks -> 16 * 32,768 * (4 + 4) bytes = 4 MiB, held by every GPU that
//              loads the library until the process exits;
//   * address: in host code the array's NAME is a host-side stand-in, not a device address; the device address
//              comes from cudaGetSymbolAddress, asked once per device with that device current (below);
//   * sharing: every call (all handles, all streams) uses the same array, and it keeps the last call's contents.
//              Nothing in it has to start at zero: every chunk's CTAs write each of their nq * k slots (lists
//              padded with idx = -1, d2 = +inf) before the merge kernel, ordered behind them on the handle's
//              stream, reads exactly those slots.  Two calls that overlap on the GPU (different streams, or
//              different host threads on different streams) RACE on the array: serialize them.
namespace {
constexpr int64_t kStaticMaxResults = BGR_KNN_STATIC_MAX_RESULTS;
constexpr int kStaticMaxChunks = 16;
constexpr size_t kStaticHalf = (size_t)kStaticMaxChunks * kStaticMaxResults * sizeof(int32_t);   // idx half, then d2 half
static_assert(sizeof(int32_t) == sizeof(float), "the two halves have equal size");
}  // namespace
static __device__ __align__(256) unsigned char g_knn_static_scratch[2 * kStaticHalf];

namespace {
// device address of g_knn_static_scratch on the CURRENT device (== device, checked by the caller)
int knn_static_scratch(int device, unsigned char** out) {
    static std::atomic<unsigned char*> addr[kMaxDevices];
    BGR_ARG(device >= 0 && device < kMaxDevices, "device %d outside the static scratch table (%d devices)", device, kMaxDevices);
    unsigned char* p = addr[device].load(std::memory_order_acquire);
    if (!p) {
        void* sym = nullptr;
        BGR_CUDA(cudaGetSymbolAddress(&sym, g_knn_static_scratch));
        if (!sym) {
            set_error("cudaGetSymbolAddress returned a null address for the static kNN scratch");
            return BGR_ERR_CUDA;
        }
        p = static_cast<unsigned char*>(sym);
        addr[device].store(p, std::memory_order_release);
    }
    *out = p;
    return BGR_OK;
}
}  // namespace

extern "C" int bgr_nns_knn_scan_static_dev(bgr_nns* h, const float* d_q, int64_t nq, int k, unsigned flags,
                                           float max_radius, int32_t* d_idx, float* d_d2, int* chunks_used) {
    BGR_CHECK(check_knn_args(h, nq, k, 0.f, flags, max_radius));
    BGR_ARG(nq == 0 || (d_q && d_idx && d_d2), "null query/result buffer");
    BGR_ARG(nq * (int64_t)k <= kStaticMaxResults,
            "nq * k = %lld results exceed the static scratch limit of %lld (BGR_KNN_STATIC_MAX_RESULTS): split the "
            "batch or use bgr_nns_knn_dev", (long long)(nq * (int64_t)k), (long long)kStaticMaxResults);
    BGR_CHECK(check_device(h->device));
    if (chunks_used) *chunks_used = 0;
    if (nq == 0) return BGR_OK;
    unsigned char* scratch = nullptr;
    BGR_CHECK(knn_static_scratch(h->device, &scratch));
    int R, chunks;
    int64_t per;
    plan_scan(nq, h->n, &R, &chunks, &per);
    if (chunks > kStaticMaxChunks) {
        const int64_t tiles = div_up(h->n, kTile);
        per = div_up(tiles, (int64_t)kStaticMaxChunks) * kTile;
        chunks = (int)div_up(tiles * kTile, per);
    }
    BruteArgs a;
    a.x = h->x.p; a.y = h->y.p; a.z = h->z.p; a.f4 = h->f4.p; a.n = h->n; a.n_dev = nullptr;
    a.q = d_q; a.nq = nq; a.k = k; a.flags = flags;
    a.max_radius = max_radius; a.max_radii = nullptr;
    a.chunk_nodes = per;
    a.index_base = h->index_base;
    int32_t* part_idx = reinterpret_cast<int32_t*>(scratch);
    float* part_d2 = reinterpret_cast<float*>(scratch + kStaticHalf);
    if (chunks == 1) {
        a.idx = d_idx; a.d2 = d_d2;          // a single chunk has no partial lists: straight into the caller's rows
    } else {
        a.idx = part_idx; a.d2 = part_d2;
    }
    switch (R) {
        case 8: BGR_CHECK(launch_brute<8>(h, a, chunks)); break;
        case 4: BGR_CHECK(launch_brute<4>(h, a, chunks)); break;
        case 2: BGR_CHECK(launch_brute<2>(h, a, chunks)); break;
        default: BGR_CHECK(launch_brute<1>(h, a, chunks)); break;
    }
    h->n_brute++;
    if (chunks > 1) {
        knn_merge_kernel<<<(unsigned)div_up(nq, 8), 256, 0, h->stream>>>(part_idx, part_d2, chunks, nq * k, nq, k, d_idx, d_d2);
        BGR_LAUNCH_CHECK();
    }
    if (chunks_used) *chunks_used = chunks;
    return BGR_OK;
}
