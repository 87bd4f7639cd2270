// ORACLE (test infrastructure) and timed CPU baseline ("port").
// Restatement of libnabo 1.0.7's default search object, the one the reference builds with
// Nabo::NNSearchF::createKDTreeLinearHeap (MS:1306) and queries with knn (MS:1312):
//   KDTreeUnbalancedPtInLeavesImplicitBoundsStackOpt<float, IndexHeapBruteForceVector>
// libnabo (pinned by abb_manipulator/CMakeLists.txt:66, libnabo.so.1.0.7) is an
// un-vendored dependency: its published algorithm is restated here —
//   build : sliding-midpoint split on the widest dimension of the implicit cell bounds,
//           points stored in leaf buckets (bucketSize 8), left child at pos+1;
//   query : depth-first, near side first, far side iff rd<=maxR2 && rd*maxErr2 < head,
//           leaf test dist<=maxR2 && dist<head && (self || dist>FLT_EPSILON);
//   heap  : k-entry ascending vector, insert shifts entries with value strictly greater.
// SURVEY App. A.1 recalls a median (nth_element) split instead; both are recollections,
// PARITY IS UNPINNED.  The split rule only changes visit order, i.e. which index wins
// among exactly tied distances, which tests canonicalise and count.
#include "../oracle.h"
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <limits>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

struct orc_nabo {
    struct Node {
        uint32_t dim;    // 0..2 split dim, 3 = leaf
        uint32_t child;  // right child (internal) or bucket size (leaf)
        union {
            float cut;
            uint32_t bucket;
        };
    };
    struct Bucket {
        float p[3];
        int32_t index;
    };
    std::vector<float> cloud;  // 3xN col-major copy
    int64_t n = 0;
    int bucket_size = 8;
    std::vector<Node> nodes;
    std::vector<Bucket> buckets;

    unsigned build(int32_t* first, int32_t* last, const float mn[3], const float mx[3]);
};

unsigned orc_nabo::build(int32_t* first, int32_t* last, const float mn[3], const float mx[3]) {
    const int count = (int)(last - first);
    const unsigned pos = (unsigned)nodes.size();
    if (count <= bucket_size) {
        const uint32_t start = (uint32_t)buckets.size();
        for (int i = 0; i < count; ++i) {
            const int32_t ix = first[i];
            Bucket b;
            b.p[0] = cloud[3 * (size_t)ix];
            b.p[1] = cloud[3 * (size_t)ix + 1];
            b.p[2] = cloud[3 * (size_t)ix + 2];
            b.index = ix;
            buckets.push_back(b);
        }
        Node nd;
        nd.dim = 3;
        nd.child = (uint32_t)count;
        nd.bucket = start;
        nodes.push_back(nd);
        return pos;
    }
    // widest dimension of the (implicit) cell, first max wins
    unsigned cd = 0;
    float best = 0.f;
    for (unsigned d = 0; d < 3; ++d) {
        const float w = mx[d] - mn[d];
        if (w > best) { best = w; cd = d; }
    }
    const float ideal = (mx[cd] + mn[cd]) / 2;
    float lo = cloud[3 * (size_t)first[0] + cd], hi = lo;
    for (int i = 1; i < count; ++i) {
        const float v = cloud[3 * (size_t)first[i] + cd];
        lo = std::min(lo, v);
        hi = std::max(hi, v);
    }
    float cut = ideal;
    if (ideal < lo) cut = lo;
    else if (ideal > hi) cut = hi;

    auto C = [&](int i) { return cloud[3 * (size_t)first[i] + cd]; };
    int l = 0, r = count - 1;
    while (true) {
        while (l < count && C(l) < cut) ++l;
        while (r >= 0 && C(r) >= cut) --r;
        if (l > r) break;
        std::swap(first[l], first[r]);
        ++l; --r;
    }
    const int br1 = l;
    r = count - 1;
    while (true) {
        while (l < count && C(l) <= cut) ++l;
        while (r >= br1 && C(r) > cut) --r;
        if (l > r) break;
        std::swap(first[l], first[r]);
        ++l; --r;
    }
    const int br2 = l;
    int left;
    if (ideal < lo) left = 1;
    else if (ideal > hi) left = count - 1;
    else if (br1 > count / 2) left = br1;
    else if (br2 < count / 2) left = br2;
    else left = count / 2;

    float lmx[3] = {mx[0], mx[1], mx[2]};
    lmx[cd] = cut;
    float rmn[3] = {mn[0], mn[1], mn[2]};
    rmn[cd] = cut;
    Node nd;
    nd.dim = cd;
    nd.child = 0;
    nd.cut = cut;
    nodes.push_back(nd);
    build(first, first + left, mn, lmx);
    const unsigned rc = build(first + left, last, rmn, mx);
    nodes[pos].child = rc;
    return pos;
}

namespace {

struct Heap {
    int k;
    float* v;
    int32_t* ix;
    inline float head() const { return v[k - 1]; }
    inline void replace_head(int32_t index, float value) {
        int i;
        for (i = k - 1; i > 0; --i) {
            if (v[i - 1] > value) { v[i] = v[i - 1]; ix[i] = ix[i - 1]; }
            else break;
        }
        v[i] = value;
        ix[i] = index;
    }
};

template <bool SELF>
unsigned long recurse(const orc_nabo* t, const float* q, unsigned n, float rd, Heap& h,
                      float off[3], float max_err2, float max_r2) {
    const orc_nabo::Node& nd = t->nodes[n];
    if (nd.dim == 3) {
        const orc_nabo::Bucket* b = &t->buckets[nd.bucket];
        for (uint32_t i = 0; i < nd.child; ++i, ++b) {
            float dist = 0;
            for (int d = 0; d < 3; ++d) {
                const float diff = q[d] - b->p[d];
                dist += diff * diff;
            }
            if (dist <= max_r2 && dist < h.head() && (SELF || dist > FLT_EPSILON))
                h.replace_head(b->index, dist);
        }
        return nd.child;
    }
    unsigned long touched = 0;
    const unsigned cd = nd.dim;
    const float old_off = off[cd];
    const float new_off = q[cd] - nd.cut;
    const unsigned near_c = new_off > 0 ? nd.child : n + 1;
    const unsigned far_c = new_off > 0 ? n + 1 : nd.child;
    touched += recurse<SELF>(t, q, near_c, rd, h, off, max_err2, max_r2);
    rd += -old_off * old_off + new_off * new_off;
    if (rd <= max_r2 && rd * max_err2 < h.head()) {
        off[cd] = new_off;
        touched += recurse<SELF>(t, q, far_c, rd, h, off, max_err2, max_r2);
        off[cd] = old_off;
    }
    return touched;
}

}  // namespace

extern "C" {

orc_nabo* orc_nabo_create(const float* cloud, int64_t n, int bucket_size) {
    if (n <= 0 || bucket_size < 2) return nullptr;
    orc_nabo* t = new orc_nabo;
    t->n = n;
    t->bucket_size = bucket_size;
    t->cloud.assign(cloud, cloud + 3 * n);
    float mn[3], mx[3];
    for (int d = 0; d < 3; ++d) mn[d] = mx[d] = cloud[d];
    for (int64_t i = 1; i < n; ++i)
        for (int d = 0; d < 3; ++d) {
            mn[d] = std::min(mn[d], cloud[3 * i + d]);
            mx[d] = std::max(mx[d], cloud[3 * i + d]);
        }
    std::vector<int32_t> pts((size_t)n);
    for (int64_t i = 0; i < n; ++i) pts[(size_t)i] = (int32_t)i;
    t->nodes.reserve((size_t)(n / bucket_size * 3 + 16));
    t->buckets.reserve((size_t)n);
    t->build(pts.data(), pts.data() + n, mn, mx);
    return t;
}

void orc_nabo_destroy(orc_nabo* t) { delete t; }

int64_t orc_nabo_knn(const orc_nabo* t, const float* q, int64_t nq, int k, float eps,
                     unsigned flags, float max_radius, int32_t* idx, float* d2, int threads) {
    if (!t || k <= 0 || (flags & ~3u)) return -1;
    const bool self = (flags & ORC_ALLOW_SELF_MATCH) != 0;
    if (self ? (k > t->n) : (k > t->n - 1)) return -2;  // libnabo checkSizesKnn throws
    const float max_r2 = max_radius * max_radius;
    const float max_err2 = (1 + eps) * (1 + eps);
    const float inf = std::numeric_limits<float>::infinity();
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
    int64_t touched = 0;
#pragma omp parallel for reduction(+ : touched) schedule(guided, 32)
    for (int64_t c = 0; c < nq; ++c) {
        Heap h{k, d2 + (size_t)c * k, idx + (size_t)c * k};
        for (int j = 0; j < k; ++j) { h.v[j] = inf; h.ix[j] = -1; }
        float off[3] = {0.f, 0.f, 0.f};
        if (self) touched += (int64_t)recurse<true>(t, q + 3 * c, 0, 0.f, h, off, max_err2, max_r2);
        else touched += (int64_t)recurse<false>(t, q + 3 * c, 0, 0.f, h, off, max_err2, max_r2);
    }
    return touched;
}

}  // extern "C"
