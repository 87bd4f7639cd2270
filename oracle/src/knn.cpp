// ORACLE (test infrastructure).  Brute-force kNN with libnabo 1.0.7 leaf arithmetic and
// admission rule (call site MS:1306-1312), the reference's literal
// Norm/Nearest (MS:130-154) and steer step (MS:1233-1253).
// libnabo is not present in /root/reference or this image: parity vs the real library is
// UNPINNED; the canonical tie rule (ascending (d2, index)) is the contract BASELINE.json states.
#include "../oracle.h"
#include <atomic>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <limits>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

struct Ent {
    float d;
    int32_t i;
};
inline bool ent_less(float d, int32_t i, const Ent& e) { return d < e.d || (d == e.d && i < e.i); }

}  // namespace

extern "C" {

int orc_knn_brute(const float* cloud, int64_t n, const float* q, int64_t nq, int k,
                  unsigned flags, float max_radius, const float* max_radii, int32_t* idx,
                  float* d2, int threads) {
    if (k <= 0 || n < 0 || nq < 0 || (flags & ~3u)) return -1;
    const bool self = (flags & ORC_ALLOW_SELF_MATCH) != 0;
    const float inf = std::numeric_limits<float>::infinity();
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
#pragma omp parallel
    {
        std::vector<Ent> heap((size_t)k);
#pragma omp for schedule(dynamic, 16)
        for (int64_t c = 0; c < nq; ++c) {
            const float qx = q[3 * c], qy = q[3 * c + 1], qz = q[3 * c + 2];
            const float mr = max_radii ? max_radii[c] : max_radius;
            const float mr2 = mr * mr;  // libnabo: maxRadius2 = maxRadius*maxRadius
            for (int j = 0; j < k; ++j) heap[j] = {inf, -1};
            int filled = 0;
            for (int64_t p = 0; p < n; ++p) {
                // libnabo leaf loop: dist = 0; for d: diff = q[d]-p[d]; dist += diff*diff;
                const float dx = qx - cloud[3 * p], dy = qy - cloud[3 * p + 1],
                            dz = qz - cloud[3 * p + 2];
                float dist = dx * dx;
                dist = dist + dy * dy;
                dist = dist + dz * dz;
                if (!(dist <= mr2)) continue;
                if (!(dist < inf)) continue;  // head starts at +inf, admission is strict <
                if (!self && !(dist > FLT_EPSILON)) continue;
                const int32_t pi = (int32_t)p;
                if (filled == k && !ent_less(dist, pi, heap[k - 1])) continue;
                int pos = filled < k ? filled : k - 1;
                while (pos > 0 && ent_less(dist, pi, heap[pos - 1])) {
                    heap[pos] = heap[pos - 1];
                    --pos;
                }
                heap[pos] = {dist, pi};
                if (filled < k) ++filled;
            }
            for (int j = 0; j < k; ++j) {
                idx[(size_t)c * k + j] = heap[j].i;
                d2[(size_t)c * k + j] = heap[j].d;
            }
        }
    }
    return 0;
}

// MS:130-134: x2-x1 in float, pow(.,2) and the sum and pow(.,0.5) in double (C pow
// overloads promote), order (x^2 + y^2) + z^2, result truncated to float on return.
// The product computes the root with a correctly rounded sqrt (CUDA __dsqrt_rn); glibc's pow is
// within 1 ulp but not correctly rounded by contract, so the two float results can differ only when
// the double root sits within an ulp of a float rounding boundary.  Every call checks it and counts
// the cases (orc_norm_pow_sqrt_mismatches; also the steer magnitude MS:1246 and the edge length MS:399): the tests assert the count stayed 0, so the agreement of
// the two forms on the tested inputs is checked, not assumed.
static std::atomic<long long> g_pow_sqrt_mismatch{0};
float orc_pow_root(double s) {
    const float r = (float)std::pow(s, 0.5);
    const float q = (float)std::sqrt(s);
    if (r != q && !(r != r && q != q)) g_pow_sqrt_mismatch.fetch_add(1, std::memory_order_relaxed);   // NaN == NaN here
    return r;
}
float orc_norm(const float* a, const float* b) {
    const float dx = b[0] - a[0], dy = b[1] - a[1], dz = b[2] - a[2];
    const double s = (std::pow((double)dx, 2) + std::pow((double)dy, 2)) + std::pow((double)dz, 2);
    return orc_pow_root(s);
}
long long orc_norm_pow_sqrt_mismatches(void) { return g_pow_sqrt_mismatch.load(); }

// MS:138-154: min = Norm(V[0],x); for j: if (Norm(V[j],x) <= min) {min=..; min_index=j;}
int orc_nearest(const float* V, int64_t n, const float* q, int64_t nq, int32_t* out_idx,
                float* out_dist, int threads) {
    if (n <= 0) return -1;
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
#pragma omp parallel for schedule(static)
    for (int64_t c = 0; c < nq; ++c) {
        const float* x = q + 3 * c;
        float mn = orc_norm(V, x);
        int32_t mi = 0;
        for (int64_t j = 0; j < n; ++j) {
            const float t = orc_norm(V + 3 * j, x);
            if (t <= mn) {
                mn = t;
                mi = (int32_t)j;
            }
        }
        out_idx[c] = mi;
        if (out_dist) out_dist[c] = mn;
    }
    return 0;
}

// MS:1233-1253.  Note the magnitude sums z^2 + y^2 + x^2 (MS:1246), unlike Norm.
int orc_steer(const float* nearest, const float* rnd, float eta, float* out) {
    if (orc_norm(nearest, rnd) <= eta) {
        out[0] = rnd[0]; out[1] = rnd[1]; out[2] = rnd[2];
        return 1;
    }
    const float dx = rnd[0] - nearest[0], dy = rnd[1] - nearest[1], dz = rnd[2] - nearest[2];
    const float mag = orc_pow_root((std::pow((double)dz, 2) + std::pow((double)dy, 2)) + std::pow((double)dx, 2));
    const float cx = dx / mag, cy = dy / mag, cz = dz / mag;
    const float ex = eta * cx, ey = eta * cy, ez = eta * cz;
    out[0] = nearest[0] + ex;
    out[1] = nearest[1] + ey;
    out[2] = nearest[2] + ez;
    return 0;
}

}  // extern "C"
