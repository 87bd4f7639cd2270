// Compile-and-run test of include/bgrrt/nabo_shim.hpp, written the way the reference planner
// uses libnabo and its two validity functions (rrt_star_milestones.cpp:1277-1315, 1270, 589).
// No Eigen in this image: Mat is the minimal column-major matrix the shim's templates need
// (data()/rows()/cols()) — Eigen::MatrixXf/VectorXf/VectorXi expose the same three members.
// exit 0: ran on a GPU and matched a scalar check; exit 3: no CUDA device and the shim threw.
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <vector>

#include "bgrrt/nabo_shim.hpp"

template <typename T>
struct Mat {
    std::vector<T> v;
    long r, c;
    Mat(long rows, long cols) : v((size_t)(rows * cols)), r(rows), c(cols) {}
    T* data() { return v.data(); }
    const T* data() const { return v.data(); }
    long rows() const { return r; }
    long cols() const { return c; }
    T& operator()(long i, long j) { return v[(size_t)(j * r + i)]; }
    T operator()(long i, long j) const { return v[(size_t)(j * r + i)]; }
};

int main() {
    const int N = 500, K = 5;
    Mat<float> M(3, N);
    unsigned s = 12345u;
    for (int j = 0; j < N; ++j)
        for (int i = 0; i < 3; ++i) {
            s = s * 1664525u + 1013904223u;
            M(i, j) = (float)(s >> 8) * (1.0f / 16777216.0f) * 3.2f;
        }
    Mat<float> q(3, 1);
    q(0, 0) = 1.1f; q(1, 0) = 2.2f; q(2, 0) = 0.7f;
    Mat<int> indices(K, 1);
    Mat<float> dists2(K, 1);
    try {
        Nabo::NNSearchF* nns = Nabo::NNSearchF::createKDTreeLinearHeap(M);      // MS:1306
        nns->knn(q, indices, dists2, K);                                        // MS:1312
        // scalar check: libnabo leaf arithmetic, ascending, ties by index
        std::vector<int> best;
        std::vector<float> bd;
        for (int kk = 0; kk < K; ++kk) {
            int bi = -1;
            float bv = INFINITY;
            for (int j = 0; j < N; ++j) {
                bool used = false;
                for (int u : best) used |= (u == j);
                if (used) continue;
                const float dx = q(0, 0) - M(0, j), dy = q(1, 0) - M(1, j), dz = q(2, 0) - M(2, j);
                volatile float a = dx * dx, b = dy * dy, c = dz * dz;
                volatile float ab = a + b;
                const float d = ab + c;
                if (d > FLT_EPSILON && d < bv) { bv = d; bi = j; }
            }
            best.push_back(bi);
            bd.push_back(bv);
        }
        int bad = 0;
        for (int kk = 0; kk < K; ++kk) bad += (indices(kk, 0) != best[kk]) || (dists2(kk, 0) != bd[kk]);
        // libnabo's error conditions surface as std::runtime_error
        int threw = 0;
        try { Mat<int> i2(K + 1, 1); nns->knn(q, i2, dists2, K); } catch (const std::runtime_error&) { ++threw; }
        try { nns->knn(q, indices, dists2, K, 0.f, 4u); } catch (const std::runtime_error&) { ++threw; }
        try { Mat<int> i3(N, 1); Mat<float> d3(N, 1); nns->knn(q, i3, d3, N); } catch (const std::runtime_error&) { ++threw; }
        float one[3] = {0.5f, 0.5f, 0.5f};
        nns->appendPoint(one);
        float nrm = 0.f;
        const int ni = nns->nearest(one, &nrm);
        delete nns;                                                             // MS:1315

        const double origin[3] = {0, 0, 0};
        const int dims[3] = {64, 64, 64};
        bgrrt::CollisionWorld world(0.05, origin, dims);
        std::vector<bgrrt::CollisionWorld::Box> boxes;
        boxes.push_back({{1.025, 1.025, 1.025, 0.05, 0.97, 0.5}});
        world.setBoxes(boxes);
        const bool in_hit = world.isStateValid({1.0f, 1.0f, 1.0f});             // true == collision
        const bool in_free = world.isStateValid({2.0f, 2.0f, 2.0f});
        const bool line_hit = world.isLineValid({0.5f, 1.025f, 1.025f}, {1.5f, 1.025f, 1.025f});
        const bool line_free = world.isLineValid({0.5f, 2.0f, 2.0f}, {1.5f, 2.0f, 2.0f});
        // path post-processing the way main() uses it (MS:1352-1356): PrunePlan -> SmoothenPlan -> find_labels
        std::vector<float> path = {0.5f, 2.0f, 2.0f, 0.9f, 2.1f, 2.0f, 1.2f, 2.0f, 2.2f, 1.5f, 2.0f, 2.0f, 2.5f, 2.2f, 2.1f};
        const std::vector<int32_t> kept = world.PrunePlan(path);               // free space: start -> goal
        std::vector<double> pd(path.begin(), path.end());
        const std::vector<double> smooth = bgrrt::SmoothenPlan(pd);             // > 3 waypoints: cubic over the first four
        std::vector<int32_t> lab_idx;
        const std::vector<float> labels = world.find_labels(pd, 0.1, 3, &lab_idx);
        // one occupied depth-16 leaf at key (32768+20, 32768+20, 32768+20) as a full-message stream
        std::vector<int8_t> msg;
        auto put = [&](float v, unsigned char mask) {
            const unsigned char* b = reinterpret_cast<const unsigned char*>(&v);
            for (int i = 0; i < 4; ++i) msg.push_back((int8_t)b[i]);
            msg.push_back((int8_t)mask);
        };
        put(1.f, 1u << 7);                                                       // root -> +x+y+z octant
        for (int d = 1; d < 16; ++d) {
            const int bit = (20 >> (15 - d)) & 1;                                // key offset 20 = 0b10100 from the octant's corner
            put(1.f, (unsigned char)(1u << (bit ? 7 : 0)));
        }
        put(1.f, 0);
        const std::vector<bgrrt::CollisionWorld::Box> mb = bgrrt::CollisionWorld::boxesFromOctomapMsg(msg.data(), msg.size(), false, 0.05);
        world.refreshBoxes(mb);
        const bool msg_hit = world.isStateValid({1.025f, 1.025f, 1.025f});
        const bool old_gone = !world.isStateValid({1.3f, 1.3f, 1.3f});
        const bool post_ok = kept.size() == 2 && kept[0] == 0 && kept[1] == 4 && smooth.size() == 33 && smooth[0] == 0.5 &&
                             labels.size() == 9 && lab_idx.size() == 3 && mb.size() == 1 && std::fabs(mb[0][0] - 1.025) < 1e-12 &&
                             msg_hit && old_gone;
        const bool ok = bad == 0 && threw == 3 && ni == N && nrm == 0.f && in_hit && !in_free && line_hit && !line_free && post_ok;
        std::printf("shim test: knn mismatches=%d threw=%d nearest=%d hit=%d/%d line=%d/%d -> %s\n", bad, threw, ni,
                    (int)in_hit, (int)in_free, (int)line_hit, (int)line_free, ok ? "ok" : "FAIL");
        return ok ? 0 : 1;
    } catch (const std::runtime_error& e) {
        std::printf("runtime_error: %s\n", e.what());
        return std::string(e.what()).find("[CUDA]") != std::string::npos ? 3 : 2;
    }
}
