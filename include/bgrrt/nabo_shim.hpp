// nabo_shim.hpp — C++11 header-only mirror of the reference planner's call surface over the
// C ABI in bgrrt.h.  Host code stays in C++ (the planner's language); everything here forwards
// to libbgrrt.so (hand-written sm_100a CUDA) — there is no CPU fallback: every call throws
// std::runtime_error carrying bgr_last_error() when the library reports an error (no CUDA
// device included).
//
//  * Nabo::NNSearchF — the subset of libnabo 1.0.7's NearestNeighbourSearch<float> the
//    reference uses (abb_manipulator/src/rrt_star_milestones.cpp:1306 createKDTreeLinearHeap,
//    :1312 knn, :1315 delete) plus the matrix and per-query-maxRadii overloads, with libnabo's
//    argument meaning, defaults and std::runtime_error conditions (checkSizesKnn).  It is
//    templated on the matrix/vector types (anything with data()/rows()/cols(), column-major:
//    Eigen::MatrixXf / VectorXf / VectorXi drop in unchanged), so it compiles with or without
//    Eigen.  Extension: append() replaces the per-iteration rebuild of MS:1277-1306.
//  * bgrrt::CollisionWorld — isStateValid (MS:310-351) / isLineValid (MS:385-471) with the
//    reference's polarity (true == IN COLLISION) and generateBoxesFromOctomap's box layout
//    (MS:288-306: fcl::OcTree::toBoxes() rows {x,y,z,size,occupancy,threshold}).
#ifndef BGRRT_NABO_SHIM_HPP
#define BGRRT_NABO_SHIM_HPP

#include <array>
#include <cstdint>
#include <limits>
#include <stdexcept>
#include <string>
#include <vector>

#include "../bgrrt.h"

namespace bgrrt {

inline void check(int rc, const char* what) {
    if (rc != BGR_OK)
        throw std::runtime_error(std::string("libbgrrt ") + what + ": " + bgr_last_error() +
                                 (rc == BGR_ERR_CUDA ? " [CUDA]" : ""));
}

}  // namespace bgrrt

namespace Nabo {

// libnabo: NearestNeighbourSearch<T>::SearchOptionFlags / CreationOptionFlags
struct NNSearchF {
    typedef int Index;
    enum SearchOptionFlags { ALLOW_SELF_MATCH = 1, SORT_RESULTS = 2 };
    enum CreationOptionFlags { TOUCH_STATISTICS = 1 };

    const Index dim;
    const unsigned creationOptionFlags;

    // libnabo: static NearestNeighbourSearch* createKDTreeLinearHeap(const CloudType& cloud,
    //          const Index dim = max, const unsigned creationOptionFlags = 0, ...)
    // cloud: dim x N, one point per column, column-major.  The cloud is COPIED into the
    // append-only HBM node store (libnabo keeps a reference and forbids resizing it).
    template <typename Matrix>
    static NNSearchF* createKDTreeLinearHeap(const Matrix& cloud,
                                             const Index dim = std::numeric_limits<Index>::max(),
                                             const unsigned creationOptionFlags = 0) {
        const Index d = dim < (Index)cloud.rows() ? dim : (Index)cloud.rows();
        if (d <= 0) throw std::runtime_error("Your space must have at least one dimension");
        if ((Index)cloud.rows() != d)
            throw std::runtime_error("libbgrrt: dim < cloud.rows() (strided clouds) is not supported");
        bgr_nns* h = nullptr;
        bgrrt::check(bgr_nns_create(cloud.data(), (int64_t)cloud.cols(), d, creationOptionFlags, &h),
                     "createKDTreeLinearHeap");
        return new NNSearchF(h, d, creationOptionFlags);
    }
    // the other libnabo factories resolve to the same exact search
    template <typename Matrix>
    static NNSearchF* createBruteForce(const Matrix& cloud, const Index dim = std::numeric_limits<Index>::max(),
                                       const unsigned creationOptionFlags = 0) {
        NNSearchF* s = createKDTreeLinearHeap(cloud, dim, creationOptionFlags);
        bgr_nns_set_method(s->h_, BGR_METHOD_BRUTE);
        return s;
    }
    template <typename Matrix>
    static NNSearchF* createKDTreeTreeHeap(const Matrix& cloud, const Index dim = std::numeric_limits<Index>::max(),
                                           const unsigned creationOptionFlags = 0) {
        return createKDTreeLinearHeap(cloud, dim, creationOptionFlags);
    }

    // libnabo: unsigned long knn(const Vector& query, IndexVector& indices, Vector& dists2,
    //          const Index k = 1, const T epsilon = 0, const unsigned optionFlags = 0,
    //          const T maxRadius = inf) const          (the call at MS:1312)
    // Also accepts dim x Q query matrices with k x Q result matrices (libnabo's matrix overload).
    template <typename Q, typename IV, typename DV>
    unsigned long knn(const Q& query, IV& indices, DV& dists2, const Index k = 1, const float epsilon = 0.f,
                      const unsigned optionFlags = 0,
                      const float maxRadius = std::numeric_limits<float>::infinity()) const {
        checkSizesKnn(query, indices, dists2, k, optionFlags);
        uint64_t touched = 0;
        bgrrt::check(bgr_nns_knn(h_, query.data(), (int64_t)query.cols(), k, epsilon, optionFlags, maxRadius, nullptr,
                                 indices.data(), dists2.data(), &touched), "knn");
        return (creationOptionFlags & TOUCH_STATISTICS) ? (unsigned long)touched : 0ul;
    }
    // libnabo: knn(query, indices, dists2, const Vector& maxRadii, k, epsilon, optionFlags)
    template <typename Q, typename IV, typename DV, typename RV>
    unsigned long knn(const Q& query, IV& indices, DV& dists2, const RV& maxRadii, const Index k,
                      const float epsilon = 0.f, const unsigned optionFlags = 0) const {
        checkSizesKnn(query, indices, dists2, k, optionFlags);
        if ((int64_t)maxRadii.rows() * (int64_t)maxRadii.cols() != (int64_t)query.cols())
            throw std::runtime_error("Maximum radii vector has not the same length as the query matrix has columns");
        uint64_t touched = 0;
        bgrrt::check(bgr_nns_knn(h_, query.data(), (int64_t)query.cols(), k, epsilon, optionFlags,
                                 std::numeric_limits<float>::infinity(), maxRadii.data(), indices.data(),
                                 dists2.data(), &touched), "knn");
        return (creationOptionFlags & TOUCH_STATISTICS) ? (unsigned long)touched : 0ul;
    }

    // ---- extensions over libnabo (the append-only store; reference Nearest())
    template <typename Matrix>
    void append(const Matrix& pts) {   // dim x M
        if ((Index)pts.rows() != dim) throw std::runtime_error("append: wrong point dimension");
        bgrrt::check(bgr_nns_append(h_, pts.data(), (int64_t)pts.cols()), "append");
    }
    void appendPoint(const float* xyz) { bgrrt::check(bgr_nns_append(h_, xyz, 1), "append"); }
    int64_t size() const { return bgr_nns_size(h_); }
    // Nearest(V, x) of MS:138-154: index of the nearest node (last index among float-equal
    // minima) and the reference's mixed-precision Norm.
    int nearest(const float xyz[3], float* norm = nullptr) const {
        int32_t i = -1;
        bgrrt::check(bgr_nns_nearest(h_, xyz, 1, &i, norm), "nearest");
        return i;
    }
    bgr_nns* handle() const { return h_; }

    ~NNSearchF() { bgr_nns_destroy(h_); }

private:
    NNSearchF(bgr_nns* h, Index d, unsigned f) : dim(d), creationOptionFlags(f), h_(h) {}
    NNSearchF(const NNSearchF&);
    NNSearchF& operator=(const NNSearchF&);
    bgr_nns* h_;

    // libnabo nabo.cpp checkSizesKnn: same conditions, same exception type
    template <typename Q, typename IV, typename DV>
    void checkSizesKnn(const Q& query, const IV& indices, const DV& dists2, const Index k,
                       const unsigned optionFlags) const {
        const int64_t n = bgr_nns_size(h_);
        if (optionFlags & ALLOW_SELF_MATCH) {
            if ((int64_t)k > n)
                throw std::runtime_error("Requesting more points (" + std::to_string(k) + ") than available in cloud (" +
                                         std::to_string(n) + ")");
        } else if ((int64_t)k > n - 1) {
            throw std::runtime_error("Requesting more points (" + std::to_string(k) +
                                     ") than available in cloud minus 1 (" + std::to_string(n - 1) +
                                     ") (as self match is forbidden)");
        }
        if ((Index)query.rows() < dim)
            throw std::runtime_error("Query has less dimensions (" + std::to_string(query.rows()) +
                                     ") than requested for cloud (" + std::to_string(dim) + ")");
        if ((Index)query.rows() != dim)
            throw std::runtime_error("libbgrrt: query.rows() > dim (strided queries) is not supported");
        if ((int64_t)indices.rows() != k || (int64_t)dists2.rows() != k)
            throw std::runtime_error("Index/distance matrix has less rows than k (" + std::to_string(k) + ")");
        if ((int64_t)indices.cols() != (int64_t)query.cols() || (int64_t)dists2.cols() != (int64_t)query.cols())
            throw std::runtime_error("Index/distance matrix has a different number of columns than query");
        if (optionFlags > (ALLOW_SELF_MATCH | SORT_RESULTS))
            throw std::runtime_error("OR-ed value of option flags (" + std::to_string(optionFlags) +
                                     ") is larger than maximal valid value (" +
                                     std::to_string(ALLOW_SELF_MATCH | SORT_RESULTS) + ")");
    }
};

}  // namespace Nabo

namespace bgrrt {

// Replaces the `octTree` global + generateBoxesFromOctomap + the two per-leaf fcl::collide
// loops.  One object per map; setBoxes on every octomap_Callback (MS:273-285), or
// updateBoxes with the leaves that changed.
class CollisionWorld {
public:
    typedef std::array<double, 6> Box;   // fcl::OcTree::toBoxes(): {x,y,z,size,occupancy,threshold}

    CollisionWorld(double res, const double origin[3], const int dims[3]) : h_(nullptr) {
        check(bgr_map_create(res, origin, dims, &h_), "map_create");
    }
    ~CollisionWorld() { bgr_map_destroy(h_); }

    void setBoxes(const std::vector<Box>& boxes) {
        check(bgr_map_set_boxes(h_, boxes.empty() ? nullptr : boxes[0].data(), (int64_t)boxes.size()), "set_boxes");
    }
    void updateBoxes(const std::vector<Box>& add, const std::vector<Box>& remove) {
        check(bgr_map_update_boxes(h_, add.empty() ? nullptr : add[0].data(), (int64_t)add.size(),
                                   remove.empty() ? nullptr : remove[0].data(), (int64_t)remove.size()), "update_boxes");
    }
    // bool isStateValid(std::vector<float> point)  MS:310 — true == collision, Sphere(0.1)
    bool isStateValid(const std::vector<float>& point, double radius = 0.1) const {
        uint8_t hit = 0;
        check(bgr_map_collide_spheres(h_, point.data(), 1, radius, &hit, nullptr), "collide_spheres");
        return hit != 0;
    }
    // bool isLineValid(std::vector<float> point1, std::vector<float> point2)  MS:385 — true ==
    // collision, Cylinder(0.03, |p2-p1|)
    bool isLineValid(const std::vector<float>& p1, const std::vector<float>& p2, double radius = 0.03,
                     int shape = BGR_SHAPE_CYLINDER) const {
        uint8_t hit = 0;
        check(bgr_map_collide_edges(h_, p1.data(), p2.data(), 1, radius, shape, &hit, nullptr), "collide_edges");
        return hit != 0;
    }
    // batched forms (m x 3 row-major = 3 x m column-major); in_band may be null
    void collideSpheres(const float* c, int64_t m, double radius, uint8_t* hit, uint8_t* in_band = nullptr) const {
        check(bgr_map_collide_spheres(h_, c, m, radius, hit, in_band), "collide_spheres");
    }
    void collideEdges(const float* p1, const float* p2, int64_t m, double radius, int shape, uint8_t* hit,
                      uint8_t* in_band = nullptr) const {
        check(bgr_map_collide_edges(h_, p1, p2, m, radius, shape, hit, in_band), "collide_edges");
    }
    // map refresh: only the leaves that differ from the current list are rasterised
    void refreshBoxes(const std::vector<Box>& boxes, int64_t* added = nullptr, int64_t* removed = nullptr) {
        check(bgr_map_refresh_boxes(h_, boxes.empty() ? nullptr : boxes[0].data(), (int64_t)boxes.size(), added, removed),
              "refresh_boxes");
    }
    // octomap_msgs::fullMsgToMap + fcl::OcTree::toBoxes (MS:279,291) without octomap/FCL: the
    // message's data bytes (binary == msg.binary, res == msg.resolution) -> toBoxes rows
    static std::vector<Box> boxesFromOctomapMsg(const int8_t* data, size_t n, bool binary, double res,
                                                double occ_threshold = 0.5, double bbox_min[3] = nullptr,
                                                double bbox_max[3] = nullptr) {
        int64_t nb = 0;
        const int fmt = binary ? BGR_OCTREE_BINARY_DATA : BGR_OCTREE_FULL_DATA;
        check(bgr_octree_to_boxes(data, (int64_t)n, fmt, res, occ_threshold, nullptr, &nb, nullptr, nullptr, nullptr), "octree_to_boxes");
        std::vector<Box> boxes((size_t)nb);
        check(bgr_octree_to_boxes(data, (int64_t)n, fmt, res, occ_threshold, nb ? boxes[0].data() : nullptr, &nb, nullptr,
                                  bbox_min, bbox_max), "octree_to_boxes");
        return boxes;
    }
    // std::vector<Pose> PrunePlan(std::vector<Pose> rrt_path)  MS:562 — waypoints as x,y,z triples;
    // returns the kept waypoint indices (all anchor-waypoint edges checked in one launch)
    std::vector<int32_t> PrunePlan(const std::vector<float>& path_xyz, double radius = 0.03,
                                   int shape = BGR_SHAPE_CYLINDER, int64_t* in_band = nullptr) const {
        const int64_t p = (int64_t)path_xyz.size() / 3;
        std::vector<int32_t> keep((size_t)p);
        int64_t nk = p;
        check(bgr_map_prune_path(h_, path_xyz.data(), p, radius, shape, keep.data(), &nk, in_band), "prune_path");
        keep.resize((size_t)nk);
        return keep;
    }
    // void find_labels(std::vector<Pose> final_path)  MS:943 — the 3 interior waypoints closest to the
    // obstacles as 9 floats (the CNN's training label); Sphere(0.1) clearance as in MS:958
    std::vector<float> find_labels(const std::vector<double>& final_path_xyz, double radius = 0.1, int n_labels = 3,
                                   std::vector<int32_t>* indices = nullptr) const {
        std::vector<float> labels((size_t)(3 * n_labels));
        std::vector<int32_t> idx((size_t)n_labels);
        check(bgr_map_find_labels(h_, final_path_xyz.data(), (int64_t)final_path_xyz.size() / 3, radius, n_labels, idx.data(),
                                  labels.data(), nullptr, nullptr), "find_labels");
        if (indices) *indices = idx;
        return labels;
    }
    // fcl::distance(Sphere(radius), every leaf) minimum per point (MS:956-975, RD:890-917); -1 = touching
    void clearance(const double* pts_xyz, int64_t m, double radius, double* dist) const {
        check(bgr_map_clearance(h_, pts_xyz, m, radius, dist), "clearance");
    }
    bgr_map* handle() const { return h_; }

private:
    CollisionWorld(const CollisionWorld&);
    CollisionWorld& operator=(const CollisionWorld&);
    bgr_map* h_;
};

// std::vector<Pose> SmoothenPlan(std::vector<Pose> rrt_path)  MS:645 — x,y,z triples in, 11 samples out
inline std::vector<double> SmoothenPlan(const std::vector<double>& path_xyz) {
    int64_t ns = 0;
    check(bgr_path_smoothen(path_xyz.data(), (int64_t)path_xyz.size() / 3, nullptr, &ns), "path_smoothen");
    std::vector<double> out((size_t)(3 * ns));
    check(bgr_path_smoothen(path_xyz.data(), (int64_t)path_xyz.size() / 3, out.data(), &ns), "path_smoothen");
    return out;
}

}  // namespace bgrrt

#endif  // BGRRT_NABO_SHIM_HPP
