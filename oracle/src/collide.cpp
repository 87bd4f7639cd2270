// ORACLE (test infrastructure).  Restatement of the reference's collision queries:
//   isStateValid  MS:310-351  fcl::collide(Sphere(0.1)@point, Box) OR-ed over every leaf box
//   isLineValid   MS:385-471  fcl::collide(Cylinder(0.03,|p2-p1|), Box) OR-ed likewise
//   generateBoxesFromOctomap MS:288-306  boxes {x,y,z,size,occupancy,threshold}
// FCL 0.5.x and libccd 2.0 are un-vendored dependencies absent from /root/reference and
// from this image; their published algorithm is restated: FCL's GJKSolver_libccd routes
// Sphere/Cylinder-vs-Box (no specialisation in 0.5) to ccdMPRPenetration with
// mpr_tolerance 1e-6, whose boolean is discoverPortal + refinePortal.
// PARITY IS UNPINNED against the real libraries.
// All arithmetic is double, compiled with -ffp-contract=off (libccd is plain C, no FMA).
#include "../oracle.h"
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

// ------------------------------------------------------------------ libccd vec3.h
const double CCD_EPS = DBL_EPSILON;
struct V3 { double v[3]; };
inline bool is_zero(double x) { return std::fabs(x) < CCD_EPS; }
inline int sgn(double x) { return is_zero(x) ? 0 : (x < 0 ? -1 : 1); }
inline bool ccd_eq(double a, double b) {
    double ab = std::fabs(a - b);
    if (ab < CCD_EPS) return true;
    a = std::fabs(a); b = std::fabs(b);
    return b > a ? ab < CCD_EPS * b : ab < CCD_EPS * a;
}
inline bool v_eq0(const V3& a) { return ccd_eq(a.v[0], 0) && ccd_eq(a.v[1], 0) && ccd_eq(a.v[2], 0); }
inline double dot(const V3& a, const V3& b) {
    double d = a.v[0] * b.v[0];
    d += a.v[1] * b.v[1];
    d += a.v[2] * b.v[2];
    return d;
}
inline V3 sub(const V3& a, const V3& b) { return {{a.v[0] - b.v[0], a.v[1] - b.v[1], a.v[2] - b.v[2]}}; }
inline V3 cross(const V3& a, const V3& b) {
    return {{a.v[1] * b.v[2] - a.v[2] * b.v[1], a.v[2] * b.v[0] - a.v[0] * b.v[2],
             a.v[0] * b.v[1] - a.v[1] * b.v[0]}};
}
inline void scale(V3& a, double k) { a.v[0] *= k; a.v[1] *= k; a.v[2] *= k; }
inline void normalize(V3& a) { scale(a, 1.0 / std::sqrt(dot(a, a))); }

// ------------------------------------------------------------------ libccd quat.h
struct Q4 { double q[4]; };  // x,y,z,w
inline void quat_rot(V3& v, const Q4& q) {  // ccdQuatRotVec (18 mul + 12 add form)
    const double x = q.q[0], y = q.q[1], z = q.q[2], w = q.q[3];
    const double vx = v.v[0], vy = v.v[1], vz = v.v[2];
    const double c1x = y * vz - z * vy + w * vx;
    const double c1y = z * vx - x * vz + w * vy;
    const double c1z = x * vy - y * vx + w * vz;
    const double c2x = y * c1z - z * c1y;
    const double c2y = z * c1x - x * c1z;
    const double c2z = x * c1y - y * c1x;
    v.v[0] = vx + 2 * c2x; v.v[1] = vy + 2 * c2y; v.v[2] = vz + 2 * c2z;
}
inline Q4 quat_inv(const Q4& q) {  // ccdQuatInvert2
    double len2 = q.q[0] * q.q[0];
    len2 += q.q[1] * q.q[1];
    len2 += q.q[2] * q.q[2];
    len2 += q.q[3] * q.q[3];
    Q4 r = q;
    if (len2 < CCD_EPS) return r;
    len2 = 1.0 / len2;
    r.q[0] = -q.q[0] * len2; r.q[1] = -q.q[1] * len2; r.q[2] = -q.q[2] * len2; r.q[3] = q.q[3] * len2;
    return r;
}

// ------------------------------------------------------------------ FCL gjk_libccd shapes
enum { T_BOX = 0, T_SPHERE = 1, T_CYL = 2, T_CAP = 3 };
struct Obj {
    int type;
    V3 pos;
    Q4 rot, rot_inv;
    double dim[3];   // box half extents
    double radius;   // sphere / cyl / cap
    double height;   // cyl / cap half length
};

void support(const Obj& o, const V3& dir_, V3& v) {
    V3 dir = dir_;
    quat_rot(dir, o.rot_inv);
    switch (o.type) {
        case T_BOX:
            v = {{sgn(dir.v[0]) * o.dim[0], sgn(dir.v[1]) * o.dim[1], sgn(dir.v[2]) * o.dim[2]}};
            break;
        case T_SPHERE: {
            v = dir;
            scale(v, o.radius);
            scale(v, 1.0 / std::sqrt(dot(dir, dir)));
            break;
        }
        case T_CYL: {
            double zdist = dir.v[0] * dir.v[0] + dir.v[1] * dir.v[1];
            zdist = std::sqrt(zdist);
            if (is_zero(zdist)) v = {{0., 0., sgn(dir.v[2]) * o.height}};
            else {
                const double rad = o.radius / zdist;
                v = {{rad * dir.v[0], rad * dir.v[1], sgn(dir.v[2]) * o.height}};
            }
            break;
        }
        default: {  // capsule
            V3 p1 = {{0, 0, o.height}}, p2 = {{0, 0, -o.height}};
            v = dir;
            scale(v, o.radius);
            for (int i = 0; i < 3; ++i) { p1.v[i] += v.v[i]; p2.v[i] += v.v[i]; }
            v = dot(dir, p1) > dot(dir, p2) ? p1 : p2;
        }
    }
    quat_rot(v, o.rot);
    v.v[0] += o.pos.v[0]; v.v[1] += o.pos.v[1]; v.v[2] += o.pos.v[2];
}

// Minkowski-difference support: __ccdSupport
inline V3 mink(const Obj& a, const Obj& b, const V3& dir) {
    V3 s1, s2, nd = dir;
    support(a, dir, s1);
    scale(nd, -1.0);
    support(b, nd, s2);
    return sub(s1, s2);
}

const int MPR_ITER_CAP = 512;  // libccd loops are unbounded; cap shared with the CUDA path

// returns 1 = intersect (ccdMPRPenetration()==0), 0 = miss.  *iters counts support calls.
int mpr(const Obj& o1, const Obj& o2, double tol, int* iters) {
    int it = 0;
    V3 v0 = sub(o1.pos, o2.pos), v1, v2, v3, v4, dir, va, vb;
    if (v_eq0(v0)) v0.v[0] += CCD_EPS * 10.0;
    // ---- discoverPortal
    dir = v0; scale(dir, -1.0); normalize(dir);
    v1 = mink(o1, o2, dir); ++it;
    double d = dot(v1, dir);
    int res = 0;
    if (is_zero(d) || d < 0) { res = -1; goto done; }
    dir = cross(v0, v1);
    if (is_zero(dot(dir, dir))) { res = v_eq0(v1) ? 1 : 2; goto done; }
    normalize(dir);
    v2 = mink(o1, o2, dir); ++it;
    d = dot(v2, dir);
    if (is_zero(d) || d < 0) { res = -1; goto done; }
    va = sub(v1, v0); vb = sub(v2, v0);
    dir = cross(va, vb); normalize(dir);
    d = dot(dir, v0);
    if (d > 0) { std::swap(v1, v2); scale(dir, -1.0); }
    for (;;) {
        if (it >= MPR_ITER_CAP) { res = -1; goto done; }
        v3 = mink(o1, o2, dir); ++it;
        d = dot(v3, dir);
        if (is_zero(d) || d < 0) { res = -1; goto done; }
        bool cont = false;
        va = cross(v1, v3);
        d = dot(va, v0);
        if (d < 0 && !is_zero(d)) { v2 = v3; cont = true; }
        if (!cont) {
            va = cross(v3, v2);
            d = dot(va, v0);
            if (d < 0 && !is_zero(d)) { v1 = v3; cont = true; }
        }
        if (!cont) break;
        va = sub(v1, v0); vb = sub(v2, v0);
        dir = cross(va, vb); normalize(dir);
    }
    // ---- refinePortal
    for (;;) {
        va = sub(v2, v1); vb = sub(v3, v1);
        dir = cross(va, vb); normalize(dir);
        d = dot(dir, v1);
        if (is_zero(d) || d > 0) { res = 0; goto done; }            // portalEncapsulesOrigin
        if (it >= MPR_ITER_CAP) { res = -1; goto done; }
        v4 = mink(o1, o2, dir); ++it;
        const double d4 = dot(v4, dir);
        if (!(is_zero(d4) || d4 > 0)) { res = -1; goto done; }      // portalCanEncapsuleOrigin
        {                                                            // portalReachTolerance
            const double dv1 = dot(v1, dir), dv2 = dot(v2, dir), dv3 = dot(v3, dir);
            double m = std::min(std::min(d4 - dv1, d4 - dv2), d4 - dv3);
            if (ccd_eq(m, tol) || m < tol) { res = -1; goto done; }
        }
        {                                                            // expandPortal
            const V3 v4v0 = cross(v4, v0);
            if (dot(v1, v4v0) > 0) {
                if (dot(v2, v4v0) > 0) v1 = v4; else v3 = v4;
            } else {
                if (dot(v3, v4v0) > 0) v2 = v4; else v1 = v4;
            }
        }
    }
done:
    if (iters) *iters = it;
    return res >= 0 ? 1 : 0;
}

Obj make_box(const double c[3], double size) {
    Obj o{};
    o.type = T_BOX;
    o.pos = {{c[0], c[1], c[2]}};
    o.rot = {{0, 0, 0, 1}};
    o.rot_inv = quat_inv(o.rot);
    o.dim[0] = o.dim[1] = o.dim[2] = size / 2;
    return o;
}
Obj make_sphere(const double c[3], double r) {
    Obj o{};
    o.type = T_SPHERE;
    o.pos = {{c[0], c[1], c[2]}};
    o.rot = {{0, 0, 0, 1}};
    o.rot_inv = quat_inv(o.rot);
    o.radius = r;
    return o;
}

// The reference's f32 transform setup for the edge shape, MS:396-437:
//   centre  = (p1+p2)/2 in float                                 (MS:396)
//   lz      = float(pow(dz^2 + dy^2 + dx^2, 0.5)), doubles inside (MS:399)
//   q       = Eigen::Quaternionf::setFromTwoVectors((0,0,1), p2-p1) in float (MS:429-432)
//   tf1.setQuatRotation(Quaternion3f(w,x,y,z)) widened to double  (MS:435)
// Eigen: v1 = b/sqrt(b.squaredNorm()) (fixed-size-3 redux order x^2+(y^2+z^2)),
// c = v1.z, axis = (0,0,1) x v1 = (-v1.y, v1.x, 0), s = sqrt((1+c)*2), vec = axis*(1/s),
// w = s*0.5.  For c < -1+1e-5 Eigen falls back to a JacobiSVD null-space axis (a unit vector
// perpendicular to both; its SIGN is an implementation detail of the SVD, and the wrong sign
// mirrors the edge about z by up to ~9e-3 rad): here the axis normalize(-vy, vx, 0), which
// reproduces the edge direction.  The reference's result in this branch is therefore unpinned.
// p1 == p2: Eigen yields NaN; defined here as identity rotation, lz = 0.
void edge_frame_f32(const float p1[3], const float p2[3], float c[3], float q[4], float* lz) {
    c[0] = (p1[0] + p2[0]) / 2; c[1] = (p1[1] + p2[1]) / 2; c[2] = (p1[2] + p2[2]) / 2;
    const float dx = p2[0] - p1[0], dy = p2[1] - p1[1], dz = p2[2] - p1[2];
    *lz = orc_pow_root((std::pow((double)dz, 2) + std::pow((double)dy, 2)) + std::pow((double)dx, 2));
    const float n2 = dx * dx + (dy * dy + dz * dz);
    if (!(n2 > 0.f)) { q[0] = q[1] = q[2] = 0.f; q[3] = 1.f; return; }
    const float nrm = std::sqrt(n2);
    const float vx = dx / nrm, vy = dy / nrm, vz = dz / nrm;
    const float cc = vz;
    if (cc < -1.f + 1e-5f) {
        // rotation axis = the unit vector perpendicular to both, on the right-handed side, so that
        // the rotated z axis IS the edge direction (the shapes are symmetric about their axis: only
        // that direction matters); exactly antiparallel: (1,0,0)
        const float c2 = std::max(cc, -1.f);
        const float w2 = (1.f + c2) * 0.5f;
        const float sv = std::sqrt(1.f - w2);
        const float an2 = vy * vy + vx * vx;
        float ax = 1.f, ay = 0.f;
        if (an2 > 0.f) {
            const float inva = 1.f / std::sqrt(an2);
            ax = -vy * inva;
            ay = vx * inva;
        }
        q[0] = ax * sv; q[1] = ay * sv; q[2] = 0.f; q[3] = std::sqrt(w2);
        return;
    }
    const float s = std::sqrt((1.f + cc) * 2.f);
    const float invs = 1.f / s;
    q[0] = -vy * invs; q[1] = vx * invs; q[2] = 0.f * invs; q[3] = s * 0.5f;
}

Obj make_edge(const float p1[3], const float p2[3], double r, int shape) {
    float c[3], q[4], lz;
    edge_frame_f32(p1, p2, c, q, &lz);
    Obj o{};
    o.type = shape == 1 ? T_CAP : T_CYL;
    o.pos = {{(double)c[0], (double)c[1], (double)c[2]}};
    o.rot = {{(double)q[0], (double)q[1], (double)q[2], (double)q[3]}};
    o.rot_inv = quat_inv(o.rot);
    o.radius = r;
    o.height = (double)lz / 2;
    return o;
}

// ------------------------------------------------------------------ independent gaps
// Not MPR: closed form for the sphere, nested golden-section minimisation of a convex
// function over the box for the edge shapes.  gap(x) for a flat-capped cylinder is the
// minimal uniform inflation t of (radius, half-length) that makes x belong to it:
//   g(x) = max(rho(x) - r, |a(x)| - h),  gap = min over the box  (<=0 iff they intersect;
//   for gap>0 the Euclidean separation lies in [gap, sqrt(2) gap]).  Capsule:
//   g(x) = dist(x, axis segment) - r, the exact signed separation while positive.
struct Frame {
    double c[3], u[3], h, r;
    int cap;
};
Frame frame_of(const Obj& o) {
    Frame f;
    V3 z = {{0, 0, 1}};
    quat_rot(z, o.rot);
    const double n = std::sqrt(dot(z, z));
    for (int i = 0; i < 3; ++i) { f.c[i] = o.pos.v[i]; f.u[i] = z.v[i] / n; }
    f.h = o.height; f.r = o.radius; f.cap = o.type == T_CAP;
    return f;
}
inline double gfun(const Frame& f, double x, double y, double z) {
    const double wx = x - f.c[0], wy = y - f.c[1], wz = z - f.c[2];
    const double a = wx * f.u[0] + wy * f.u[1] + wz * f.u[2];
    if (f.cap) {
        const double t = std::max(-f.h, std::min(f.h, a));
        const double ex = wx - t * f.u[0], ey = wy - t * f.u[1], ez = wz - t * f.u[2];
        return std::sqrt(ex * ex + ey * ey + ez * ez) - f.r;
    }
    const double ex = wx - a * f.u[0], ey = wy - a * f.u[1], ez = wz - a * f.u[2];
    const double rho = std::sqrt(ex * ex + ey * ey + ez * ez);
    return std::max(rho - f.r, std::fabs(a) - f.h);
}
template <class F>
double golden(F&& fn, double lo, double hi) {
    const double g = 0.6180339887498949;
    double a = lo, b = hi, x1 = b - g * (b - a), x2 = a + g * (b - a);
    double f1 = fn(x1), f2 = fn(x2);
    for (int i = 0; i < 80; ++i) {
        if (f1 <= f2) { b = x2; x2 = x1; f2 = f1; x1 = b - g * (b - a); f1 = fn(x1); }
        else { a = x1; x1 = x2; f1 = f2; x2 = a + g * (b - a); f2 = fn(x2); }
    }
    return std::min(std::min(f1, f2), std::min(fn(lo), fn(hi)));
}
double gap_edge_box(const Frame& f, const double bc[3], double size) {
    const double e = size / 2;
    return golden([&](double x) {
        return golden([&](double y) {
            return golden([&](double z) { return gfun(f, x, y, z); }, bc[2] - e, bc[2] + e);
        }, bc[1] - e, bc[1] + e);
    }, bc[0] - e, bc[0] + e);
}
inline double gap_sphere_box(const double c[3], double r, const double bc[3], double size) {
    const double e = size / 2;
    double d2 = 0;
    for (int i = 0; i < 3; ++i) {
        const double t = std::max(0.0, std::fabs(c[i] - bc[i]) - e);
        d2 += t * t;
    }
    return std::sqrt(d2) - r;
}
// cheap certified bounds on gap_edge_box from the box's inscribed/circumscribed balls
inline void gap_bounds(const Frame& f, const double bc[3], double size, double* lo, double* hi) {
    const double g = gfun(f, bc[0], bc[1], bc[2]);
    // g is 1-Lipschitz in the max-norm sense used here up to sqrt(2)
    *hi = g;
    *lo = g - size * 0.5 * 1.7320508075688774 * 1.0000001;
}

}  // namespace

extern "C" {

void orc_edge_frame(const float p1[3], const float p2[3], double out[8]) {
    float c[3], q[4], lz;
    edge_frame_f32(p1, p2, c, q, &lz);
    for (int i = 0; i < 3; ++i) out[i] = c[i];
    for (int i = 0; i < 4; ++i) out[3 + i] = q[i];
    out[7] = lz;
}

int orc_mpr_sphere_box(const double c[3], double r, const double box_c[3], double box_size,
                       int* iters) {
    return mpr(make_sphere(c, r), make_box(box_c, box_size), 1e-6, iters);
}
int orc_mpr_cyl_box(const float p1[3], const float p2[3], double r, int shape,
                    const double box_c[3], double box_size, int* iters) {
    return mpr(make_edge(p1, p2, r, shape), make_box(box_c, box_size), 1e-6, iters);
}
double orc_sep_sphere_box(const double c[3], double r, const double box_c[3], double box_size) {
    return gap_sphere_box(c, r, box_c, box_size);
}
double orc_sep_cyl_box(const float p1[3], const float p2[3], double r, double dr, int shape,
                       const double box_c[3], double box_size) {
    Obj o = make_edge(p1, p2, r, shape);
    Frame f = frame_of(o);
    f.r += dr;
    if (!f.cap) f.h += dr;
    return gap_edge_box(f, box_c, box_size);
}

// mode 2 = "library-best" CPU baseline: what a maintainer gets from an FCL
// broadphase instead of the reference's loop over every leaf — the boxes are bucketed once into a
// coarse uniform grid (8 finest cells per bucket side, CSR lists, cached per box array); an edge
// visits the buckets its axis-aligned bounding box overlaps, culls by bounding sphere and runs the
// same MPR on what is left.  Conservative (a colliding box overlaps the AABB), so the booleans equal
// modes 0 / 1 (tests/test_oracle.py).
namespace {
struct EdgeBuckets {
    const double* boxes = nullptr;
    int64_t nb = -1;
    double cell = 0, org[3] = {0, 0, 0};
    int g[3] = {0, 0, 0};
    double max_half = 0;                 // largest box half size (buckets hold box CENTRES)
    std::vector<int64_t> start;
    std::vector<int32_t> items;
};
EdgeBuckets g_buckets;

void build_buckets(const double* boxes, int64_t nb) {
    EdgeBuckets& B = g_buckets;
    if (B.boxes == boxes && B.nb == nb) return;
    B.boxes = boxes; B.nb = nb;
    double mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY}, smin = INFINITY;
    B.max_half = 0;
    for (int64_t b = 0; b < nb; ++b) {
        for (int a = 0; a < 3; ++a) { mn[a] = std::min(mn[a], boxes[6 * b + a]); mx[a] = std::max(mx[a], boxes[6 * b + a]); }
        smin = std::min(smin, boxes[6 * b + 3]);
        B.max_half = std::max(B.max_half, 0.5 * boxes[6 * b + 3]);
    }
    if (nb == 0) { B.g[0] = B.g[1] = B.g[2] = 0; B.start.assign(1, 0); B.items.clear(); return; }
    B.cell = 8.0 * smin;
    int64_t cells = 1;
    for (int a = 0; a < 3; ++a) {
        B.org[a] = mn[a];
        B.g[a] = (int)std::floor((mx[a] - mn[a]) / B.cell) + 1;
        cells *= B.g[a];
    }
    auto cell_of = [&](const double* c) {
        int64_t id = 0;
        for (int a = 2; a >= 0; --a) {
            int v = (int)std::floor((c[a] - B.org[a]) / B.cell);
            v = std::min(std::max(v, 0), B.g[a] - 1);
            id = id * B.g[a] + v;
        }
        return id;
    };
    B.start.assign((size_t)cells + 1, 0);
    for (int64_t b = 0; b < nb; ++b) B.start[(size_t)cell_of(boxes + 6 * b) + 1]++;
    for (int64_t c = 0; c < cells; ++c) B.start[(size_t)c + 1] += B.start[(size_t)c];
    B.items.resize((size_t)nb);
    std::vector<int64_t> cur(B.start.begin(), B.start.end() - 1);
    for (int64_t b = 0; b < nb; ++b) B.items[(size_t)cur[(size_t)cell_of(boxes + 6 * b)]++] = (int32_t)b;
}
}  // namespace

int orc_collide_spheres(const double* boxes, int64_t nb, const float* centers, int64_t m,
                        double radius, int mode, double band_tol, uint8_t* hit, uint8_t* band,
                        int threads) {
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
    if (mode == 2) build_buckets(boxes, nb);
#pragma omp parallel for schedule(dynamic, 8)
    for (int64_t i = 0; i < m; ++i) {
        // MS:319-322: Vec3f(point[0],point[1],point[2]) widens the floats; Sphere(0.1)
        const double c[3] = {(double)centers[3 * i], (double)centers[3 * i + 1], (double)centers[3 * i + 2]};
        const Obj s = make_sphere(c, radius);
        bool flag = false;
        if (mode == 2) {
            // bucket-grid broadphase (see build_buckets): the boxes whose centre bucket overlaps the
            // sphere's AABB grown by the largest box half size, same bounding-sphere cull, same MPR;
            // conservative, so the boolean equals modes 0 / 1.  The band uses the same candidates (a
            // box outside them is farther than 1e-3 m from the sphere).
            const EdgeBuckets& B = g_buckets;
            double g = INFINITY;
            int lo[3], hi[3];
            bool empty = nb == 0;
            for (int a = 0; a < 3 && !empty; ++a) {
                const double ext = radius + B.max_half + 1e-3;
                lo[a] = std::max(0, (int)std::floor((c[a] - ext - B.org[a]) / B.cell));
                hi[a] = std::min(B.g[a] - 1, (int)std::floor((c[a] + ext - B.org[a]) / B.cell));
                empty = lo[a] > hi[a];
            }
            for (int z = empty ? 1 : lo[2]; !empty && z <= hi[2]; ++z)
                for (int y = lo[1]; y <= hi[1]; ++y) {
                    const int64_t row = ((int64_t)z * B.g[1] + y) * B.g[0];
                    for (int64_t k = B.start[(size_t)(row + lo[0])]; k < B.start[(size_t)(row + hi[0] + 1)]; ++k) {
                        const double* bx = boxes + 6 * (int64_t)B.items[(size_t)k];
                        const double R = radius + bx[3] * 0.8660254037844387 + 1e-3;
                        const double dx = bx[0] - c[0], dy = bx[1] - c[1], dz = bx[2] - c[2];
                        if (dx * dx + dy * dy + dz * dz > R * R) continue;
                        if (band) g = std::min(g, gap_sphere_box(c, radius, bx, bx[3]));
                        if (!flag && mpr(s, make_box(bx, bx[3]), 1e-6, nullptr)) flag = true;
                    }
                }
            hit[i] = flag ? 1 : 0;
            if (band) band[i] = std::fabs(g) <= band_tol ? 1 : 0;
            continue;
        }
        for (int64_t b = 0; b < nb; ++b) {
            const double* bx = boxes + 6 * b;
            if (mode == 1) {
                const double R = radius + bx[3] * 0.8660254037844387 + 1e-3;
                const double dx = bx[0] - c[0], dy = bx[1] - c[1], dz = bx[2] - c[2];
                if (dx * dx + dy * dy + dz * dz > R * R) continue;
            }
            if (mpr(s, make_box(bx, bx[3]), 1e-6, nullptr)) { flag = true; break; }   // MS:331-340
        }
        hit[i] = flag ? 1 : 0;
        if (band) {
            double g = INFINITY;
            for (int64_t b = 0; b < nb; ++b) g = std::min(g, gap_sphere_box(c, radius, boxes + 6 * b, boxes[6 * b + 3]));
            band[i] = std::fabs(g) <= band_tol ? 1 : 0;
        }
    }
    return 0;
}

int orc_collide_edges(const double* boxes, int64_t nb, const float* p1, const float* p2,
                      int64_t m, double radius, int shape, int mode, double band_tol,
                      uint8_t* hit, uint8_t* band, int threads) {
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
    if (mode == 2) {
        build_buckets(boxes, nb);
        const EdgeBuckets& B = g_buckets;
#pragma omp parallel for schedule(dynamic, 16)
        for (int64_t i = 0; i < m; ++i) {
            const Obj e = make_edge(p1 + 3 * i, p2 + 3 * i, radius, shape);
            const Frame f = frame_of(e);
            const double reach = std::sqrt(e.height * e.height + radius * radius) + radius;
            bool flag = false;
            int lo[3], hi[3];
            bool empty = nb == 0;
            for (int a = 0; a < 3 && !empty; ++a) {
                // AABB of the (capsule hull of the) shape, grown by the largest box half size
                const double ext = f.h * std::fabs(f.u[a]) + radius + B.max_half + 1e-3;
                lo[a] = std::max(0, (int)std::floor((f.c[a] - ext - B.org[a]) / B.cell));
                hi[a] = std::min(B.g[a] - 1, (int)std::floor((f.c[a] + ext - B.org[a]) / B.cell));
                empty = lo[a] > hi[a];
            }
            for (int z = empty ? 1 : lo[2]; !empty && z <= hi[2] && !flag; ++z)
                for (int y = lo[1]; y <= hi[1] && !flag; ++y) {
                    const int64_t row = ((int64_t)z * B.g[1] + y) * B.g[0];
                    for (int64_t k = B.start[(size_t)(row + lo[0])]; k < B.start[(size_t)(row + hi[0] + 1)]; ++k) {
                        const double* bx = boxes + 6 * (int64_t)B.items[(size_t)k];
                        const double R = reach + bx[3] * 0.8660254037844387 + 1e-3;
                        const double dx = bx[0] - e.pos.v[0], dy = bx[1] - e.pos.v[1], dz = bx[2] - e.pos.v[2];
                        if (dx * dx + dy * dy + dz * dz > R * R) continue;
                        // tighter cull: distance of the box centre to the axis segment
                        const double al = dx * f.u[0] + dy * f.u[1] + dz * f.u[2];
                        const double t = std::max(-f.h, std::min(f.h, al));
                        const double ex = dx - t * f.u[0], ey = dy - t * f.u[1], ez = dz - t * f.u[2];
                        const double R2 = radius + bx[3] * 0.8660254037844387 + 1e-3;
                        if (ex * ex + ey * ey + ez * ez > R2 * R2) continue;
                        if (mpr(e, make_box(bx, bx[3]), 1e-6, nullptr)) { flag = true; break; }
                    }
                }
            hit[i] = flag ? 1 : 0;
            if (band) band[i] = 0;
        }
        return 0;
    }
#pragma omp parallel for schedule(dynamic, 4)
    for (int64_t i = 0; i < m; ++i) {
        const Obj e = make_edge(p1 + 3 * i, p2 + 3 * i, radius, shape);
        const double reach = std::sqrt(e.height * e.height + radius * radius) + radius;
        bool flag = false;
        for (int64_t b = 0; b < nb; ++b) {
            const double* bx = boxes + 6 * b;
            if (mode == 1) {
                const double R = reach + bx[3] * 0.8660254037844387 + 1e-3;
                const double dx = bx[0] - e.pos.v[0], dy = bx[1] - e.pos.v[1], dz = bx[2] - e.pos.v[2];
                if (dx * dx + dy * dy + dz * dz > R * R) continue;
            }
            if (mpr(e, make_box(bx, bx[3]), 1e-6, nullptr)) { flag = true; break; }   // MS:451-462
        }
        hit[i] = flag ? 1 : 0;
        if (band) {
            const Frame f = frame_of(e);
            double lo_min = INFINITY;
            bool deep = false;
            std::vector<int64_t> near;
            for (int64_t b = 0; b < nb && !deep; ++b) {
                double lo, hi;
                gap_bounds(f, boxes + 6 * b, boxes[6 * b + 3], &lo, &hi);
                if (hi < -band_tol) deep = true;
                else if (lo <= band_tol) near.push_back(b);
                lo_min = std::min(lo_min, lo);
            }
            uint8_t bd = 0;
            if (!deep) {
                double g = INFINITY;
                for (int64_t b : near) {
                    g = std::min(g, gap_edge_box(f, boxes + 6 * b, boxes[6 * b + 3]));
                    if (g < -band_tol) break;
                }
                bd = std::fabs(g) <= band_tol ? 1 : 0;
            }
            band[i] = bd;
        }
    }
    return 0;
}

}  // extern "C"
