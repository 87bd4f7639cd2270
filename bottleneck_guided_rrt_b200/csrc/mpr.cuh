// Device narrow phase: FCL-0.5 / libccd-2.0 MPR boolean for {Sphere, Cylinder, Capsule} vs an
// axis-aligned leaf box, in double, with the operation order of the published algorithm
// (ccdMPRPenetration = discoverPortal + refinePortal, mpr_tolerance 1e-6; FCL support
// functions supportCyl / supportCap / supportBox / supportSphere).  Reference call sites:
// fcl::collide at MS:331 (sphere) and MS:451 (cylinder).  The translation unit is compiled
// with -fmad=false: libccd is plain C built without FMA contraction, and the boolean near
// the tolerance band depends on every rounding.
// The leaf box has identity rotation, for which ccdQuatRotVec is exactly the identity, so
// the box rotations are elided.
#pragma once
#include <cfloat>
#include "common.cuh"

namespace bgr {

constexpr double kCcdEps = DBL_EPSILON;
constexpr int kMprIterCap = 512;   // libccd's loops are unbounded; cap shared with the oracle

struct D3 {
    double x, y, z;
};
__device__ __forceinline__ bool ccd_is_zero(double v) { return fabs(v) < kCcdEps; }
__device__ __forceinline__ double ccd_sign(double v) { return ccd_is_zero(v) ? 0.0 : (v < 0 ? -1.0 : 1.0); }
__device__ __forceinline__ bool ccd_eq(double a, double b) {
    double ab = fabs(a - b);
    if (ab < kCcdEps) return true;
    a = fabs(a);
    b = fabs(b);
    return b > a ? ab < kCcdEps * b : ab < kCcdEps * a;
}
__device__ __forceinline__ bool d3_eq0(const D3& a) { return ccd_eq(a.x, 0) && ccd_eq(a.y, 0) && ccd_eq(a.z, 0); }
__device__ __forceinline__ double d3_dot(const D3& a, const D3& b) {
    double d = a.x * b.x;
    d += a.y * b.y;
    d += a.z * b.z;
    return d;
}
__device__ __forceinline__ D3 d3_sub(const D3& a, const D3& b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ D3 d3_cross(const D3& a, const D3& b) {
    return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
__device__ __forceinline__ void d3_scale(D3& a, double k) {
    a.x *= k;
    a.y *= k;
    a.z *= k;
}
__device__ __forceinline__ void d3_normalize(D3& a) { d3_scale(a, 1.0 / sqrt(d3_dot(a, a))); }

struct Quat {
    double x, y, z, w;
};
__device__ __forceinline__ void quat_rot(D3& v, const Quat& q) {   // ccdQuatRotVec
    const double vx = v.x, vy = v.y, vz = v.z;
    const double c1x = q.y * vz - q.z * vy + q.w * vx;
    const double c1y = q.z * vx - q.x * vz + q.w * vy;
    const double c1z = q.x * vy - q.y * vx + q.w * vz;
    const double c2x = q.y * c1z - q.z * c1y;
    const double c2y = q.z * c1x - q.x * c1z;
    const double c2z = q.x * c1y - q.y * c1x;
    v.x = vx + 2 * c2x;
    v.y = vy + 2 * c2y;
    v.z = vz + 2 * c2z;
}
__device__ __forceinline__ Quat quat_inv(const Quat& q) {          // ccdQuatInvert2
    double len2 = q.x * q.x;
    len2 += q.y * q.y;
    len2 += q.z * q.z;
    len2 += q.w * q.w;
    if (len2 < kCcdEps) return q;
    len2 = 1.0 / len2;
    return {-q.x * len2, -q.y * len2, -q.z * len2, q.w * len2};
}

enum { kShapeCyl = 0, kShapeCap = 1, kShapeSphere = 2 };

struct ShapeObj {
    D3 pos;
    Quat rot, rot_inv;
    double radius, height;   // height = half length
    int type;
};

__device__ __forceinline__ D3 support_shape(const ShapeObj& o, const D3& dir_) {
    D3 dir = dir_, v;
    quat_rot(dir, o.rot_inv);
    if (o.type == kShapeCyl) {
        double zdist = dir.x * dir.x + dir.y * dir.y;
        zdist = sqrt(zdist);
        if (ccd_is_zero(zdist)) v = {0., 0., ccd_sign(dir.z) * o.height};
        else {
            const double rad = o.radius / zdist;
            v = {rad * dir.x, rad * dir.y, ccd_sign(dir.z) * o.height};
        }
    } else if (o.type == kShapeCap) {
        D3 p1 = {0, 0, o.height}, p2 = {0, 0, -o.height};
        v = dir;
        d3_scale(v, o.radius);
        p1.x += v.x; p1.y += v.y; p1.z += v.z;
        p2.x += v.x; p2.y += v.y; p2.z += v.z;
        v = d3_dot(dir, p1) > d3_dot(dir, p2) ? p1 : p2;
    } else {
        v = dir;
        d3_scale(v, o.radius);
        d3_scale(v, 1.0 / sqrt(d3_dot(dir, dir)));
    }
    quat_rot(v, o.rot);
    v.x += o.pos.x;
    v.y += o.pos.y;
    v.z += o.pos.z;
    return v;
}

// Minkowski support shape (-) box, __ccdSupport: support1(dir) - support2(-dir)
__device__ __forceinline__ D3 mink_support(const ShapeObj& o, const D3& bc, double e, const D3& dir) {
    const D3 s1 = support_shape(o, dir);
    const D3 nd = {dir.x * -1.0, dir.y * -1.0, dir.z * -1.0};
    const D3 s2 = {ccd_sign(nd.x) * e + bc.x, ccd_sign(nd.y) * e + bc.y, ccd_sign(nd.z) * e + bc.z};
    return d3_sub(s1, s2);
}

// returns 1 = intersect, 0 = separate; *capped set when the iteration cap fired
static __device__ __noinline__ int mpr_shape_box(const ShapeObj& o, const D3& bc, double e, double tol, int* capped) {
    int it = 0;
    D3 v0 = d3_sub(o.pos, bc), v1, v2, v3, v4, dir, va, vb;
    if (d3_eq0(v0)) v0.x += kCcdEps * 10.0;
    dir = v0;
    d3_scale(dir, -1.0);
    d3_normalize(dir);
    v1 = mink_support(o, bc, e, dir); ++it;
    double d = d3_dot(v1, dir);
    if (ccd_is_zero(d) || d < 0) return 0;
    dir = d3_cross(v0, v1);
    if (ccd_is_zero(d3_dot(dir, dir))) return 1;   // origin on v1 or on the v0-v1 segment
    d3_normalize(dir);
    v2 = mink_support(o, bc, e, dir); ++it;
    d = d3_dot(v2, dir);
    if (ccd_is_zero(d) || d < 0) return 0;
    va = d3_sub(v1, v0);
    vb = d3_sub(v2, v0);
    dir = d3_cross(va, vb);
    d3_normalize(dir);
    d = d3_dot(dir, v0);
    if (d > 0) {
        const D3 t = v1; v1 = v2; v2 = t;
        d3_scale(dir, -1.0);
    }
    for (;;) {
        if (it >= kMprIterCap) { *capped = 1; return 0; }
        v3 = mink_support(o, bc, e, dir); ++it;
        d = d3_dot(v3, dir);
        if (ccd_is_zero(d) || d < 0) return 0;
        bool cont = false;
        va = d3_cross(v1, v3);
        d = d3_dot(va, v0);
        if (d < 0 && !ccd_is_zero(d)) { v2 = v3; cont = true; }
        if (!cont) {
            va = d3_cross(v3, v2);
            d = d3_dot(va, v0);
            if (d < 0 && !ccd_is_zero(d)) { v1 = v3; cont = true; }
        }
        if (!cont) break;
        va = d3_sub(v1, v0);
        vb = d3_sub(v2, v0);
        dir = d3_cross(va, vb);
        d3_normalize(dir);
    }
    for (;;) {
        va = d3_sub(v2, v1);
        vb = d3_sub(v3, v1);
        dir = d3_cross(va, vb);
        d3_normalize(dir);
        d = d3_dot(dir, v1);
        if (ccd_is_zero(d) || d > 0) return 1;                       // portalEncapsulesOrigin
        if (it >= kMprIterCap) { *capped = 1; return 0; }
        v4 = mink_support(o, bc, e, dir); ++it;
        const double d4 = d3_dot(v4, dir);
        if (!(ccd_is_zero(d4) || d4 > 0)) return 0;                  // portalCanEncapsuleOrigin
        {
            const double dv1 = d3_dot(v1, dir), dv2 = d3_dot(v2, dir), dv3 = d3_dot(v3, dir);
            const double m = fmin(fmin(d4 - dv1, d4 - dv2), d4 - dv3);
            if (ccd_eq(m, tol) || m < tol) return 0;                 // portalReachTolerance
        }
        {
            const D3 v4v0 = d3_cross(v4, v0);                        // expandPortal
            if (d3_dot(v1, v4v0) > 0) {
                if (d3_dot(v2, v4v0) > 0) v1 = v4; else v3 = v4;
            } else {
                if (d3_dot(v3, v4v0) > 0) v2 = v4; else v1 = v4;
            }
        }
    }
}

// The reference's f32 transform setup for the edge shape (MS:396-437): Eigen quaternion
// setFromTwoVectors((0,0,1), p2-p1) written out.
__device__ __forceinline__ ShapeObj make_edge_obj(const float p1[3], const float p2[3], double radius, int shape) {
    ShapeObj o;
    const float cx = __fdiv_rn(__fadd_rn(p1[0], p2[0]), 2.f), cy = __fdiv_rn(__fadd_rn(p1[1], p2[1]), 2.f),
                cz = __fdiv_rn(__fadd_rn(p1[2], p2[2]), 2.f);
    const float dx = __fsub_rn(p2[0], p1[0]), dy = __fsub_rn(p2[1], p1[1]), dz = __fsub_rn(p2[2], p1[2]);
    const double s2 = ((double)dz * (double)dz + (double)dy * (double)dy) + (double)dx * (double)dx;
    const float lz = __double2float_rn(sqrt(s2));
    float qx, qy, qz, qw;
    const float n2 = __fadd_rn(__fmul_rn(dx, dx), __fadd_rn(__fmul_rn(dy, dy), __fmul_rn(dz, dz)));
    if (!(n2 > 0.f)) {
        qx = qy = qz = 0.f;
        qw = 1.f;
    } else {
        const float nrm = __fsqrt_rn(n2);
        const float vx = __fdiv_rn(dx, nrm), vy = __fdiv_rn(dy, nrm), vz = __fdiv_rn(dz, nrm);
        if (vz < -1.f + 1e-5f) {
            // Eigen takes a JacobiSVD null-space axis here (perpendicular to both vectors, sign
            // unspecified); the right-handed one, normalize(-vy, vx, 0), reproduces the edge direction
            // (the shapes are symmetric about their axis).  Exactly antiparallel: (1,0,0).
            const float c2 = fmaxf(vz, -1.f);
            const float w2 = __fmul_rn(__fadd_rn(1.f, c2), 0.5f);
            const float sv = __fsqrt_rn(__fsub_rn(1.f, w2));
            const float an2 = __fadd_rn(__fmul_rn(vy, vy), __fmul_rn(vx, vx));
            float ax = 1.f, ay = 0.f;
            if (an2 > 0.f) {
                const float inva = __fdiv_rn(1.f, __fsqrt_rn(an2));
                ax = __fmul_rn(-vy, inva);
                ay = __fmul_rn(vx, inva);
            }
            qx = __fmul_rn(ax, sv);
            qy = __fmul_rn(ay, sv);
            qz = 0.f;
            qw = __fsqrt_rn(w2);
        } else {
            const float s = __fsqrt_rn(__fmul_rn(__fadd_rn(1.f, vz), 2.f));
            const float invs = __fdiv_rn(1.f, s);
            qx = __fmul_rn(-vy, invs);
            qy = __fmul_rn(vx, invs);
            qz = __fmul_rn(0.f, invs);
            qw = __fmul_rn(s, 0.5f);
        }
    }
    o.pos = {(double)cx, (double)cy, (double)cz};
    o.rot = {(double)qx, (double)qy, (double)qz, (double)qw};
    o.rot_inv = quat_inv(o.rot);
    o.radius = radius;
    o.height = (double)lz / 2;
    o.type = shape == BGR_SHAPE_CAPSULE ? kShapeCap : kShapeCyl;
    return o;
}

}  // namespace bgr
