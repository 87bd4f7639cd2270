"""ctypes/numpy binding of the CPU oracle (liboracle.so).  TEST INFRASTRUCTURE ONLY.

Every function here restates a reference function; see oracle/oracle.h and oracle/src/*.cpp
for the reference file:line each follows.  PARITY IS UNPINNED against the real
libnabo / FCL / libccd (absent here); MT19937 is pinned against the reference's mtrand.cpp.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

ALLOW_SELF_MATCH = 1
SORT_RESULTS = 2


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    if force or not os.path.exists(so):
        subprocess.check_call(["make", "-C", _HERE], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        vp, i64, i32, f32, f64, u32 = C.c_void_p, C.c_int64, C.c_int, C.c_float, C.c_double, C.c_uint32
        L.orc_mt_new_seed.restype = vp
        L.orc_mt_new_seed.argtypes = [u32]
        L.orc_mt_new_array.restype = vp
        L.orc_mt_new_array.argtypes = [vp, i32]
        L.orc_mt_free.argtypes = [vp]
        L.orc_mt_u32.restype = u32
        L.orc_mt_u32.argtypes = [vp]
        L.orc_mt_real.restype = f64
        L.orc_mt_real.argtypes = [vp]
        L.orc_mt_std_uniform.restype = f64
        L.orc_mt_std_uniform.argtypes = [vp, f64, f64]
        L.orc_mt_fill_u32.argtypes = [vp, vp, i64]
        L.orc_mt_fill_real.argtypes = [vp, vp, i64]
        L.orc_knn_brute.argtypes = [vp, i64, vp, i64, i32, C.c_uint, f32, vp, vp, vp, i32]
        L.orc_nabo_create.restype = vp
        L.orc_nabo_create.argtypes = [vp, i64, i32]
        L.orc_nabo_destroy.argtypes = [vp]
        L.orc_nabo_knn.restype = i64
        L.orc_nabo_knn.argtypes = [vp, vp, i64, i32, f32, C.c_uint, f32, vp, vp, i32]
        L.orc_nearest.argtypes = [vp, i64, vp, i64, vp, vp, i32]
        L.orc_norm.restype = f32
        L.orc_norm.argtypes = [vp, vp]
        L.orc_steer.argtypes = [vp, vp, f32, vp]
        L.orc_collide_spheres.argtypes = [vp, i64, vp, i64, f64, i32, f64, vp, vp, i32]
        L.orc_collide_edges.argtypes = [vp, i64, vp, vp, i64, f64, i32, i32, f64, vp, vp, i32]
        L.orc_mpr_sphere_box.argtypes = [vp, f64, vp, f64, vp]
        L.orc_mpr_cyl_box.argtypes = [vp, vp, f64, i32, vp, f64, vp]
        L.orc_sep_sphere_box.restype = f64
        L.orc_sep_sphere_box.argtypes = [vp, f64, vp, f64]
        L.orc_sep_cyl_box.restype = f64
        L.orc_sep_cyl_box.argtypes = [vp, vp, f64, f64, i32, vp, f64]
        L.orc_edge_frame.argtypes = [vp, vp, vp]
        L.orc_grid_to_boxes.restype = i64
        L.orc_grid_to_boxes.argtypes = [vp, i32, i32, i32, f64, vp, i32, vp, i64]
        L.orc_make_map.argtypes = [vp, i32, f64, u32, i32, i32, i32, f64]
        L.orc_planner_create.restype = vp
        L.orc_planner_create.argtypes = [vp, vp, i64, vp, i64, vp, i32]
        L.orc_planner_destroy.argtypes = [vp]
        L.orc_planner_set_collision_mode.argtypes = [vp, i32]
        L.orc_norm_pow_sqrt_mismatches.restype = C.c_longlong
        L.orc_planner_step.argtypes = [vp, vp, vp]
        L.orc_planner_run.argtypes = [vp, i64, i64]
        L.orc_planner_size.restype = i64
        L.orc_planner_size.argtypes = [vp]
        L.orc_planner_get_tree.argtypes = [vp, vp, vp]
        L.orc_planner_stats.argtypes = [vp, vp]
        _LIB = L
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


class MT:
    """MT19937 stream (reference mtrand.cpp semantics)."""

    def __init__(self, seed=5489, key=None):
        L = lib()
        if key is not None:
            k = np.ascontiguousarray(key, dtype=np.uint32)
            self._h = L.orc_mt_new_array(_p(k), len(k))
        else:
            self._h = L.orc_mt_new_seed(int(seed) & 0xFFFFFFFF)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_mt_free(self._h)
            self._h = None

    def u32(self, n=None):
        if n is None:
            return lib().orc_mt_u32(self._h)
        out = np.empty(n, dtype=np.uint32)
        lib().orc_mt_fill_u32(self._h, _p(out), n)
        return out

    def real(self, n=None):
        if n is None:
            return lib().orc_mt_real(self._h)
        out = np.empty(n, dtype=np.float64)
        lib().orc_mt_fill_real(self._h, _p(out), n)
        return out

    def std_uniform(self, a, b):
        return lib().orc_mt_std_uniform(self._h, a, b)


def points(seed, n, scale):
    """MTRand [0,1) double -> * scale -> f32, xyz interleaved (n,3)."""
    return (MT(seed).real(3 * n) * scale).astype(np.float32).reshape(n, 3)


def knn_brute(cloud, q, k, flags=0, max_radius=np.inf, max_radii=None, threads=0):
    """cloud (N,3) f32, q (Q,3) f32 -> idx (Q,k) i32, d2 (Q,k) f32 (libnabo kxQ col-major)."""
    cloud, q = _f32(cloud), _f32(q)
    idx = np.empty((q.shape[0], k), dtype=np.int32)
    d2 = np.empty((q.shape[0], k), dtype=np.float32)
    mr = _f32(max_radii) if max_radii is not None else None
    rc = lib().orc_knn_brute(_p(cloud), cloud.shape[0], _p(q), q.shape[0], k, flags,
                             np.float32(max_radius), _p(mr), _p(idx), _p(d2), threads)
    if rc:
        raise ValueError("orc_knn_brute rc=%d" % rc)
    return idx, d2


class NaboKDTree:
    def __init__(self, cloud, bucket_size=8):
        self.cloud = _f32(cloud)
        self._h = lib().orc_nabo_create(_p(self.cloud), self.cloud.shape[0], bucket_size)
        if not self._h:
            raise ValueError("orc_nabo_create failed")

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_nabo_destroy(self._h)
            self._h = None

    def knn(self, q, k, eps=0.0, flags=0, max_radius=np.inf, threads=0):
        q = _f32(q)
        idx = np.empty((q.shape[0], k), dtype=np.int32)
        d2 = np.empty((q.shape[0], k), dtype=np.float32)
        t = lib().orc_nabo_knn(self._h, _p(q), q.shape[0], k, np.float32(eps), flags,
                               np.float32(max_radius), _p(idx), _p(d2), threads)
        if t < 0:
            raise RuntimeError("libnabo would throw: rc=%d" % t)
        self.touched = t
        return idx, d2


def nearest(V, q, threads=0):
    V, q = _f32(V), _f32(q)
    idx = np.empty(q.shape[0], dtype=np.int32)
    dist = np.empty(q.shape[0], dtype=np.float32)
    rc = lib().orc_nearest(_p(V), V.shape[0], _p(q), q.shape[0], _p(idx), _p(dist), threads)
    if rc:
        raise ValueError("orc_nearest rc=%d" % rc)
    return idx, dist


def norm(a, b):
    a, b = _f32(a), _f32(b)
    return float(lib().orc_norm(_p(a), _p(b)))


def pow_sqrt_mismatches():
    """calls of the oracle's Norm / steer / edge length in this process where float(pow(s, 0.5)) (the
    reference's form) differed from float(sqrt(s)) (correctly rounded, what the CUDA path computes)"""
    return int(lib().orc_norm_pow_sqrt_mismatches())


def steer(nearest_pt, rnd, eta=0.15):
    a, b = _f32(nearest_pt), _f32(rnd)
    out = np.empty(3, dtype=np.float32)
    kept = lib().orc_steer(_p(a), _p(b), np.float32(eta), _p(out))
    return out, bool(kept)


def collide_spheres(boxes, centers, radius=0.1, mode=1, band_tol=1e-6, want_band=True, threads=0):
    boxes = np.ascontiguousarray(boxes, dtype=np.float64).reshape(-1, 6)
    c = _f32(centers).reshape(-1, 3)
    hit = np.zeros(c.shape[0], dtype=np.uint8)
    band = np.zeros(c.shape[0], dtype=np.uint8) if want_band else None
    lib().orc_collide_spheres(_p(boxes), boxes.shape[0], _p(c), c.shape[0], radius, mode,
                              band_tol, _p(hit), _p(band), threads)
    return hit, band


def collide_edges(boxes, p1, p2, radius=0.03, shape=0, mode=1, band_tol=1e-6, want_band=True, threads=0):
    boxes = np.ascontiguousarray(boxes, dtype=np.float64).reshape(-1, 6)
    p1, p2 = _f32(p1).reshape(-1, 3), _f32(p2).reshape(-1, 3)
    hit = np.zeros(p1.shape[0], dtype=np.uint8)
    band = np.zeros(p1.shape[0], dtype=np.uint8) if want_band else None
    lib().orc_collide_edges(_p(boxes), boxes.shape[0], _p(p1), _p(p2), p1.shape[0], radius,
                            shape, mode, band_tol, _p(hit), _p(band), threads)
    return hit, band


def mpr_sphere_box(c, r, box_c, box_size):
    c = np.ascontiguousarray(c, dtype=np.float64)
    b = np.ascontiguousarray(box_c, dtype=np.float64)
    it = C.c_int(0)
    h = lib().orc_mpr_sphere_box(_p(c), r, _p(b), box_size, C.addressof(it))
    return bool(h), it.value


def mpr_cyl_box(p1, p2, r, box_c, box_size, shape=0):
    p1, p2 = _f32(p1), _f32(p2)
    b = np.ascontiguousarray(box_c, dtype=np.float64)
    it = C.c_int(0)
    h = lib().orc_mpr_cyl_box(_p(p1), _p(p2), r, shape, _p(b), box_size, C.addressof(it))
    return bool(h), it.value


def sep_sphere_box(c, r, box_c, box_size):
    c = np.ascontiguousarray(c, dtype=np.float64)
    b = np.ascontiguousarray(box_c, dtype=np.float64)
    return lib().orc_sep_sphere_box(_p(c), r, _p(b), box_size)


def sep_cyl_box(p1, p2, r, box_c, box_size, dr=0.0, shape=0):
    p1, p2 = _f32(p1), _f32(p2)
    b = np.ascontiguousarray(box_c, dtype=np.float64)
    return lib().orc_sep_cyl_box(_p(p1), _p(p2), r, dr, shape, _p(b), box_size)


def edge_frame(p1, p2):
    p1, p2 = _f32(p1), _f32(p2)
    out = np.empty(8, dtype=np.float64)
    lib().orc_edge_frame(_p(p1), _p(p2), _p(out))
    return out


def make_map(n, res=0.05, seed=1, n_walls=3, wall_thick=2, door_cells=6, occupancy=0.10):
    occ = np.zeros((n, n, n), dtype=np.uint8)  # [z][y][x]
    lib().orc_make_map(_p(occ), n, res, seed, n_walls, wall_thick, door_cells, occupancy)
    return occ


def grid_to_boxes(occ, res=0.05, origin=(0.0, 0.0, 0.0), max_level=2):
    occ = np.ascontiguousarray(occ, dtype=np.uint8)
    nz, ny, nx = occ.shape
    o = np.ascontiguousarray(origin, dtype=np.float64)
    n = lib().orc_grid_to_boxes(_p(occ), nx, ny, nz, res, _p(o), max_level, None, 0)
    boxes = np.empty((n, 6), dtype=np.float64)
    lib().orc_grid_to_boxes(_p(occ), nx, ny, nz, res, _p(o), max_level, _p(boxes), n)
    return boxes


class RefMT:
    """The REFERENCE's own mtrand.cpp (oracle/_ref/libmtrand_ref.so); one global stream."""

    def __init__(self):
        so = os.path.join(_HERE, "_ref", "libmtrand_ref.so")
        if not os.path.exists(so):
            raise FileNotFoundError(so)
        self.L = C.CDLL(so)
        self.L.ref_mt_u32.restype = C.c_uint32
        self.L.ref_mt_real.restype = C.c_double
        self.L.ref_mt_seed.argtypes = [C.c_uint32]
        self.L.ref_mt_seed_array.argtypes = [C.c_void_p, C.c_int]

    def seed(self, s):
        self.L.ref_mt_seed(s)

    def seed_array(self, key):
        k = np.ascontiguousarray(key, dtype=np.uint32)
        self.L.ref_mt_seed_array(_p(k), len(k))

    def u32(self):
        return self.L.ref_mt_u32()

    def real(self):
        return self.L.ref_mt_real()


class Planner:
    """CPU restatement of the planner loop (oracle/src/planner.cpp).  `params` is any ctypes
    struct with the orc_plan_params layout (the product's PlannerParams has it)."""

    def __init__(self, params, boxes, table, milestones=None, collision_mode=1):
        b = np.ascontiguousarray(boxes, dtype=np.float64).reshape(-1, 6)
        t = _f32(table).reshape(-1, 3)
        m = _f32(milestones).reshape(-1, 3) if milestones is not None else np.zeros((0, 3), np.float32)
        self._keep = (b, t, m)
        self._h = lib().orc_planner_create(C.byref(params), _p(b), b.shape[0], _p(t), t.shape[0], _p(m), m.shape[0])
        if not self._h:
            raise ValueError("orc_planner_create failed")
        lib().orc_planner_set_collision_mode(self._h, int(collision_mode))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_planner_destroy(self._h)
            self._h = None

    def __len__(self):
        return int(lib().orc_planner_size(self._h))

    def step(self):
        a, g = C.c_int(0), C.c_int(0)
        lib().orc_planner_step(self._h, C.byref(a), C.byref(g))
        return a.value, bool(g.value)

    def run(self, max_nodes, max_batches=1 << 40):
        return bool(lib().orc_planner_run(self._h, max_nodes, max_batches))

    def tree(self):
        n = len(self)
        xyz = np.empty((n, 3), dtype=np.float32)
        par = np.empty(n, dtype=np.int32)
        lib().orc_planner_get_tree(self._h, _p(xyz), _p(par))
        return xyz, par

    def stats(self):
        s = np.zeros(8, dtype=np.int64)
        lib().orc_planner_stats(self._h, _p(s))
        return dict(zip(("batches", "candidates", "rejected_state", "rejected_edge", "improved_parent", "rewired",
                         "in_band", "goal_candidate"), (int(v) for v in s)))
