"""ctypes binding of libbgrrt.so + host-side mirror of the reference call surface.

``NNSearchF`` mirrors libnabo's ``Nabo::NNSearchF`` as the reference uses it
(rrt_star_milestones.cpp:1306-1315): ``createKDTreeLinearHeap(cloud)`` / ``knn(query,
indices, dists2, k, eps, flags, maxRadius)`` with libnabo's argument meaning and error
behaviour (``RuntimeError`` where libnabo throws ``std::runtime_error``).
``OccupancyMap`` mirrors ``isStateValid`` (:310-351) / ``isLineValid`` (:385-471): True means
IN COLLISION, as in the reference.
"""
import ctypes as C
import os

import numpy as np

ALLOW_SELF_MATCH = 1
SORT_RESULTS = 2
KNN_STATIC_MAX_RESULTS = 32768   # BGR_KNN_STATIC_MAX_RESULTS (bgr_nns_knn_scan_static_dev)
TOUCH_STATISTICS = 1
METHOD_AUTO, METHOD_BRUTE, METHOD_GRID, METHOD_GRID_GENERAL = 0, 1, 2, 3
SHAPE_CYLINDER, SHAPE_CAPSULE = 0, 1

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class BgrError(RuntimeError):
    def __init__(self, rc, msg):
        super().__init__("libbgrrt rc=%d: %s" % (rc, msg))
        self.rc = rc


def library_path():
    return os.path.join(_HERE, "libbgrrt.so")


def lib():
    """Load libbgrrt.so.  Raises if the CUDA extension is missing (no fallback)."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = library_path()
    if not os.path.exists(path):
        raise ImportError(
            "libbgrrt.so not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C bottleneck_guided_rrt_b200/csrc` (there is no CPU fallback)")
    L = C.CDLL(path)
    vp, i64, i32, f32, f64, u32 = C.c_void_p, C.c_int64, C.c_int, C.c_float, C.c_double, C.c_uint
    L.bgr_last_error.restype = C.c_char_p
    L.bgr_version.restype = i32
    L.bgr_launch_count.restype = C.c_uint64
    L.bgr_device_count.argtypes = [vp]
    L.bgr_device_info.argtypes = [vp, vp, vp]
    L.bgr_microbench_l2.argtypes = [i64, i32, i32, vp]
    L.bgr_nns_create.argtypes = [vp, i64, i32, u32, vp]
    L.bgr_nns_create_dev.argtypes = [vp, i64, i32, u32, vp, vp]
    L.bgr_nns_append.argtypes = [vp, vp, i64]
    L.bgr_nns_append_dev.argtypes = [vp, vp, i64]
    L.bgr_nns_size.restype = i64
    L.bgr_nns_size.argtypes = [vp]
    L.bgr_nns_set_stream.argtypes = [vp, vp]
    L.bgr_nns_set_method.argtypes = [vp, i32]
    L.bgr_nns_set_index_base.argtypes = [vp, i64]
    L.bgr_nns_build_index.argtypes = [vp]
    L.bgr_nns_knn.argtypes = [vp, vp, i64, i32, f32, u32, f32, vp, vp, vp, vp]
    L.bgr_nns_knn_dev.argtypes = [vp, vp, i64, i32, f32, u32, f32, vp, vp, vp, vp]
    L.bgr_nns_nearest.argtypes = [vp, vp, i64, vp, vp]
    L.bgr_nns_nearest_dev.argtypes = [vp, vp, i64, vp, vp]
    L.bgr_nns_debug_stats.argtypes = [vp, vp]
    L.bgr_nns_knn_skip_dev.argtypes = [vp, vp, i64, i32, C.c_uint, vp, vp, vp]
    L.bgr_nns_knn_scan_static_dev.argtypes = [vp, vp, i64, i32, C.c_uint, f32, vp, vp, vp]
    L.bgr_nns_set_tile.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int]
    L.bgr_nns_tile_stats.argtypes = [vp, vp]
    L.bgr_nns_destroy.argtypes = [vp]
    L.bgr_knn_merge_dev.argtypes = [vp, vp, i32, i64, i32, vp, vp, vp]
    L.bgr_knn_merge_strided_dev.argtypes = [vp, vp, i32, i64, i64, i32, vp, vp, vp]
    L.bgr_peer_handle_bytes.restype = i32
    L.bgr_peer_create.argtypes = [i32, i32, i64, vp]
    L.bgr_peer_export.argtypes = [vp, vp]
    L.bgr_peer_open.argtypes = [vp, i32, vp]
    L.bgr_peer_slab.argtypes = [vp, vp, vp]
    L.bgr_peer_result.argtypes = [vp, vp, vp]
    L.bgr_peer_merge_dev.argtypes = [vp, i64, i32, vp]
    L.bgr_peer_status.argtypes = [vp, vp, vp]
    L.bgr_peer_destroy.argtypes = [vp]
    L.bgr_map_create.argtypes = [f64, vp, vp, vp]
    L.bgr_map_set_stream.argtypes = [vp, vp]
    L.bgr_map_set_boxes.argtypes = [vp, vp, i64]
    L.bgr_map_update_boxes.argtypes = [vp, vp, i64, vp, i64]
    L.bgr_map_occupied_cells.argtypes = [vp, vp]
    L.bgr_map_download_bitmap.argtypes = [vp, vp, vp]
    L.bgr_map_collide_spheres.argtypes = [vp, vp, i64, f64, vp, vp]
    L.bgr_map_collide_spheres_dev.argtypes = [vp, vp, i64, f64, vp, vp]
    L.bgr_map_collide_edges.argtypes = [vp, vp, vp, i64, f64, i32, vp, vp]
    L.bgr_map_collide_edges_dev.argtypes = [vp, vp, vp, i64, f64, i32, vp, vp]
    L.bgr_map_last_stats.argtypes = [vp, vp]
    L.bgr_map_set_timing.argtypes = [vp, i32]
    L.bgr_map_set_gate.argtypes = [vp, i32]
    L.bgr_map_last_gate_stats.argtypes = [vp, vp]
    L.bgr_map_dilation_stats.argtypes = [vp, vp]
    L.bgr_map_download_gate.argtypes = [vp, i32, vp, vp]
    L.bgr_map_last_kernel_ms.argtypes = [vp, vp]
    L.bgr_map_destroy.argtypes = [vp]
    L.bgr_map_refresh_boxes.argtypes = [vp, vp, i64, vp, vp]
    L.bgr_octree_to_boxes.argtypes = [vp, i64, i32, f64, f64, vp, vp, vp, vp, vp]
    L.bgr_map_prune_path.argtypes = [vp, vp, i64, f64, i32, vp, vp, vp]
    L.bgr_path_smoothen.argtypes = [vp, i64, vp, vp]
    L.bgr_map_clearance.argtypes = [vp, vp, i64, f64, vp]
    L.bgr_map_find_labels.argtypes = [vp, vp, i64, f64, i32, vp, vp, vp, vp]
    L.bgr_planner_create.argtypes = [vp, vp, vp, i64, vp, i32, vp]
    L.bgr_planner_set_stream.argtypes = [vp, vp]
    L.bgr_planner_step.argtypes = [vp, vp, vp]
    L.bgr_planner_run_many.argtypes = [vp, i32, i64, i64, vp]
    L.bgr_planner_run.argtypes = [vp, i64, i64, vp]
    L.bgr_planner_size.restype = i64
    L.bgr_planner_size.argtypes = [vp]
    L.bgr_planner_get_tree.argtypes = [vp, vp, vp]
    L.bgr_planner_get_path.argtypes = [vp, vp, vp]
    L.bgr_planner_stats.argtypes = [vp, vp]
    L.bgr_planner_persist_stats.argtypes = [vp, vp]
    L.bgr_planner_reset.argtypes = [vp, vp, vp, u32]
    L.bgr_planner_run_queue.argtypes = [vp, i32, vp, vp, vp, i64, i64, i64, vp, vp]
    L.bgr_planner_nns.argtypes = [vp, vp]
    L.bgr_planner_destroy.argtypes = [vp]
    _LIB = L
    return L


def microbench_l2(nbytes, mode, reps=10):
    """bgr_microbench_l2: measured L2 read bandwidth (GB/s) of the current device"""
    v = C.c_double(0.0)
    rc = lib().bgr_microbench_l2(int(nbytes), int(mode), int(reps), C.byref(v))
    if rc != 0:
        raise BgrError(rc, lib().bgr_last_error().decode("utf-8", "replace"))
    return v.value


PLAN_EDGE_CHECK, PLAN_REWIRE, PLAN_QUIRKS, PLAN_NO_GOAL, PLAN_DEDUP, PLAN_NO_PERSIST, PLAN_NO_GRAPH = 1, 2, 4, 8, 16, 32, 64
PLAN_MODE_RS, PLAN_MODE_MS = 0, 1


class PlannerParams(C.Structure):
    """bgr_planner_params (include/bgrrt.h).
    Defaults are the reference's compile-time constants."""
    _fields_ = [("start", C.c_float * 3), ("goal", C.c_float * 3), ("eta", C.c_float), ("goal_tol", C.c_float),
                ("goal_chance", C.c_float), ("mode", C.c_int32), ("k", C.c_int32), ("batch", C.c_int32),
                ("edge_shape", C.c_int32), ("flags", C.c_uint32), ("seed", C.c_uint32), ("pad_", C.c_uint32),
                ("state_radius", C.c_double), ("edge_radius", C.c_double)]

    def __init__(self, start, goal, mode=PLAN_MODE_MS, k=5, batch=1, flags=PLAN_QUIRKS, seed=5489, eta=0.15,
                 goal_tol=0.4, goal_chance=0.1, state_radius=0.1, edge_radius=0.03, edge_shape=SHAPE_CYLINDER):
        super().__init__()
        self.start[:] = [float(v) for v in start]
        self.goal[:] = [float(v) for v in goal]
        self.eta, self.goal_tol, self.goal_chance = eta, goal_tol, goal_chance
        self.mode, self.k, self.batch, self.edge_shape = mode, k, batch, edge_shape
        self.flags, self.seed = flags, seed
        self.state_radius, self.edge_radius = state_radius, edge_radius


def _check(rc):
    if rc != 0:
        raise BgrError(rc, lib().bgr_last_error().decode("utf-8", "replace"))


def _ptr(a):
    """host numpy array / raw int device pointer / torch tensor -> void*"""
    if a is None:
        return None
    if isinstance(a, int):
        return C.c_void_p(a)
    if isinstance(a, np.ndarray):
        return a.ctypes.data_as(C.c_void_p)
    if hasattr(a, "data_ptr"):
        return C.c_void_p(a.data_ptr())
    raise TypeError(type(a))


def _host_f32(a, cols=3):
    a = np.ascontiguousarray(a, dtype=np.float32)
    if a.ndim == 1:
        a = a.reshape(-1, cols)
    return a


def _is_dev(a):
    return hasattr(a, "data_ptr") and getattr(a, "is_cuda", False)


class NNSearchF:
    """libnabo-shaped search object over the append-only HBM node store.

    Points are passed as (N, 3) float32 arrays — the memory image of libnabo's 3xN
    column-major ``Eigen::MatrixXf`` (one point per column).
    """

    ALLOW_SELF_MATCH = ALLOW_SELF_MATCH
    SORT_RESULTS = SORT_RESULTS
    TOUCH_STATISTICS = TOUCH_STATISTICS

    def __init__(self, handle):
        self._h = handle

    # Nabo::NNSearchF::createKDTreeLinearHeap(cloud, dim, creationOptionFlags)
    @classmethod
    def createKDTreeLinearHeap(cls, cloud, dim=None, creationOptionFlags=0, stream=None):
        h = C.c_void_p()
        if _is_dev(cloud):
            n = cloud.shape[0]
            _check(lib().bgr_nns_create_dev(_ptr(cloud), n, dim or 3, creationOptionFlags,
                                            C.c_void_p(stream or 0), C.byref(h)))
        else:
            cloud = _host_f32(cloud)
            _check(lib().bgr_nns_create(_ptr(cloud), cloud.shape[0], dim or cloud.shape[1],
                                        creationOptionFlags, C.byref(h)))
            if stream:
                _check(lib().bgr_nns_set_stream(h, C.c_void_p(stream)))
        return cls(h)

    create = createKDTreeLinearHeap

    def __del__(self):
        self.close()

    def close(self):
        if getattr(self, "_h", None):
            lib().bgr_nns_destroy(self._h)
            self._h = None

    def __len__(self):
        return int(lib().bgr_nns_size(self._h))

    def append(self, pts):
        if _is_dev(pts):
            _check(lib().bgr_nns_append_dev(self._h, _ptr(pts), pts.shape[0]))
        else:
            pts = _host_f32(pts)
            _check(lib().bgr_nns_append(self._h, _ptr(pts), pts.shape[0]))

    def set_stream(self, stream):
        _check(lib().bgr_nns_set_stream(self._h, C.c_void_p(stream or 0)))

    def set_method(self, method):
        _check(lib().bgr_nns_set_method(self._h, method))

    def set_index_base(self, base):
        _check(lib().bgr_nns_set_index_base(self._h, base))

    def build_index(self):
        _check(lib().bgr_nns_build_index(self._h))

    def knn_skip(self, d_query, k, d_skip, d_indices, d_dists2, optionFlags=0):
        """bgr_nns_knn_skip_dev: CUDA tensors only; rows of queries with a non-zero skip byte are left untouched"""
        _check(lib().bgr_nns_knn_skip_dev(self._h, _ptr(d_query), d_query.shape[0], k, optionFlags, _ptr(d_skip),
                                          _ptr(d_indices), _ptr(d_dists2)))

    def knn_scan_static(self, d_query, k, d_indices, d_dists2, optionFlags=0, maxRadius=float("inf")):
        """bgr_nns_knn_scan_static_dev: the allocation-free plain scan (CUDA tensors only; nq * k <=
        KNN_STATIC_MAX_RESULTS; partial lists in the library's static device array).  Returns the number of
        node chunks the scan used (1 = the static scratch was not needed)."""
        chunks = C.c_int(0)
        _check(lib().bgr_nns_knn_scan_static_dev(self._h, _ptr(d_query), d_query.shape[0], k, optionFlags, maxRadius,
                                                 _ptr(d_indices), _ptr(d_dists2), C.byref(chunks)))
        return chunks.value

    def set_tile(self, tx=0, ty=0, tz=0, warps=0):
        """tile kNN tuning (bgr_nns_set_tile): all 0 = automatic, tx < 0 = tile search off"""
        _check(lib().bgr_nns_set_tile(self._h, tx, ty, tz, warps))

    def tile_stats(self):
        """bgr_nns_tile_stats: tile searches so far, queries the last one handed back, its tile shape / count"""
        s = np.zeros(4, dtype=np.uint64)
        _check(lib().bgr_nns_tile_stats(self._h, _ptr(s)))
        sh = int(s[2])
        return {"tile_searches": int(s[0]), "fallback_queries_last": int(s[1]),
                "tile": (sh & 255, (sh >> 8) & 255, (sh >> 16) & 255), "warps": sh >> 24, "tiles": int(s[3])}

    def knn(self, query, k=1, epsilon=0.0, optionFlags=0, maxRadius=float("inf"), maxRadii=None,
            indices=None, dists2=None, want_dists2=True):
        """Returns (indices (Q,k) int32, dists2 (Q,k) float32, touched).  Host arrays in,
        host arrays out; CUDA tensors in, CUDA tensors out (asynchronous on the stream)."""
        touched = C.c_uint64(0)
        if _is_dev(query):
            import torch
            nq = query.shape[0]
            if indices is None:
                indices = torch.empty((nq, k), dtype=torch.int32, device=query.device)
            if dists2 is None:
                dists2 = torch.empty((nq, k), dtype=torch.float32, device=query.device)
            if maxRadii is not None and not _is_dev(maxRadii):      # per-query radii must live where the queries do
                maxRadii = torch.from_numpy(np.ascontiguousarray(maxRadii, dtype=np.float32)).to(query.device)
            _check(lib().bgr_nns_knn_dev(self._h, _ptr(query), nq, k, epsilon, optionFlags, maxRadius,
                                         _ptr(maxRadii), _ptr(indices), _ptr(dists2), C.byref(touched)))
            return indices, dists2, touched.value
        q = _host_f32(query)
        nq = q.shape[0]
        if indices is None:
            indices = np.empty((nq, k), dtype=np.int32)
        if dists2 is None and want_dists2:
            dists2 = np.empty((nq, k), dtype=np.float32)
        # libnabo checkSizesKnn: result matrices must be k x Q
        if indices.shape != (nq, k) or (dists2 is not None and dists2.shape != (nq, k)):
            raise RuntimeError("Index/distance matrix has wrong shape: expected %dx%d" % (k, nq))
        mr = None
        if maxRadii is not None:
            mr = np.ascontiguousarray(maxRadii, dtype=np.float32)
            if mr.shape[0] != nq:
                raise RuntimeError("Maximum radii vector has %d rows != %d queries" % (mr.shape[0], nq))
        _check(lib().bgr_nns_knn(self._h, _ptr(q), nq, k, epsilon, optionFlags, maxRadius, _ptr(mr),
                                 _ptr(indices), _ptr(dists2), C.byref(touched)))
        return indices, dists2, touched.value

    def debug_stats(self):
        """which code paths this handle has run (bgr_nns_debug_stats): test / measurement aid"""
        return _nns_debug(self._h)

    def nearest(self, query, idx=None, dist=None):
        """Batched reference Nearest(): returns (index (Q,) int32, Norm (Q,) float32)."""
        if _is_dev(query):
            import torch
            nq = query.shape[0]
            if idx is None:
                idx = torch.empty(nq, dtype=torch.int32, device=query.device)
            if dist is None:
                dist = torch.empty(nq, dtype=torch.float32, device=query.device)
            _check(lib().bgr_nns_nearest_dev(self._h, _ptr(query), nq, _ptr(idx), _ptr(dist)))
            return idx, dist
        q = _host_f32(query)
        nq = q.shape[0]
        idx = np.empty(nq, dtype=np.int32) if idx is None else idx
        dist = np.empty(nq, dtype=np.float32) if dist is None else dist
        _check(lib().bgr_nns_nearest(self._h, _ptr(q), nq, _ptr(idx), _ptr(dist)))
        return idx, dist


_NNS_DEBUG_KEYS = ("index_builds", "grid_knn_launches", "grid_knn_piled_launches", "heavy_queries_last_launch",
                   "grid_nearest_launches", "scan_nearest_launches", "brute_knn_launches", "nodes_indexed")


def _nns_debug(handle):
    s = np.zeros(8, dtype=np.uint64)
    _check(lib().bgr_nns_debug_stats(handle, _ptr(s)))
    return dict(zip(_NNS_DEBUG_KEYS, (int(v) for v in s)))


def knn_merge_dev(idx_slabs, d2_slabs, idx_out, d2_out, stream=0):
    """Merge G per-shard (Q,k) top-k slabs (CUDA tensors (G,Q,k), each slab contiguous, any
    stride between slabs) into one, by (d2, index)."""
    g, nq, k = idx_slabs.shape
    assert idx_slabs.stride(1) == k and idx_slabs.stride(2) == 1 and idx_slabs.stride() == d2_slabs.stride()
    _check(lib().bgr_knn_merge_strided_dev(_ptr(idx_slabs), _ptr(d2_slabs), g, idx_slabs.stride(0), nq, k,
                                           _ptr(idx_out), _ptr(d2_out), C.c_void_p(stream or 0)))
    return idx_out, d2_out


class _DevView:
    """raw device memory as a __cuda_array_interface__ object (torch.as_tensor wraps it, no copy)"""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 2, "strides": None}


class PeerExchange:
    """bgr_peer_*: the sharded search's exchange step over NVLink peer memory (CUDA IPC windows,
    one merge kernel with peer loads/stores instead of all-gather + merge).  `exchange_handles`
    is a callable that takes this rank's handle (bytes) and returns the list of all ranks'
    handles (e.g. torch.distributed.all_gather_object)."""

    def __init__(self, rank, world, max_entries, exchange_handles, device=None):
        h = C.c_void_p()
        _check(lib().bgr_peer_create(rank, world, max_entries, C.byref(h)))
        self._h, self.rank, self.world, self.device = h, rank, world, device
        nb = lib().bgr_peer_handle_bytes()
        buf = C.create_string_buffer(nb)
        _check(lib().bgr_peer_export(self._h, buf))
        handles = exchange_handles(bytes(buf.raw))
        for p, hb in enumerate(handles):
            if p != rank:
                _check(lib().bgr_peer_open(self._h, p, C.create_string_buffer(hb, nb)))

    def __del__(self):
        self.close()

    def close(self):
        if getattr(self, "_h", None):
            lib().bgr_peer_destroy(self._h)
            self._h = None

    def _region(self, fn):
        pi, pd = C.c_void_p(), C.c_void_p()
        _check(fn(self._h, C.byref(pi), C.byref(pd)))
        return pi.value, pd.value

    def slab_ptrs(self):
        return self._region(lib().bgr_peer_slab)

    def result(self, nq, k):
        """(idx (Q,k) int32, d2 (Q,k) float32) CUDA tensors viewing the window's result region"""
        import torch
        pi, pd = self._region(lib().bgr_peer_result)
        dev = self.device if self.device is not None else torch.device("cuda", torch.cuda.current_device())
        return (torch.as_tensor(_DevView(pi, (nq, k), "<i4"), device=dev),
                torch.as_tensor(_DevView(pd, (nq, k), "<f4"), device=dev))

    def merge(self, nq, k, stream=0):
        _check(lib().bgr_peer_merge_dev(self._h, nq, k, C.c_void_p(stream or 0)))

    def status(self, stream=0):
        e = C.c_int(0)
        _check(lib().bgr_peer_status(self._h, C.c_void_p(stream or 0), C.byref(e)))
        return e.value


class OccupancyMap:
    """Octomap occupied leaves flattened to HBM bit grids; isStateValid / isLineValid batched."""

    def __init__(self, res, origin, dims, stream=None):
        o = np.ascontiguousarray(origin, dtype=np.float64)
        d = np.ascontiguousarray(dims, dtype=np.int32)
        h = C.c_void_p()
        _check(lib().bgr_map_create(res, _ptr(o), _ptr(d), C.byref(h)))
        self._h = h
        self.res, self.origin, self.dims = res, o, d
        if stream:
            self.set_stream(stream)

    def __del__(self):
        self.close()

    def close(self):
        if getattr(self, "_h", None):
            lib().bgr_map_destroy(self._h)
            self._h = None

    def set_stream(self, stream):
        _check(lib().bgr_map_set_stream(self._h, C.c_void_p(stream or 0)))

    def set_boxes(self, boxes):
        b = np.ascontiguousarray(boxes, dtype=np.float64).reshape(-1, 6)
        _check(lib().bgr_map_set_boxes(self._h, _ptr(b), b.shape[0]))

    def update_boxes(self, add=None, remove=None):
        a = np.ascontiguousarray(add if add is not None else np.zeros((0, 6)), dtype=np.float64).reshape(-1, 6)
        r = np.ascontiguousarray(remove if remove is not None else np.zeros((0, 6)), dtype=np.float64).reshape(-1, 6)
        _check(lib().bgr_map_update_boxes(self._h, _ptr(a), a.shape[0], _ptr(r), r.shape[0]))

    def refresh_boxes(self, boxes):
        """map refresh: diff against the current leaf list; returns (n_added, n_removed)"""
        b = np.ascontiguousarray(boxes, dtype=np.float64).reshape(-1, 6)
        a, r = C.c_int64(0), C.c_int64(0)
        _check(lib().bgr_map_refresh_boxes(self._h, _ptr(b), b.shape[0], C.byref(a), C.byref(r)))
        return a.value, r.value

    @classmethod
    def from_octomap(cls, data, fmt=2, res=0.0, occ_threshold=0.5, stream=None):
        """OccupancyMap over the bounding box of a serialised Octomap's occupied leaves
        (origin snapped down to a multiple of 32 cells from the key origin).  Returns (map, boxes)."""
        boxes, res, lo, hi = octree_to_boxes(data, fmt, res, occ_threshold)
        blk = 32 * res
        lo_c = np.floor(np.asarray(lo) / blk + 1e-9).astype(np.int64) * 32
        hi_c = np.ceil(np.asarray(hi) / res - 1e-9).astype(np.int64)
        dims = np.maximum(hi_c - lo_c, 1)
        m = cls(res, lo_c * res, dims, stream=stream)
        m.set_boxes(boxes)
        return m, boxes

    def occupied_cells(self):
        c = C.c_int64(0)
        _check(lib().bgr_map_occupied_cells(self._h, C.byref(c)))
        return c.value

    def download_bitmap(self):
        n = C.c_int64(0)
        _check(lib().bgr_map_download_bitmap(self._h, None, C.byref(n)))
        out = np.empty(n.value, dtype=np.uint64)
        _check(lib().bgr_map_download_bitmap(self._h, _ptr(out), C.byref(n)))
        return out

    def collide_spheres(self, centers, radius=0.1, want_band=True, hit=None, band=None):
        if _is_dev(centers):
            import torch
            n = centers.shape[0]
            hit = torch.empty(n, dtype=torch.uint8, device=centers.device) if hit is None else hit
            if want_band and band is None:
                band = torch.empty(n, dtype=torch.uint8, device=centers.device)
            _check(lib().bgr_map_collide_spheres_dev(self._h, _ptr(centers), n, radius, _ptr(hit),
                                                     _ptr(band) if want_band else None))
            return hit, band
        c = _host_f32(centers)
        n = c.shape[0]
        hit = np.empty(n, dtype=np.uint8) if hit is None else hit
        if want_band and band is None:
            band = np.empty(n, dtype=np.uint8)
        _check(lib().bgr_map_collide_spheres(self._h, _ptr(c), n, radius, _ptr(hit), _ptr(band) if want_band else None))
        return hit, band

    def collide_edges(self, p1, p2, radius=0.03, shape=SHAPE_CYLINDER, want_band=True, hit=None, band=None):
        if _is_dev(p1):
            import torch
            n = p1.shape[0]
            hit = torch.empty(n, dtype=torch.uint8, device=p1.device) if hit is None else hit
            if want_band and band is None:
                band = torch.empty(n, dtype=torch.uint8, device=p1.device)
            _check(lib().bgr_map_collide_edges_dev(self._h, _ptr(p1), _ptr(p2), n, radius, shape, _ptr(hit),
                                                   _ptr(band) if want_band else None))
            return hit, band
        a, b = _host_f32(p1), _host_f32(p2)
        n = a.shape[0]
        hit = np.empty(n, dtype=np.uint8) if hit is None else hit
        if want_band and band is None:
            band = np.empty(n, dtype=np.uint8)
        _check(lib().bgr_map_collide_edges(self._h, _ptr(a), _ptr(b), n, radius, shape, _ptr(hit),
                                           _ptr(band) if want_band else None))
        return hit, band

    def set_timing(self, on=True):
        _check(lib().bgr_map_set_timing(self._h, 1 if on else 0))

    def last_kernel_ms(self):
        t = np.zeros(4, dtype=np.float32)
        _check(lib().bgr_map_last_kernel_ms(self._h, _ptr(t)))
        return {"edges_sweep_kernel": float(t[0]), "edges_narrow_kernel": float(t[1]),
                "collide_edges_kernel": float(t[2]), "edges_finalize_kernel": float(t[3])}

    def last_stats(self):
        s = np.zeros(4, dtype=np.uint64)
        _check(lib().bgr_map_last_stats(self._h, _ptr(s)))
        return {"work_items": int(s[0]), "candidates": int(s[1]), "narrow_tests": int(s[2]), "iter_cap": int(s[3])}

    def set_gate(self, on):
        """sweep gate of collide_edges on/off (results are identical; measurement / test aid)"""
        _check(lib().bgr_map_set_gate(self._h, 1 if on else 0))

    def gate_stats(self):
        s = np.zeros(2, dtype=np.uint64)
        _check(lib().bgr_map_last_gate_stats(self._h, _ptr(s)))
        return {"rect_items": int(s[0]), "dilation_cells": int(s[1])}

    def download_gate(self, d):
        """level 0 dilated by d cells (the gate map of the queries), brick words as download_bitmap"""
        nw = C.c_int64(0)
        _check(lib().bgr_map_download_gate(self._h, d, None, C.byref(nw)))
        out = np.empty(nw.value, dtype=np.uint64)
        _check(lib().bgr_map_download_gate(self._h, d, _ptr(out), C.byref(nw)))
        return out

    def dilation_stats(self):
        """bgr_map_dilation_stats: full three-pass dilations / incremental updates of a cached gate map so far"""
        s = np.zeros(2, dtype=np.uint64)
        _check(lib().bgr_map_dilation_stats(self._h, _ptr(s)))
        return {"full": int(s[0]), "incremental": int(s[1])}

    # ---- path post-processing (PrunePlan MS:562-643, find_labels MS:943-1011)
    def prune_path(self, path, radius=0.03, shape=SHAPE_CYLINDER):
        """returns (kept waypoint indices int32, number of consulted in-band edges)"""
        p = _host_f32(path)
        keep = np.empty(p.shape[0], dtype=np.int32)
        nk, nb = C.c_int64(p.shape[0]), C.c_int64(0)
        _check(lib().bgr_map_prune_path(self._h, _ptr(p), p.shape[0], radius, shape, _ptr(keep), C.byref(nk), C.byref(nb)))
        return keep[:nk.value].copy(), nb.value

    def clearance(self, pts, radius=0.1):
        p = np.ascontiguousarray(pts, dtype=np.float64).reshape(-1, 3)
        d = np.empty(p.shape[0], dtype=np.float64)
        _check(lib().bgr_map_clearance(self._h, _ptr(p), p.shape[0], radius, _ptr(d)))
        return d

    def find_labels(self, final_path, radius=0.1, n_labels=3):
        """returns (indices into final_path[1:-1], labels (3*n_labels,) float32, clearances float32, ties)"""
        p = np.ascontiguousarray(final_path, dtype=np.float64).reshape(-1, 3)
        idx = np.empty(n_labels, dtype=np.int32)
        lab = np.empty(3 * n_labels, dtype=np.float32)
        clr = np.empty(max(p.shape[0] - 2, 0), dtype=np.float32)
        ties = C.c_int(0)
        _check(lib().bgr_map_find_labels(self._h, _ptr(p), p.shape[0], radius, n_labels, _ptr(idx), _ptr(lab), _ptr(clr),
                                         C.byref(ties)))
        return idx, lab, clr, ties.value

    # reference-shaped single queries (True = in collision)
    def isStateValid(self, point):
        hit, _ = self.collide_spheres(np.asarray(point, dtype=np.float32).reshape(1, 3), 0.1, want_band=False)
        return bool(hit[0])

    def isLineValid(self, point1, point2):
        hit, _ = self.collide_edges(np.asarray(point1, dtype=np.float32).reshape(1, 3),
                                    np.asarray(point2, dtype=np.float32).reshape(1, 3), 0.03, SHAPE_CYLINDER,
                                    want_band=False)
        return bool(hit[0])


OCTREE_FULL_DATA, OCTREE_BINARY_DATA, OCTREE_FILE = 0, 1, 2


def octree_to_boxes(data, fmt=OCTREE_FILE, res=0.0, occ_threshold=0.5):
    """bgr_octree_to_boxes: serialised Octomap (bytes) -> (boxes (B,6) float64, res, bbox_min, bbox_max)"""
    buf = bytes(data)
    nb, r = C.c_int64(0), C.c_double(0.0)
    lo, hi = np.zeros(3), np.zeros(3)
    _check(lib().bgr_octree_to_boxes(buf, len(buf), fmt, res, occ_threshold, None, C.byref(nb), C.byref(r), _ptr(lo), _ptr(hi)))
    boxes = np.empty((nb.value, 6), dtype=np.float64)
    _check(lib().bgr_octree_to_boxes(buf, len(buf), fmt, res, occ_threshold, _ptr(boxes), C.byref(nb), C.byref(r), _ptr(lo), _ptr(hi)))
    return boxes, r.value, lo, hi


def smoothen_plan(path):
    """SmoothenPlan (MS:645-799): (P,3) float64 waypoints -> (11,3) float64 samples."""
    p = np.ascontiguousarray(path, dtype=np.float64).reshape(-1, 3)
    n = C.c_int64(0)
    _check(lib().bgr_path_smoothen(_ptr(p), p.shape[0], None, C.byref(n)))
    out = np.empty((n.value, 3), dtype=np.float64)
    _check(lib().bgr_path_smoothen(_ptr(p), p.shape[0], _ptr(out), C.byref(n)))
    return out


class Planner:
    """Batched extension loop of the reference planner (rrt_star_milestones.cpp:1160-1406 /
    rrt_star.cpp:1140-1300) on the device: bgr_planner_*.  `omap` is an OccupancyMap."""

    def __init__(self, params, omap, table, milestones=None, stream=None):
        t = _host_f32(table)
        m = _host_f32(milestones) if milestones is not None else np.zeros((0, 3), np.float32)
        h = C.c_void_p()
        _check(lib().bgr_planner_create(C.byref(params), omap._h, _ptr(t), t.shape[0], _ptr(m) if len(m) else None,
                                        m.shape[0], C.byref(h)))
        self._h, self._map, self.params = h, omap, params
        if stream:
            _check(lib().bgr_planner_set_stream(self._h, C.c_void_p(stream)))

    def __del__(self):
        self.close()

    def close(self):
        if getattr(self, "_h", None):
            lib().bgr_planner_destroy(self._h)
            self._h = None

    def __len__(self):
        return int(lib().bgr_planner_size(self._h))

    def step(self):
        a, g = C.c_int(0), C.c_int(0)
        _check(lib().bgr_planner_step(self._h, C.byref(a), C.byref(g)))
        return a.value, bool(g.value)

    def run(self, max_nodes, max_batches=1 << 40):
        g = C.c_int(0)
        _check(lib().bgr_planner_run(self._h, max_nodes, max_batches, C.byref(g)))
        return bool(g.value)

    @staticmethod
    def run_many(planners, max_nodes, max_batches=1 << 40):
        """bgr_planner_run_many: independent queries stepped round-robin, one stream + one map
        handle per planner.  Returns the list of goal flags."""
        n = len(planners)
        arr = (C.c_void_p * n)(*[p._h for p in planners])
        flags = (C.c_int * n)()
        _check(lib().bgr_planner_run_many(arr, n, max_nodes, max_batches, flags))
        return [bool(f) for f in flags]

    def reset(self, start, goal, seed):
        """bgr_planner_reset: same map / table / milestones, new query"""
        a = np.ascontiguousarray(start, dtype=np.float32)
        b = np.ascontiguousarray(goal, dtype=np.float32)
        _check(lib().bgr_planner_reset(self._h, _ptr(a), _ptr(b), int(seed)))

    @staticmethod
    def run_queue(planners, starts, goals, seeds, max_nodes, max_batches=1 << 40):
        """bgr_planner_run_queue: len(starts) queries through the given planners (one worker thread each).
        Returns (reached (Q,) bool, nodes (Q,) int64)."""
        n = len(planners)
        arr = (C.c_void_p * n)(*[p._h for p in planners])
        st = np.ascontiguousarray(starts, dtype=np.float32).reshape(-1, 3)
        gl = np.ascontiguousarray(goals, dtype=np.float32).reshape(-1, 3)
        sd = np.ascontiguousarray(seeds, dtype=np.uint32)
        nq = st.shape[0]
        reached = np.zeros(nq, dtype=np.int32)
        nodes = np.zeros(nq, dtype=np.int64)
        _check(lib().bgr_planner_run_queue(arr, n, _ptr(st), _ptr(gl), _ptr(sd), nq, max_nodes, max_batches, _ptr(reached), _ptr(nodes)))
        return reached.astype(bool), nodes

    def tree(self):
        n = len(self)
        xyz = np.empty((n, 3), dtype=np.float32)
        par = np.empty(n, dtype=np.int32)
        _check(lib().bgr_planner_get_tree(self._h, _ptr(xyz), _ptr(par)))
        return xyz, par

    def path(self):
        n = C.c_int64(0)
        _check(lib().bgr_planner_get_path(self._h, None, C.byref(n)))
        out = np.empty(n.value, dtype=np.int32)
        _check(lib().bgr_planner_get_path(self._h, _ptr(out), C.byref(n)))
        return out

    def nns_debug_stats(self):
        """code paths the planner's own node store has run (bgr_planner_nns + bgr_nns_debug_stats)"""
        h = C.c_void_p()
        _check(lib().bgr_planner_nns(self._h, C.byref(h)))
        return _nns_debug(h)

    def persist_stats(self):
        """bgr_planner_persist_stats: launches of the persistent small-batch kernel and the steps they ran"""
        s = np.zeros(3, dtype=np.uint64)
        _check(lib().bgr_planner_persist_stats(self._h, _ptr(s)))
        return {"launches": int(s[0]), "steps": int(s[1]), "graph_steps": int(s[2])}

    def stats(self):
        s = np.zeros(8, dtype=np.int64)
        _check(lib().bgr_planner_stats(self._h, _ptr(s)))
        return dict(zip(("batches", "candidates", "rejected_state", "rejected_edge", "improved_parent", "rewired",
                         "in_band", "goal_candidate"), (int(v) for v in s)))
