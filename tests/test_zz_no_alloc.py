"""GPU: bgr_nns_knn_scan_static_dev (the allocation-free brute-force kNN scan, include/bgrrt.h) against
the existing entry point bgr_nns_knn_dev, bit-exact (the bar of tests/test_gpu_knn.py: integer indices and
f32 distance bits), on two calls in a row — the second with other queries, another batch size and another
k, so that anything left in the shared static scratch by the first would show.

No counterpart in the reference.  This file is named to be
collected LAST."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _both(bgr, nns, q, k, flags=0, max_radius=float("inf")):
    """(existing entry point on the plain scan, existing entry point on the grid, new entry point, chunks)"""
    import torch
    dq = torch.from_numpy(q).cuda()
    out = []
    for method in (bgr.METHOD_BRUTE, bgr.METHOD_GRID):
        nns.set_method(method)
        i, d, _ = nns.knn(dq, k, 0.0, flags, max_radius)
        out.append((i.cpu().numpy(), d.cpu().numpy()))
    # poisoned result rows: every slot must be written by the call
    si = torch.full((q.shape[0], k), -7, dtype=torch.int32, device="cuda")
    sd = torch.full((q.shape[0], k), -1.0, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    chunks = nns.knn_scan_static(dq, k, si, sd, flags, max_radius)
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    return out[0], out[1], (si.cpu().numpy(), sd.cpu().numpy()), chunks, free0 - free1


def _same(a, b):
    return np.array_equal(a[0], b[0]) and np.array_equal(a[1].view(np.uint32), b[1].view(np.uint32))


def test_static_scan_matches_knn_dev_twice(bgr, oracle):
    pts = oracle.points(11, 120000, 25.6)
    nns = bgr.NNSearchF.createKDTreeLinearHeap(pts)
    try:
        # ---- call 1: 1,024 queries, k = 16 (the planner's batch)
        q1 = oracle.points(12, 1024, 25.6)
        brute, grid, static, chunks, grew = _both(bgr, nns, q1, 16)
        assert chunks > 1, "the scan ran in one chunk: the static scratch was not exercised"
        assert _same(static, brute), "static scan differs from bgr_nns_knn_dev (plain scan)"
        assert _same(static, grid), "static scan differs from bgr_nns_knn_dev (grid)"
        oi, od = oracle.knn_brute(pts, q1, 16, 0, np.inf, None)
        assert _same(static, (oi, od)), "static scan differs from the oracle"
        assert grew == 0, "device memory in use grew by %d bytes across the call" % grew
        # ---- call 2, straight after: other queries (half of them nodes: self-match rule), other size, other k,
        #      a radius limit that leaves rows partly empty (idx = -1, d2 = +inf)
        q2 = np.concatenate([oracle.points(13, 350, 25.6), pts[5000:5349]]).astype(np.float32)
        flags = bgr.ALLOW_SELF_MATCH | bgr.SORT_RESULTS
        brute, grid, static, chunks2, grew = _both(bgr, nns, q2, 10, flags, 0.35)
        assert chunks2 > 1
        assert _same(static, brute), "second call differs from bgr_nns_knn_dev (plain scan)"
        assert _same(static, grid), "second call differs from bgr_nns_knn_dev (grid)"
        assert (static[0] == -1).any() and (static[0] >= 0).any(), "the radius limit should leave some slots empty"
        assert np.array_equal(static[0][350:, 0], np.arange(5000, 5349, dtype=np.int32))
        assert grew == 0, "device memory in use grew by %d bytes across the call" % grew
        # ---- and the first batch once more: same answer as before
        _, _, again, _, _ = _both(bgr, nns, q1, 16)
        assert _same(again, (oi, od))
    finally:
        nns.close()


def test_static_scan_shared_by_handles_and_limits(bgr, oracle):
    import torch
    pts_a = oracle.points(21, 60000, 25.6)
    pts_b = oracle.points(22, 90000, 12.8)
    a = bgr.NNSearchF.createKDTreeLinearHeap(pts_a)
    b = bgr.NNSearchF.createKDTreeLinearHeap(pts_b)
    try:
        q = oracle.points(23, 512, 12.8)
        for nns in (a, b, a):             # the same array serves every handle, one call after the other
            brute, _, static, chunks, _ = _both(bgr, nns, q, 8)
            assert chunks > 1 and _same(static, brute)
        # the stated maximum: nq * k == limit passes, one row more is refused with a clear error
        k = 16
        nq = bgr.KNN_STATIC_MAX_RESULTS // k
        qmax = oracle.points(24, nq + 1, 25.6)
        brute, _, static, _, _ = _both(bgr, a, qmax[:nq], k)
        assert _same(static, brute)
        dq = torch.from_numpy(qmax).cuda()
        si = torch.empty((nq + 1, k), dtype=torch.int32, device="cuda")
        sd = torch.empty((nq + 1, k), dtype=torch.float32, device="cuda")
        with pytest.raises(bgr.BgrError, match="static scratch limit"):
            a.knn_scan_static(dq, k, si, sd)
        # argument rules of the existing entry point hold here too
        with pytest.raises(bgr.BgrError):
            a.knn_scan_static(dq[:4], 33, si, sd)
        # an empty batch is a no-op
        assert a.knn_scan_static(dq[:0], k, si[:0], sd[:0]) == 0
    finally:
        a.close()
        b.close()
