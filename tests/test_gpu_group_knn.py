"""GPU parity: grid_knn_group_kernel<LQ> (a query per group of 16 / 8 lanes, grid_group.cuh) against the CPU oracle
and against the warp-per-query kernel (BGR_KNN_GROUP=0), bit-exact.  Reference surface: Nabo::NNSearchF::knn
(rrt_star_milestones.cpp:1309-1312)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _knn(bgr, nns, q, k, flags=0, max_radius=np.inf, max_radii=None, lanes=16):
    old = os.environ.get("BGR_KNN_GROUP")
    os.environ["BGR_KNN_GROUP"] = str(lanes)
    try:
        nns.set_method(bgr.METHOD_GRID)
        idx, d2, _ = nns.knn(q, k, 0.0, flags, max_radius, max_radii)
    finally:
        if old is None:
            del os.environ["BGR_KNN_GROUP"]
        else:
            os.environ["BGR_KNN_GROUP"] = old
    return idx, d2


def _same(a, b):
    return np.array_equal(a[0], b[0]) and np.array_equal(a[1].view(np.uint32), b[1].view(np.uint32))


@pytest.mark.parametrize("n,nq,k", [(200000, 4097, 16), (60000, 1001, 8), (30000, 3, 16), (30000, 2, 1), (90000, 555, 32),
                                    (120000, 2048, 24)])
@pytest.mark.parametrize("lanes", [16, 8])
def test_group_matches_oracle_and_warp_kernel(bgr, oracle, lanes, n, nq, k):
    pts = oracle.points(31, n, 25.6)
    q = oracle.points(32, nq, 25.6)
    nns = bgr.NNSearchF.createKDTreeLinearHeap(pts)
    try:
        pair = _knn(bgr, nns, q, k, lanes=lanes)
        warp = _knn(bgr, nns, q, k, lanes=0)
        assert _same(pair, warp), "group kernel differs from the warp-per-query kernel"
        oi, od = oracle.knn_brute(pts, q, k, 0, np.inf, None)
        assert _same(pair, (oi, od)), "group kernel differs from the oracle"
    finally:
        nns.close()


@pytest.mark.parametrize("lanes", [16, 8])
def test_group_self_match_radius_and_ties(bgr, oracle, lanes):
    # queries = nodes (self-match rule), a radius limit that leaves rows partly empty, per-query radii
    pts = oracle.points(33, 80000, 25.6)
    q = np.concatenate([pts[1000:2001], oracle.points(34, 700, 25.6)]).astype(np.float32)
    nns = bgr.NNSearchF.createKDTreeLinearHeap(pts)
    try:
        for flags in (0, bgr.ALLOW_SELF_MATCH | bgr.SORT_RESULTS):
            for mr in (np.inf, 0.45, 0.05):
                pair = _knn(bgr, nns, q, 12, flags, mr, lanes=lanes)
                oi, od = oracle.knn_brute(pts, q, 12, flags, mr, None)
                assert _same(pair, (oi, od)), (flags, mr)
        radii = np.abs(oracle.points(35, len(q), 1.0)[:, 0]).astype(np.float32) * np.float32(0.6)
        pair = _knn(bgr, nns, q, 10, 0, np.inf, radii, lanes=lanes)
        oi, od = oracle.knn_brute(pts, q, 10, 0, np.inf, radii)
        assert _same(pair, (oi, od))
    finally:
        nns.close()
    # integer lattice: exactly equal distances among the first k ranks are ordered by index
    g = np.arange(14, dtype=np.float32)
    lat = np.stack(np.meshgrid(g, g, g, indexing="ij"), -1).reshape(-1, 3)
    lat = lat[np.random.default_rng(5).permutation(len(lat))]
    ql = np.concatenate([lat[:301] + np.float32(0.5), lat[301:500], lat[500:600] + np.float32(0.25)])
    nns = bgr.NNSearchF.createKDTreeLinearHeap(lat)
    try:
        for k in (6, 16, 27):
            for flags in (0, bgr.ALLOW_SELF_MATCH):
                pair = _knn(bgr, nns, ql, k, flags, lanes=lanes)
                oi, od = oracle.knn_brute(lat, ql, k, flags, np.inf, None)
                assert _same(pair, (oi, od)), (k, flags)
    finally:
        nns.close()


@pytest.mark.parametrize("lanes", [16, 8])
def test_group_sparse_queries_far_outside_and_skip(bgr, oracle, lanes, monkeypatch):
    import torch
    # a cloud in one corner: most queries find an empty ring-1 block or need ring 2 (general search per half)
    pts = (oracle.points(36, 50000, 6.4)).astype(np.float32)
    q = oracle.points(37, 999, 25.6)
    q[::7] = np.float32(1e6)                         # far outside the grid
    nns = bgr.NNSearchF.createKDTreeLinearHeap(pts)
    try:
        pair = _knn(bgr, nns, q, 16, lanes=lanes)
        oi, od = oracle.knn_brute(pts, q, 16, 0, np.inf, None)
        assert _same(pair, (oi, od))
        # skip flags: rows of skipped queries stay untouched, whichever half they sit in
        skip = (np.arange(len(q)) % 3 == 1).astype(np.uint8)
        di = torch.full((len(q), 16), -9, dtype=torch.int32, device="cuda")
        dd = torch.full((len(q), 16), -2.0, dtype=torch.float32, device="cuda")
        nns.set_method(bgr.METHOD_GRID)
        monkeypatch.setenv("BGR_KNN_GROUP", str(lanes))
        nns.knn_skip(torch.from_numpy(q).cuda(), 16, torch.from_numpy(skip).cuda(), di, dd)
        ri, rd = di.cpu().numpy(), dd.cpu().numpy()
        keep = skip == 0
        assert np.array_equal(ri[keep], oi[keep]) and np.array_equal(rd[keep].view(np.uint32), od[keep].view(np.uint32))
        assert (ri[~keep] == -9).all() and (rd[~keep] == -2.0).all()
    finally:
        nns.close()
