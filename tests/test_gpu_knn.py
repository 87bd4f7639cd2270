"""GPU parity: bgr_nns_* (through the C ABI) vs the CPU oracle, bit-exact.

Reference surface: Nabo::NNSearchF::createKDTreeLinearHeap / knn (rrt_star_milestones.cpp:
1306-1312) and Nearest() (:138-154)."""
import numpy as np
import pytest

from conftest import canon_equal

pytestmark = pytest.mark.gpu


def _check(B, O, pts, q, k, method, flags=0, max_radius=np.inf, max_radii=None):
    nns = B.NNSearchF.createKDTreeLinearHeap(pts)
    nns.set_method(method)
    idx, d2, _ = nns.knn(q, k, 0.0, flags, max_radius, max_radii)
    oi, od = O.knn_brute(pts, q, k, flags, max_radius, max_radii)
    assert np.array_equal(d2.view(np.uint32), od.view(np.uint32)), "distances differ"
    assert np.array_equal(idx, oi), "indices differ"
    nns.close()


@pytest.mark.parametrize("method", [1, 2, 3])
@pytest.mark.parametrize("n,nq,k", [(5000, 300, 10), (40000, 1000, 16), (33, 7, 32), (1500, 64, 1), (70000, 333, 5)])
def test_knn_matches_oracle(bgr, oracle, method, n, nq, k):
    pts = oracle.points(2, n, 25.6)
    q = oracle.points(3, nq, 25.6)
    _check(bgr, oracle, pts, q, k, method)


@pytest.mark.parametrize("method", [1, 2, 3])
def test_knn_self_match_rule(bgr, oracle, method):
    """queries = nodes: without ALLOW_SELF_MATCH points with d2 <= FLT_EPSILON are skipped."""
    pts = oracle.points(2, 30000, 25.6)
    q = pts[:2000].copy()
    _check(bgr, oracle, pts, q, 16, method, flags=0)
    _check(bgr, oracle, pts, q, 16, method, flags=bgr.ALLOW_SELF_MATCH | bgr.SORT_RESULTS)
    nns = bgr.NNSearchF.createKDTreeLinearHeap(pts)
    nns.set_method(method)
    idx, d2, _ = nns.knn(q, 4, 0.0, bgr.ALLOW_SELF_MATCH)
    assert np.array_equal(idx[:, 0], np.arange(2000, dtype=np.int32))
    assert (d2[:, 0] == 0).all()
    idx, d2, _ = nns.knn(q, 4, 0.0, 0)
    assert (idx[:, 0] != np.arange(2000)).all() and (d2[:, 0] > np.finfo(np.float32).eps).all()


@pytest.mark.parametrize("method", [1, 2, 3])
def test_knn_lattice_ties(bgr, oracle, method):
    """integer lattice: many exactly equal distances; ties must come out in index order."""
    g = np.arange(12, dtype=np.float32)
    pts = np.stack(np.meshgrid(g, g, g, indexing="ij"), -1).reshape(-1, 3)
    rng = np.random.default_rng(7)
    pts = pts[rng.permutation(len(pts))]
    q = np.concatenate([pts[:200] + np.float32(0.5), pts[200:300]])
    for k in (6, 16, 27):
        _check(bgr, oracle, pts, q, k, method, flags=bgr.ALLOW_SELF_MATCH)
        _check(bgr, oracle, pts, q, k, method, flags=0)


@pytest.mark.parametrize("method", [1, 2, 3])
def test_knn_radius(bgr, oracle, method):
    pts = oracle.points(2, 20000, 10.0)
    q = oracle.points(3, 500, 10.0)
    _check(bgr, oracle, pts, q, 16, method, max_radius=0.35)
    radii = (oracle.MT(11).real(500) * 0.6).astype(np.float32)
    _check(bgr, oracle, pts, q, 8, method, max_radii=radii)
    # unfilled slots: idx -1, d2 +inf
    nns = bgr.NNSearchF.createKDTreeLinearHeap(pts)
    nns.set_method(method)
    idx, d2, _ = nns.knn(q, 16, 0.0, 0, 0.05)
    assert ((idx == -1) == np.isinf(d2)).all() and (idx == -1).any()


@pytest.mark.parametrize("method", [1, 2, 3])
def test_knn_append_only_store(bgr, oracle, method):
    """create + repeated append == one-shot create (replaces the reference's per-iteration rebuild)."""
    pts = oracle.points(5, 50000, 5.0)
    q = oracle.points(6, 400, 5.0)
    nns = bgr.NNSearchF.createKDTreeLinearHeap(pts[:20000])
    nns.set_method(method)
    nns.knn(q, 10)                         # builds the index over the first part
    for a, b in ((20000, 20001), (20001, 23000), (23000, 50000)):
        nns.append(pts[a:b])
        idx, d2, _ = nns.knn(q, 10)
        oi, od = oracle.knn_brute(pts[:b], q, 10)
        assert canon_equal(idx, d2, oi, od)
    assert len(nns) == 50000


def test_knn_clustered_and_outside_queries(bgr, oracle):
    """non-uniform node density (RRT-like clusters) and queries far outside the node AABB."""
    rng = np.random.default_rng(3)
    centers = rng.uniform(0, 20, (12, 3))
    pts = (centers[rng.integers(0, 12, 60000)] + rng.normal(0, 0.15, (60000, 3))).astype(np.float32)
    q = np.concatenate([rng.uniform(-30, 50, (300, 3)), pts[:300] + 0.01]).astype(np.float32)
    for method in (1, 2, 3):
        _check(bgr, oracle, pts, q, 16, method)


@pytest.mark.parametrize("flags", [0, 1, 2, 3])
def test_knn_piled_up_nodes(bgr, oracle, flags):
    """a planner tree holds thousands of coincident copies of a reached milestone: the queries whose
    ring-1 block is such a pile go through grid_knn_heavy_kernel (one CTA per query over the raw
    store).  Rows must equal the oracle's, with and without self matches (every copy is a self
    match of a query at the pile), sorted or not, for queries at, next to and far from the piles,
    and when more queries are heavy than the heavy list holds (the rest take the ring search)."""
    rng = np.random.default_rng(11)
    base = rng.uniform(0, 12.8, (60000, 3)).astype(np.float32)
    piles = np.array([[3.0, 3.0, 3.0], [9.5, 2.25, 7.0], [6.4, 6.4, 6.4]], np.float32)
    pts = np.concatenate([base, np.repeat(piles[:1], 30000, 0), np.repeat(piles[1:2], 9000, 0),
                          np.repeat(piles[2:], 5000, 0)])
    pts = pts[rng.permutation(len(pts))]
    near = (piles[rng.integers(0, 3, 400)] + rng.normal(0, 0.05, (400, 3))).astype(np.float32)
    q = np.concatenate([piles, np.repeat(piles, 4, 0), near, rng.uniform(0, 12.8, (300, 3)).astype(np.float32)])
    for method in (2, 3):
        _check(bgr, oracle, pts, q, 16, method, flags=flags)
    # appended tail on top of the piled index, k = 32, a radius limit
    nns = bgr.NNSearchF.createKDTreeLinearHeap(pts)
    nns.set_method(2)
    nns.knn(q[:8], 4)
    extra = np.concatenate([np.repeat(piles[:1], 500, 0), rng.uniform(0, 12.8, (700, 3)).astype(np.float32)])
    nns.append(extra)
    allp = np.concatenate([pts, extra])
    idx, d2, _ = nns.knn(q, 32, 0.0, flags, 0.3)
    oi, od = oracle.knn_brute(allp, q, 32, flags, 0.3)
    assert np.array_equal(d2.view(np.uint32), od.view(np.uint32)) and np.array_equal(idx, oi)
    if flags == 0:
        # Nearest over the piles: the copies take the coincident-node shortcut (equal Norm: last index wins)
        for method in (1, 2):
            nns.set_method(method)
            ni, nd = nns.nearest(q)
            ri, rd = oracle.nearest(allp, q)
            assert np.array_equal(ni, ri) and np.array_equal(nd.view(np.uint32), rd.view(np.uint32))
    nns.close()


def test_knn_errors_like_libnabo(bgr, oracle):
    pts = oracle.points(2, 10, 1.0)
    nns = bgr.NNSearchF.createKDTreeLinearHeap(pts)
    q = pts[:2]
    with pytest.raises(RuntimeError):   # k > N-1 without self match (why the reference guards V.size()>5)
        nns.knn(q, 10)
    nns.knn(q, 10, 0.0, bgr.ALLOW_SELF_MATCH)
    with pytest.raises(RuntimeError):
        nns.knn(q, 11, 0.0, bgr.ALLOW_SELF_MATCH)
    with pytest.raises(RuntimeError):   # unknown flag bits
        nns.knn(q, 2, 0.0, 4)
    with pytest.raises(RuntimeError):   # approximate search is not offered
        nns.knn(q, 2, 0.5)
    with pytest.raises(RuntimeError):
        bgr.NNSearchF.createKDTreeLinearHeap(np.zeros((4, 2), np.float32), dim=2)


def test_knn_device_pointers_and_stream(bgr, oracle):
    import torch
    pts = oracle.points(2, 30000, 25.6)
    q = oracle.points(3, 1000, 25.6)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        dp = torch.from_numpy(pts).cuda()
        dq = torch.from_numpy(q).cuda()
        nns = bgr.NNSearchF.createKDTreeLinearHeap(dp, stream=st.cuda_stream)
        idx, d2, _ = nns.knn(dq, 16)
    st.synchronize()
    oi, od = oracle.knn_brute(pts, q, 16)
    assert canon_equal(idx.cpu().numpy(), d2.cpu().numpy(), oi, od)


def test_knn_merge_shards(bgr, oracle):
    """index-range shards + merge == single search (the multi-GPU exchange step, on one GPU)."""
    import torch
    pts = oracle.points(5, 40000, 25.6)
    q = oracle.points(6, 700, 25.6)
    g, k = 4, 16
    dq = torch.from_numpy(q).cuda()
    slabs_i, slabs_d = [], []
    for s in range(g):
        lo, hi = s * 10000, (s + 1) * 10000
        nns = bgr.NNSearchF.createKDTreeLinearHeap(pts[lo:hi])
        nns.set_index_base(lo)
        i, d, _ = nns.knn(dq, k)
        slabs_i.append(i)
        slabs_d.append(d)
    torch.cuda.synchronize()
    si, sd = torch.stack(slabs_i), torch.stack(slabs_d)
    oi_t = torch.empty_like(slabs_i[0])
    od_t = torch.empty_like(slabs_d[0])
    bgr.knn_merge_dev(si, sd, oi_t, od_t)
    torch.cuda.synchronize()
    oi, od = oracle.knn_brute(pts, q, k)
    assert canon_equal(oi_t.cpu().numpy(), od_t.cpu().numpy(), oi, od)


@pytest.mark.parametrize("n,nq", [(1, 5), (777, 100), (5000, 2000), (120000, 512)])
def test_nearest_matches_reference_literal(bgr, oracle, n, nq):
    pts = oracle.points(2, n, 3.2)
    q = oracle.points(3, nq, 3.2)
    nns = bgr.NNSearchF.createKDTreeLinearHeap(pts)
    idx, dist = nns.nearest(q)
    oi, od = oracle.nearest(pts, q)
    assert np.array_equal(idx, oi)
    assert np.array_equal(dist.view(np.uint32), od.view(np.uint32))


def test_nearest_last_index_wins_ties(bgr, oracle):
    pts = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [1, 0, 0], [0, 0, 0]], dtype=np.float32)
    q = np.array([[0, 0, 0], [1, 0, 0], [0.5, 0, 0]], dtype=np.float32)
    nns = bgr.NNSearchF.createKDTreeLinearHeap(pts)
    idx, _ = nns.nearest(q)
    oi, _ = oracle.nearest(pts, q)
    assert idx.tolist() == oi.tolist() == [4, 3, 4]


@pytest.mark.parametrize("method", [2, 3])
def test_nearest_grid_equals_plain_scan(bgr, oracle, method):
    """the grid Nearest (large trees) returns the plain scan's answer bit for bit: random and
    clustered trees, lattice points with equal float Norms (last index wins), duplicated nodes,
    queries outside the cloud / infinite / NaN, and an unsorted tail after appends."""
    mt = oracle.MT(71)
    uni = oracle.points(2, 60000, 25.6)
    clus = (np.array([5.0, 6.0, 7.0]) + (mt.real(3 * 30000).reshape(-1, 3) - 0.5) * 0.3).astype(np.float32)
    lat = np.stack(np.meshgrid(np.arange(20), np.arange(20), np.arange(20), indexing="ij"), -1).reshape(-1, 3).astype(np.float32)
    pts = np.concatenate([uni, clus, lat, lat[:500], uni[:1000]])
    q = np.concatenate([oracle.points(3, 3000, 25.6), clus[:500] + 0.01, lat[:300] + 0.5, lat[:300],
                        np.array([[-50, 3, 3], [1e6, 1e6, 1e6], [np.inf, 0, 0], [np.nan, 1, 1]], dtype=np.float32)]).astype(np.float32)
    ref = bgr.NNSearchF.createKDTreeLinearHeap(pts)
    ref.set_method(bgr.METHOD_BRUTE)
    ri, rd = ref.nearest(q)
    oi, od = oracle.nearest(pts, q[:600])
    assert np.array_equal(ri[:600], oi) and np.array_equal(rd[:600].view(np.uint32), od.view(np.uint32))
    nns = bgr.NNSearchF.createKDTreeLinearHeap(pts[:80000])
    nns.set_method(method)
    nns.build_index()
    nns.append(pts[80000:])                      # unsorted tail: a small batch does not trigger a rebuild
    sub = np.r_[0:40, 3000:3020, len(q) - 8:len(q)]
    si, sd = nns.nearest(q[sub])
    assert np.array_equal(si, ri[sub]) and np.array_equal(sd.view(np.uint32), rd[sub].view(np.uint32))
    gi, gd = nns.nearest(q)
    assert np.array_equal(gi, ri), np.nonzero(gi != ri)[0][:10]
    assert np.array_equal(gd.view(np.uint32), rd.view(np.uint32))
    nns.build_index()
    gi, gd = nns.nearest(q)
    assert np.array_equal(gi, ri) and np.array_equal(gd.view(np.uint32), rd.view(np.uint32))


def test_knn_full_size_properties(bgr, oracle):
    """BASELINE config 2 (1M nodes, 64k queries, k=16): grid == brute on the GPU, and an
    oracle check on a 256-query subsample; sortedness; distances recomputed on the host."""
    n, nq, k = 1_000_000, 65_536, 16
    pts = oracle.points(2, n, 25.6)
    q = oracle.points(3, nq, 25.6)
    nns = bgr.NNSearchF.createKDTreeLinearHeap(pts)
    nns.set_method(bgr.METHOD_GRID)
    gi, gd, _ = nns.knn(q, k)
    assert (np.diff(gd, axis=1) >= 0).all()
    nns.set_method(bgr.METHOD_GRID_GENERAL)          # ring-expansion search only: same answer
    ri, rd, _ = nns.knn(q, k)
    assert canon_equal(gi, gd, ri, rd)
    sub = slice(0, 4096)
    nns.set_method(bgr.METHOD_BRUTE)
    bi, bd, _ = nns.knn(q[sub], k)
    assert canon_equal(gi[sub], gd[sub], bi, bd)
    oi, od = oracle.knn_brute(pts, q[:256], k)
    assert canon_equal(gi[:256], gd[:256], oi, od)
    # d2 recomputed in numpy f32 with the same operation order
    d = q[:, None, :] - pts[gi]
    r = (d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1]) + d[..., 2] * d[..., 2]
    assert np.array_equal(r.astype(np.float32).view(np.uint32), gd.view(np.uint32))
    # every query against the oracle's libnabo kd-tree restatement (its order inside a run of
    # exactly equal d2 is the kd-tree visit order: compare d2 bit for bit, indices as sets per run)
    ti, td = oracle.NaboKDTree(pts).knn(q, k)
    assert np.array_equal(td.view(np.uint32), gd.view(np.uint32))
    diff = np.nonzero((ti != gi).any(axis=1))[0]
    for r in diff:
        for v in np.unique(gd[r]):
            sel = gd[r] == v
            assert set(ti[r][sel]) == set(gi[r][sel]) or sel[-1], "row %d differs beyond tie order" % r
    # the host call above was chunked (D2H of chunk c overlaps the kernel of chunk c+1): the
    # one-launch device call returns the same bytes
    import torch
    nns.set_method(bgr.METHOD_GRID)
    di, dd, _ = nns.knn(torch.from_numpy(q).cuda(), k)
    torch.cuda.synchronize()
    assert np.array_equal(di.cpu().numpy(), gi) and np.array_equal(dd.cpu().numpy().view(np.uint32), gd.view(np.uint32))


def test_knn_full_size_second_pass_self_match(bgr, oracle):
    """BASELINE config 2, second pass (libnabo call site MS:1309-1312): queries = the
    first 65,536 NODES of the 1 M-node tree, once with SORT_RESULTS|ALLOW_SELF_MATCH (every query finds
    itself first, d2 = 0) and once with flags = 0 (the d2 <= FLT_EPSILON exclusion drops the node itself):
    every row against the oracle's libnabo kd-tree restatement, a 256-row subsample against the oracle's
    brute force, grid == ring search == device call."""
    import torch
    n, nq, k = 1_000_000, 65_536, 16
    pts = oracle.points(2, n, 25.6)
    q = np.ascontiguousarray(pts[:nq])
    nns = bgr.NNSearchF.createKDTreeLinearHeap(pts)
    tree = oracle.NaboKDTree(pts)
    for flags in (bgr.SORT_RESULTS | bgr.ALLOW_SELF_MATCH, 0):
        nns.set_method(bgr.METHOD_GRID)
        gi, gd, _ = nns.knn(q, k, optionFlags=flags)
        assert (np.diff(gd, axis=1) >= 0).all() and (gi >= 0).all()
        if flags & bgr.ALLOW_SELF_MATCH:
            assert np.array_equal(gi[:, 0], np.arange(nq, dtype=np.int32)) and (gd[:, 0] == 0).all()
        else:
            assert (gi != np.arange(nq, dtype=np.int32)[:, None]).all() and (gd > np.finfo(np.float32).eps).all()
        nns.set_method(bgr.METHOD_GRID_GENERAL)
        ri, rd, _ = nns.knn(q, k, optionFlags=flags)
        assert canon_equal(gi, gd, ri, rd)
        oi, od = oracle.knn_brute(pts, q[:256], k, flags)
        assert canon_equal(gi[:256], gd[:256], oi, od)
        ti, td = tree.knn(q, k, flags=flags)
        assert np.array_equal(td.view(np.uint32), gd.view(np.uint32))
        for r in np.nonzero((ti != gi).any(axis=1))[0]:
            for v in np.unique(gd[r]):
                sel = gd[r] == v
                assert set(ti[r][sel]) == set(gi[r][sel]) or sel[-1], "row %d differs beyond tie order" % r
        nns.set_method(bgr.METHOD_GRID)
        di, dd, _ = nns.knn(torch.from_numpy(q).cuda(), k, optionFlags=flags)
        torch.cuda.synchronize()
        assert np.array_equal(di.cpu().numpy(), gi) and np.array_equal(dd.cpu().numpy().view(np.uint32), gd.view(np.uint32))


def test_knn_pinned_result_buffers_written_by_the_kernel(bgr, oracle):
    """host call with page-locked result buffers: the kernel writes them directly (no staging copy); same rows as
    the pageable-buffer call and the oracle, with and without the distance array, grid and brute force"""
    import torch
    pts = oracle.points(2, 80_000, 10.0)
    q = oracle.points(3, 20_000, 10.0)
    oi, od = oracle.knn_brute(pts, q[:256], 16)
    nns = bgr.NNSearchF.createKDTreeLinearHeap(pts)
    for method in (bgr.METHOD_GRID, bgr.METHOD_BRUTE):
        nns.set_method(method)
        qp = q if method == bgr.METHOD_GRID else q[:2048]
        ri, rd, _ = nns.knn(qp, 16)                                     # pageable numpy arrays: staged path
        ih = torch.full((len(qp), 16), -9, dtype=torch.int32).pin_memory()
        dh = torch.full((len(qp), 16), -9.0, dtype=torch.float32).pin_memory()
        nns.knn(qp, 16, indices=ih.numpy(), dists2=dh.numpy())
        assert canon_equal(ih.numpy(), dh.numpy(), ri, rd)
        assert canon_equal(ih.numpy()[:256], dh.numpy()[:256], oi, od)
        ih.fill_(-9)
        nns.knn(qp, 16, indices=ih.numpy(), want_dists2=False)
        assert np.array_equal(ih.numpy(), ri)


def test_nearest_scan_edge_cases(bgr, oracle):
    """the plain Nearest scan runs an f32 filter before the reference's mixed
    f32 / f64 arithmetic: it must never lose the reference's answer — large coordinates (coarse filter), queries on
    nodes, a hair away from them, far, astronomically far, infinite and NaN; duplicated nodes (last index wins)"""
    rng = np.random.default_rng(13)
    base = oracle.points(2, 40_000, 6.4)
    pts = np.concatenate([base + np.float32(500.0), base[:5000] + np.float32(500.0), base[:3000] * np.float32(1e-3)]).astype(np.float32)
    q = np.concatenate([
        pts[rng.integers(0, len(pts), 600)],
        pts[rng.integers(0, len(pts), 600)] + rng.normal(0, 1e-4, (600, 3)).astype(np.float32),
        oracle.points(3, 600, 6.4) + np.float32(500.0),
        rng.uniform(-2000, 2000, (300, 3)).astype(np.float32),
        np.array([[1e6, -1e6, 3], [1e12, 0, 0], [1e18, 1e18, -1e18], [3e25, 0, 1], [np.inf, 0, 0], [0, -np.inf, 5], [np.nan, 1, 1],
                  [0, 0, 0], [500, 500, 500]], dtype=np.float32)]).astype(np.float32)
    nns = bgr.NNSearchF.createKDTreeLinearHeap(pts)
    nns.set_method(bgr.METHOD_BRUTE)
    gi, gd = nns.nearest(q)
    oi, od = oracle.nearest(pts, q)
    ok = ~np.isnan(q).any(axis=1)       # a NaN query compares false with everything: the library reports "no node" (-1)
    assert np.array_equal(gi[ok], oi[ok]), np.nonzero(gi != oi)[0][:10]
    assert np.array_equal(gd[ok].view(np.uint32), od[ok].view(np.uint32))
    assert (gi[~ok] == -1).all()
