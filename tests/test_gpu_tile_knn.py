"""GPU parity of the tile kNN (csrc/tile_knn.cuh: one lane per query out of shared-memory tiles; an option,
off by default) against the warp-per-query grid search, the brute-force scan and the CPU
oracle: bit-exact distances, identical indices.  Reference surface: Nabo::NNSearchF::knn
(rrt_star_milestones.cpp:1309-1312)."""
import numpy as np
import pytest

from conftest import canon_equal

pytestmark = pytest.mark.gpu


def _both(B, pts, q, k, flags=0, max_radius=np.inf, max_radii=None, tile=None, index_base=0):
    """one device-pointer search through the tile path and one with the tile path off, same handle"""
    import torch
    nns = B.NNSearchF.createKDTreeLinearHeap(pts)
    nns.set_method(B.METHOD_GRID)
    if index_base:
        nns.set_index_base(index_base)
    dq = torch.from_numpy(np.ascontiguousarray(q, dtype=np.float32)).cuda()
    nns.set_tile(*(tile or (0, 0, 0, 0)))      # (0,0,0,0) = tile size chosen from the query density
    ti, td, _ = nns.knn(dq, k, 0.0, flags, max_radius, max_radii)
    torch.cuda.synchronize()
    st = nns.tile_stats()
    nns.set_tile(-1)
    ri, rd, _ = nns.knn(dq, k, 0.0, flags, max_radius, max_radii)
    torch.cuda.synchronize()
    assert nns.tile_stats()["tile_searches"] == st["tile_searches"], "tile search ran although it was turned off"
    nns.close()
    return ti.cpu().numpy(), td.cpu().numpy(), ri.cpu().numpy(), rd.cpu().numpy(), st


@pytest.mark.parametrize("k", [1, 5, 10, 16])
def test_tile_knn_uniform(bgr, oracle, k):
    pts = oracle.points(2, 200_000, 12.8)
    q = oracle.points(3, 20_000, 12.8)
    ti, td, ri, rd, st = _both(bgr, pts, q, k)
    assert st["tile_searches"] == 1 and st["fallback_queries_last"] < len(q) // 8, st
    assert canon_equal(ti, td, ri, rd)
    oi, od = oracle.knn_brute(pts, q[:512], k)
    assert canon_equal(ti[:512], td[:512], oi, od)


@pytest.mark.parametrize("tile", [(1, 1, 1, 1), (2, 2, 2, 2), (4, 4, 3, 1), (3, 5, 2, 3), (8, 6, 6, 4), (8, 1, 1, 1), (1, 6, 6, 2)])
def test_tile_knn_forced_tile_shapes(bgr, oracle, tile):
    pts = oracle.points(4, 150_000, 10.0)
    q = oracle.points(5, 30_000, 10.0)
    ti, td, ri, rd, st = _both(bgr, pts, q, 16, tile=tile)
    assert st["tile_searches"] == 1 and st["tile"] == tile[:3] and st["warps"] == tile[3], st
    assert canon_equal(ti, td, ri, rd)


def test_tile_knn_self_match_rule(bgr, oracle):
    """queries = nodes: d2 <= FLT_EPSILON excluded unless ALLOW_SELF_MATCH"""
    pts = oracle.points(2, 120_000, 10.0)
    q = pts[:16_000].copy()
    for flags in (0, bgr.ALLOW_SELF_MATCH | bgr.SORT_RESULTS):
        ti, td, ri, rd, st = _both(bgr, pts, q, 16, flags=flags)
        assert st["tile_searches"] == 1
        assert canon_equal(ti, td, ri, rd)
        if flags:
            assert np.array_equal(ti[:, 0], np.arange(len(q), dtype=np.int32)) and (td[:, 0] == 0).all()
        else:
            assert (ti != np.arange(len(q), dtype=np.int32)[:, None]).all()
        oi, od = oracle.knn_brute(pts, q[:256], 16, flags)
        assert canon_equal(ti[:256], td[:256], oi, od)


def test_tile_knn_radius_limits(bgr, oracle):
    """maxRadius / per-query maxRadii: rows shorter than k are padded with -1 / inf"""
    pts = oracle.points(2, 100_000, 10.0)
    q = oracle.points(3, 8_192, 10.0)
    for mr in (0.05, 0.2, 1.0):
        ti, td, ri, rd, st = _both(bgr, pts, q, 16, max_radius=mr)
        assert st["tile_searches"] == 1
        assert canon_equal(ti, td, ri, rd)
        oi, od = oracle.knn_brute(pts, q[:300], 16, 0, mr)
        assert canon_equal(ti[:300], td[:300], oi, od)
    rng = np.random.default_rng(5)
    radii = rng.choice(np.array([0.0, 0.03, 0.1, 0.3, np.inf], dtype=np.float32), len(q))
    ti, td, ri, rd, st = _both(bgr, pts, q, 10, max_radii=radii)
    assert canon_equal(ti, td, ri, rd)
    oi, od = oracle.knn_brute(pts, q[:300], 10, 0, np.inf, radii[:300])
    assert canon_equal(ti[:300], td[:300], oi, od)
    assert (ti[radii == 0] == -1).all() and np.isinf(td[radii == 0]).all()


def test_tile_knn_odd_queries(bgr, oracle):
    """queries far outside the node box, on its faces, NaN and infinite coordinates"""
    pts = oracle.points(2, 100_000, 10.0)
    rng = np.random.default_rng(9)
    q = oracle.points(3, 6_000, 10.0)
    q[:500] = rng.uniform(-30, 40, (500, 3))
    q[500:600, 0] = 0.0
    q[600:700, 2] = 10.0
    q[700:720] = np.nan
    q[720:730, 1] = np.inf
    q[730:740, 0] = -np.inf
    ti, td, ri, rd, st = _both(bgr, pts, q, 16)
    assert st["tile_searches"] == 1
    assert canon_equal(ti, td, ri, rd)
    oi, od = oracle.knn_brute(pts, q[:800], 16)
    assert canon_equal(ti[:800], td[:800], oi, od)


def test_tile_knn_nonuniform_density(bgr, oracle):
    """a density gradient: the tiles of the dense corner overflow the staging buffer and are handed back;
    clustered queries overflow their tile lists"""
    rng = np.random.default_rng(11)
    pts = (rng.uniform(0, 1, (250_000, 3)) ** 2.0 * 10.0).astype(np.float32)
    q = np.concatenate([(rng.uniform(0, 1, (20_000, 3)) ** 2.0 * 10.0), rng.normal(5.0, 0.05, (6_000, 3)),
                        rng.uniform(0, 10, (6_000, 3))]).astype(np.float32)
    ti, td, ri, rd, st = _both(bgr, pts, q, 16)
    assert canon_equal(ti, td, ri, rd)
    if st["tile_searches"]:
        assert st["fallback_queries_last"] > 0
    oi, od = oracle.knn_brute(pts, q[::64], 16)
    assert canon_equal(ti[::64], td[::64], oi, od)


def test_tile_knn_lattice_ties(bgr, oracle):
    """integer lattice: every query has many exactly equal distances; the 26-bit tie rule must hand them back
    and the rows must still come out in (d2, index) order"""
    g = np.arange(40, dtype=np.float32)
    pts = np.stack(np.meshgrid(g, g, g, indexing="ij"), -1).reshape(-1, 3)
    rng = np.random.default_rng(7)
    pts = pts[rng.permutation(len(pts))]
    q = np.concatenate([pts[:3000] + np.float32(0.5), pts[3000:5000], pts[5000:6000] + np.float32(0.25)])
    for k, flags in ((6, 0), (16, 0), (16, bgr.ALLOW_SELF_MATCH)):
        ti, td, ri, rd, st = _both(bgr, pts, q, k, flags=flags)
        assert canon_equal(ti, td, ri, rd)
        oi, od = oracle.knn_brute(pts, q[::16], k, flags)
        assert canon_equal(ti[::16], td[::16], oi, od)


def test_tile_knn_repeated_calls_and_index_base(bgr, oracle):
    """the per-tile counters and the two fallback counters are left clean by every call: several searches with
    different batches on one handle, and shard-style global indices"""
    import torch
    pts = oracle.points(2, 300_000, 12.8)
    nns = bgr.NNSearchF.createKDTreeLinearHeap(pts)
    nns.set_method(bgr.METHOD_GRID)
    nns.set_tile(0)
    nns.set_index_base(1_000_000)
    ref = bgr.NNSearchF.createKDTreeLinearHeap(pts)
    ref.set_method(bgr.METHOD_GRID)
    ref.set_tile(-1)
    ref.set_index_base(1_000_000)
    for rep, (seed, nq, k) in enumerate([(3, 40_000, 16), (4, 9_000, 16), (5, 40_000, 10), (6, 2_048, 16), (7, 65_536, 16), (3, 40_000, 16)]):
        dq = torch.from_numpy(oracle.points(seed, nq, 12.8)).cuda()
        ti, td, _ = nns.knn(dq, k)
        ri, rd, _ = ref.knn(dq, k)
        torch.cuda.synchronize()
        assert nns.tile_stats()["tile_searches"] == rep + 1
        assert torch.equal(ti, ri) and torch.equal(td.view(torch.int32), rd.view(torch.int32)), "call %d" % rep
        assert int(ti.min()) >= 1_000_000
    assert ref.tile_stats()["tile_searches"] == 0


def test_tile_knn_host_call_and_skip_flags(bgr, oracle):
    """the host-pointer call (H2D, search, D2H) and the planner's skip-flag variant give the same rows"""
    import torch
    pts = oracle.points(2, 200_000, 12.8)
    q = oracle.points(3, 30_000, 12.8)
    nns = bgr.NNSearchF.createKDTreeLinearHeap(pts)
    nns.set_method(bgr.METHOD_GRID)
    nns.set_tile(0)
    hi, hd, _ = nns.knn(q, 16)
    assert nns.tile_stats()["tile_searches"] >= 1
    oi, od = oracle.knn_brute(pts, q[:256], 16)
    assert canon_equal(hi[:256], hd[:256], oi, od)
    nns.set_tile(-1)
    ri, rd, _ = nns.knn(q, 16)
    assert canon_equal(hi, hd, ri, rd)
    nns.set_tile(0)
    if hasattr(nns, "knn_skip"):
        skip = (np.arange(len(q)) % 3 == 0).astype(np.uint8)
        di = torch.full((len(q), 16), -7, dtype=torch.int32, device="cuda")
        dd = torch.full((len(q), 16), -7.0, dtype=torch.float32, device="cuda")
        nns.knn_skip(torch.from_numpy(q).cuda(), 16, torch.from_numpy(skip).cuda(), di, dd)
        torch.cuda.synchronize()
        di, dd = di.cpu().numpy(), dd.cpu().numpy()
        assert (di[skip == 1] == -7).all() and (dd[skip == 1] == -7.0).all()
        assert canon_equal(di[skip == 0], dd[skip == 0], hi[skip == 0], hd[skip == 0])


def test_tile_knn_falls_back_where_it_does_not_apply(bgr, oracle):
    """k > 16, small batches and argument validation: the warp-per-query search answers, set_tile rejects bad sizes"""
    import torch
    pts = oracle.points(2, 60_000, 8.0)
    q = oracle.points(3, 5_000, 8.0)
    nns = bgr.NNSearchF.createKDTreeLinearHeap(pts)
    nns.set_method(bgr.METHOD_GRID)
    nns.set_tile(0)
    for k, nq in ((20, 5000), (16, 1000)):
        i, d, _ = nns.knn(torch.from_numpy(q[:nq]).cuda(), k)
        torch.cuda.synchronize()
        oi, od = oracle.knn_brute(pts, q[:256], k)
        assert canon_equal(i.cpu().numpy()[:256], d.cpu().numpy()[:256], oi, od)
    assert nns.tile_stats()["tile_searches"] == 0
    for bad in ((9, 1, 1, 1), (1, 7, 1, 1), (2, 2, 0, 1), (2, 2, 2, 5)):
        with pytest.raises(Exception):
            nns.set_tile(*bad)
