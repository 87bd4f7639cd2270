"""GPU parity of the path post-processing (C ABI) vs the oracle restatement (oracle/path.py):
PrunePlan rrt_star_milestones.cpp:562-609 (all-pairs edge batch + the reference's greedy walk),
find_labels :943-996 (clearance = min sphere-leaf distance, 3 smallest)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _map(B, O, n=64, seed=1, max_level=2):
    occ = O.make_map(n, seed=seed)
    boxes = O.grid_to_boxes(occ, 0.05, (0, 0, 0), max_level)
    m = B.OccupancyMap(0.05, (0, 0, 0), (n, n, n))
    m.set_boxes(boxes)
    return occ, boxes, m


def _free_points(occ, rng, n, res=0.05):
    out = []
    while len(out) < n:
        p = rng.uniform(0.1, occ.shape[0] * res - 0.1, 3)
        c = np.floor(p / res).astype(int)
        if not occ[max(c[2] - 1, 0):c[2] + 2, max(c[1] - 1, 0):c[1] + 2, max(c[0] - 1, 0):c[0] + 2].any():
            out.append(p)
    return np.asarray(out, dtype=np.float32)


@pytest.mark.parametrize("seed", [0, 1, 2, 3, 4, 5])
def test_prune_path_matches_reference_walk(bgr, oracle, seed):
    from oracle import path as P
    occ, boxes, m = _map(bgr, oracle, 64, 1 + seed % 2)
    rng = np.random.default_rng(seed)
    # a path of short hops (mostly free) so that the walk usually terminates; both sides must agree either way
    start = _free_points(occ, rng, 1)[0]
    pts = [start]
    while len(pts) < 14:
        q = pts[-1] + rng.normal(0, 0.25, 3).astype(np.float32)
        if (q > 0.1).all() and (q < 3.1).all():
            pts.append(q.astype(np.float32))
    path = np.asarray(pts, dtype=np.float32)
    try:
        ref, ref_band = P.prune_plan(boxes, path)
        ref_err = None
    except RuntimeError as e:
        ref, ref_band, ref_err = None, 0, e
    try:
        got, got_band = m.prune_path(path)
        got_err = None
    except bgr.BgrError as e:
        got, got_band, got_err = None, 0, e
    if ref_band or got_band:
        pytest.skip("a consulted edge lies in the MPR tolerance band (reported, not compared)")
    assert (ref_err is None) == (got_err is None), (ref_err, got_err)
    if ref_err is None:
        assert list(ref) == list(got)
        assert got[0] == 0 and path[got[-1]][0] == path[-1][0]


def test_prune_path_trivial_and_quirk(bgr, oracle):
    occ, boxes, m = _map(bgr, oracle)
    empty = bgr.OccupancyMap(0.05, (0, 0, 0), (64, 64, 64))
    empty.set_boxes(np.zeros((0, 6)))
    path = np.array([[0.2, 0.2, 0.2], [0.5, 0.4, 0.3], [1.0, 1.0, 1.0], [2.0, 2.5, 3.0]], dtype=np.float32)
    keep, _ = empty.prune_path(path)
    assert list(keep) == [0, 3]                       # free space: straight to the goal
    keep, _ = empty.prune_path(path[:1])
    assert list(keep) == [0]
    # MS:579 compares x only: a waypoint sharing the goal's x ends the walk as soon as it is kept
    q = path.copy()
    q[0, 0] = q[3, 0]
    keep, _ = empty.prune_path(q)
    assert list(keep) == [0]


def test_clearance_and_labels_match_oracle(bgr, oracle):
    from oracle import path as P
    occ, boxes, m = _map(bgr, oracle, 64, 1, 2)
    rng = np.random.default_rng(7)
    pts = rng.uniform(0.05, 3.15, (40, 3))
    got = m.clearance(pts, 0.1)
    sub = boxes
    ref = P.clearance(sub, pts, 0.1)
    touching = np.abs(ref) < 1e-9
    assert np.array_equal(got[~touching] == -1.0, ref[~touching] == -1.0)
    ok = (ref != -1.0) & ~touching
    assert np.allclose(got[ok], ref[ok], rtol=0, atol=1e-12)
    path = np.vstack([[0.1, 0.1, 0.1], pts[:12], [3.1, 3.1, 3.1]])
    gi, gl, gc, ties = m.find_labels(path, 0.1, 3)
    ri, rl, rc = P.find_labels(boxes, path, 0.1, 3)
    assert np.allclose(gc, rc, rtol=0, atol=1e-6)
    if ties == 0 and len(set(np.sort(rc)[:4])) == 4:
        assert list(gi) == list(ri) and np.array_equal(gl, rl)
    # empty map: the reference's initial min_dist
    empty = bgr.OccupancyMap(0.05, (0, 0, 0), (64, 64, 64))
    empty.set_boxes(np.zeros((0, 6)))
    assert np.array_equal(empty.clearance(pts[:3], 0.1), [1000.0, 1000.0, 1000.0])
