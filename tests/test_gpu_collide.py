"""GPU parity: bgr_map_* (through the C ABI) vs the CPU oracle's FCL/libccd-MPR restatement.

Reference surface: generateBoxesFromOctomap (rrt_star_milestones.cpp:288-306),
isStateValid (:310-351), isLineValid (:385-471).  Booleans must be identical except inside
the 1e-6 m MPR tolerance band, which the library flags (in_band) and the test counts."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _map(B, O, n=64, seed=1, max_level=2, **kw):
    occ = O.make_map(n, seed=seed, **kw)
    boxes = O.grid_to_boxes(occ, 0.05, (0, 0, 0), max_level)
    m = B.OccupancyMap(0.05, (0, 0, 0), (n, n, n))
    m.set_boxes(boxes)
    return occ, boxes, m


def _bits_to_occ(words, n):
    nb = (n + 3) // 4
    w = words.reshape(nb, nb, nb)  # [bz][by][bx]
    occ = np.zeros((n, n, n), dtype=np.uint8)
    for z in range(4):
        for y in range(4):
            for x in range(4):
                bit = (z << 4) | (y << 2) | x
                occ[z::4, y::4, x::4] = ((w >> np.uint64(bit)) & np.uint64(1)).astype(np.uint8)[: (n - z + 3) // 4, : (n - y + 3) // 4, : (n - x + 3) // 4]
    return occ


@pytest.mark.parametrize("max_level", [0, 2, 4])
def test_flatten_boxes_to_bitmap(bgr, oracle, max_level):
    occ, boxes, m = _map(bgr, oracle, 64, 1, max_level)
    assert m.occupied_cells() == int(occ.sum())
    assert np.array_equal(_bits_to_occ(m.download_bitmap(), 64), occ)


def test_incremental_update(bgr, oracle):
    occ, boxes, m = _map(bgr, oracle, 64, 1, 2)
    occ2 = oracle.make_map(64, seed=2)
    b2 = oracle.grid_to_boxes(occ2, 0.05, (0, 0, 0), 2)
    m.update_boxes(add=None, remove=boxes)
    assert m.occupied_cells() == 0
    m.update_boxes(add=b2)
    assert np.array_equal(_bits_to_occ(m.download_bitmap(), 64), occ2)
    # partial refresh: drop a tenth of the leaves, add them back
    part = b2[::10]
    m.update_boxes(remove=part)
    assert m.occupied_cells() == int(occ2.sum()) - int(round(((part[:, 3] / 0.05) ** 3).sum()))
    m.update_boxes(add=part)
    assert np.array_equal(_bits_to_occ(m.download_bitmap(), 64), occ2)


def _dilate_np(occ, d):
    """box dilation by d cells (Chebyshev), zero outside the map"""
    out = occ.astype(bool)
    for axis in range(3):
        acc = out.copy()
        for s in range(1, d + 1):
            sl_a = [slice(None)] * 3
            sl_b = [slice(None)] * 3
            sl_a[axis], sl_b[axis] = slice(s, None), slice(None, -s)
            acc[tuple(sl_a)] |= out[tuple(sl_b)]
            acc[tuple(sl_b)] |= out[tuple(sl_a)]
        out = acc
    return out.astype(np.uint8)


@pytest.mark.parametrize("n", [100, 128])
def test_gate_maps_follow_updates_incrementally(bgr, oracle, n):
    """north_star (1) "rebuilt incrementally when the map updates", for the gate maps too: after update_boxes /
    refresh_boxes the cached dilations (d = 1 for the 0.03 m edges, d = 3 for the 0.1 m state sphere) are
    brought up to date by re-dilating only the bricks near the changed leaves — compared cell by cell with a
    numpy dilation of the new occupancy, and counted (no full re-dilation after the first one)."""
    occ = oracle.make_map(n, seed=3, n_walls=2, wall_thick=2, door_cells=4)
    boxes = oracle.grid_to_boxes(occ, 0.05, (0, 0, 0), 3)
    m = bgr.OccupancyMap(0.05, (0, 0, 0), (n, n, n))
    m.set_boxes(boxes)
    for d in (1, 3):
        assert np.array_equal(_bits_to_occ(m.download_gate(d), n), _dilate_np(occ, d))
    assert m.dilation_stats() == {"full": 2, "incremental": 0}
    rng = np.random.default_rng(5)
    cur = occ.copy()
    for step in range(6):
        # remove some leaves (any level), add single cells and a few 2^3 / 4^3 blocks, some at the map border
        ridx = rng.choice(len(boxes), 20, replace=False) if step % 2 == 0 else np.zeros(0, dtype=np.int64)
        rem = boxes[ridx]
        for b in rem:
            s_ = int(round(b[3] / 0.05))
            x0, y0, z0 = (int(round(b[a] / 0.05 - s_ / 2)) for a in range(3))
            cur[z0:z0 + s_, y0:y0 + s_, x0:x0 + s_] = 0
        boxes = np.delete(boxes, ridx, axis=0)
        add = []
        for _ in range(20):
            L = int(rng.integers(0, 3))
            s_ = 1 << L
            c = (rng.integers(0, (n - s_) // s_ + 1, 3) * s_) if rng.random() < 0.8 else np.array([0, n - s_ - (n - s_) % s_, 0])
            x0, y0, z0 = int(c[0]), int(c[1]), int(c[2])
            if cur[z0:z0 + s_, y0:y0 + s_, x0:x0 + s_].any():
                continue
            cur[z0:z0 + s_, y0:y0 + s_, x0:x0 + s_] = 1
            add.append([(x0 + s_ / 2) * 0.05, (y0 + s_ / 2) * 0.05, (z0 + s_ / 2) * 0.05, s_ * 0.05, 1.0, 0.5])
        add = np.asarray(add, dtype=np.float64).reshape(-1, 6)
        m.update_boxes(add if len(add) else None, rem if len(rem) else None)
        boxes = np.concatenate([boxes, add])
        assert np.array_equal(_bits_to_occ(m.download_bitmap(), n), cur)
        for d in (1, 3):
            assert np.array_equal(_bits_to_occ(m.download_gate(d), n), _dilate_np(cur, d)), "step %d d %d" % (step, d)
    st = m.dilation_stats()
    assert st["full"] == 2 and st["incremental"] >= 12, st
    # refresh_boxes (whole-tree republish, diffed) keeps them current the same way
    m2 = bgr.OccupancyMap(0.05, (0, 0, 0), (n, n, n))
    m2.set_boxes(oracle.grid_to_boxes(occ, 0.05, (0, 0, 0), 3))
    m2.download_gate(1)
    occ_b = occ.copy()
    occ_b[n // 2:n // 2 + 6, 3:9, 5:11] ^= 1
    m2.refresh_boxes(oracle.grid_to_boxes(occ_b, 0.05, (0, 0, 0), 3))
    assert np.array_equal(_bits_to_occ(m2.download_gate(1), n), _dilate_np(occ_b, 1))
    assert m2.dilation_stats()["full"] == 1 and m2.dilation_stats()["incremental"] >= 1


def test_box_validation(bgr, oracle):
    m = bgr.OccupancyMap(0.05, (0, 0, 0), (64, 64, 64))
    with pytest.raises(RuntimeError):
        m.set_boxes(np.array([[0.03, 0.025, 0.025, 0.05, 1, 0.5]]))      # off-grid centre
    with pytest.raises(RuntimeError):
        m.set_boxes(np.array([[0.025, 0.025, 0.025, 0.07, 1, 0.5]]))     # size not res*2^L
    with pytest.raises(RuntimeError):
        m.set_boxes(np.array([[4.025, 0.025, 0.025, 0.05, 1, 0.5]]))     # outside
    m.set_boxes(np.array([[0.025, 0.025, 0.025, 0.05, 0.3, 0.5]]))       # below threshold: ignored
    assert m.occupied_cells() == 0
    # empty map => free (the reference returns a stale flag there; documented divergence)
    assert not m.isStateValid([1.0, 1.0, 1.0]) and not m.isLineValid([0.1, 0.1, 0.1], [3, 3, 3])


def _points_near_obstacles(O, occ, n, seed, spread):
    zz, yy, xx = np.nonzero(occ)
    mt = O.MT(seed)
    pick = (mt.real(n) * len(zz)).astype(np.int64)
    c = np.stack([xx[pick], yy[pick], zz[pick]], 1).astype(np.float64) * 0.05
    return (c + (mt.real(3 * n).reshape(n, 3) - 0.5) * spread).astype(np.float32)


@pytest.mark.parametrize("max_level", [0, 2])
def test_spheres_match_mpr_oracle(bgr, oracle, max_level):
    occ, boxes, m = _map(bgr, oracle, 64, 1, max_level)
    pts = np.concatenate([oracle.points(21, 3000, 3.2), _points_near_obstacles(oracle, occ, 3000, 22, 0.5)])
    hit, band = m.collide_spheres(pts, 0.1)
    ohit, oband = oracle.collide_spheres(boxes, pts, 0.1, mode=1)
    diff = hit != ohit
    assert not (diff & (band == 0)).any(), "mismatch outside the tolerance band: %d" % int((diff & (band == 0)).sum())
    assert np.array_equal(band, oband)
    # the sphere path is a closed-form distance, not FCL's MPR: inside the band (|gap| <= 1e-6) the two
    # may answer differently; counted and reported, never silently dropped
    print("spheres: %d in band, %d of them answered differently from the MPR restatement" % (int(band.sum()), int((diff & (band != 0)).sum())))
    assert 0.2 < hit.mean() < 0.98


def test_spheres_known_answers(bgr, oracle):
    m = bgr.OccupancyMap(0.05, (0, 0, 0), (64, 64, 64))
    box = np.array([[1.025, 1.025, 1.025, 0.05, 1.0, 0.5]])
    m.set_boxes(box)
    for sep, want in ((1e-3, 0), (1e-5, 0), (-1e-5, 1), (-1e-3, 1)):
        for d in np.eye(3):
            c = np.array([1.025, 1.025, 1.025]) + d * (0.025 + 0.1 + sep)
            c32 = c.astype(np.float32)
            gap = oracle.sep_sphere_box(c32.astype(np.float64), 0.1, box[0, :3], 0.05)
            hit, band = m.collide_spheres(c32.reshape(1, 3), 0.1)
            assert hit[0] == (1 if gap < 0 else 0)
            assert band[0] == (1 if abs(gap) <= 1e-6 else 0)
            if abs(sep) >= 1e-4:
                assert hit[0] == want == int(oracle.mpr_sphere_box(c32.astype(np.float64), 0.1, box[0, :3], 0.05)[0])


def _edges(O, occ, n, seed, lo=0.15, hi=3.0, extent=3.2):
    mt = O.MT(seed)
    p1 = _points_near_obstacles(O, occ, n, seed + 1, 1.0)
    u = mt.real(2 * n).reshape(n, 2)
    ct = 2 * u[:, 0] - 1
    st = np.sqrt(1 - ct * ct)
    ph = 2 * np.pi * u[:, 1]
    d = np.stack([st * np.cos(ph), st * np.sin(ph), ct], 1)
    ln = lo * (hi / lo) ** mt.real(n)
    p2 = (p1 + d * ln[:, None]).astype(np.float32)
    return p1, p2


@pytest.mark.parametrize("shape", [0, 1])
@pytest.mark.parametrize("max_level", [0, 2])
def test_edges_match_mpr_oracle(bgr, oracle, shape, max_level):
    occ, boxes, m = _map(bgr, oracle, 64, 1, max_level)
    p1, p2 = _edges(oracle, occ, 4000, 31)
    # axis-aligned, diagonal and -z (Eigen's antiparallel branch) edges
    extra1 = np.array([[0.3, 0.3, 0.3], [0.3, 0.3, 0.3], [0.3, 0.3, 2.9], [1.0, 1.0, 1.0]], dtype=np.float32)
    extra2 = np.array([[2.9, 0.3, 0.3], [2.9, 2.9, 2.9], [0.3, 0.3, 0.3], [1.0, 1.0, 1.0]], dtype=np.float32)
    p1, p2 = np.concatenate([p1, extra1]), np.concatenate([p2, extra2])
    hit, band = m.collide_edges(p1, p2, 0.03, shape)
    ohit, oband = oracle.collide_edges(boxes, p1, p2, 0.03, shape, mode=1)
    st = m.last_stats()
    assert st["iter_cap"] == 0
    diff = hit != ohit
    assert not (diff & (band == 0)).any(), "mismatch outside the band: %s" % np.nonzero(diff & (band == 0))[0][:10]
    assert not ((oband == 1) & (band == 0)).any(), "oracle-band contact not flagged"
    assert diff.sum() <= band.sum()
    # the band flag must not hide a real divergence: the device narrow phase is libccd's MPR in the
    # oracle's operation order (no FMA contraction), so flagged pairs agree as well unless the
    # iteration cap fired (none here)
    masked = int((diff & (band != 0)).sum())
    print("edges shape %d: %d in band, %d oracle-band, %d mismatches masked by the band flag" % (shape, int(band.sum()), int(oband.sum()), masked))
    assert masked == 0, "in-band mismatches: %s" % np.nonzero(diff & (band != 0))[0][:10]
    assert 0.1 < hit.mean() < 0.95


def test_edges_mode0_equals_mode1_oracle(oracle):
    """the oracle's conservative culling never changes the reference-literal loop's result"""
    occ = oracle.make_map(32, seed=4, n_walls=2)
    boxes = oracle.grid_to_boxes(occ, 0.05, (0, 0, 0), 2)
    p1, p2 = _edges(oracle, occ, 300, 41, 0.1, 1.5, 1.6)
    a, _ = oracle.collide_edges(boxes, p1, p2, 0.03, 0, mode=0, want_band=False)
    b, _ = oracle.collide_edges(boxes, p1, p2, 0.03, 0, mode=1, want_band=False)
    assert np.array_equal(a, b)


def test_edges_known_answers(bgr, oracle):
    """cylinder vs a single leaf at exact separations."""
    m = bgr.OccupancyMap(0.05, (0, 0, 0), (64, 64, 64))
    box = np.array([[1.025, 1.025, 1.025, 0.05, 1.0, 0.5]])
    m.set_boxes(box)
    for sep in (1e-3, 1e-5, -1e-5, -1e-3):
        # side contact: vertical edge passing the box at radial gap `sep`
        x = 1.025 + 0.025 + 0.03 + sep
        p1 = np.array([[x, 1.025, 0.5]], dtype=np.float32)
        p2 = np.array([[x, 1.025, 1.5]], dtype=np.float32)
        hit, band = m.collide_edges(p1, p2, 0.03, 0)
        ohit, oband = oracle.collide_edges(box, p1, p2, 0.03, 0, mode=0)
        assert hit[0] == ohit[0] or band[0]
        if abs(sep) >= 1e-4:
            assert hit[0] == (1 if sep < 0 else 0) and band[0] == 0
        # cap contact: flat cap facing the box (a capsule would reach further)
        z = 1.025 - 0.025 - sep
        p1 = np.array([[1.025, 1.025, 0.2]], dtype=np.float32)
        p2 = np.array([[1.025, 1.025, z]], dtype=np.float32)
        hit, band = m.collide_edges(p1, p2, 0.03, 0)
        ohit, _ = oracle.collide_edges(box, p1, p2, 0.03, 0, mode=0)
        assert hit[0] == ohit[0] or band[0]
        if abs(sep) >= 1e-4:
            assert hit[0] == (1 if sep < 0 else 0)
    # flat-capped cylinder vs capsule: a gap of 0.02 in front of the cap is free for the
    # cylinder (parity shape) and a hit for the capsule (conservative superset)
    p1 = np.array([[1.025, 1.025, 0.2]], dtype=np.float32)
    p2 = np.array([[1.025, 1.025, 0.98]], dtype=np.float32)
    assert m.collide_edges(p1, p2, 0.03, 0)[0][0] == 0
    assert m.collide_edges(p1, p2, 0.03, 1)[0][0] == 1
    assert m.isLineValid(p1[0], p2[0]) is False


def test_edges_device_pointers(bgr, oracle):
    import torch
    occ, boxes, m = _map(bgr, oracle, 64, 1, 2)
    p1, p2 = _edges(oracle, occ, 3000, 51)
    h0, b0 = m.collide_edges(p1, p2)
    d1, d2 = torch.from_numpy(p1).cuda(), torch.from_numpy(p2).cuda()
    h1, b1 = m.collide_edges(d1, d2)
    torch.cuda.synchronize()
    assert np.array_equal(h0, h1.cpu().numpy()) and np.array_equal(b0, b1.cpu().numpy())


def test_edges_large_map_properties(bgr, oracle):
    """256^3 map: order independence, idempotence, reversal symmetry outside the band."""
    occ = oracle.make_map(256, seed=9, n_walls=4, wall_thick=4, door_cells=6)
    boxes = oracle.grid_to_boxes(occ, 0.05, (0, 0, 0), 3)
    m = bgr.OccupancyMap(0.05, (0, 0, 0), (256, 256, 256))
    m.set_boxes(boxes)
    assert m.occupied_cells() == int(occ.sum())
    mt = oracle.MT(77)
    p1 = (mt.real(3 * 100000).reshape(-1, 3) * 12.8).astype(np.float32)
    p2, _ = _edges(oracle, occ, 100000, 78, extent=12.8)
    p2 = (p1 + (p2 - p2.mean(0)) * 0.2).astype(np.float32)
    h, b = m.collide_edges(p1, p2)
    h2, b2 = m.collide_edges(p1, p2)
    assert np.array_equal(h, h2) and np.array_equal(b, b2)
    perm = np.random.default_rng(1).permutation(len(p1))
    hp, bp = m.collide_edges(p1[perm], p2[perm])
    assert np.array_equal(hp, h[perm]) and np.array_equal(bp, b[perm])
    hr, br = m.collide_edges(p2, p1)
    bad = (hr != h) & (b == 0) & (br == 0)
    assert not bad.any()
    # oracle on a subsample
    sub = slice(0, 400)
    oh, ob = oracle.collide_edges(boxes, p1[sub], p2[sub], 0.03, 0, mode=1, want_band=False)
    assert not ((oh != h[sub]) & (b[sub] == 0)).any()


@pytest.mark.parametrize("shape", [0, 1])
@pytest.mark.parametrize("radius", [0.03, 0.0499, 0.07])
def test_edges_gate_is_exact(bgr, oracle, radius, shape):
    """the dilated-map gate of the sweep (d = 1 and 2 cells here) never changes hit / in_band:
    gate on == gate off, including edges that leave the map, hug its border, are shorter than
    a cell or longer than the 128 slabs the gate tracks, and after an incremental map update."""
    occ = oracle.make_map(192, seed=12, n_walls=3, wall_thick=3, door_cells=5)
    boxes = oracle.grid_to_boxes(occ, 0.05, (0, 0, 0), 3)
    m = bgr.OccupancyMap(0.05, (0, 0, 0), (192, 192, 192))
    m.set_boxes(boxes)
    n = 60000
    p1, p2 = _edges(oracle, occ, n, 91, lo=0.01, hi=4.0, extent=9.6)
    mt = oracle.MT(92)
    # a third of the edges start up to 0.1 m outside the map / on its border planes
    shift = (mt.real(3 * n).reshape(n, 3) * 0.25 - 0.1).astype(np.float32)
    side = mt.real(n)
    k = n // 3
    p1[:k] = np.where(side[:k, None] < 0.5, shift[:k], 9.6 - shift[:k] * 0.5) * np.array([1, 0, 0], np.float32) + p1[:k] * np.array([0, 1, 1], np.float32)
    # very long edges (> 128 slabs)
    p2[k:k + 2000] = (p1[k:k + 2000] + (p2[k:k + 2000] - p1[k:k + 2000]) * 3.0).astype(np.float32)
    m.set_gate(True)
    h1, b1 = m.collide_edges(p1, p2, radius, shape)
    g = m.gate_stats()
    st = m.last_stats()
    assert g["dilation_cells"] in {0.03: (1,), 0.0499: (1, 2), 0.07: (2,)}[radius]
    assert 0 < g["rect_items"] < st["work_items"]
    m.set_gate(False)
    h0, b0 = m.collide_edges(p1, p2, radius, shape)
    assert m.gate_stats()["rect_items"] == m.last_stats()["work_items"]
    assert np.array_equal(h1, h0) and np.array_equal(b1, b0)
    assert 0.05 < h0.mean() < 0.95
    # incremental update: the dilated map must follow the level-0 change
    add = np.array([[4.825 + 0.05 * i, 4.825, 4.825, 0.05, 1.0, 0.5] for i in range(4)])
    rem = boxes[:5000]
    m.update_boxes(add, rem)
    h0u, b0u = m.collide_edges(p1, p2, radius, shape)
    m.set_gate(True)
    h1u, b1u = m.collide_edges(p1, p2, radius, shape)
    assert np.array_equal(h1u, h0u) and np.array_equal(b1u, b0u)
    assert not np.array_equal(h0u, h0)


@pytest.mark.parametrize("radius", [0.03, 0.1, 0.149])
def test_spheres_gate_is_exact(bgr, oracle, radius):
    """the dilated-map gate of the sphere check never changes hit / in_band (gate on == off),
    including centres outside the map and after an incremental update."""
    occ, boxes, m = _map(bgr, oracle, 96, 5, 3, n_walls=2, wall_thick=3, door_cells=5)
    mt = oracle.MT(55)
    pts = np.concatenate([_points_near_obstacles(oracle, occ, 40000, 56, 0.6),
                          (mt.real(3 * 20000).reshape(-1, 3) * 5.2 - 0.2).astype(np.float32)])
    m.set_gate(True)
    h1, b1 = m.collide_spheres(pts, radius)
    m.set_gate(False)
    h0, b0 = m.collide_spheres(pts, radius)
    assert np.array_equal(h1, h0) and np.array_equal(b1, b0)
    assert 0.05 < h0.mean() < 0.95
    m.update_boxes(np.array([[2.425, 2.425, 2.425, 0.05, 1.0, 0.5]]), boxes[:3000])
    h0u, b0u = m.collide_spheres(pts, radius)
    m.set_gate(True)
    h1u, b1u = m.collide_spheres(pts, radius)
    assert np.array_equal(h1u, h0u) and np.array_equal(b1u, b0u)
    assert not np.array_equal(h0u, h0)


@pytest.mark.parametrize("radius", [0.0, 0.03, 0.05, 0.1, 0.149, 0.2])
def test_spheres_lane_form_equals_warp_forms(bgr, oracle, radius, monkeypatch):
    """the lane-per-sphere kernel (large batches of small spheres) writes the same hit / in_band bytes as the two
    warp forms — gate on and off, centres outside / on the border of the map, after an incremental update — and
    agrees with the oracle's MPR loop outside the band.  r = 0.2 spans more cells than the lane form takes: the
    request falls back to the warp form."""
    occ, boxes, m = _map(bgr, oracle, 96, 5, 3, n_walls=2, wall_thick=3, door_cells=5)
    mt = oracle.MT(77)
    pts = np.concatenate([_points_near_obstacles(oracle, occ, 30000, 78, 0.6),
                          (mt.real(3 * 20000).reshape(-1, 3) * 5.4 - 0.3).astype(np.float32),
                          np.array([[0.0, 0.0, 0.0], [4.8, 4.8, 4.8], [-0.05, 2.0, 2.0], [4.85, 2.0, 2.0], [2.4, 2.4, 1e6]], dtype=np.float32)])
    res = {}
    for gate in (True, False):
        m.set_gate(gate)
        for form in ("lane", "warp32", "warp1"):
            monkeypatch.setenv("BGR_SPHERES", form)
            res[(gate, form)] = m.collide_spheres(pts, radius)
    ref = res[(True, "warp32")]
    for key, (h, b) in res.items():
        assert np.array_equal(h, ref[0]) and np.array_equal(b, ref[1]), key
    if radius > 0:
        assert 0.02 < ref[0].mean() < 0.98
    sub = np.arange(0, len(pts), 37)
    oh, _ = oracle.collide_spheres(boxes, pts[sub], radius, mode=1, want_band=False)
    assert not ((oh != ref[0][sub]) & (ref[1][sub] == 0)).any()
    m.set_gate(True)
    m.update_boxes(np.array([[2.425, 2.425, 2.425, 0.05, 1.0, 0.5]]), boxes[:3000])
    upd = {}
    for form in ("lane", "warp32"):
        monkeypatch.setenv("BGR_SPHERES", form)
        upd[form] = m.collide_spheres(pts, radius)
    assert np.array_equal(upd["lane"][0], upd["warp32"][0]) and np.array_equal(upd["lane"][1], upd["warp32"][1])
    if radius > 0:
        assert not np.array_equal(upd["lane"][0], ref[0])


def test_edges_chunked_host_call_equals_device_call(bgr, oracle):
    """a large host-pointer batch is cut into chunks whose H2D copies overlap the sweep of the
    previous chunk: same hit / in_band bytes and same counters as the one-launch device call."""
    import torch
    occ, boxes, m = _map(bgr, oracle, 128, 3, 2, n_walls=3, wall_thick=3, door_cells=5)
    p1, p2 = _edges(oracle, occ, 200001, 61, lo=0.05, hi=2.0, extent=6.4)
    h0, b0 = m.collide_edges(p1, p2)
    st0 = m.last_stats()
    d1, d2 = torch.from_numpy(p1).cuda(), torch.from_numpy(p2).cuda()
    h1, b1 = m.collide_edges(d1, d2)
    torch.cuda.synchronize()
    st1 = m.last_stats()
    assert np.array_equal(h0, h1.cpu().numpy()) and np.array_equal(b0, b1.cpu().numpy())
    assert st0["work_items"] == st1["work_items"] and st0["candidates"] == st1["candidates"]
    hp = torch.from_numpy(p1).pin_memory().numpy(), torch.from_numpy(p2).pin_memory().numpy()
    h2, b2 = m.collide_edges(hp[0], hp[1])
    assert np.array_equal(h0, h2) and np.array_equal(b0, b2)


def test_edges_full_size_properties(bgr, oracle):
    """BASELINE config 3 at full size (1,000,000 cylinder edges vs the 512^3 map, ~6.5 M leaf
    boxes): gate on == gate off, chunked host call == device call, reversal symmetry outside the
    band, and the oracle's booleans for all 1,000,000 edges."""
    import torch
    from bottleneck_guided_rrt_b200 import synth
    occ = synth.make_map(512, 4, n_walls=8, wall_thick=4, door_cells=5)
    boxes = synth.grid_to_boxes(occ, 0.05, (0, 0, 0), 3)
    m = bgr.OccupancyMap(0.05, (0, 0, 0), (512, 512, 512))
    m.set_boxes(boxes)
    assert m.occupied_cells() == int(occ.sum())
    p1, p2 = synth.edges(4, 1_000_000, occ)
    d1, d2 = torch.from_numpy(p1).cuda(), torch.from_numpy(p2).cuda()
    h, b = m.collide_edges(d1, d2)
    torch.cuda.synchronize()
    h, b = h.cpu().numpy(), b.cpu().numpy()
    st, g = m.last_stats(), m.gate_stats()
    assert st["iter_cap"] == 0 and 0 < g["rect_items"] < st["work_items"] // 4
    assert 0.3 < h.mean() < 0.7 and b.sum() < 1000
    m.set_gate(False)
    h0, b0 = m.collide_edges(d1, d2)
    torch.cuda.synchronize()
    assert np.array_equal(h, h0.cpu().numpy()) and np.array_equal(b, b0.cpu().numpy())
    m.set_gate(True)
    hh, bh = m.collide_edges(p1, p2)                       # host pointers: 4 chunks
    assert np.array_equal(h, hh) and np.array_equal(b, bh)
    hr, br = m.collide_edges(d2, d1)
    torch.cuda.synchronize()
    hr, br = hr.cpu().numpy(), br.cpu().numpy()
    assert not ((hr != h) & (b == 0) & (br == 0)).any()
    # the oracle on EVERY edge (its bucketed broadphase in front of the same MPR: mode 2, equal to the
    # reference-literal loop by tests/test_oracle.py) and the literal loop itself on a subsample
    oh, _ = oracle.collide_edges(boxes, p1, p2, 0.03, 0, mode=2, want_band=False)
    bad = (oh != h) & (b == 0)
    assert not bad.any(), np.nonzero(bad)[0][:10]
    assert (oh != h).sum() <= b.sum()
    assert int(((oh != h) & (b != 0)).sum()) == 0, "in-band mismatches hidden by the band flag"
    sub = np.arange(0, 1_000_000, 10000)
    ol, _ = oracle.collide_edges(boxes, p1[sub], p2[sub], 0.03, 0, mode=1, want_band=False)
    assert np.array_equal(ol, oh[sub])


def test_spheres_full_size_properties(bgr, oracle, monkeypatch):
    """1,000,000 state checks (r = 0.1) against the 512^3 map: gate on == gate off, the lane-per-sphere form ==
    the warp forms, chunk / permutation independence, and the oracle's MPR loop on a subsample."""
    import torch
    from bottleneck_guided_rrt_b200 import synth
    occ = synth.make_map(512, 4, n_walls=8, wall_thick=4, door_cells=5)
    boxes = synth.grid_to_boxes(occ, 0.05, (0, 0, 0), 3)
    m = bgr.OccupancyMap(0.05, (0, 0, 0), (512, 512, 512))
    m.set_boxes(boxes)
    pts = synth.points(21, 1_000_000, 25.6)
    dp = torch.from_numpy(pts).cuda()
    h, b = m.collide_spheres(dp, 0.1)
    torch.cuda.synchronize()
    h, b = h.cpu().numpy(), b.cpu().numpy()
    assert 0.3 < h.mean() < 0.99
    m.set_gate(False)
    h0, b0 = m.collide_spheres(dp, 0.1)
    torch.cuda.synchronize()
    assert np.array_equal(h, h0.cpu().numpy()) and np.array_equal(b, b0.cpu().numpy())
    m.set_gate(True)
    perm = np.random.default_rng(3).permutation(len(pts))[:100000]       # another batch size, another order
    hp, bp = m.collide_spheres(pts[perm], 0.1)
    assert np.array_equal(hp, h[perm]) and np.array_equal(bp, b[perm])
    hs, bs = m.collide_spheres(pts[perm[:10000]], 0.1)                   # small batch: a warp per sphere
    assert np.array_equal(hs, h[perm[:10000]]) and np.array_equal(bs, b[perm[:10000]])
    monkeypatch.setenv("BGR_SPHERES", "warp32")                          # the 32-spheres-per-warp form on the full batch
    hw, bw = m.collide_spheres(dp, 0.1)
    torch.cuda.synchronize()
    assert np.array_equal(hw.cpu().numpy(), h) and np.array_equal(bw.cpu().numpy(), b)
    monkeypatch.delenv("BGR_SPHERES")
    hh, bh = m.collide_spheres(pts, 0.1)                                 # host buffers: chunked, copies overlapped
    assert np.array_equal(hh, h) and np.array_equal(bh, b)
    sub = np.arange(0, 1_000_000, 8000)
    oh, _ = oracle.collide_spheres(boxes, pts[sub], 0.1, mode=1, want_band=False)
    assert not ((oh != h[sub]) & (b[sub] == 0)).any()


def test_octomap_ingestion_and_refresh(bgr, oracle):
    """serialised Octomap (full message data, MS:279) -> toBoxes rows -> bitmap; then a map refresh
    that rasterises only the changed leaves (bgr_map_refresh_boxes)."""
    from oracle import octree_io as T
    occ = oracle.make_map(64, seed=1)
    res, key0 = 0.05, (T.ORIGIN + 32, T.ORIGIN, T.ORIGIN - 64)
    root, _ = T.tree_from_grid(occ, key0)
    m, boxes = bgr.OccupancyMap.from_octomap(T.write_full(root), bgr.OCTREE_FULL_DATA, res)
    assert m.occupied_cells() == int(occ.sum())
    off = np.round((np.asarray([(k - T.ORIGIN) * res for k in key0]) - m.origin) / res).astype(int)
    got = _bits_to_occ(m.download_bitmap(), int(m.dims.max()))[: m.dims[2], : m.dims[1], : m.dims[0]] if len(set(m.dims)) == 1 else None
    full = np.zeros((m.dims[2], m.dims[1], m.dims[0]), dtype=np.uint8)
    full[off[2]:off[2] + 64, off[1]:off[1] + 64, off[0]:off[0] + 64] = occ
    if got is not None:
        assert np.array_equal(got, full)
    # collision answers through the ingested map == through the directly flattened grid
    origin = [(k - T.ORIGIN) * res for k in key0]
    m2 = bgr.OccupancyMap(res, origin, (64, 64, 64))
    m2.set_boxes(oracle.grid_to_boxes(occ, res, origin, 2))
    rng = np.random.default_rng(3)
    p1 = (rng.uniform(0.2, 3.0, (4000, 3)) + origin).astype(np.float32)
    p2 = (p1 + rng.normal(0, 0.3, p1.shape)).astype(np.float32)
    p2 = np.clip(p2, np.asarray(origin, np.float32) + 0.1, np.asarray(origin, np.float32) + 3.1)
    h1, b1 = m.collide_edges(p1, p2)
    h2, b2 = m2.collide_edges(p1, p2)
    same = (b1 == 0) & (b2 == 0)
    assert np.array_equal(h1[same], h2[same]) and same.mean() > 0.99
    # refresh: second map differs in a few blobs only
    occ2 = occ.copy()
    occ2[10:14, 10:14, 10:14] = 1
    occ2[40:44, 40:48, 40:44] = 0
    root2, _ = T.tree_from_grid(occ2, key0)
    boxes2 = bgr.octree_to_boxes(T.write_full(root2), bgr.OCTREE_FULL_DATA, res)[0]
    added, removed = m.refresh_boxes(boxes2)
    assert 0 < added + removed < len(boxes2) // 4
    assert m.occupied_cells() == int(occ2.sum())
    fresh = bgr.OccupancyMap(res, m.origin, m.dims)
    fresh.set_boxes(boxes2)
    assert np.array_equal(m.download_bitmap(), fresh.download_bitmap())
    h3, b3 = m.collide_edges(p1, p2)
    h4, b4 = fresh.collide_edges(p1, p2)
    assert np.array_equal(h3, h4) and np.array_equal(b3, b4)
    assert m.refresh_boxes(boxes2) == (0, 0)
