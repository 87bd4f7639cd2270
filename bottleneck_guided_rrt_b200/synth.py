"""Synthetic workloads of the named BASELINE.json shapes, numpy only.

These are input generators for the benchmark, the smoke test and the examples — not part
of the oracle.  The sample stream is MT19937 seeded like the reference's ``MTRand``
(abb_manipulator/src/mtrand.cpp:21-31), drawn as ``MTRand::operator()`` does
(include/mtrand.h:103-104): u32 * 2^-32.
"""
import numpy as np


def mt_u32(seed, n):
    bg = np.random.MT19937()
    bg._legacy_seeding(int(seed) & 0xFFFFFFFF)
    return bg.random_raw(int(n)).astype(np.uint32)


def mt_real(seed, n):
    return mt_u32(seed, n).astype(np.float64) * (1.0 / 4294967296.0)


def points(seed, n, scale):
    """n points U[0,scale)^3 as float32 (n,3): MTRand double -> *scale -> f32."""
    return (mt_real(seed, 3 * n) * scale).astype(np.float32).reshape(n, 3)


def make_map(n, seed=1, n_walls=3, wall_thick=2, door_cells=6, occupancy=0.10):
    """n^3 occupancy grid [z][y][x] (uint8): wall slabs normal to x, each pierced by one
    door_cells^2 door, plus clutter cubes (side 1..4 cells) up to ~`occupancy`; a small cube
    is kept free at both corners (start / goal regions)."""
    r = mt_real(seed, 2 * n_walls + 8_000_000)
    occ = np.zeros((n, n, n), dtype=np.uint8)
    p = 0
    for w in range(n_walls):
        x0 = (w + 1) * n // (n_walls + 1) - wall_thick // 2
        dy = int(r[p] * (n - door_cells)); dz = int(r[p + 1] * (n - door_cells)); p += 2
        sl = occ[:, :, max(0, x0):min(n, x0 + wall_thick)]
        sl[:] = 1
        sl[dz:dz + door_cells, dy:dy + door_cells, :] = 0
    total = n ** 3
    target = int(occupancy * total)
    m = min(6, n // 4)
    while True:
        filled = int(occ.sum())
        if filled >= target:
            break
        need = max(64, int((target - filled) / 20))
        if p + 4 * need > len(r):
            r = mt_real(seed + 7919 + p, 4 * need + 1024)
            p = 0
        s = 1 + (r[p:p + need] * 4).astype(np.int64)
        x0 = (r[p + need:p + 2 * need] * (n - s + 1)).astype(np.int64)
        y0 = (r[p + 2 * need:p + 3 * need] * (n - s + 1)).astype(np.int64)
        z0 = (r[p + 3 * need:p + 4 * need] * (n - s + 1)).astype(np.int64)
        p += 4 * need
        for side in (1, 2, 3, 4):
            sel = s == side
            if not sel.any():
                continue
            for dz in range(side):
                for dy in range(side):
                    for dx in range(side):
                        occ[z0[sel] + dz, y0[sel] + dy, x0[sel] + dx] = 1
        occ[:m, :m, :m] = 0
        occ[n - m:, n - m:, n - m:] = 0
    return occ


def grid_to_boxes(occ, res=0.05, origin=(0.0, 0.0, 0.0), max_level=2):
    """fcl::OcTree::toBoxes() view of an occupancy grid: rows {x,y,z,size,occupancy,threshold}
    with octree-aligned pruning (a full aligned 2^L block becomes one leaf, L <= max_level)."""
    occ = np.ascontiguousarray(occ, dtype=bool)
    nz, ny, nx = occ.shape
    side = 1
    while side < max(nx, ny, nz):
        side *= 2
    full = [np.zeros((side, side, side), dtype=bool)]
    full[0][:nz, :ny, :nx] = occ
    L = 0
    while L < max_level and full[L].shape[0] > 1:
        f = full[L]
        s = f.shape[0] // 2
        f8 = f.reshape(s, 2, s, 2, s, 2)
        full.append(f8.all(axis=(1, 3, 5)))
        L += 1
    out = []
    for L in range(len(full)):
        f = full[L]
        if L + 1 < len(full):
            parent = np.repeat(np.repeat(np.repeat(full[L + 1], 2, 0), 2, 1), 2, 2)
            f = f & ~parent
        z, y, x = np.nonzero(f)
        if len(z) == 0:
            continue
        sz = res * (1 << L)
        b = np.empty((len(z), 6), dtype=np.float64)
        half = 0.5 * (1 << L)
        b[:, 0] = origin[0] + ((x.astype(np.float64) * (1 << L)) + half) * res
        b[:, 1] = origin[1] + ((y.astype(np.float64) * (1 << L)) + half) * res
        b[:, 2] = origin[2] + ((z.astype(np.float64) * (1 << L)) + half) * res
        b[:, 3] = sz
        b[:, 4] = 0.97
        b[:, 5] = 0.5
        out.append(b)
    return np.concatenate(out) if out else np.zeros((0, 6))


def edges(seed, n, occ, res=0.05, lo=0.15, hi=3.0):
    """n edges: p1 uniform in free cells, direction uniform on the sphere, length
    log-uniform in [lo, hi] (eta-steps to prune-length).  Returns float32 (n,3),(n,3)."""
    nz, ny, nx = occ.shape
    r = mt_real(seed, 8 * n)
    cand = np.stack([r[:2 * n] * nx, r[2 * n:4 * n] * ny, r[4 * n:6 * n] * nz], 1)
    ci = cand.astype(np.int64)
    free = occ[ci[:, 2], ci[:, 1], ci[:, 0]] == 0
    p1 = cand[free][:n]
    if len(p1) < n:
        p1 = np.concatenate([p1, cand[:n - len(p1)]])
    p1 = p1 * res
    u = r[6 * n:8 * n].reshape(n, 2)
    ct = 2 * u[:, 0] - 1
    st = np.sqrt(np.maximum(0.0, 1 - ct * ct))
    ph = 2 * np.pi * u[:, 1]
    d = np.stack([st * np.cos(ph), st * np.sin(ph), ct], 1)
    ln = lo * (hi / lo) ** mt_real(seed + 1, n)
    p2 = p1 + d * ln[:, None]
    return p1.astype(np.float32), p2.astype(np.float32)
