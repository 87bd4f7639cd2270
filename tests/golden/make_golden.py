#!/usr/bin/env python
"""Regenerates the committed golden fixtures under tests/golden/.  Run in the BUILD container
(it needs /root/reference for two of the fixtures); the fixtures, not this script, travel to
the GPU box.

  mt19937.json      outputs of the REFERENCE's own abb_manipulator/src/mtrand.cpp, compiled
                    where it lies into oracle/_ref/libmtrand_ref.so (oracle/Makefile `ref`).
  milestones.npz    the reference's labelled_data/{outputs,start_end_points}/*.npy (202 CNN
                    milestone triples + start/goal pairs) packed verbatim — the "3 CNN
                    milestone points" consumed unchanged by the planner (MS:249-271, 1213-1215).
  knn_numpy.npz     libnabo leaf arithmetic restated INDEPENDENTLY of oracle/ in numpy float32
                    (every op rounded, no FMA): d2 = ((dx*dx)+dy*dy)+dz*dz, admission rule,
                    ascending (d2, index) order — plus Nearest()/Norm() in the reference's
                    mixed precision (MS:130-154).
  collide_indep.npz sphere/cylinder-vs-box separation distances from solvers that share no
                    code with oracle/: closed-form point-AABB distance (sphere) and a bounded
                    L-BFGS-B minimisation of the convex point-to-cylinder distance over the
                    box (cylinder).  Labels: hit if sep < 1e-7, miss if sep > 1e-5, else band.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference"


def gen_mt():
    from oracle import oracle as O
    r = O.RefMT()
    out = {"source": "abb_manipulator/src/mtrand.cpp + include/mtrand.h compiled unmodified (oracle/_ref)"}
    r.seed_array([0x123, 0x234, 0x345, 0x456])
    out["init_by_array_0x123_0x234_0x345_0x456_u32_first16"] = [int(r.u32()) for _ in range(16)]
    r.seed_array([0x123, 0x234, 0x345, 0x456])
    out["init_by_array_real_first8_hex"] = [float(r.real()).hex() for _ in range(8)]
    r.seed(5489)
    v = [int(r.u32()) for _ in range(10000)]
    out["seed5489_u32_first8"] = v[:8]
    out["seed5489_u32_10000th"] = v[9999]
    seeds = {}
    for s in (0, 1, 2, 3, 4, 5, 6, 100, 4195, 0xFFFFFFFF):
        r.seed(s)
        a = [int(r.u32()) for _ in range(700)]     # crosses the 624-word regeneration boundary
        seeds[str(s)] = {"first4": a[:4], "at623_626": a[623:627], "xor700": int(np.bitwise_xor.reduce(np.array(a, dtype=np.uint32)))}
    out["seeds"] = seeds
    with open(os.path.join(HERE, "mt19937.json"), "w") as f:
        json.dump(out, f, indent=1)


def gen_milestones():
    d = os.path.join(REF, "labelled_data")
    ids = sorted(os.path.splitext(x)[0] for x in os.listdir(os.path.join(d, "outputs")))
    ms = np.stack([np.load(os.path.join(d, "outputs", i + ".npy")).reshape(9) for i in ids]).astype(np.float32)
    se = np.stack([np.load(os.path.join(d, "start_end_points", i + ".npy")).reshape(6) for i in ids]).astype(np.float64)
    occ_cells = np.array([int(np.count_nonzero(np.load(os.path.join(d, "inputs", i + ".npy")))) for i in ids], dtype=np.int32)
    np.savez_compressed(os.path.join(HERE, "milestones.npz"), ids=np.array(ids), milestones=ms, start_end=se,
                        occupied_cells=occ_cells)


def np_knn(pts, q, k, self_match, max_r2):
    """independent numpy-f32 restatement; returns idx (Q,k) i32, d2 (Q,k) f32."""
    f = np.float32
    idx = np.full((len(q), k), -1, dtype=np.int32)
    d2o = np.full((len(q), k), np.inf, dtype=np.float32)
    for i in range(len(q)):
        dx = (q[i, 0] - pts[:, 0]).astype(f)
        dy = (q[i, 1] - pts[:, 1]).astype(f)
        dz = (q[i, 2] - pts[:, 2]).astype(f)
        d2 = ((dx * dx).astype(f) + (dy * dy).astype(f)).astype(f)
        d2 = (d2 + (dz * dz).astype(f)).astype(f)
        ok = d2 <= max_r2[i]
        if not self_match:
            ok &= d2 > np.finfo(np.float32).eps
        cand = np.nonzero(ok)[0]
        order = np.lexsort((cand, d2[cand]))[:k]
        idx[i, :len(order)] = cand[order]
        d2o[i, :len(order)] = d2[cand][order]
    return idx, d2o


def np_nearest(V, q):
    """MS:130-154: dx in f32; pow/sum/sqrt in f64; result f32; `<=` keeps the LAST index."""
    idx = np.empty(len(q), dtype=np.int32)
    dist = np.empty(len(q), dtype=np.float32)
    for i in range(len(q)):
        d = (V - q[i]).astype(np.float32).astype(np.float64)          # x2 - x1 in float
        s = (d[:, 0] ** 2 + d[:, 1] ** 2) + d[:, 2] ** 2
        n = np.sqrt(s).astype(np.float32)
        m = n.min()
        idx[i] = np.nonzero(n == m)[0][-1]
        dist[i] = m
    return idx, dist


def gen_knn():
    from bottleneck_guided_rrt_b200 import synth
    pts = synth.points(2, 3000, 3.2)
    q = np.concatenate([synth.points(3, 80, 3.2), pts[:16]])                    # 16 self queries
    g = np.arange(7, dtype=np.float32)
    lat = np.stack(np.meshgrid(g, g, g, indexing="ij"), -1).reshape(-1, 3)
    lat = lat[np.random.default_rng(7).permutation(len(lat))]
    lq = np.concatenate([lat[:24] + np.float32(0.5), lat[24:48]])
    inf = np.full(len(q), np.inf, dtype=np.float32)
    out = {"pts": pts, "q": q, "lat": lat, "lq": lq}
    out["idx_k16"], out["d2_k16"] = np_knn(pts, q, 16, False, inf)
    out["idx_k16_self"], out["d2_k16_self"] = np_knn(pts, q, 16, True, inf)
    r = np.float32(0.3)
    out["idx_k8_r03"], out["d2_k8_r03"] = np_knn(pts, q, 8, False, np.full(len(q), r * r, dtype=np.float32))
    linf = np.full(len(lq), np.inf, dtype=np.float32)
    out["lat_idx_k27"], out["lat_d2_k27"] = np_knn(lat, lq, 27, False, linf)
    out["lat_idx_k6_self"], out["lat_d2_k6_self"] = np_knn(lat, lq, 6, True, linf)
    out["near_idx"], out["near_dist"] = np_nearest(pts, q)
    out["lat_near_idx"], out["lat_near_dist"] = np_nearest(lat, lq)
    np.savez_compressed(os.path.join(HERE, "knn_numpy.npz"), **out)


def cyl_box_sep(p1, p2, r, bc, size):
    from scipy.optimize import minimize
    p1, p2 = np.asarray(p1, np.float64), np.asarray(p2, np.float64)
    c = 0.5 * (p1 + p2)
    ax = p2 - p1
    h = 0.5 * np.linalg.norm(ax)
    u = ax / (2 * h)

    def f(y):
        w = y - c
        a = w @ u
        e = w - a * u
        rho = np.linalg.norm(e)
        da, dq = max(0.0, abs(a) - h), max(0.0, rho - r)
        val = da * da + dq * dq
        g = np.zeros(3)
        if da > 0:
            g += 2 * da * np.sign(a) * u
        if dq > 0:
            g += 2 * dq * e / rho
        return val, g

    lo, hi = bc - size / 2, bc + size / 2
    best = np.inf
    for y0 in (np.clip(c, lo, hi), np.clip(p1, lo, hi), np.clip(p2, lo, hi)):
        res = minimize(f, y0, jac=True, bounds=list(zip(lo, hi)), method="L-BFGS-B",
                       options={"ftol": 1e-30, "gtol": 1e-14, "maxiter": 500})
        best = min(best, res.fun)
    return float(np.sqrt(max(best, 0.0)))


def gen_collide():
    rng = np.random.default_rng(11)
    n = 600
    bc = rng.uniform(0.5, 2.5, (n, 3))
    size = 0.05 * (2.0 ** rng.integers(0, 4, n))
    # spheres: centres near the box surface
    sc = (bc + rng.normal(0, 1, (n, 3)) * (size[:, None] * 0.5 + 0.1)).astype(np.float32)
    t = np.maximum(0.0, np.abs(sc.astype(np.float64) - bc) - size[:, None] / 2)
    s_sep = np.maximum(0.0, np.sqrt((t ** 2).sum(1)) - 0.1)
    # cylinders: edges passing near the box
    d = rng.normal(0, 1, (n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    ln = 0.15 * (3.0 / 0.15) ** rng.uniform(0, 1, n)
    off = rng.normal(0, 1, (n, 3)) * (size[:, None] * 0.5 + 0.03)
    mid = bc + off + d * (rng.uniform(-0.6, 0.6, (n, 1)) * ln[:, None])
    p1 = (mid - 0.5 * d * ln[:, None]).astype(np.float32)
    p2 = (mid + 0.5 * d * ln[:, None]).astype(np.float32)
    c_sep = np.array([cyl_box_sep(p1[i], p2[i], 0.03, bc[i], size[i]) for i in range(n)])
    np.savez_compressed(os.path.join(HERE, "collide_indep.npz"), box_c=bc, box_size=size, sphere_c=sc,
                        sphere_r=0.1, sphere_sep=s_sep, p1=p1, p2=p2, cyl_r=0.03, cyl_sep=c_sep)


if __name__ == "__main__":
    what = sys.argv[1:] or ["mt", "milestones", "knn", "collide"]
    for w in what:
        {"mt": gen_mt, "milestones": gen_milestones, "knn": gen_knn, "collide": gen_collide}[w]()
        print("wrote", w)
