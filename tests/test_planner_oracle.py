"""CPU suite: the oracle's planner loop against a literal pure-Python walk of the reference's
sequential iteration (rrt_star.cpp RS:1140-1300 with batch = 1) built from the oracle's own
primitives — checks the loop STRUCTURE (draw order, V.size() > K guard, parent rule, `found`
quirk, goal node), not the primitives, which test_oracle.py pins."""
import numpy as np


def _chain_cost(O, xyz, par, P, child):
    cost = np.float32(0) + np.float32(O.norm(xyz[P], child))
    cur, l, prev = par[P], 0, 0
    while cur > 0:
        cost = np.float32(cost + np.float32(O.norm(xyz[cur], xyz[P] if l == 0 else xyz[prev])))
        l, prev, cur = 1, cur, par[cur]
    return cost


def test_sequential_loop_literal(oracle, bgr):
    from bottleneck_guided_rrt_b200 import synth
    O = oracle
    occ = synth.make_map(64, 1, occupancy=0.0)
    boxes = synth.grid_to_boxes(occ, 0.05, (0, 0, 0), 2)
    table = synth.points(9, 5000, 3.2)
    start, goal = (0.2, 0.2, 0.2), (1.0, 0.6, 0.5)
    P = bgr.PlannerParams(start, goal, mode=bgr.PLAN_MODE_RS, k=5, batch=1, flags=bgr.PLAN_QUIRKS, seed=99, goal_chance=0.2)
    op = O.Planner(P, boxes, table)
    op.run(200, 500)
    oxyz, opar = op.tree()

    mt = O.MT(99)
    xyz, par = [np.array(start, np.float32)], [-1]
    goal_f = np.array(goal, np.float32)
    for _ in range(500):
        if len(xyz) >= 200:
            break
        rc = np.float32(mt.std_uniform(0.0, 1.0))
        ri = int(np.float32(int(mt.std_uniform(0.0, float(len(table))))))
        prand = table[ri] if rc > np.float32(0.2) else goal_f
        V = np.stack(xyz)
        ni, _ = O.nearest(V, prand[None])
        ni = int(ni[0])
        pnew, found = O.steer(V[ni], prand, 0.15)
        hit, _ = O.collide_spheres(boxes, pnew[None], 0.1, want_band=False)
        if hit[0]:
            continue
        parent = ni
        if len(xyz) > 5:
            idx, _ = O.knn_brute(V, pnew[None], 5)
            lowest, pc = _chain_cost(O, V, par, ni, pnew), -5
            for ind in idx[0]:
                t = _chain_cost(O, V, par, int(ind), pnew)
                if t < lowest:
                    pc, lowest = int(ind), t
            if pc > 0:
                parent = pc
        xyz.append(pnew)
        par.append(ni if found else parent)          # MS:1236-1238 quirk
        if np.float32(O.norm(goal_f, pnew)) < np.float32(0.4):
            par.append(len(xyz) - 1)
            xyz.append(goal_f)
            break
    assert len(xyz) == len(opar)
    assert np.array_equal(np.stack(xyz).view(np.uint32), oxyz.view(np.uint32))
    assert par == opar.tolist()


def test_batched_rewiring_stays_a_tree(oracle, bgr):
    """beyond-reference rewiring, many proposals applied per batch: the reference's cost skips the
    edge into the root (cost_wiring stops at index <= 0), so cost alone would let a depth-1 node
    adopt its own descendant, and two proposals of one batch could close a loop between them.  The
    psi rule (true path length must drop too) keeps every batch acyclic: every node still reaches
    the root, in fewer steps than there are nodes."""
    from bottleneck_guided_rrt_b200 import synth
    occ = synth.make_map(64, 1, occupancy=0.0)
    boxes = synth.grid_to_boxes(occ, 0.05, (0, 0, 0), 3)
    table = synth.points(9, 20000, 3.2)
    P = bgr.PlannerParams((0.2, 0.2, 0.2), (3.0, 3.0, 3.0), mode=bgr.PLAN_MODE_RS, k=12, batch=512,
                          flags=bgr.PLAN_REWIRE | bgr.PLAN_NO_GOAL, seed=5)
    op = oracle.Planner(P, boxes, table)
    op.run(12000, 60)
    xyz, par = op.tree()
    st = op.stats()
    assert st["rewired"] > 100 and par[0] == -1
    depth = np.zeros(len(par), dtype=np.int64)
    cur = np.arange(len(par))
    for _ in range(len(par)):
        alive = cur > 0
        if not alive.any():
            break
        depth[alive] += 1
        cur = np.where(alive, par[np.maximum(cur, 0)], cur)
    assert not (cur > 0).any(), "a parent chain does not reach the root: rewiring closed a cycle"
    assert depth.max() < len(par)


def test_dedup_keeps_one_candidate_per_fixed_sample(oracle, bgr):
    """ORC_PLAN_DEDUP (beyond reference): the goal / milestone samples of one batch are the same
    point against the same snapshot; with the flag each batch appends at most one node per fixed
    sample, so nodes of one batch no longer coincide, and the table-sample nodes are the ones the
    plain run would have appended in its first batch."""
    from bottleneck_guided_rrt_b200 import synth
    occ = synth.make_map(64, 1, occupancy=0.0)
    boxes = synth.grid_to_boxes(occ, 0.05, (0, 0, 0), 3)
    table = synth.points(9, 20000, 3.2)
    ms = np.array([[1.0, 0.5, 1.5], [2.0, 1.0, 2.0], [2.5, 2.0, 1.0]], np.float32)   # not collinear with start -> goal

    def grow(flags, batches):
        P = bgr.PlannerParams((0.2, 0.2, 0.2), (3.0, 3.0, 3.0), mode=bgr.PLAN_MODE_MS, k=8, batch=256,
                              flags=flags | bgr.PLAN_NO_GOAL, seed=17)
        op = oracle.Planner(P, boxes, table, ms)
        sizes = [len(op)]
        for _ in range(batches):
            op.step()
            sizes.append(len(op))
        return op, sizes

    op, sizes = grow(bgr.PLAN_DEDUP, 12)
    xyz, _ = op.tree()
    def copies(x, sz):
        return sum((b - a) - np.unique(x[a:b].view(np.uint32).reshape(-1, 3), axis=0).shape[0]
                   for a, b in zip(sz[:-1], sz[1:]))

    plain, psizes = grow(0, 12)
    pxyz, _ = plain.tree()
    # what is left are table entries drawn twice in one batch (256 draws of 20 000: rare)
    assert copies(xyz, sizes) <= 3 and copies(pxyz, psizes) > 300
    # same draws: the first batch differs only by the dropped copies
    first = set(map(tuple, xyz[sizes[0]:sizes[1]].view(np.uint32).reshape(-1, 3).tolist()))
    pfirst = set(map(tuple, pxyz[psizes[0]:psizes[1]].view(np.uint32).reshape(-1, 3).tolist()))
    assert first == pfirst and sizes[1] < psizes[1]
    assert op.stats()["rejected_state"] > plain.stats()["rejected_state"]


def test_collision_mode_2_grows_the_same_tree(oracle, bgr):
    """the bucket-grid collision mode of the oracle planner (used for the 256^3 / batch-4096 parity run)
    only changes how the leaf boxes are visited: same booleans, so the same tree and the same counters
    (edge band flags aside, which mode 2 does not compute)."""
    from bottleneck_guided_rrt_b200 import synth
    occ = synth.make_map(64, 1, occupancy=0.03)
    boxes = synth.grid_to_boxes(occ, 0.05, (0, 0, 0), 3)
    table = synth.points(9, 20000, 3.2)
    ms = np.array([[1.0, 0.5, 1.5], [2.0, 1.0, 2.0], [2.5, 2.0, 1.0]], np.float32)
    trees = []
    pow_sqrt_before = oracle.pow_sqrt_mismatches()
    for mode in (1, 2):
        P = bgr.PlannerParams((0.2, 0.2, 0.2), (3.0, 3.0, 3.0), mode=bgr.PLAN_MODE_MS, k=8, batch=128,
                              flags=bgr.PLAN_EDGE_CHECK | bgr.PLAN_REWIRE | bgr.PLAN_NO_GOAL, seed=23)
        op = oracle.Planner(P, boxes, table, ms, collision_mode=mode)
        op.run(1500, 20)
        st = op.stats()
        st.pop("in_band")
        trees.append((op.tree(), st))
    (x1, p1), s1 = trees[0]
    (x2, p2), s2 = trees[1]
    assert np.array_equal(x1.view(np.uint32), x2.view(np.uint32)) and np.array_equal(p1, p2) and s1 == s2
    assert s1["rejected_state"] > 0 and s1["improved_parent"] > 0 and len(p1) > 500
    assert oracle.pow_sqrt_mismatches() == pow_sqrt_before
