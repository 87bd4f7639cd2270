"""GPU parity: bgr_planner_* (through the C ABI) vs the CPU restatement of the planner loop
(oracle/src/planner.cpp).  Reference: rrt_star_milestones.cpp main() MS:1160-1406, cost_wiring
MS:802-831, extend_tree MS:834-853; rrt_star.cpp RS:1145-1178.  Trees must be identical: node
coordinates bit for bit, parent indices equal, same acceptance / rejection counts."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def world(bgr, oracle):
    from bottleneck_guided_rrt_b200 import synth
    occ = synth.make_map(64, 1, occupancy=0.0)          # 3 wall slabs with doors, no clutter
    boxes = synth.grid_to_boxes(occ, 0.05, (0, 0, 0), 2)
    m = bgr.OccupancyMap(0.05, (0, 0, 0), (64, 64, 64))
    m.set_boxes(boxes)
    table = synth.points(9, 50000, 3.2)
    ms = np.load(os.path.join(GOLD, "milestones.npz"))["milestones"][3].reshape(3, 3) * np.float32(1.5)
    return m, boxes, table, ms.astype(np.float32)


def _first_band_batch(gp, max_nodes, max_batches):
    """steps a GPU planner and returns the index of the first batch that saw a collision result inside
    the MPR tolerance band (None: no such batch before the run ended)"""
    b = 0
    while b < max_batches and len(gp) < max_nodes:
        _, goal = gp.step()
        if gp.stats()["in_band"]:
            return b
        if goal:
            break
        b += 1
    return None


def _assert_same_tree(gp, op, g_goal, o_goal):
    gs, os_ = gp.stats(), op.stats()
    gx, gpar = gp.tree()
    ox, opar = op.tree()
    assert len(gp) == len(op) and g_goal == o_goal
    assert np.array_equal(gx.view(np.uint32), ox.view(np.uint32)), "node coordinates differ"
    assert np.array_equal(gpar, opar), "parents differ"
    for key in ("batches", "candidates", "rejected_state", "rejected_edge", "improved_parent", "rewired", "goal_candidate"):
        assert gs[key] == os_[key], (key, gs, os_)
    # the tree is a tree: every parent precedes... (rewiring may point forward) -> walk to the root instead
    depth_guard = len(gpar) + 1
    for i in np.random.default_rng(0).integers(0, len(gpar), 50):
        cur, steps = int(i), 0
        while cur > 0 and steps < depth_guard:
            cur, steps = int(gpar[cur]), steps + 1
        assert cur == 0 or gpar[cur] == -1
    return gs


def _compare(bgr, oracle, world, max_nodes, max_batches, collision_mode=1, **kw):
    """GPU planner vs the oracle's.  A collision result inside the 1e-6 m MPR tolerance band may
    legitimately differ between the two, after which the trees diverge: every batch BEFORE the first
    such event (found with a scouting run of the GPU planner) is compared, bit for bit."""
    m, boxes, table, ms = world
    P = bgr.PlannerParams((0.2, 0.2, 0.2), (3.0, 3.0, 3.0), **kw)
    first = _first_band_batch(bgr.Planner(P, m, table, ms), max_nodes, max_batches)
    if first is not None:
        if first == 0:
            pytest.skip("the very first batch holds an in-band collision result: nothing to compare")
        max_batches = first
    gp = bgr.Planner(P, m, table, ms)
    op = oracle.Planner(P, boxes, table, ms, collision_mode=collision_mode)
    pow_sqrt_before = oracle.pow_sqrt_mismatches()       # process-wide counter: other tests may have fed it odd inputs
    g_goal = gp.run(max_nodes, max_batches)
    o_goal = op.run(max_nodes, max_batches)
    assert gp.stats()["in_band"] == 0
    gs = _assert_same_tree(gp, op, g_goal, o_goal)
    assert oracle.pow_sqrt_mismatches() == pow_sqrt_before, "the oracle met a Norm where float(pow(s,0.5)) != float(sqrt(s))"
    return gp, gs


def test_reference_sequential_loop_rs(bgr, oracle, world):
    """batch = 1, RS sampler (10 % goal bias), K = 5, quirks on: the reference's own loop."""
    gp, gs = _compare(bgr, oracle, world, 400, 1500, mode=bgr.PLAN_MODE_RS, k=5, batch=1, flags=bgr.PLAN_QUIRKS, seed=5489)
    assert gs["improved_parent"] > 0 and gs["rejected_state"] > 0


def test_reference_sequential_loop_ms(bgr, oracle, world):
    """batch = 1, MS sampler (0.8 / 0.5 thresholds + 3 milestones)."""
    _compare(bgr, oracle, world, 300, 1200, mode=bgr.PLAN_MODE_MS, k=5, batch=1, flags=bgr.PLAN_QUIRKS, seed=7)


@pytest.mark.parametrize("batch,k", [(16, 5), (256, 10), (1024, 16)])
def test_batched_step(bgr, oracle, world, batch, k):
    _compare(bgr, oracle, world, 3000, 40, mode=bgr.PLAN_MODE_RS, k=k, batch=batch, flags=0, seed=11)


@pytest.mark.parametrize("flags", [1, 2, 3, 3 | 8])
def test_batched_edge_checks_and_rewiring(bgr, oracle, world, flags):
    gp, gs = _compare(bgr, oracle, world, 2500, 30, mode=bgr.PLAN_MODE_MS, k=10, batch=256, flags=flags, seed=3)
    if flags & 2:
        assert gs["rewired"] > 0
    if flags & 1:
        assert gs["rejected_edge"] >= 0


@pytest.mark.parametrize("mode,flags", [(0, 16), (1, 16), (1, 16 | 3), (1, 16 | 3 | 8)])
def test_batched_dedup_of_goal_and_milestone_samples(bgr, oracle, world, mode, flags):
    # BGR_PLAN_DEDUP: one candidate per distinct goal / milestone sample of a batch, oracle in lockstep
    gp, gs = _compare(bgr, oracle, world, 2500, 30, mode=mode, k=10, batch=256, flags=flags, seed=3)
    x, _ = gp.tree()
    # no two nodes appended in the same batch coincide: distinct coordinates >= nodes - (goal node copies)
    uniq = np.unique(x.view(np.uint32).reshape(-1, 3), axis=0).shape[0]
    plain = bgr.Planner(bgr.PlannerParams((0.2, 0.2, 0.2), (3.0, 3.0, 3.0), mode=mode, k=10, batch=256,
                                          flags=flags & ~16, seed=3), world[0], world[2], world[3])
    plain.run(2500, 30)
    px, _ = plain.tree()
    puniq = np.unique(px.view(np.uint32).reshape(-1, 3), axis=0).shape[0]
    assert uniq / len(x) > puniq / len(px), (uniq, len(x), puniq, len(px))
    assert gs["rejected_state"] > plain.stats()["rejected_state"]


def test_goal_reached_and_path(bgr, oracle, world):
    """run to the goal: goal node appended with the reaching node as parent (MS:1346-1347);
    BuildPlan walks parents back to the root (MS:156-171)."""
    m, boxes, table, ms = world
    P = bgr.PlannerParams((0.2, 0.2, 0.2), (0.9, 0.5, 0.4), mode=bgr.PLAN_MODE_RS, k=5, batch=8, flags=0, seed=21, goal_chance=0.3)
    gp = bgr.Planner(P, m, table, ms)
    op = oracle.Planner(P, boxes, table, ms)
    assert gp.run(5000, 3000) and op.run(5000, 3000)
    gx, gpar = gp.tree()
    ox, opar = op.tree()
    assert np.array_equal(gx.view(np.uint32), ox.view(np.uint32)) and np.array_equal(gpar, opar)
    assert np.array_equal(gx[-1], np.array([0.9, 0.5, 0.4], np.float32)) and gpar[-1] == len(gpar) - 2
    path = gp.path()
    assert path[0] == 0 and path[-1] == len(gpar) - 1
    assert all(gpar[path[i + 1]] == path[i] for i in range(len(path) - 1))
    # a finished planner does nothing more
    assert gp.step() == (0, True)


def test_planner_argument_errors(bgr, world):
    m, boxes, table, ms = world
    with pytest.raises(bgr.BgrError):
        bgr.Planner(bgr.PlannerParams((0, 0, 0), (1, 1, 1), k=40), m, table, ms)
    with pytest.raises(bgr.BgrError):
        bgr.Planner(bgr.PlannerParams((0, 0, 0), (1, 1, 1), mode=bgr.PLAN_MODE_MS), m, table, None)
    with pytest.raises(bgr.BgrError):
        bgr.Planner(bgr.PlannerParams((0, 0, 0), (1, 1, 1), flags=128), m, table, ms)


def test_run_many_equals_run(bgr, oracle):
    """bgr_planner_run_many (independent queries stepped round-robin on their own streams / map
    handles) grows exactly the trees bgr_planner_run grows one query at a time."""
    import torch
    from bottleneck_guided_rrt_b200 import synth
    occ = synth.make_map(64, 1, occupancy=0.02)
    boxes = synth.grid_to_boxes(occ, 0.05, (0, 0, 0), 2)
    table = synth.points(9, 50000, 3.2)
    ms = np.array([[0.8, 1.6, 1.6], [1.6, 1.6, 1.6], [2.4, 1.6, 1.6]], dtype=np.float32)
    goals = [(3.0, 3.0, 3.0), (2.9, 0.3, 2.8), (0.3, 2.9, 0.4), (3.0, 0.2, 0.2)]

    def make(i, stream=None):
        m = bgr.OccupancyMap(0.05, (0, 0, 0), (64, 64, 64), stream=stream)
        m.set_boxes(boxes)
        P = bgr.PlannerParams((0.2, 0.2, 0.2), goals[i], mode=bgr.PLAN_MODE_MS, k=8, batch=64,
                              flags=bgr.PLAN_EDGE_CHECK | bgr.PLAN_REWIRE, seed=100 + i)
        return bgr.Planner(P, m, table, ms, stream=stream)

    streams = [torch.cuda.Stream() for _ in goals]
    many = [make(i, s.cuda_stream) for i, s in enumerate(streams)]
    flags = bgr.Planner.run_many(many, 3000, 400)
    torch.cuda.synchronize()
    for i in range(len(goals)):
        one = make(i)
        g = one.run(3000, 400)
        assert g == flags[i] and len(one) == len(many[i])
        ax, ap = one.tree()
        bx, bp = many[i].tree()
        assert np.array_equal(ax.view(np.uint32), bx.view(np.uint32)) and np.array_equal(ap, bp)
        assert one.stats() == many[i].stats()
    with pytest.raises(RuntimeError):
        bgr.Planner.run_many([many[0], many[0]], 10, 1)


def test_unfilled_neighbour_slots_are_skipped(bgr, oracle, world):
    """Every sample is the same point P within eta of the start: from the second batch on the tree is the
    root plus copies of P, the self-match rule (d2 <= FLT_EPSILON) leaves ONE admissible neighbour and
    K-1 unfilled slots (index -1).  The parent choice must ignore them (their cost entries are stale or
    uninitialised): identical to the oracle, and identical between two fresh planners."""
    m, boxes, _, _ = world
    P0 = np.array([[0.28, 0.22, 0.2]], dtype=np.float32)
    ms = np.repeat(P0, 3, axis=0)
    trees = []
    for flags in (8, 8 | 3):
        for rep in range(2):
            P = bgr.PlannerParams((0.2, 0.2, 0.2), tuple(P0[0]), mode=bgr.PLAN_MODE_MS, k=5, batch=8, flags=flags, seed=17)
            gp = bgr.Planner(P, m, P0, ms)
            if rep == 0:
                # poison the scratch a fresh planner would otherwise find zeroed: a batch on a real tree first
                warm = bgr.Planner(bgr.PlannerParams((0.2, 0.2, 0.2), (3.0, 3.0, 3.0), mode=bgr.PLAN_MODE_RS, k=5, batch=8,
                                                     flags=8, seed=3), m, world[2], world[3])
                warm.run(200, 40)
            op = oracle.Planner(P, boxes, P0, ms)
            gp.run(100, 6)
            op.run(100, 6)
            assert gp.stats()["in_band"] == 0
            _assert_same_tree(gp, op, False, False)
            trees.append(gp.tree())
        assert np.array_equal(trees[-1][1], trees[-2][1])
    x, par = trees[0]
    assert len(x) == 1 + 6 * 8 and (par[1:] >= 0).all()


def test_config3_shape_parity(bgr, oracle):
    """BASELINE configs[3] shape: 256^3 narrow-passage map, MS sampler with the 3 door-centre
    milestones, batch 4096, K = 16, edge checks + rewiring (the planner shape bench.py times), grown to
    ~65 k nodes — where the tree is piled up at the milestones, so grid_knn_kernel<PILED>,
    grid_knn_heavy_kernel, the tail-work index rebuilds and the distinct-sample Nearest run.  Trees bit
    for bit against oracle/src/planner.cpp (bucket-grid collision mode: same booleans) on every batch
    before the first in-band collision event."""
    from bottleneck_guided_rrt_b200 import synth
    occ = synth.make_map(256, 5, n_walls=3, wall_thick=4, door_cells=5)
    boxes = synth.grid_to_boxes(occ, 0.05, (0, 0, 0), 3)
    m = bgr.OccupancyMap(0.05, (0, 0, 0), (256,) * 3)
    m.set_boxes(boxes)
    ms = []
    for w in range(3):
        x0 = (w + 1) * 256 // 4
        zz, yy = np.nonzero(occ[:, :, x0] == 0)
        ms.append([(x0 + 0.5) * 0.05, (yy.mean() + 0.5) * 0.05, (zz.mean() + 0.5) * 0.05])
    ms = np.asarray(ms, dtype=np.float32)
    table = synth.points(9, 200_000, 12.8)
    P = bgr.PlannerParams((0.2, 0.2, 0.2), (12.6, 12.6, 12.6), mode=bgr.PLAN_MODE_MS, k=16, batch=4096,
                          flags=bgr.PLAN_EDGE_CHECK | bgr.PLAN_REWIRE | bgr.PLAN_NO_GOAL, seed=5489)
    n_batches = 24
    scout = bgr.Planner(P, m, table, ms)
    heavy_max, first = 0, None
    for b in range(n_batches):
        scout.step()
        heavy_max = max(heavy_max, scout.nns_debug_stats()["heavy_queries_last_launch"])
        if scout.stats()["in_band"]:
            first = b
            break
    dbg = scout.nns_debug_stats()
    if first is None:
        # the shape really exercises what it is here for
        assert len(scout) > 40_000
        assert dbg["index_builds"] >= 3 and dbg["grid_knn_piled_launches"] > 0 and heavy_max > 0, dbg
    else:
        assert first >= 8, "in-band event in batch %d: too early to exercise the crowded-index paths" % first
        n_batches = first
    gp = bgr.Planner(P, m, table, ms)
    op = oracle.Planner(P, boxes, table, ms, collision_mode=2)
    pow_sqrt_before = oracle.pow_sqrt_mismatches()
    gp.run(1 << 30, n_batches)
    op.run(1 << 30, n_batches)
    assert gp.stats()["in_band"] == 0
    gs = _assert_same_tree(gp, op, False, False)
    assert gs["improved_parent"] > 0 and gs["rejected_state"] > 0 and gs["candidates"] == n_batches * 4096
    assert oracle.pow_sqrt_mismatches() == pow_sqrt_before


def _same_gpu_trees(a, b):
    ax, ap = a.tree()
    bx, bp = b.tree()
    assert len(a) == len(b)
    assert np.array_equal(ax.view(np.uint32), bx.view(np.uint32)) and np.array_equal(ap, bp)
    sa, sb = a.stats(), b.stats()
    for key in ("batches", "candidates", "rejected_state", "rejected_edge", "improved_parent", "rewired", "in_band", "goal_candidate"):
        assert sa[key] == sb[key], (key, sa, sb)


@pytest.mark.parametrize("mode,k,batch,flags,seed", [(0, 5, 1, 4, 5489), (1, 5, 1, 4, 7), (0, 10, 4, 0, 11), (1, 10, 3, 2, 3),
                                                     (1, 8, 2, 16 | 2, 5), (0, 5, 1, 8, 1), (1, 16, 4, 2 | 8, 9), (1, 10, 8, 2, 13), (0, 6, 7, 0, 2)])
def test_persistent_steps_equal_host_driven_steps(bgr, oracle, world, mode, k, batch, flags, seed):
    """small batches without edge checks run whole steps inside one launch (plan_persist_kernel): the tree, the
    statistics and the sample stream position are those of the host-driven steps (BGR_PLAN_NO_PERSIST) and of the
    oracle, also when run() and step() calls are mixed (the step() path prepares the next step's words ahead)."""
    m, boxes, table, ms = world
    P = bgr.PlannerParams((0.2, 0.2, 0.2), (3.0, 3.0, 3.0), mode=mode, k=k, batch=batch, flags=flags, seed=seed)
    Q = bgr.PlannerParams((0.2, 0.2, 0.2), (3.0, 3.0, 3.0), mode=mode, k=k, batch=batch, flags=flags | bgr.PLAN_NO_PERSIST, seed=seed)
    a, b = bgr.Planner(P, m, table, ms), bgr.Planner(Q, m, table, ms)
    op = oracle.Planner(P, boxes, table, ms)
    ga, gb, go = a.run(150, 400), b.run(150, 400), op.run(150, 400)
    assert a.persist_stats()["launches"] >= 1 and a.persist_stats()["steps"] == a.stats()["batches"]
    assert b.persist_stats()["launches"] == 0
    assert ga == gb
    _same_gpu_trees(a, b)
    if a.stats()["in_band"] == 0:
        _assert_same_tree(a, op, ga, go)
    # mixed: single steps (host-driven, words prepared ahead), then persistent runs again
    for _ in range(7):
        if not ga:
            ga = a.step()[1]
            gb = b.step()[1]
    if not ga:
        ga, gb = a.run(400, 300), b.run(400, 300)
        assert ga == gb
    for _ in range(3):
        if not ga:
            ga = a.step()[1]
            gb = b.step()[1]
    if not ga:
        ga, gb = a.run(len(a) + 40, 10_000), b.run(len(b) + 40, 10_000)
    assert ga == gb
    _same_gpu_trees(a, b)
    assert 0 < a.persist_stats()["steps"] <= a.stats()["batches"]


def test_persistent_steps_reach_the_goal(bgr, oracle, world):
    """batch 1 to the goal: goal cut, goal node and path as the host-driven loop's"""
    m, boxes, table, ms = world
    for seed in (3, 4):
        P = bgr.PlannerParams((0.2, 0.2, 0.2), (0.6, 0.6, 0.6), mode=bgr.PLAN_MODE_RS, k=5, batch=1, flags=0, seed=seed, goal_tol=0.2)
        Q = bgr.PlannerParams((0.2, 0.2, 0.2), (0.6, 0.6, 0.6), mode=bgr.PLAN_MODE_RS, k=5, batch=1, flags=bgr.PLAN_NO_PERSIST, seed=seed, goal_tol=0.2)
        a, b = bgr.Planner(P, m, table, ms), bgr.Planner(Q, m, table, ms)
        ga, gb = a.run(4000, 20_000), b.run(4000, 20_000)
        assert ga and gb
        _same_gpu_trees(a, b)
        assert np.array_equal(a.path(), b.path())
        assert a.persist_stats()["launches"] >= 1


def test_persistent_steps_large_tree(bgr, oracle, world):
    """batch 4 to 24,000 nodes: past the size where the host-driven path switches from the plain scans to the grid
    index (16,384 nodes), the persistent kernel keeps scanning — same tree"""
    m, boxes, table, ms = world
    P = bgr.PlannerParams((0.2, 0.2, 0.2), (3.0, 3.0, 3.0), mode=bgr.PLAN_MODE_RS, k=10, batch=4, flags=bgr.PLAN_REWIRE | bgr.PLAN_NO_GOAL, seed=21)
    Q = bgr.PlannerParams((0.2, 0.2, 0.2), (3.0, 3.0, 3.0), mode=bgr.PLAN_MODE_RS, k=10, batch=4, flags=bgr.PLAN_REWIRE | bgr.PLAN_NO_GOAL | bgr.PLAN_NO_PERSIST, seed=21)
    a, b = bgr.Planner(P, m, table, ms), bgr.Planner(Q, m, table, ms)
    a.run(24_000)
    b.run(24_000)
    _same_gpu_trees(a, b)
    assert a.persist_stats()["steps"] == a.stats()["batches"] and a.stats()["rewired"] > 0


def test_reset_and_query_queue_equal_fresh_planners(bgr, oracle, world):
    """bgr_planner_reset re-arms a planner for another query (same map / table / milestones): the tree it then grows
    is the one a freshly created planner grows, bit for bit — host-driven batches, persistent small batches, and
    through bgr_planner_run_queue (workers take the next query as they finish one)."""
    import torch
    from bottleneck_guided_rrt_b200 import synth
    m, boxes, table, ms = world
    starts = np.array([[0.2, 0.2, 0.2], [0.3, 2.9, 0.4], [0.25, 1.5, 2.8], [0.5, 0.5, 3.0], [0.3, 0.3, 1.5], [0.2, 3.0, 3.0]], dtype=np.float32)
    goals = np.array([[3.0, 3.0, 3.0], [2.9, 0.3, 2.9], [3.0, 1.6, 0.3], [2.8, 2.8, 0.3], [3.0, 0.4, 1.4], [3.0, 0.2, 0.2]], dtype=np.float32)
    seeds = np.array([11, 12, 13, 14, 15, 16], dtype=np.uint32)
    for batch, k, flags, cap in ((64, 10, 3, 3000), (2, 8, 2, 600)):
        fresh = []
        for q in range(len(seeds)):
            P = bgr.PlannerParams(tuple(starts[q]), tuple(goals[q]), mode=bgr.PLAN_MODE_MS, k=k, batch=batch, flags=flags, seed=int(seeds[q]))
            pl = bgr.Planner(P, m, table, ms)
            g = pl.run(cap, 200)
            fresh.append((g, pl.tree(), pl.stats()))
        # one planner, re-armed for every query (after a different query ran on it)
        P0 = bgr.PlannerParams((1.0, 1.0, 1.0), (2.0, 2.0, 2.0), mode=bgr.PLAN_MODE_MS, k=k, batch=batch, flags=flags, seed=99)
        pl = bgr.Planner(P0, m, table, ms)
        pl.run(cap // 2, 50)
        for q in range(len(seeds)):
            pl.reset(starts[q], goals[q], seeds[q])
            assert len(pl) == 1
            g = pl.run(cap, 200)
            x, par = pl.tree()
            assert g == fresh[q][0] and np.array_equal(x.view(np.uint32), fresh[q][1][0].view(np.uint32)) and np.array_equal(par, fresh[q][1][1])
            assert pl.stats() == fresh[q][2]
        # the queue: 2 workers (own map handle + stream each), 6 queries
        streams = [torch.cuda.Stream() for _ in range(2)]
        maps = []
        for s_ in streams:
            mm = bgr.OccupancyMap(0.05, (0, 0, 0), (64, 64, 64), stream=s_.cuda_stream)
            mm.set_boxes(boxes)
            maps.append(mm)
        workers = [bgr.Planner(P0, maps[i], table, ms, stream=streams[i].cuda_stream) for i in range(2)]
        reached, nodes = bgr.Planner.run_queue(workers, starts, goals, seeds, cap, 200)
        torch.cuda.synchronize()
        for q in range(len(seeds)):
            assert bool(reached[q]) == fresh[q][0] and int(nodes[q]) == len(fresh[q][1][1]), (q, reached, nodes)


def test_persistent_steps_respect_max_batches(bgr, oracle, world):
    """run(max_nodes, max_batches) stops after exactly max_batches steps in the persistent kernel too, and a
    sequence of short runs leaves the tree and the sample stream where the host-driven steps leave them"""
    m, boxes, table, ms = world
    P = bgr.PlannerParams((0.2, 0.2, 0.2), (3.0, 3.0, 3.0), mode=bgr.PLAN_MODE_MS, k=6, batch=3, flags=bgr.PLAN_REWIRE | bgr.PLAN_NO_GOAL, seed=17)
    Q = bgr.PlannerParams((0.2, 0.2, 0.2), (3.0, 3.0, 3.0), mode=bgr.PLAN_MODE_MS, k=6, batch=3, flags=bgr.PLAN_REWIRE | bgr.PLAN_NO_GOAL | bgr.PLAN_NO_PERSIST, seed=17)
    a, b = bgr.Planner(P, m, table, ms), bgr.Planner(Q, m, table, ms)
    total = 0
    for nb in (1, 2, 3, 5, 1, 7, 40, 1, 1, 13):
        a.run(1 << 30, nb)
        b.run(1 << 30, nb)
        total += nb
        assert a.stats()["batches"] == total == b.stats()["batches"]
        assert len(a) == len(b)
    _same_gpu_trees(a, b)
    # node limit inside a launch: stops at the first step that reaches it
    a.run(len(a) + 5, 1000)
    b.run(len(b) + 5, 1000)
    _same_gpu_trees(a, b)
    assert a.persist_stats()["steps"] == a.stats()["batches"]


def test_run_queue_corner_cases(bgr, oracle, world):
    import torch
    m, boxes, table, ms = world
    P0 = bgr.PlannerParams((1.0, 1.0, 1.0), (2.0, 2.0, 2.0), mode=bgr.PLAN_MODE_RS, k=5, batch=1, flags=0, seed=1)
    w = bgr.Planner(P0, m, table, ms)
    reached, nodes = bgr.Planner.run_queue([w], np.zeros((0, 3), np.float32), np.zeros((0, 3), np.float32), np.zeros(0, np.uint32), 100)
    assert len(reached) == 0 and len(nodes) == 0
    starts = np.array([[0.2, 0.2, 0.2], [0.3, 0.3, 0.3]], np.float32)
    goals = np.array([[0.5, 0.5, 0.5], [0.6, 0.4, 0.5]], np.float32)
    reached, nodes = bgr.Planner.run_queue([w], starts, goals, np.array([5, 6], np.uint32), 2000, 5000)
    for q in range(2):
        P = bgr.PlannerParams(tuple(starts[q]), tuple(goals[q]), mode=bgr.PLAN_MODE_RS, k=5, batch=1, flags=0, seed=5 + q)
        f = bgr.Planner(P, m, table, ms)
        assert f.run(2000, 5000) == bool(reached[q]) and len(f) == int(nodes[q])
    with pytest.raises(Exception):
        bgr.Planner.run_queue([w, w], starts, goals, np.array([5, 6], np.uint32), 100)      # shared handle


@pytest.mark.parametrize("mode,k,batch,flags,seed", [(0, 10, 256, 0, 11), (1, 10, 256, 3, 3), (1, 16, 1024, 3 | 16, 5), (1, 8, 64, 1, 9),
                                                     (0, 5, 16, 2 | 4, 2), (1, 10, 300, 3 | 8, 4)])
def test_graph_replayed_steps_equal_host_driven_steps(bgr, oracle, mode, k, batch, flags, seed):
    """a planner on a stream of its own replays its step as one CUDA graph launch while the tree is below 16,384 nodes
    (the tree size is read from the counters block, Nearest / kNN are the plain scans laid out for the cap): same
    tree, statistics and sample-stream position as the launch-by-launch steps (BGR_PLAN_NO_GRAPH) and the oracle,
    across the switch to host-driven steps at the cap, with run() and step() mixed, and after a map update"""
    import torch
    from bottleneck_guided_rrt_b200 import synth
    occ = synth.make_map(64, 1, occupancy=0.0)
    boxes = synth.grid_to_boxes(occ, 0.05, (0, 0, 0), 2)
    table = synth.points(9, 50000, 3.2)
    ms = np.load(os.path.join(GOLD, "milestones.npz"))["milestones"][3].reshape(3, 3).astype(np.float32) * np.float32(1.5)
    planners = []
    for extra in (0, bgr.PLAN_NO_GRAPH):
        st = torch.cuda.Stream()
        m = bgr.OccupancyMap(0.05, (0, 0, 0), (64, 64, 64), stream=st.cuda_stream)
        m.set_boxes(boxes)
        P = bgr.PlannerParams((0.2, 0.2, 0.2), (3.0, 3.0, 3.0), mode=mode, k=k, batch=batch, flags=flags | extra | bgr.PLAN_NO_PERSIST, seed=seed)
        planners.append((bgr.Planner(P, m, table, ms, stream=st.cuda_stream), m, st))
    (a, ma, _), (b, mb, _) = planners
    P0 = bgr.PlannerParams((0.2, 0.2, 0.2), (3.0, 3.0, 3.0), mode=mode, k=k, batch=batch, flags=flags, seed=seed)
    op = oracle.Planner(P0, boxes, table, ms)
    ga, gb, go = a.run(3000, 12), b.run(3000, 12), op.run(3000, 12)
    torch.cuda.synchronize()
    assert a.persist_stats()["graph_steps"] >= 5 and b.persist_stats()["graph_steps"] == 0
    assert ga == gb
    _same_gpu_trees(a, b)
    if a.stats()["in_band"] == 0:
        _assert_same_tree(a, op, ga, go)
    for _ in range(3):
        if not ga:
            ga, gb = a.step()[1], b.step()[1]
    # a map update between steps: the captured step must not keep using what it captured
    add = np.array([[1.025 + 0.05 * i, 1.025, 1.025, 0.05, 1.0, 0.5] for i in range(6)])
    ma.update_boxes(add, boxes[:200])
    mb.update_boxes(add, boxes[:200])
    if not ga:
        ga, gb = a.run(20_000, 40), b.run(20_000, 40)       # across the 16,384-node cap for the large batches
    torch.cuda.synchronize()
    assert ga == gb
    _same_gpu_trees(a, b)


def test_graph_replay_with_rearmed_planner(bgr, oracle, world):
    """bgr_planner_reset changes the goal the captured kernels carry by value: the step is captured again"""
    import torch
    m0, boxes, table, ms = world
    st = torch.cuda.Stream()
    m = bgr.OccupancyMap(0.05, (0, 0, 0), (64, 64, 64), stream=st.cuda_stream)
    m.set_boxes(boxes)
    P = bgr.PlannerParams((0.2, 0.2, 0.2), (3.0, 3.0, 3.0), mode=bgr.PLAN_MODE_MS, k=10, batch=128, flags=3, seed=1)
    w = bgr.Planner(P, m, table, ms, stream=st.cuda_stream)
    w.run(2000, 10)
    for q, (s_, g_, seed) in enumerate([((0.3, 2.9, 0.4), (2.9, 0.3, 2.9), 12), ((0.25, 1.5, 2.8), (3.0, 1.6, 0.3), 13)]):
        w.reset(s_, g_, seed)
        gw = w.run(2500, 30)
        Pq = bgr.PlannerParams(s_, g_, mode=bgr.PLAN_MODE_MS, k=10, batch=128, flags=3 | bgr.PLAN_NO_GRAPH, seed=seed)
        f = bgr.Planner(Pq, m0, table, ms)
        gf = f.run(2500, 30)
        torch.cuda.synchronize()
        assert gw == gf
        _same_gpu_trees(w, f)
    assert w.persist_stats()["graph_steps"] > 10


def test_config4_query_shape_parity(bgr, oracle):
    """BASELINE configs[4](i) shape: one start/goal query on the 256^3 query map of bench.py (0.4 m doors, light
    clutter), MS sampler with the door-centre milestones, batch 1024, K = 16, edge checks + rewiring +
    BGR_PLAN_DEDUP, the planner on a stream of its own (so its steps are replayed as a CUDA graph while the tree is
    small, then host-driven): tree bit for bit against oracle/src/planner.cpp on every batch before the first
    in-band collision event, goal flag included."""
    import torch
    from bottleneck_guided_rrt_b200 import synth
    import bench
    occ = synth.make_map(256, 5, n_walls=3, wall_thick=4, door_cells=bench.QUERY_DOOR_CELLS, occupancy=bench.QUERY_OCCUPANCY)
    boxes = synth.grid_to_boxes(occ, 0.05, (0, 0, 0), 3)
    ms = bench.door_milestones(occ, 256)
    table = synth.points(9, 200_000, 12.8)
    st = torch.cuda.Stream()
    m = bgr.OccupancyMap(0.05, (0, 0, 0), (256,) * 3, stream=st.cuda_stream)
    m.set_boxes(boxes)
    cand = synth.points(77, 4000, 12.8)
    hit, _ = m.collide_spheres(cand, 0.1)
    free = cand[np.asarray(hit) == 0]
    flags = bgr.PLAN_EDGE_CHECK | bgr.PLAN_REWIRE | bgr.PLAN_DEDUP
    compared = 0
    for qi in (0, 3):
        P = bgr.PlannerParams(tuple(free[2 * qi]), tuple(free[2 * qi + 1]), mode=bgr.PLAN_MODE_MS, k=16, batch=1024, flags=flags, seed=100 + qi)
        n_batches = 40
        first = _first_band_batch(bgr.Planner(P, m, table, ms, stream=st.cuda_stream), 1 << 30, n_batches)
        if first is not None:
            n_batches = first
        if n_batches < 5:
            continue
        gp = bgr.Planner(P, m, table, ms, stream=st.cuda_stream)
        op = oracle.Planner(P, boxes, table, ms, collision_mode=2)
        before = oracle.pow_sqrt_mismatches()
        g_goal, o_goal = gp.run(1 << 30, n_batches), op.run(1 << 30, n_batches)
        torch.cuda.synchronize()
        assert gp.stats()["in_band"] == 0
        gs = _assert_same_tree(gp, op, g_goal, o_goal)
        assert gs["rejected_state"] > 0 and len(gp) > 1000
        assert gp.persist_stats()["graph_steps"] > 0
        assert oracle.pow_sqrt_mismatches() == before
        compared += 1
    assert compared >= 1
