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


def _compare(bgr, oracle, world, max_nodes, max_batches, **kw):
    m, boxes, table, ms = world
    P = bgr.PlannerParams((0.2, 0.2, 0.2), (3.0, 3.0, 3.0), **kw)
    gp = bgr.Planner(P, m, table, ms)
    op = oracle.Planner(P, boxes, table, ms)
    g_goal = gp.run(max_nodes, max_batches)
    o_goal = op.run(max_nodes, max_batches)
    gs, os_ = gp.stats(), op.stats()
    if gs["in_band"] or os_["in_band"]:
        pytest.skip("a collision result fell inside the 1e-6 m MPR tolerance band (%d/%d): trees may "
                    "legitimately diverge" % (gs["in_band"], os_["in_band"]))
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
        bgr.Planner(bgr.PlannerParams((0, 0, 0), (1, 1, 1), flags=64), m, table, ms)


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
