"""CPU tests of the path post-processing restatement (oracle/path.py) and of the library's host
arithmetic for SmoothenPlan (bgr_path_smoothen runs on the host: no GPU needed).
Reference: rrt_star_milestones.cpp BuildPlan :156-171, SmoothenPlan :645-799."""
import numpy as np
import pytest


def test_build_plan_walks_parents():
    from oracle import path as P
    par = np.array([-1, 0, 1, 1, 3, 2], dtype=np.int32)
    assert list(P.build_plan(par)) == [0, 1, 2, 5]
    assert list(P.build_plan(np.array([-1]))) == [0]


def test_smoothen_known_answers():
    from oracle import path as P
    # P == 2: straight line, 11 samples, float t
    s = P.smoothen_plan(np.array([[0, 0, 0], [1, 2, 4.0]]))
    assert s.shape == (11, 3) and np.array_equal(s[0], [0, 0, 0])
    assert np.allclose(s[5], [0.5, 1.0, 2.0], atol=1e-6) and np.allclose(s[10], [1, 2, 4], atol=1e-6)
    # P == 3: the quadratic Bezier passes through the middle waypoint at t = 0.5 (control = 2*p2 - (p1+p3)/2)
    p = np.array([[0, 0, 0], [1, 1, 0.5], [2, 0, 1.0]])
    s = P.smoothen_plan(p)
    assert s.shape == (11, 3) and np.allclose(s[5], p[1], atol=1e-6) and np.allclose(s[10], p[2], atol=1e-5)
    # P > 3: cubic Bezier over the FIRST FOUR waypoints, t = 0 .. 0.9999999999999999 (11 samples)
    p = np.array([[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0], [9, 9, 9.0]])
    s = P.smoothen_plan(p)
    assert s.shape == (11, 3) and np.array_equal(s[0], p[0]) and np.allclose(s[10], p[3], atol=1e-12)
    assert np.allclose(s[5], [0.75, 0.5, 0.0], atol=1e-12)           # B(0.5) of that control polygon
    with pytest.raises(ValueError):
        P.smoothen_plan(np.zeros((1, 3)))


@pytest.mark.parametrize("n", [2, 3, 4, 9])
def test_library_smoothen_matches_oracle_bit_exact(n):
    import bottleneck_guided_rrt_b200 as B
    from oracle import path as P
    rng = np.random.default_rng(n)
    for _ in range(100):
        pts = rng.uniform(-3, 3, (n, 3)).astype(np.float32).astype(np.float64)
        a, b = B.smoothen_plan(pts), P.smoothen_plan(pts)
        assert a.shape == b.shape and np.array_equal(a.view(np.uint64), b.view(np.uint64))
    with pytest.raises(RuntimeError):
        B.smoothen_plan(np.zeros((1, 3)))


def test_clearance_and_labels_oracle():
    from oracle import path as P
    boxes = np.array([[0.5, 0.5, 0.5, 1.0, 1.0, 0.5], [3.0, 0.5, 0.5, 1.0, 1.0, 0.5], [9.0, 9.0, 9.0, 1.0, 0.2, 0.5]])
    pts = np.array([[1.6, 0.5, 0.5], [2.0, 0.5, 0.5], [1.05, 0.5, 0.5]])
    d = P.clearance(boxes, pts, 0.1)
    assert np.allclose(d[:2], [0.5, 0.4]) and d[2] == -1.0           # third sphere penetrates; the unoccupied box is ignored
    path = np.array([[0, 5, 5], [1.6, 0.5, 0.5], [2.0, 0.5, 0.5], [1.3, 0.5, 0.5], [2.2, 3, 3], [5, 5, 5.0]])
    idx, labels, clr = P.find_labels(boxes, path, 0.1, 3)
    assert list(idx) == [2, 1, 0] and labels.shape == (9,) and np.allclose(labels[:3], [1.3, 0.5, 0.5])
