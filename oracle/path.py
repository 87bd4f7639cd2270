"""CPU restatement of the reference's path post-processing.  TEST INFRASTRUCTURE ONLY.

Follows abb_manipulator/src/rrt_star_milestones.cpp ("MS"): BuildPlan MS:156-171,
PrunePlan MS:562-609, SmoothenPlan MS:645-799, find_labels MS:943-996.  Plain Python loops
over the oracle's collision restatement — small cases only.  PARITY UNPINNED (the reference has
no tests or golden vectors for these functions; FCL/libccd are restated, see oracle.py).
"""
import math

import numpy as np

from . import oracle as O

F = np.float32


def build_plan(parent):
    """BuildPlan MS:159-170: walk parents from the last node while index > 0, then push 0."""
    path, cur = [], len(parent) - 1
    while cur > 0:
        path.insert(0, cur)
        cur = int(parent[cur])
    path.insert(0, 0)
    return np.asarray(path, dtype=np.int32)


def prune_plan(boxes, path, radius=0.03, shape=0):
    """PrunePlan MS:562-609.  path: (P,3) float32 waypoints (geometry_msgs doubles that came
    from floats).  Returns (kept indices, consulted in-band count) or raises RuntimeError where
    the reference would loop forever (no collision-free successor from the current anchor)."""
    path = np.asarray(path, dtype=np.float32).reshape(-1, 3)
    p = len(path)
    kept, cur, n_band = [0], 0, 0
    # MS:579 `while (pruned.back().x != rrt_path.back().x)`: the x coordinate alone ends the loop
    while float(path[kept[-1]][0]) != float(path[p - 1][0]):
        found = False
        for it in range(p - 1, cur, -1):                       # MS:584
            hit, band = O.collide_edges(boxes, path[cur:cur + 1], path[it:it + 1], radius, shape, mode=1)
            n_band += int(band[0])
            if not hit[0]:                                     # MS:591 line is safe
                kept.append(it)
                cur = it
                found = True
                break
        if not found:
            raise RuntimeError("PrunePlan: no collision-free successor from waypoint %d" % cur)
    return np.asarray(kept, dtype=np.int32), n_band


def smoothen_plan(path):
    """SmoothenPlan MS:645-799.  path: (P,3) float64 (geometry_msgs::Pose positions).
    P == 3: quadratic Bezier through a float control point, float t = 0, 0.1, ... < 1.1;
    P > 3: cubic Bezier over the FIRST FOUR waypoints only, double t = 0, 0.1, ... < 1.0;
    P == 2: straight line resampled, float t.  Returns (S,3) float64."""
    path = np.asarray(path, dtype=np.float64).reshape(-1, 3)
    n = len(path)
    if n < 2:
        raise ValueError("SmoothenPlan reads rrt_path[1]: needs >= 2 waypoints")
    out = []
    if n == 3:
        p1, p2, p3 = (path[i].astype(F) for i in range(3))      # Eigen Vector3f
        c1 = np.array([F(F(p2[j] * F(2)) - F(F(p1[j] + p3[j]) / F(2))) for j in range(3)], dtype=F)   # MS:671-673
        t = F(0)
        while float(t) < 1.1:                                    # MS:676 float t, double 1.1
            omt = F(F(1) - t)
            row = []
            for j in range(3):
                a = math.pow(float(omt), 2) * float(p1[j])                  # double
                b = F(F(F(omt * F(2)) * t) * c1[j])                          # float chain
                c = F(F(t * t) * p3[j])
                row.append((a + float(b)) + float(c))
            out.append(row)
            t = F(float(t) + 0.1)                                # t = t + 0.1 (double add, float store)
    elif n > 3:
        x = [[float(F(path[i][j])) for i in range(4)] for j in range(3)]   # vector<float> MS:721-723
        t = 0.0
        while t < 1.0:                                           # MS:725 double t
            row = []
            for j in range(3):
                v = math.pow(1 - t, 3) * x[j][0] + 3 * t * math.pow(1 - t, 2) * x[j][1] \
                    + 3 * math.pow(t, 2) * (1 - t) * x[j][2] + math.pow(t, 3) * x[j][3]
                row.append(v)
            out.append(row)
            t += 0.1
    else:
        p1, p2 = path[0].astype(F), path[1].astype(F)
        u1 = (p2 - p1).astype(F)                                 # MS:761, not normalised
        t = F(0)
        while float(t) < 1.1:
            out.append([float(F(p1[j] + F(u1[j] * t))) for j in range(3)])   # MS:768
            t = F(float(t) + 0.1)
    return np.asarray(out, dtype=np.float64)


def clearance(boxes, pts, radius=0.1):
    """find_labels MS:956-975: per point, min over ALL leaf boxes of fcl::distance(Sphere(r), Box)
    (GJK: -1 when the shapes intersect), starting from min_dist = 1000."""
    boxes = np.asarray(boxes, dtype=np.float64).reshape(-1, 6)
    pts = np.asarray(pts, dtype=np.float64).reshape(-1, 3)
    out = np.empty(len(pts), dtype=np.float64)
    for i, c in enumerate(pts):
        best = 1000.0
        for b in boxes:
            if b[4] < b[5]:
                continue
            d = O.sep_sphere_box(c, radius, b[:3], b[3])
            d = -1.0 if d <= 0.0 else d
            if d < best:
                best = d
        out[i] = best
    return out


def find_labels(boxes, final_path, radius=0.1, n_labels=3):
    """find_labels MS:943-996: drop the start and the goal waypoint (MS:950-952 "Removing the
    start point and the end point"; the reference writes erase(end()), which is undefined
    behaviour — the stated intent, which is also what libstdc++ does for a non-trivially-copyable
    element type, is restated), clearance per remaining waypoint stored as float, the n_labels
    waypoints with the smallest clearance -> labels (n_labels x 3 floats).
    Ordering: by (float clearance, index) — std::sort is not stable; ties are reported by the
    caller.  Returns (indices into the trimmed path, labels, clearances float32)."""
    pts = np.asarray(final_path, dtype=np.float64).reshape(-1, 3)[1:-1]
    d = clearance(boxes, pts, radius).astype(F)
    order = sorted(range(len(d)), key=lambda i: (float(d[i]), i))[:n_labels]
    labels = np.asarray([pts[i] for i in order], dtype=F).reshape(-1)
    return np.asarray(order, dtype=np.int32), labels, d
