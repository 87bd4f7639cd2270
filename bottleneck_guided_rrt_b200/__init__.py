"""bottleneck_guided_rrt_b200 — B200-native hot paths of the bottleneck-guided RRT* planner.

The product is the C-ABI shared library ``libbgrrt.so`` (include/bgrrt.h): hand-written
sm_100a CUDA kernels behind the planner's own call surface (libnabo-style NNSearch and
isStateValid / isLineValid).  This package is the thin Python binding used by the tests,
the benchmark and the multi-GPU launcher; importing it FAILS LOUDLY if the CUDA library
has not been built — there is no CPU fallback anywhere in the product path.
"""
from .capi import (  # noqa: F401
    ALLOW_SELF_MATCH,
    SORT_RESULTS,
    KNN_STATIC_MAX_RESULTS,
    TOUCH_STATISTICS,
    METHOD_AUTO,
    METHOD_BRUTE,
    METHOD_GRID,
    METHOD_GRID_GENERAL,
    SHAPE_CYLINDER,
    SHAPE_CAPSULE,
    BgrError,
    NNSearchF,
    OccupancyMap,
    lib,
    library_path,
    knn_merge_dev,
    smoothen_plan,
    octree_to_boxes,
    OCTREE_FULL_DATA,
    OCTREE_BINARY_DATA,
    OCTREE_FILE,
    PeerExchange,
    Planner,
    PlannerParams,
    PLAN_EDGE_CHECK,
    PLAN_REWIRE,
    PLAN_QUIRKS,
    PLAN_NO_GOAL,
    PLAN_DEDUP,
    PLAN_NO_PERSIST,
    PLAN_NO_GRAPH,
    PLAN_MODE_RS,
    PLAN_MODE_MS,
    microbench_l2,
)
