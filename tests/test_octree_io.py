"""OcTree ingestion (bgr_octree_to_boxes, host code: no GPU needed) against the plain-Python
restatement of the octomap 1.8 stream formats (oracle/octree_io.py) and hand-built streams.
Reference call sites: octomap_msgs::fullMsgToMap rrt_star_milestones.cpp:279, toBoxes :291."""
import struct

import numpy as np
import pytest


def _grid(seed, n=32, p=0.08):
    rng = np.random.default_rng(seed)
    occ = (rng.random((n, n, n)) < p).astype(np.uint8)
    occ[:, 8:16, 8:16] = 1          # a pruned slab (8x8 columns): multi-level leaves
    occ[16:32, 16:32, 16:32] = 1    # one 16^3 block -> a level-4 leaf
    return occ


def _sorted_rows(b):
    b = np.asarray(b).reshape(-1, 6)
    return b[np.lexsort((b[:, 0], b[:, 1], b[:, 2], b[:, 3]))]


def test_hand_built_full_stream():
    import bottleneck_guided_rrt_b200 as B
    # root -> child 7 (+x+y+z) ... 15 more levels of child 0 -> leaf at key (32768, 32768, 32768), log-odds 2.0
    data = b""
    for d in range(16):
        data += struct.pack("<fB", 2.0, (1 << 7) if d == 0 else 1)
    data += struct.pack("<fB", 2.0, 0)
    boxes, res, lo, hi = B.octree_to_boxes(data, B.OCTREE_FULL_DATA, 0.1)
    assert boxes.shape == (1, 6)
    assert np.allclose(boxes[0, :4], [0.05, 0.05, 0.05, 0.1], rtol=0, atol=1e-15)
    assert abs(boxes[0, 4] - (1 - 1 / (1 + np.exp(2.0)))) < 1e-15 and boxes[0, 5] == 0.5
    assert np.allclose(lo, [0, 0, 0]) and np.allclose(hi, [0.1, 0.1, 0.1])
    # a free leaf (log-odds < 0) is not a box; a depth-1 leaf is a 3276.8 m cube centred at -1638.4
    free = struct.pack("<fB", -1.0, 1) + struct.pack("<fB", -1.0, 0)
    assert B.octree_to_boxes(free, B.OCTREE_FULL_DATA, 0.1)[0].shape == (0, 6)
    big = struct.pack("<fB", 1.0, 1) + struct.pack("<fB", 1.0, 0)
    b = B.octree_to_boxes(big, B.OCTREE_FULL_DATA, 0.1)[0]
    assert np.allclose(b[0, :4], [-1638.4, -1638.4, -1638.4, 3276.8])
    # empty message
    assert B.octree_to_boxes(b"", B.OCTREE_FULL_DATA, 0.1)[0].shape == (0, 6)


def test_hand_built_binary_stream():
    import bottleneck_guided_rrt_b200 as B
    # root: child 7 inner; then 14 levels: child 0 inner; last level: child 0 occupied (01 -> value 2), child 1 free (value 1)
    data = bytes([0, 3 << 6])
    for _ in range(14):
        data += bytes([3, 0])
    data += bytes([2 | (1 << 2), 0])
    boxes, _, _, _ = B.octree_to_boxes(data, B.OCTREE_BINARY_DATA, 0.05)
    assert boxes.shape == (1, 6) and np.allclose(boxes[0, :4], [0.025, 0.025, 0.025, 0.05], rtol=0, atol=1e-15)
    assert abs(boxes[0, 4] - 0.971) < 1e-6


@pytest.mark.parametrize("seed,free", [(1, False), (2, True)])
def test_streams_against_python_restatement(oracle, seed, free):
    import bottleneck_guided_rrt_b200 as B
    from oracle import octree_io as T
    occ = _grid(seed)
    res, key0 = 0.05, (T.ORIGIN + 64, T.ORIGIN - 32, T.ORIGIN)
    root, n_nodes = T.tree_from_grid(occ, key0, free_value=T.CLAMP_MIN if free else None)
    full, binary = T.write_full(root), T.write_binary(root)
    origin = tuple((k - T.ORIGIN) * res for k in key0)
    want = _sorted_rows(oracle.grid_to_boxes(occ, res, origin, 5))
    for fmt, data, reader in ((B.OCTREE_FULL_DATA, full, T.read_full), (B.OCTREE_BINARY_DATA, binary, T.read_binary)):
        got, _, lo, hi = B.octree_to_boxes(data, fmt, res)
        ref = reader(data, res)
        assert got.shape == ref.shape and np.array_equal(got.view(np.uint64), ref.view(np.uint64))   # same rows, same order
        g = _sorted_rows(got)
        assert g.shape == want.shape and np.allclose(g[:, :4], want[:, :4], rtol=0, atol=1e-9)        # = the pruned leaf set of the grid
        assert int(round(((got[:, 3] / res) ** 3).sum())) == int(occ.sum())
        assert np.all(lo >= np.asarray(origin) - 1e-9) and np.all(hi <= np.asarray(origin) + 32 * res + 1e-9)
    # whole files, incl. a ColorOcTree payload
    for binary_file, data in ((False, full), (True, binary)):
        fb = T.file_bytes(data, res, n_nodes, binary_file)
        got, r, _, _ = B.octree_to_boxes(fb, B.OCTREE_FILE)
        assert r == res and np.array_equal(got, B.octree_to_boxes(data, B.OCTREE_BINARY_DATA if binary_file else B.OCTREE_FULL_DATA, res)[0])
    col = b"".join(full[i:i + 4] + b"\x10\x20\x30" + full[i + 4:i + 5] for i in range(0, len(full), 5))
    got, _, _, _ = B.octree_to_boxes(T.file_bytes(col, res, n_nodes, False, "ColorOcTree"), B.OCTREE_FILE)
    assert np.array_equal(got, B.octree_to_boxes(full, B.OCTREE_FULL_DATA, res)[0])


def test_bad_streams(oracle):
    import bottleneck_guided_rrt_b200 as B
    from oracle import octree_io as T
    root, n = T.tree_from_grid(_grid(3), (T.ORIGIN, T.ORIGIN, T.ORIGIN))
    full = T.write_full(root)
    with pytest.raises(RuntimeError):
        B.octree_to_boxes(full[:-3], B.OCTREE_FULL_DATA, 0.05)            # truncated
    with pytest.raises(RuntimeError):
        B.octree_to_boxes(full, B.OCTREE_FULL_DATA, 0.0)                   # no resolution
    with pytest.raises(RuntimeError):
        B.octree_to_boxes(b"# something else\n", B.OCTREE_FILE)
    with pytest.raises(RuntimeError):
        B.octree_to_boxes(T.file_bytes(full, 0.05, n, False, "CountingOcTree"), B.OCTREE_FILE)
