"""Octomap stream writer + reader in plain Python.  TEST INFRASTRUCTURE ONLY.

Restates the published octomap 1.8 serialisation (OcTreeBaseImpl::writeData/readData,
OcTree::writeBinaryData/readBinaryData, AbstractOcTree::write headers) that the reference
consumes through octomap_msgs::fullMsgToMap (rrt_star_milestones.cpp:279) and
fcl::OcTree::toBoxes (:291).  octomap itself is absent from this image and from
/root/reference: PARITY UNPINNED against real .ot/.bt files; the writer and the reader here
are written independently of the library's parser and check it (and each other).
"""
import math
import struct

import numpy as np

DEPTH = 16
ORIGIN = 32768
CLAMP_MAX = np.float32(math.log(0.971 / (1 - 0.971)))      # OcTree default clamping_thres_max (log-odds)
CLAMP_MIN = np.float32(math.log(0.1192 / (1 - 0.1192)))


def _prob(l):
    return 1.0 - (1.0 / (1.0 + math.exp(float(l))))


class Node:
    __slots__ = ("value", "kids")

    def __init__(self, value=None):
        self.value, self.kids = value, None


def tree_from_grid(occ, key0, free_value=None, occ_value=CLAMP_MAX):
    """occ[z][y][x] (uint8) placed with its (0,0,0) cell at key key0 (3 ints, x y z) -> pruned
    octree (root Node, n_nodes).  Occupied cells become leaves of log-odds occ_value; free cells
    become leaves of free_value (None: unknown, no node).  Eight equal leaf children collapse into
    their parent (OcTree::prune)."""
    nz, ny, nx = occ.shape
    pts = np.argwhere(np.ones_like(occ, dtype=bool) if free_value is not None else occ > 0)   # (z, y, x)
    vals = np.where(occ[pts[:, 0], pts[:, 1], pts[:, 2]] > 0, occ_value, free_value if free_value is not None else occ_value)
    keys = pts[:, ::-1] + np.asarray(key0)[None, :]
    root = Node()
    count = [1]

    def insert(node, depth, k, v):
        for d in range(depth, DEPTH):
            bit = DEPTH - 1 - d
            i = ((k[0] >> bit) & 1) | (((k[1] >> bit) & 1) << 1) | (((k[2] >> bit) & 1) << 2)
            if node.kids is None:
                node.kids = [None] * 8
            if node.kids[i] is None:
                node.kids[i] = Node()
                count[0] += 1
            node = node.kids[i]
        node.value = np.float32(v)

    for k, v in zip(keys.tolist(), vals.tolist()):
        insert(root, 0, k, v)

    def prune(node):
        if node.kids is None:
            return
        for c in node.kids:
            if c is not None:
                prune(c)
        ks = node.kids
        if all(c is not None and c.kids is None for c in ks) and all(ks[0].value == c.value for c in ks):
            node.value, node.kids = ks[0].value, None
            count[0] -= 8
        else:
            node.value = max(c.value for c in ks if c is not None)      # inner node = max of children

    prune(root)
    return root, count[0]


def write_full(root):
    """OcTreeBaseImpl::writeData: depth-first, node = f32 value + u8 child mask."""
    out = bytearray()
    stack = [root]
    while stack:
        n = stack.pop()
        mask = 0
        if n.kids is not None:
            for i, c in enumerate(n.kids):
                if c is not None:
                    mask |= 1 << i
        out += struct.pack("<fB", float(n.value), mask)
        if n.kids is not None:
            stack.extend(c for c in reversed(n.kids) if c is not None)
    return bytes(out)


def write_binary(root, occ_thres_log=0.0):
    """OcTree::writeBinaryData: node = 2 bytes, 2 bits per child: (bit 2i, bit 2i+1) =
    (1,0) free leaf, (0,1) occupied leaf, (1,1) inner, (0,0) unknown; inner children follow."""
    out = bytearray()
    if root.kids is None:
        raise ValueError("binary stream needs an inner root")
    stack = [root]
    while stack:
        n = stack.pop()
        b = [0, 0]
        inner = []
        for i, c in enumerate(n.kids):
            if c is None:
                continue
            if c.kids is not None:
                code = 3
                inner.append(c)
            elif float(c.value) >= occ_thres_log:
                code = 2                                     # bit 2i = 0, bit 2i+1 = 1
            else:
                code = 1
            b[i >> 2] |= code << (2 * (i & 3))
        out += bytes(b)
        stack.extend(reversed(inner))
    return bytes(out)


def file_bytes(data, res, n_nodes, binary, tree_id="OcTree"):
    head = "# Octomap OcTree binary file\n" if binary else "# Octomap OcTree file\n"
    head += "# (feel free to add / change comments, but leave the first line as it is!)\n#\n"
    head += "id %s\nsize %d\nres %s\ndata\n" % (tree_id, n_nodes, repr(float(res)))
    return head.encode() + data


def _box(depth, k, res, occ, thr):
    cells = 1 << (DEPTH - depth)
    size = res * float(cells)
    return [(math.floor((k[a] - ORIGIN) / float(cells)) + 0.5) * size for a in range(3)] + [size, occ, thr]


def read_full(data, res, thr=0.5, payload=0):
    """toBoxes rows of the occupied leaves of a full stream, leaf-iterator (depth-first) order."""
    rows, pos = [], 0
    stack = [(0, (0, 0, 0))]
    while stack:
        depth, k = stack.pop()
        value, = struct.unpack_from("<f", data, pos)
        mask = data[pos + 4 + payload]
        pos += 5 + payload
        if mask == 0:
            occ = _prob(np.float32(value))
            if occ >= thr:
                rows.append(_box(depth, k, res, occ, thr))
            continue
        half = 1 << (DEPTH - depth - 1)
        kids = [(depth + 1, (k[0] + (half if i & 1 else 0), k[1] + (half if i & 2 else 0), k[2] + (half if i & 4 else 0)))
                for i in range(8) if (mask >> i) & 1]
        stack.extend(reversed(kids))
    assert pos == len(data), "trailing bytes"
    return np.asarray(rows, dtype=np.float64).reshape(-1, 6)


def read_binary(data, res, thr=0.5):
    rows, pos = [], 0
    occ_leaf = _prob(CLAMP_MAX)

    def node(depth, k):
        nonlocal pos
        b0, b1 = data[pos], data[pos + 1]
        pos += 2
        half = 1 << (DEPTH - depth - 1)
        for i in range(8):
            code = ((b0 if i < 4 else b1) >> (2 * (i & 3))) & 3
            ck = (k[0] + (half if i & 1 else 0), k[1] + (half if i & 2 else 0), k[2] + (half if i & 4 else 0))
            if code == 2 and occ_leaf >= thr:
                rows.append(_box(depth + 1, ck, res, occ_leaf, thr))
            elif code == 3:
                node(depth + 1, ck)

    if len(data):
        node(0, (0, 0, 0))
    assert pos == len(data), "trailing bytes"
    return np.asarray(rows, dtype=np.float64).reshape(-1, 6)
