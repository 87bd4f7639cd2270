"""CPU suite: the N>1 host logic (index-range shards -> local top-k with global indices ->
all_gather -> merge by (d2, index)) over torch.distributed `gloo`, world_size 2 and 3, with the
oracle standing in for the per-rank CUDA search (the GPU kernels are covered by -m gpu)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, nq, k, flags, q_out):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from bottleneck_guided_rrt_b200.dist import ShardedNNSearch, merge_slabs_host, shard_range
    from oracle import oracle as O
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        pts = O.points(5, n, 10.0)
        q = O.points(6, nq, 10.0)
        lo, hi = shard_range(n, rank, world)

        class CpuShard:                      # same surface as NNSearchF, oracle arithmetic
            def __init__(self, p):
                self.p, self.base = np.ascontiguousarray(p), 0

            def set_index_base(self, b):
                self.base = b

            def knn(self, query, kk, eps, fl, max_radius, indices=None, dists2=None):
                kq = min(kk, len(self.p) - (0 if fl & 1 else 0))
                i, d = O.knn_brute(self.p, query.numpy(), kk, fl | 0, max_radius) if len(self.p) > kk else (None, None)
                assert i is not None and kq == kk
                i = np.where(i >= 0, i + self.base, -1).astype(np.int32)
                if indices is not None:          # the sharded search hands in its packed slab
                    indices.copy_(torch.from_numpy(i))
                    dists2.copy_(torch.from_numpy(d))
                    return indices, dists2, 0
                return torch.from_numpy(i), torch.from_numpy(d), 0

        def merge(si, sd, oi, od):
            mi, md = merge_slabs_host(si.numpy(), sd.numpy())
            oi.copy_(torch.from_numpy(mi.astype(np.int32)))
            od.copy_(torch.from_numpy(md))

        s = ShardedNNSearch(pts[lo:hi], lo, nns_factory=CpuShard, merge=merge)
        idx, d2 = s.knn(torch.from_numpy(q), k, flags)
        oi, od = O.knn_brute(pts, q, k, flags)
        ok = np.array_equal(idx.numpy(), oi) and np.array_equal(d2.numpy().view(np.uint32), od.view(np.uint32))
        # lattice with many ties: the merged order must still be by (d2, index), independent of sharding
        g = np.arange(9, dtype=np.float32)
        lat = np.stack(np.meshgrid(g, g, g, indexing="ij"), -1).reshape(-1, 3)
        lat = lat[np.random.default_rng(1).permutation(len(lat))]
        lq = lat[:40] + np.float32(0.5)
        lo2, hi2 = shard_range(len(lat), rank, world)
        s2 = ShardedNNSearch(lat[lo2:hi2], lo2, nns_factory=CpuShard, merge=merge)
        i2, dd2 = s2.knn(torch.from_numpy(lq), k, flags)
        oi2, od2 = O.knn_brute(lat, lq, k, flags)
        ok = ok and np.array_equal(i2.numpy(), oi2) and np.array_equal(dd2.numpy().view(np.uint32), od2.view(np.uint32))
        q_out.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_knn_gloo(world):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    port = _free_port()
    q_out = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, 6000, 64, 16, 0, q_out)) for r in range(world)]
    for p in procs:
        p.start()
    res = []
    import queue
    import time
    t0 = time.time()
    while len(res) < world and time.time() - t0 < 180:
        try:
            res.append(q_out.get(timeout=1))
        except queue.Empty:
            if any(p.exitcode not in (None, 0) for p in procs):
                break
    for p in procs:
        p.join(30)
        if p.is_alive():
            p.kill()
    assert sorted(r for r, _ in res) == list(range(world))
    assert all(ok for _, ok in res), res


def test_shard_ranges_cover_and_merge_rule():
    from bottleneck_guided_rrt_b200.dist import merge_slabs_host, shard_range
    for n in (1, 7, 1000, 50_000_000):
        for w in (1, 2, 3, 8):
            r = [shard_range(n, i, w) for i in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n and all(r[i][1] == r[i + 1][0] for i in range(w - 1))
    # two shards with a d2 tie across shards and an empty (-1,+inf) tail
    idx = np.array([[[5, 9, -1]], [[2, 7, 8]]], dtype=np.int32)
    d2 = np.array([[[1.0, 2.0, np.inf]], [[1.0, 2.0, 3.0]]], dtype=np.float32)
    mi, md = merge_slabs_host(idx, d2)
    assert mi.tolist() == [[2, 5, 7]] and md.tolist() == [[1.0, 1.0, 2.0]]
