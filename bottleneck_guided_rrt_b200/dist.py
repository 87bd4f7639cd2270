"""Multi-GPU forms of the hot path, one process per GPU over torch.distributed.

* ``ShardedNNSearch`` — very large node sets split by contiguous index range: every rank
  searches its shard (global indices via the handle's index base), ONE ``all_gather`` of the
  per-rank (Q,k) slabs over NCCL/NVLink (gloo on CPU hosts for the tests), then the merge
  kernel ``bgr_knn_merge_dev`` keeps the k best by (d2, index) — the same tie rule as the
  single-GPU search, so the result does not depend on the number of shards.
* Independent planning queries / edge batches need no collective: ranks work on their own
  slice (``shard_range``) — "replicas only".
"""
import numpy as np


def shard_range(n, rank, world):
    """contiguous index range [lo, hi) of rank among world (GPU g owns [g*N/G, (g+1)*N/G))."""
    return (n * rank) // world, (n * (rank + 1)) // world


class ShardedNNSearch:
    def __init__(self, local_points, index_base, group=None, stream=None, nns_factory=None, merge=None,
                 exchange="nccl", max_entries=None):
        """local_points: this rank's shard ((n_local,3) float32 host array or CUDA tensor).
        nns_factory / merge are injectable so the host logic can be tested on CPU (gloo).
        exchange: "nccl" = one all_gather_into_tensor of the packed (idx,d2) slabs + the merge
        kernel; "peer" = bgr_peer_* (CUDA IPC windows, ONE merge kernel doing peer loads of this
        rank's query slice and peer stores of the merged lists; needs max_entries >= k*Q)."""
        import torch.distributed as dist
        self.dist = dist
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.exchange = exchange if self.world > 1 else "none"
        self.stream = stream or 0
        self.peer = None
        if self.exchange == "peer":
            from .capi import PeerExchange

            def ship(mine):
                allh = [None] * self.world
                dist.all_gather_object(allh, mine, group=group)
                return allh

            assert max_entries, "exchange='peer' needs max_entries (>= k*Q)"
            self.peer = PeerExchange(self.rank, self.world, int(max_entries), ship)
            dist.barrier(group=group)      # every window is mapped before anyone signals into it
        if nns_factory is None:
            from .capi import NNSearchF
            nns_factory = lambda pts: NNSearchF.createKDTreeLinearHeap(pts, stream=stream)  # noqa: E731
        if merge is None:
            from .capi import knn_merge_dev
            merge = lambda si, sd, oi, od: knn_merge_dev(si, sd, oi, od, stream or 0)  # noqa: E731
        self.nns = nns_factory(local_points)
        self.nns.set_index_base(int(index_base))
        self.merge = merge
        self._slabs = None

    def knn(self, query, k, optionFlags=0, maxRadius=float("inf"), events=None):
        """query: replicated (Q,3) CUDA tensor.  Returns merged (idx, d2) on every rank.
        events: optional 3 torch.cuda.Event objects recorded on the current stream before the local
        search, after it, and after the exchange + merge (measurement aid: local vs exchange time)."""
        import torch
        nq = query.shape[0]

        def mark(i):
            if events is not None:
                events[i].record(torch.cuda.current_stream())

        mark(0)
        if self.exchange == "peer":
            pi, pd = self.peer.slab_ptrs()
            self.nns.knn(query, k, 0.0, optionFlags, maxRadius, indices=pi, dists2=pd)
            mark(1)
            self.peer.merge(nq, k, self.stream)
            mark(2)
            if self._slabs is None or self._slabs[0] != (nq, k):
                self._slabs = ((nq, k),) + tuple(self.peer.result(nq, k))    # views of the window, made once
            return self._slabs[1], self._slabs[2]
        if self.world == 1:
            idx, d2, _ = self.nns.knn(query, k, 0.0, optionFlags, maxRadius)
            mark(1)
            mark(2)
            return idx, d2
        key = (nq, k, query.device)
        if self._slabs is None or self._slabs[0] != key:
            # packed slab: [0] = indices, [1] = d2 bits, so ONE all-gather moves both
            self._slabs = (key,
                           torch.empty((2, nq, k), dtype=torch.int32, device=query.device),
                           torch.empty((self.world, 2, nq, k), dtype=torch.int32, device=query.device),
                           torch.empty((nq, k), dtype=torch.int32, device=query.device),
                           torch.empty((nq, k), dtype=torch.float32, device=query.device))
        _, loc, slabs, oi, od = self._slabs
        self.nns.knn(query, k, 0.0, optionFlags, maxRadius, indices=loc[0], dists2=loc[1].view(torch.float32))
        mark(1)
        # concatenated-rows form (G*2*Q, k): accepted by both NCCL and gloo
        self.dist.all_gather_into_tensor(slabs.view(self.world * 2 * nq, k), loc.view(2 * nq, k), group=self.group)
        self.merge(slabs[:, 0], slabs[:, 1].view(torch.float32), oi, od)
        mark(2)
        return oi, od


def merge_slabs_host(idx_slabs, d2_slabs):
    """numpy statement of the merge rule, used by the CPU (gloo) tests of the host logic:
    k best of G*k candidates per query by (d2 bits, index), empty slots (-1, +inf) last."""
    g, nq, k = idx_slabs.shape
    i = np.transpose(idx_slabs, (1, 0, 2)).reshape(nq, g * k)
    d = np.transpose(d2_slabs, (1, 0, 2)).reshape(nq, g * k)
    key = (d.view(np.uint32).astype(np.uint64) << np.uint64(32)) | i.astype(np.uint32).astype(np.uint64)
    key[i < 0] = np.uint64(0xFFFFFFFFFFFFFFFF)
    order = np.argsort(key, axis=1, kind="stable")[:, :k]
    return np.take_along_axis(i, order, 1), np.take_along_axis(d, order, 1)
