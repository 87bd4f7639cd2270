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
    def __init__(self, local_points, index_base, group=None, stream=None, nns_factory=None, merge=None):
        """local_points: this rank's shard ((n_local,3) float32 host array or CUDA tensor).
        nns_factory / merge are injectable so the host logic can be tested on CPU (gloo)."""
        import torch.distributed as dist
        self.dist = dist
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
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

    def knn(self, query, k, optionFlags=0, maxRadius=float("inf")):
        """query: replicated (Q,3) CUDA tensor.  Returns merged (idx, d2) on every rank."""
        import torch
        idx, d2, _ = self.nns.knn(query, k, 0.0, optionFlags, maxRadius)
        if self.world == 1:
            return idx, d2
        nq = idx.shape[0]
        key = (nq, k, idx.device)
        if self._slabs is None or self._slabs[0] != key:
            self._slabs = (key,
                           torch.empty((self.world, nq, k), dtype=idx.dtype, device=idx.device),
                           torch.empty((self.world, nq, k), dtype=d2.dtype, device=d2.device),
                           torch.empty_like(idx), torch.empty_like(d2))
        _, si, sd, oi, od = self._slabs
        # concatenated-rows form (G*Q, k): accepted by both NCCL and gloo
        self.dist.all_gather_into_tensor(si.view(self.world * nq, k), idx.contiguous(), group=self.group)
        self.dist.all_gather_into_tensor(sd.view(self.world * nq, k), d2.contiguous(), group=self.group)
        self.merge(si, sd, oi, od)
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
