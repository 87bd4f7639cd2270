"""Multi-GPU worker (launched by torch.distributed.run, one process per GPU): the index-range
sharded kNN with both exchange forms (NCCL all-gather + merge kernel, peer-memory merge
kernel) against the oracle and against the unsharded search.  Prints one JSON line on rank 0.
Used by tests/test_gpu_multi.py and by hand:
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tests/mgpu_worker.py"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    import bottleneck_guided_rrt_b200 as B
    from bottleneck_guided_rrt_b200 import synth
    from bottleneck_guided_rrt_b200.dist import ShardedNNSearch, shard_range
    from oracle import oracle as O

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    stream = torch.cuda.Stream(device=dev)
    out = {"world": world, "ok": True}
    with torch.cuda.stream(stream):
        n, nq, k = 300_000, 4096, 16
        pts = synth.points(11, n, 20.0)
        pts[1000:1200] = pts[:200]                      # duplicates across the set: (d2, index) ties across shards
        q = synth.points(12, nq, 20.0)
        q[:64] = pts[:64]                               # self matches (d2 = 0 rule)
        lo, hi = shard_range(n, rank, world)
        dq = torch.from_numpy(q).to(dev)
        oi, od = O.knn_brute(pts, q, k)
        for exch in ("nccl", "peer"):
            s = ShardedNNSearch(torch.from_numpy(pts[lo:hi]).to(dev), lo, stream=stream.cuda_stream, exchange=exch,
                                max_entries=65536 * k)
            for flags in (0, B.ALLOW_SELF_MATCH):
                ri, rd = (oi, od) if flags == 0 else O.knn_brute(pts, q, k, flags)
                for rep in range(3):                    # epochs: the windows are reused
                    idx, d2 = s.knn(dq, k, flags)
                    stream.synchronize()
                    good = np.array_equal(idx.cpu().numpy(), ri) and np.array_equal(d2.cpu().numpy().view(np.uint32), rd.view(np.uint32))
                    out["ok"] = out["ok"] and bool(good)
                    out["%s_flags%d_rep%d" % (exch, flags, rep)] = bool(good)
            if exch == "peer":
                out["peer_error"] = s.peer.status(stream.cuda_stream)
                out["ok"] = out["ok"] and out["peer_error"] == 0
            # timing at the bench shape: 1M nodes per rank, 65,536 queries
            big = synth.points(20 + rank, 1_000_000, 50.0)
            bq = torch.from_numpy(synth.points(30, 65536, 50.0)).to(dev)
            sb = ShardedNNSearch(torch.from_numpy(big).to(dev), rank * 1_000_000, stream=stream.cuda_stream,
                                 exchange=exch, max_entries=65536 * k)
            sb.nns.build_index()
            for _ in range(10):
                sb.knn(bq, k)
            stream.synchronize()
            dist.barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            for _ in range(20):
                sb.knn(bq, k)
            b.record(stream)
            stream.synchronize()
            t = torch.tensor([a.elapsed_time(b) / 20], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            out["ms_%s" % exch] = float(t.item())
            res_i, _ = sb.knn(bq, k)
            stream.synchronize()
            out.setdefault("big_idx", []).append(res_i.cpu().numpy())
            del s, sb
        out["big_equal"] = bool(np.array_equal(out["big_idx"][0], out["big_idx"][1]))
        out["ok"] = out["ok"] and out["big_equal"]
        del out["big_idx"]
    flag = torch.tensor([1 if out["ok"] else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    out["ok_all_ranks"] = bool(flag.item())
    if rank == 0:
        print(json.dumps(out))
    dist.destroy_process_group()
    sys.exit(0 if out["ok_all_ranks"] else 1)


if __name__ == "__main__":
    main()
