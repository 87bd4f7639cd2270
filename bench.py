#!/usr/bin/env python
"""bench.py — kNN queries/s (+ edge collision checks/s) on N B200s, beside the CPU path.

  python bench.py --gpus N --steps K --warmup W            (torchrun launches N>1)
  python bench.py --impl reference --gpus N --steps K ...  (CPU arm: the oracle's libnabo /
                                                            FCL-MPR restatement, all host threads)

Headline workload (N=1) = BASELINE.json configs[1]: 1,000,000-node 3-D tree, 65,536 queries,
k=16, eps=0, exact.  A "step" = one pass of bgr_nns_knn over one batch of 65,536 synthetic
queries.  `value` = queries/s with inputs resident in HBM (device pointers, CUDA events on
the library's stream); `e2e` = the same call through the C ABI with HOST (pinned) buffers,
H2D + D2H inside the timed region.  The second metric of BASELINE.json (edge collision
checks/s, configs[2]: 1M cylinder edges vs a 512^3 map) is reported in "edge_checks" with
its own roofline / e2e / cpu_baseline.  For N>1 every rank runs the same per-GPU workload
on its own queries/edges (replicas, no data-path collective: "scaling": "weak"); the
index-range-sharded search with its NCCL all-gather + merge is reported in "sharded_knn".
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "kNN queries/s + edge collision checks/s at 1/2/4/8 B200 vs libnabo+FCL CPU"
N_NODES, N_QUERIES, K_NN = 1_000_000, 65_536, 16
MAP_N, N_EDGES, EDGE_R = 512, 1_000_000, 0.03
SHARD_NODES = 6_250_000           # config 5(ii): 50M nodes / 8 GPUs
FP32_NOFMA_LANES = 128            # FP32 lanes per SM; 8 non-FMA flops per (query,node) pair


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return {"hbm_gbs": float(d["hbm_gbs"]), "sm_max_mhz": float(d.get("sm_max_mhz", 1965.0)), "src": "measured"}
    return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0, "src": "fallback"}


def ncu_traffic(kernel):
    """per-launch dram__bytes_read.sum + dram__bytes_write.sum of `kernel` from the committed
    `ncu --set full` capture of this workload (profiles/traffic.json, written by
    scripts/ncu_traffic.py from the .ncu-rep), or None."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(p):
        return None
    with open(p) as f:
        d = json.load(f)
    e = d.get(kernel)
    return float(e["dram_bytes"]) if e else None


class ClockSampler:
    """samples SM clock + throttle reasons DURING the timed region (pynvml, else nvidia-smi)."""

    def __init__(self, index):
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "sw_thermal_slowdown": 0x20,
                 "hw_thermal_slowdown": 0x40, "hw_power_brake_slowdown": 0x80}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.002)

    def start(self):
        if self.nv is not None:
            self._t = threading.Thread(target=self._loop, daemon=True)
            self._t.start()

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join()
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


# ------------------------------------------------------------------------------ GPU arm
def timed_steps(torch, stream, step_fn, steps, warmup, flush, barrier, prewarm_s=0.25, prewarm_iters=None):
    """W warm-up steps, then exactly K timed steps; each step has its own CUDA event pair on
    `stream` and L2 is flushed (untimed) between steps.  Returns total ms of the K steps.
    A step that contains a collective must pass prewarm_iters: a wall-clock-bounded ramp would
    run a different number of collectives on each rank (deadlock)."""
    if prewarm_iters is not None:
        for _ in range(prewarm_iters):
            step_fn()
        stream.synchronize()
    else:
        t0 = time.perf_counter()
        while time.perf_counter() - t0 < prewarm_s:      # untimed clock ramp (idle SM clock is 120 MHz)
            step_fn()
            stream.synchronize()
    for _ in range(warmup):
        flush()
        step_fn()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    stream.synchronize()
    barrier()
    torch.cuda.synchronize()
    for a, b in evs:
        flush()
        a.record(stream)
        step_fn()
        b.record(stream)
    torch.cuda.synchronize()
    barrier()
    return sum(a.elapsed_time(b) for a, b in evs)


def run_ours(args):
    import torch
    import bottleneck_guided_rrt_b200 as B
    from bottleneck_guided_rrt_b200 import synth
    from bottleneck_guided_rrt_b200.dist import ShardedNNSearch, shard_range

    rank, world, local = dist_env()
    assert world == args.gpus, "launch with torchrun --nproc-per-node %d" % args.gpus
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    L = B.lib()
    pk = peaks()
    dev = torch.device("cuda", local)
    stream = torch.cuda.Stream(device=dev)
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def barrier():
        if dist is not None:
            dist.barrier()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def flush():
        with torch.cuda.stream(stream):
            flush_buf.fill_(rank & 0xFF)

    out = {"metric": METRIC, "unit": "queries/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
           "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic"}
    sampler = ClockSampler(local)

    with torch.cuda.stream(stream):
        # ---------------------------------------------------------------- kNN (configs[1])
        pts = synth.points(2, N_NODES, 25.6)
        q_host = synth.points(3 + rank, N_QUERIES, 25.6)
        d_pts = torch.from_numpy(pts).to(dev)
        nns = B.NNSearchF.createKDTreeLinearHeap(d_pts, creationOptionFlags=B.TOUCH_STATISTICS,
                                                 stream=stream.cuda_stream)
        nns.set_method(B.METHOD_GRID if args.method == "grid" else B.METHOD_BRUTE if args.method == "brute" else B.METHOD_AUTO)
        t0 = time.perf_counter()
        nns.build_index()
        stream.synchronize()
        build_ms = (time.perf_counter() - t0) * 1e3
        d_q = torch.from_numpy(q_host).to(dev)
        d_idx = torch.empty((N_QUERIES, K_NN), dtype=torch.int32, device=dev)
        d_d2 = torch.empty((N_QUERIES, K_NN), dtype=torch.float32, device=dev)
        _, _, pairs = nns.knn(d_q, K_NN, indices=d_idx, dists2=d_d2)      # counts evaluated pairs
        stream.synchronize()
        nns_t = B.NNSearchF.createKDTreeLinearHeap(d_pts, stream=stream.cuda_stream)  # timed handle: no statistics
        nns_t.set_method(B.METHOD_GRID if args.method == "grid" else B.METHOD_BRUTE if args.method == "brute" else B.METHOD_AUTO)
        nns_t.build_index()

        def knn_step():
            nns_t.knn(d_q, K_NN, indices=d_idx, dists2=d_d2)

        sampler.start()
        l0 = L.bgr_launch_count()
        knn_ms = max_over_ranks(timed_steps(torch, stream, knn_step, args.steps, args.warmup, flush, barrier))
        del l0
        # launches per step, counted on one isolated step
        l1 = L.bgr_launch_count()
        knn_step()
        stream.synchronize()
        per_step_launches = L.bgr_launch_count() - l1

        # e2e: host (pinned) buffers through the host-pointer C ABI, H2D + D2H inside the timing
        qh = torch.from_numpy(q_host).pin_memory()
        ih = torch.empty((N_QUERIES, K_NN), dtype=torch.int32).pin_memory()
        dh = torch.empty((N_QUERIES, K_NN), dtype=torch.float32).pin_memory()
        qh_np, ih_np, dh_np = qh.numpy(), ih.numpy(), dh.numpy()

        def knn_e2e_step():
            nns_t.knn(qh_np, K_NN, indices=ih_np, dists2=dh_np)

        e2e_ms = max_over_ranks(timed_steps(torch, stream, knn_e2e_step, args.steps, args.warmup, flush, barrier, 0.1))
        clocks = sampler.stop()
        assert np.array_equal(ih_np, d_idx.cpu().numpy()), "e2e and device-resident results differ"

        ms_per_step = knn_ms / args.steps
        value = world * N_QUERIES / (ms_per_step * 1e-3)
        pairs_per_step = float(pairs)
        sm_mhz = clocks["sm_mhz"] or pk["sm_max_mhz"]
        fp32_peak = 148 * FP32_NOFMA_LANES * pk["sm_max_mhz"] * 1e6 / 1e12     # TFLOP/s, non-FMA
        achieved = 8.0 * pairs_per_step / (ms_per_step * 1e-3) / 1e12
        alg_bytes = 16.0 * N_NODES + 12.0 * N_QUERIES + 8.0 * K_NN * N_QUERIES
        out.update({
            "value": value, "ms_per_step": ms_per_step,
            "config": {"workload": "configs[1]: batched kNN, 1,000,000-node 3-D tree (U[0,25.6)^3, MT19937 seed 2), "
                                   "65,536 queries/GPU (seed 3+rank), k=16, eps=0, exact, method=%s" % args.method,
                       "nodes": N_NODES, "queries_per_gpu": N_QUERIES, "k": K_NN,
                       "l2": "flushed between steps (256 MiB fill on the same stream, untimed)",
                       "prewarm": "0.25 s untimed clock ramp before the W warm-up steps",
                       "index_build_ms": build_ms, "parallelism": "replicas x%d" % world},
            "e2e": {"value": world * N_QUERIES / (e2e_ms / args.steps * 1e-3), "unit": "queries/s",
                    "ms_per_step": e2e_ms / args.steps,
                    "h2d_bytes_per_step": int(12 * N_QUERIES), "d2h_bytes_per_step": int(8 * K_NN * N_QUERIES)},
            "gpu_launches": int(per_step_launches * args.steps),
            "clocks": clocks,
            # top-level figures: HBM roofline of the kernel (the contract's "hbm" bound: algorithmic bytes of
            # SURVEY 8(d) / measured launch time / measured HBM peak).  The node index is L2 resident and the
            # kernel is bound by the FP32 / issue pipes, not HBM: that (non-tensor) roofline is in "fp32".
            "roofline": {"bound": "hbm", "kernel": "grid_knn_kernel" if args.method != "brute" else "knn_brute_kernel<8>",
                         "achieved": alg_bytes / (ms_per_step * 1e-3) / 1e9, "peak": pk["hbm_gbs"], "unit": "GB/s",
                         "frac": alg_bytes / (ms_per_step * 1e-3) / 1e9 / pk["hbm_gbs"], "peak_src": pk["src"],
                         "algorithmic_bytes": alg_bytes,
                         "traffic": ncu_traffic("grid_knn_kernel" if args.method != "brute" else "knn_brute_kernel"),
                         "fp32": {"achieved": achieved, "peak": fp32_peak, "unit": "TFLOP/s", "frac": achieved / fp32_peak,
                                  "peak_src": "148 SM x 128 FP32 lanes x %s sm_max_mhz (non-FMA: the bit-exact contract "
                                              "forbids FMA)" % pk["src"],
                                  "pairs_evaluated_per_step": pairs_per_step, "flops_per_pair": 8, "sm_mhz_during_run": sm_mhz},
                         "note": "node index L2 resident (DRAM traffic per L2-flushed launch = the index read once); the "
                                 "kernel is issue / dependent-load bound (ncu: profiles/), see DESIGN.md 3.2"},
        })

        # ---------------------------------------------------------------- edges (configs[2])
        if not args.skip_edges:
            occ = synth.make_map(MAP_N, 4, n_walls=8, wall_thick=4, door_cells=5)
            boxes = synth.grid_to_boxes(occ, 0.05, (0, 0, 0), 3)
            omap = B.OccupancyMap(0.05, (0, 0, 0), (MAP_N,) * 3, stream=stream.cuda_stream)
            omap.set_boxes(boxes)
            p1, p2 = synth.edges(4 + rank, N_EDGES, occ)
            d_p1, d_p2 = torch.from_numpy(p1).to(dev), torch.from_numpy(p2).to(dev)
            d_hit = torch.empty(N_EDGES, dtype=torch.uint8, device=dev)
            d_band = torch.empty(N_EDGES, dtype=torch.uint8, device=dev)

            def edge_step():
                omap.collide_edges(d_p1, d_p2, EDGE_R, B.SHAPE_CYLINDER, hit=d_hit, band=d_band)

            l1 = L.bgr_launch_count()
            edge_step()
            stream.synchronize()
            edge_launches = L.bgr_launch_count() - l1
            stats = omap.last_stats()
            stats.update(omap.gate_stats())
            edge_ms = max_over_ranks(timed_steps(torch, stream, edge_step, args.steps, args.warmup, flush, barrier))
            p1h, p2h = torch.from_numpy(p1).pin_memory(), torch.from_numpy(p2).pin_memory()
            hh, bh = torch.empty(N_EDGES, dtype=torch.uint8).pin_memory(), torch.empty(N_EDGES, dtype=torch.uint8).pin_memory()
            p1n, p2n, hn, bn = p1h.numpy(), p2h.numpy(), hh.numpy(), bh.numpy()

            def edge_e2e_step():
                omap.collide_edges(p1n, p2n, EDGE_R, B.SHAPE_CYLINDER, hit=hn, band=bn)

            edge_e2e_ms = max_over_ranks(timed_steps(torch, stream, edge_e2e_step, args.steps, args.warmup, flush, barrier, 0.1))
            ems = edge_ms / args.steps
            map_bytes = (MAP_N ** 3) / 8
            ebytes = 28.0 * N_EDGES + map_bytes
            # per-kernel durations: CUDA events recorded by the library between its launches, on its stream
            omap.set_timing(True)
            kms = {}
            for _ in range(args.steps):
                flush()
                edge_step()
                for kname, v in omap.last_kernel_ms().items():
                    kms[kname] = kms.get(kname, 0.0) + v / args.steps
            omap.set_timing(False)
            sweep_ms = kms["edges_sweep_kernel"]
            out["edge_checks"] = {
                "value": world * N_EDGES / (ems * 1e-3), "unit": "edge checks/s", "ms_per_step": ems,
                "config": {"workload": "configs[2]: 1,000,000 cylinder edges/GPU (r=0.03, L log-uniform [0.15,3] m) vs 512^3 "
                                       "occupancy map @0.05 m, %.1f%% occupied, %d leaf boxes" % (100 * occ.mean(), len(boxes)),
                           "hit_fraction": float(d_hit.float().mean().item()), "in_band": int(d_band.sum().item()),
                           "sweep": stats},
                "e2e": {"value": world * N_EDGES / (edge_e2e_ms / args.steps * 1e-3), "unit": "edge checks/s",
                        "ms_per_step": edge_e2e_ms / args.steps,
                        "h2d_bytes_per_step": int(24 * N_EDGES), "d2h_bytes_per_step": int(2 * N_EDGES)},
                "gpu_launches": int(edge_launches * args.steps),
                "roofline": {"bound": "hbm", "kernel": "edges_sweep_kernel", "achieved": ebytes / (sweep_ms * 1e-3) / 1e9,
                             "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": ebytes / (sweep_ms * 1e-3) / 1e9 / pk["hbm_gbs"],
                             "algorithmic_bytes": ebytes, "peak_src": pk["src"], "traffic": ncu_traffic("edges_sweep_kernel"),
                             "kernel_ms": kms, "kernel_share_of_step": sweep_ms / max(sum(kms.values()), 1e-9),
                             "note": "the 16 MiB bitmap is L2 resident: the sweep is bound by instruction issue and L2 latency, "
                                     "not HBM (ncu: profiles/); the HBM figure is the contract of SURVEY 8(d)",
                             "work_items_per_s": stats["work_items"] / (sweep_ms * 1e-3)},
            }

            # sphere state checks (isStateValid MS:310-351) on the same map: the edges' first endpoints
            d_shit = torch.empty(N_EDGES, dtype=torch.uint8, device=dev)
            d_sband = torch.empty(N_EDGES, dtype=torch.uint8, device=dev)

            def sphere_step():
                omap.collide_spheres(d_p1, 0.1, hit=d_shit, band=d_sband)

            sphere_ms = max_over_ranks(timed_steps(torch, stream, sphere_step, args.steps, args.warmup, flush, barrier, 0.1)) / args.steps
            # the same call with HOST buffers (pinned), copies inside the timed region
            shh, sbh = torch.empty(N_EDGES, dtype=torch.uint8).pin_memory(), torch.empty(N_EDGES, dtype=torch.uint8).pin_memory()
            shn, sbn = shh.numpy(), sbh.numpy()

            def sphere_e2e_step():
                omap.collide_spheres(p1n, 0.1, hit=shn, band=sbn)

            sphere_e2e_ms = max_over_ranks(timed_steps(torch, stream, sphere_e2e_step, args.steps, args.warmup, flush, barrier, 0.1)) / args.steps
            assert np.array_equal(shn, d_shit.cpu().numpy()) and np.array_equal(sbn, d_sband.cpu().numpy()), \
                "sphere checks: host-buffer and device-resident results differ"
            sbytes = 14.0 * N_EDGES + map_bytes
            out["state_checks"] = {
                "value": world * N_EDGES / (sphere_ms * 1e-3), "unit": "sphere state checks/s", "ms_per_step": sphere_ms,
                "config": {"workload": "1,000,000 spheres/GPU (r=0.1) vs the configs[2] map", "hit_fraction": float(d_shit.float().mean().item()),
                           "in_band": int(d_sband.sum().item())},
                "e2e": {"value": world * N_EDGES / (sphere_e2e_ms * 1e-3), "unit": "sphere state checks/s", "ms_per_step": sphere_e2e_ms,
                        "h2d_bytes_per_step": int(12 * N_EDGES), "d2h_bytes_per_step": int(2 * N_EDGES)},
                "roofline": {"bound": "hbm", "kernel": "collide_spheres_kernel", "achieved": sbytes / (sphere_ms * 1e-3) / 1e9,
                             "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": sbytes / (sphere_ms * 1e-3) / 1e9 / pk["hbm_gbs"],
                             "algorithmic_bytes": sbytes, "peak_src": pk["src"], "traffic": None}}

        # ---------------------------------------------------------------- batched planner (configs[0], configs[3])
        if not args.skip_planner:
            out["planner"] = planner_bench(torch, B, synth, stream, rank, world, max_over_ranks, barrier)

        # ---------------------------------------------------------------- sharded kNN (N>1)
        if world > 1 and not args.skip_sharded:
            lo, hi = shard_range(SHARD_NODES * world, rank, world)
            shard = torch.from_numpy(synth.points(5 + 1000 * rank, hi - lo, 102.4)).to(dev)
            sq = torch.from_numpy(synth.points(6, N_QUERIES, 102.4)).to(dev)
            sh = {"nodes_total": SHARD_NODES * world, "nodes_per_gpu": SHARD_NODES, "queries": N_QUERIES, "k": K_NN,
                  "unit": "queries/s"}
            results = {}
            for exch in ("nccl", "peer"):
                try:
                    snn = ShardedNNSearch(shard, lo, stream=stream.cuda_stream, exchange=exch, max_entries=N_QUERIES * K_NN)
                    snn.nns.build_index()

                    def shard_step():
                        snn.knn(sq, K_NN)

                    sh_ms = max_over_ranks(timed_steps(torch, stream, shard_step, args.steps, args.warmup, flush, barrier,
                                                       prewarm_iters=20))
                    ri, _ = snn.knn(sq, K_NN)
                    stream.synchronize()
                    results[exch] = ri.clone()
                    sh[exch] = {"value": N_QUERIES / (sh_ms / args.steps * 1e-3), "ms_per_step": sh_ms / args.steps}
                    if exch == "peer":
                        sh[exch]["spin_timeouts"] = snn.peer.status(stream.cuda_stream)
                    del snn
                except Exception as ex:      # reported, never hidden: the other exchange still stands
                    sh[exch] = {"error": repr(ex)[:300]}
            if len(results) == 2:
                sh["exchanges_identical"] = bool(torch.equal(results["nccl"], results["peer"]))
            best = max((v for v in (sh.get("peer"), sh.get("nccl")) if v and "value" in v), key=lambda v: v["value"], default=None)
            if best:
                sh["value"], sh["ms_per_step"] = best["value"], best["ms_per_step"]
            sh["exchange"] = {"nccl": "ONE all_gather_into_tensor of the packed (idx,d2) slabs (%d B per rank) + bgr_knn_merge_strided_dev" % (8 * K_NN * N_QUERIES),
                              "peer": "bgr_peer_merge_dev: one kernel, peer loads of this rank's query slice from all slabs over NVLink, "
                                      "merge by (d2,index), peer stores of the merged lists (%d B in + out per rank)" % (8 * K_NN * N_QUERIES * (world - 1) // world)}
            out["sharded_knn"] = sh

    # -------------------------------------------------------------------- CPU baseline (N=1, rank 0)
    if rank == 0 and world == 1 and not args.skip_cpu:
        out["cpu_baseline"] = cpu_knn_baseline(pts, q_host, budget_s=8.0)
        if "edge_checks" in out:
            out["edge_checks"]["cpu_baseline"] = cpu_edge_baseline(boxes, p1, p2, budget_s=10.0)
    if rank == 0:
        print(json.dumps(out))
    if dist is not None:
        dist.destroy_process_group()


def planner_bench(torch, B, synth, stream, rank, world, max_over_ranks, barrier):
    """Tree growth of the batched planner (bgr_planner_*): configs[0] = RRT* with 10 % goal bias on a
    64^3 map to 5,000 nodes, k=10 (batch 1 = the reference's sequential loop, and batch 256);
    configs[3] = bottleneck-guided RRT* (3 milestones) on a 256^3 narrow-passage map to 200,000 nodes,
    batch 4096, K=16, edge checks + rewiring.  One independent planning query per GPU (replicas)."""
    res = {}

    def grow(n_map, seed, walls, door, mode, k, batch, flags, max_nodes, milestones, goal):
        occ = synth.make_map(n_map, seed, n_walls=walls, wall_thick=2 if n_map == 64 else 4, door_cells=door)
        boxes = synth.grid_to_boxes(occ, 0.05, (0, 0, 0), 3)
        omap = B.OccupancyMap(0.05, (0, 0, 0), (n_map,) * 3, stream=stream.cuda_stream)
        omap.set_boxes(boxes)
        ext = n_map * 0.05
        table = synth.points(9 + rank, 200_000, ext)
        P = B.PlannerParams((0.2, 0.2, 0.2), goal, mode=mode, k=k, batch=batch, flags=flags | B.PLAN_NO_GOAL, seed=5489 + rank)
        pl = B.Planner(P, omap, table, milestones, stream=stream.cuda_stream)
        pl.run(min(max_nodes, 64 * batch + 64), 3)                    # warm-up: first launches, allocations
        stream.synchronize()
        barrier()
        n0, t0 = len(pl), time.perf_counter()
        pl.run(max_nodes)
        stream.synchronize()
        el = max_over_ranks(time.perf_counter() - t0)
        st = pl.stats()
        return {"nodes": len(pl), "seconds": el, "nodes_per_s": world * (len(pl) - n0) / el,
                "candidates_per_s": world * st["candidates"] / el, "batches": st["batches"], "batch": batch, "k": k,
                "rejected_state": st["rejected_state"], "rejected_edge": st["rejected_edge"], "rewired": st["rewired"],
                "in_band": st["in_band"], "map": "%d^3 @0.05 m, %d leaf boxes" % (n_map, len(boxes))}

    g64 = (3.0, 3.0, 3.0)
    res["config0_rrtstar_64_batch1"] = grow(64, 1, 3, 6, B.PLAN_MODE_RS, 10, 1, B.PLAN_QUIRKS, 5000, None, g64)
    res["config0_rrtstar_64_batch256"] = grow(64, 1, 3, 6, B.PLAN_MODE_RS, 10, 256, 0, 5000, None, g64)
    # the 3 milestone points = the door centres of the 3 walls (what the CNN is trained to predict)
    occ3 = synth.make_map(256, 5, n_walls=3, wall_thick=4, door_cells=5)
    ms = []
    for w in range(3):
        x0 = (w + 1) * 256 // 4
        zz, yy = np.nonzero(occ3[:, :, x0] == 0)
        ms.append([(x0 + 0.5) * 0.05, (yy.mean() + 0.5) * 0.05, (zz.mean() + 0.5) * 0.05])
    ms = np.asarray(ms, dtype=np.float32)
    # wall clock of ~860 host-driven steps: sensitive to whatever else the host is doing, so the growth is
    # repeated (same seed, same tree) and every run is listed; the headline is the fastest
    runs3 = [grow(256, 5, 3, 5, B.PLAN_MODE_MS, 16, 4096, B.PLAN_EDGE_CHECK | B.PLAN_REWIRE, 200_000, ms, (12.6, 12.6, 12.6))
             for _ in range(3)]
    best3 = min(runs3, key=lambda r: r["seconds"])
    best3["seconds_all_runs"] = [r["seconds"] for r in runs3]
    res["config3_bottleneck_256_batch4096"] = best3
    # configs[4](i): independent start/goal planning queries on a 256^3 map, each run to the goal
    # (bounded sample: 8 queries per GPU instead of 512; cap 60,000 nodes per query).  The queries
    # are stepped round-robin by bgr_planner_run_many, each on its own stream and map handle, so
    # their small per-step kernels overlap; the same queries one after the other are timed beside.
    boxes3 = synth.grid_to_boxes(occ3, 0.05, (0, 0, 0), 3)
    nq_local = 8
    qstreams = [torch.cuda.Stream() for _ in range(nq_local)]
    qmaps = []
    for qs in qstreams:
        qm = B.OccupancyMap(0.05, (0, 0, 0), (256,) * 3, stream=qs.cuda_stream)
        qm.set_boxes(boxes3)
        qmaps.append(qm)
    cand = synth.points(77 + rank, 4000, 12.8)
    chit, _ = qmaps[0].collide_spheres(cand, 0.1)
    free = cand[np.asarray(chit) == 0]
    table = synth.points(9 + rank, 200_000, 12.8)

    def make_queries():
        out = []
        for qi in range(nq_local):
            P = B.PlannerParams(tuple(free[2 * qi]), tuple(free[2 * qi + 1]), mode=B.PLAN_MODE_MS, k=16, batch=1024,
                                flags=B.PLAN_EDGE_CHECK | B.PLAN_REWIRE, seed=100 + rank * 512 + qi)
            out.append(B.Planner(P, qmaps[qi], table, ms, stream=qstreams[qi].cuda_stream))
        return out

    pls = make_queries()
    torch.cuda.synchronize()
    barrier()
    t0 = time.perf_counter()
    flags = B.Planner.run_many(pls, 60_000)
    torch.cuda.synchronize()
    el = max_over_ranks(time.perf_counter() - t0)
    nodes = sum(len(p) for p in pls)
    del pls
    pls = make_queries()
    torch.cuda.synchronize()
    barrier()
    t0 = time.perf_counter()
    for p in pls:
        p.run(60_000)
    torch.cuda.synchronize()
    el_seq = max_over_ranks(time.perf_counter() - t0)
    res["config4_independent_queries_256"] = {
        "queries": world * nq_local, "queries_per_gpu": nq_local, "seconds": el, "queries_per_s": world * nq_local / el,
        "one_after_the_other": {"seconds": el_seq, "queries_per_s": world * nq_local / el_seq},
        "reached_goal_rank0": int(sum(flags)), "mean_nodes_rank0": nodes / nq_local, "batch": 1024, "k": 16, "max_nodes": 60_000,
        "sample": "8 of the 512 queries per GPU of configs[4](i); random free start/goal pairs, 3 milestones at the door "
                  "centres; bgr_planner_run_many (round-robin steps, one stream + map handle per query)"}
    res["note"] = ("wall clock of bgr_planner_run (host loop + device kernels, one sync per batch); nodes_per_s is the "
                   "aggregate over %d independent planning queries (one per GPU)" % world)
    return res


# ------------------------------------------------------------------------------ CPU arm
def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def host_cpu_model():
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("model name"):
                    return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def cpu_knn_baseline(pts, q, budget_s):
    """the oracle's libnabo-1.0.7 kd-tree restatement (kind "port"), OpenMP over queries on all
    host cores; index prebuilt (as for the GPU arm); repeated until ~budget_s of wall time."""
    from oracle import oracle as O
    cores = host_cores()
    t0 = time.perf_counter()
    tree = O.NaboKDTree(pts)
    build_s = time.perf_counter() - t0
    tree.knn(q[:4096], K_NN, threads=cores)
    reps, t0 = 0, time.perf_counter()
    while True:
        tree.knn(q, K_NN, threads=cores)
        reps += 1
        el = time.perf_counter() - t0
        if el > budget_s or reps >= 50:
            break
    return {"value": reps * len(q) / el, "unit": "queries/s", "cores": cores, "kind": "port", "cpu_model": host_cpu_model(),
            "sample": "%d x %d queries, k=%d, on the prebuilt kd-tree of %d nodes (build %.2f s not counted); "
                      "libnabo restatement oracle/src/nabo_kdtree.cpp" % (reps, len(q), K_NN, len(pts), build_s)}


def cpu_edge_baseline(boxes, p1, p2, budget_s):
    """reference-literal loop (every leaf box, MPR, early break) with a conservative bounding
    sphere cull, on a bounded edge sample; OpenMP over edges."""
    from oracle import oracle as O
    cores = host_cores()
    n = 64
    t0 = time.perf_counter()
    O.collide_edges(boxes, p1[:n], p2[:n], EDGE_R, 0, mode=1, want_band=False, threads=cores)
    el = time.perf_counter() - t0
    n2 = int(max(n, min(4096, n * budget_s / max(el, 1e-3))))
    t0 = time.perf_counter()
    O.collide_edges(boxes, p1[:n2], p2[:n2], EDGE_R, 0, mode=1, want_band=False, threads=cores)
    el = time.perf_counter() - t0
    out = {"value": n2 / el, "unit": "edge checks/s", "cores": cores, "kind": "port",
           "sample": "%d edges vs all %d leaf boxes, linear loop + MPR as rrt_star_milestones.cpp:385-471 "
                     "(bounding-sphere cull before MPR), oracle/src/collide.cpp" % (n2, len(boxes))}
    # (c2) of SURVEY 8(d): the "library-best" CPU arrangement — a broadphase over the leaf boxes (bucket
    # grid, built once per map and not timed) in front of the same MPR, all host threads
    O.collide_edges(boxes, p1[:64], p2[:64], EDGE_R, 0, mode=2, want_band=False, threads=cores)      # builds the buckets
    n3 = 20000
    t0 = time.perf_counter()
    O.collide_edges(boxes, p1[:n3], p2[:n3], EDGE_R, 0, mode=2, want_band=False, threads=cores)
    el3 = time.perf_counter() - t0
    n4 = int(max(n3, min(len(p1), n3 * 0.5 * budget_s / max(el3, 1e-3))))
    if n4 > n3:
        t0 = time.perf_counter()
        O.collide_edges(boxes, p1[:n4], p2[:n4], EDGE_R, 0, mode=2, want_band=False, threads=cores)
        el3, n3 = time.perf_counter() - t0, n4
    out["library_best"] = {"value": n3 / el3, "unit": "edge checks/s", "cores": cores, "kind": "port",
                           "sample": "%d edges, bucket-grid broadphase over the leaf boxes (built once, untimed) + bounding "
                                     "culls + MPR (oracle mode 2: same booleans as the loop over every leaf)" % n3}
    return out


def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return
    from bottleneck_guided_rrt_b200 import synth
    from oracle import oracle as O
    cores = host_cores()
    pts = synth.points(2, N_NODES, 25.6)
    q = synth.points(3, N_QUERIES, 25.6)
    tree = O.NaboKDTree(pts)
    for _ in range(max(1, args.warmup)):
        tree.knn(q, K_NN, threads=cores)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        tree.knn(q, K_NN, threads=cores)
    el = time.perf_counter() - t0
    value = args.steps * N_QUERIES / el
    cb = {"value": value, "unit": "queries/s", "cores": cores, "kind": "port", "cpu_model": host_cpu_model(),
          "sample": "%d steps x %d queries, k=%d, prebuilt libnabo-restatement kd-tree over %d nodes" % (args.steps, N_QUERIES, K_NN, N_NODES)}
    out = {"impl": "reference", "metric": METRIC, "value": value, "unit": "queries/s", "n_gpus": args.gpus,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": el / args.steps * 1e3, "higher_is_better": True,
           "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": {"workload": "configs[1]: batched kNN, 1,000,000-node 3-D tree, 65,536 queries, k=16, eps=0 "
                                  "(CPU: libnabo 1.0.7 algorithm restated in oracle/, OpenMP over queries, %d threads)" % cores},
           "cpu_baseline": cb,
           "e2e": {"value": value, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    if not args.skip_edges:
        occ = synth.make_map(MAP_N, 4, n_walls=8, wall_thick=4, door_cells=5)
        boxes = synth.grid_to_boxes(occ, 0.05, (0, 0, 0), 3)
        p1, p2 = synth.edges(4, N_EDGES, occ)
        out["edge_checks"] = cpu_edge_baseline(boxes, p1, p2, budget_s=15.0)
    print(json.dumps(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--method", default="grid", choices=["grid", "brute", "auto"])
    ap.add_argument("--skip-edges", action="store_true")
    ap.add_argument("--skip-cpu", action="store_true")
    ap.add_argument("--skip-sharded", action="store_true")
    ap.add_argument("--skip-planner", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
