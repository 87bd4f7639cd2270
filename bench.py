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


def ncu_entry(kernel):
    """the entry of `kernel` in profiles/traffic.json (written by scripts/ncu_traffic.py from the committed
    `ncu --set full` capture of this workload): per-launch DRAM and L2 traffic, or {}."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(p):
        return {}
    with open(p) as f:
        return json.load(f).get(kernel) or {}


def ncu_traffic(kernel):
    """per-launch dram__bytes_read.sum + dram__bytes_write.sum of `kernel`, or None"""
    e = ncu_entry(kernel)
    return float(e["dram_bytes"]) if e else None


def l2_roofline(kernel, kernel_ms, peak_gbs, peak_what):
    """L2 roofline of an L2-resident kernel: bytes the L2 delivered to the SMs in one launch (ncu capture of
    this workload, profiles/traffic.json) / the launch duration measured live / the L2 read bandwidth measured
    live on this box with the access shape the kernel uses (bgr_microbench_l2)."""
    e = ncu_entry(kernel)
    b = e.get("l2_to_sm_read_bytes")
    out = {"peak": peak_gbs, "unit": "GB/s", "peak_src": "bgr_microbench_l2 on this box, this run: " + peak_what,
           "l2_to_sm_read_bytes_per_launch": b, "l2_sector_bytes_per_launch": e.get("l2_bytes"),
           "bytes_src": "profiles/traffic.json (%s)" % e.get("source") if e else None,
           "l2_hit_rate_pct": e.get("l2_hit_rate_pct"), "l1_hit_rate_pct": e.get("l1_hit_rate_pct")}
    if b and kernel_ms > 0:
        out["achieved"] = b / (kernel_ms * 1e-3) / 1e9
        out["frac"] = out["achieved"] / peak_gbs if peak_gbs else None
    else:
        out["achieved"] = out["frac"] = None
    return out


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
def timed_steps(torch, stream, step_fn, steps, warmup, flush, barrier, prewarm_s=0.25, prewarm_iters=None, run_ahead=True):
    """W warm-up steps, then exactly K timed steps; each step has its own CUDA event pair on
    `stream` and L2 is flushed (untimed) between steps.  Returns total ms of the K steps.
    A step that contains a collective must pass prewarm_iters: a wall-clock-bounded ramp would
    run a different number of collectives on each rank (deadlock).
    run_ahead: the K (flush, event, step, event) groups are all enqueued behind a ~1.5 ms spin kernel, so
    the host is done launching before the first timed kernel runs and no host launch gap (several ranks
    share the box's cores under torchrun) can sit inside an event pair; steps that synchronise inside
    (host-pointer calls) gain nothing from it."""
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
    if run_ahead:
        with torch.cuda.stream(stream):
            torch.cuda._sleep(3_000_000)
    for a, b in evs:
        flush()
        a.record(stream)
        step_fn()
        b.record(stream)
    torch.cuda.synchronize()
    barrier()
    return sum(a.elapsed_time(b) for a, b in evs)


def copy_only_ms(torch, stream, pairs, reps, barrier):
    """average ms of the given (dst, src) tensor copies alone, issued together on `stream` (events on it);
    all ranks start behind the same barrier, so at N > 1 the figure is the copy rate with every GPU's
    copies in flight at once"""
    for d, s_ in pairs:
        d.copy_(s_, non_blocking=True)
    stream.synchronize()
    barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for _ in range(reps):
        for d, s_ in pairs:
            d.copy_(s_, non_blocking=True)
    b.record(stream)
    stream.synchronize()
    return a.elapsed_time(b) / reps


def run_ours(args):
    import torch
    import bottleneck_guided_rrt_b200 as B
    from bottleneck_guided_rrt_b200 import synth

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

    def per_rank(x):
        """the value of every rank, in rank order"""
        if dist is None:
            return [x]
        t = torch.zeros(world, dtype=torch.float64, device=dev)
        t[rank] = x
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return [float(v) for v in t.tolist()]

    def flush():
        with torch.cuda.stream(stream):
            flush_buf.fill_(rank & 0xFF)

    out = {"metric": METRIC, "unit": "queries/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
           "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "parity_note": "oracle parity is UNPINNED against the real libnabo / FCL / octomap (absent offline); see DESIGN.md section 0"}
    sampler = ClockSampler(local)
    ctx = dict(torch=torch, B=B, synth=synth, L=L, pk=pk, dev=dev, stream=stream, flush=flush, barrier=barrier,
               max_over_ranks=max_over_ranks, per_rank=per_rank, rank=rank, world=world, dist=dist, args=args)

    with torch.cuda.stream(stream):
        # on-box L2 read bandwidth for the access shapes of the three L2-resident kernels (roofline.l2 peaks)
        l2_peak = {"rows_16MiB": B.microbench_l2(16 << 20, 2), "sectors_32MiB": B.microbench_l2(32 << 20, 1),
                   "stream_16MiB": B.microbench_l2(16 << 20, 0)}
        ctx["l2_peak"] = l2_peak
        out["l2_microbench_gbs"] = dict(l2_peak, note="bgr_microbench_l2 (scripts/micro/l2_read_bench.cu): L1 bypassed; rows = random "
                                        "512 B rows with LDG.128, sectors = random 32 B sectors with 8 B loads (32 B counted per load), "
                                        "stream = consecutive rows; resident buffers")
        sampler.start()
        pts, q_host = knn_section(ctx, out)
        out["clocks"] = sampler.stop()
        boxes = p1 = p2 = None
        if not args.skip_edges:
            boxes, p1, p2 = edge_section(ctx, out)
        if not args.skip_planner:
            out["planner"] = planner_bench(ctx)
        if not args.skip_sharded:
            out["sharded_knn"] = sharded_section(ctx)

    # -------------------------------------------------------------------- CPU baseline (N=1, rank 0)
    if rank == 0 and world == 1 and not args.skip_cpu:
        out["cpu_baseline"] = cpu_knn_baseline(pts, q_host, budget_s=8.0)
        if "edge_checks" in out:
            out["edge_checks"]["cpu_baseline"] = cpu_edge_baseline(boxes, p1, p2, budget_s=10.0)
        if "planner" in out:
            out["planner"]["cpu_baseline"] = cpu_planner_baseline(budget_s=12.0)
    if rank == 0:
        print(json.dumps(out))
    if dist is not None:
        dist.destroy_process_group()


def knn_section(c, out):
    """configs[1]: 1 M-node tree, 65,536 queries, k = 16 (the headline)"""
    torch, B, synth, L, pk, dev, stream = c["torch"], c["B"], c["synth"], c["L"], c["pk"], c["dev"], c["stream"]
    args, rank, world, max_over_ranks, barrier, flush = c["args"], c["rank"], c["world"], c["max_over_ranks"], c["barrier"], c["flush"]
    method = B.METHOD_GRID if args.method == "grid" else B.METHOD_BRUTE if args.method == "brute" else B.METHOD_AUTO
    pts = synth.points(2, N_NODES, 25.6)
    q_host = synth.points(3 + rank, N_QUERIES, 25.6)
    d_pts = torch.from_numpy(pts).to(dev)
    nns = B.NNSearchF.createKDTreeLinearHeap(d_pts, creationOptionFlags=B.TOUCH_STATISTICS, stream=stream.cuda_stream)
    nns.set_method(method)
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
    nns_t.set_method(method)
    nns_t.build_index()

    def knn_step():
        nns_t.knn(d_q, K_NN, indices=d_idx, dists2=d_d2)

    knn_ms_mine = timed_steps(torch, stream, knn_step, args.steps, args.warmup, flush, barrier)
    knn_ms = max_over_ranks(knn_ms_mine)
    knn_ms_ranks = [v / args.steps for v in c["per_rank"](knn_ms_mine)]
    # the exhaustive scan of the same workload (knn_brute_kernel: the FP32-pipe-bound kernel of north_star (3)):
    # 3 timed steps of ~16 ms, results compared with the grid search's
    brute = None
    if args.method != "brute" and not args.skip_brute:
        nns_t.set_method(B.METHOD_BRUTE)
        b_idx, b_d2 = torch.empty_like(d_idx), torch.empty_like(d_d2)

        def brute_step():
            nns_t.knn(d_q, K_NN, indices=b_idx, dists2=b_d2)

        brute_ms = max_over_ranks(timed_steps(torch, stream, brute_step, 3, 3, flush, barrier, 0.05)) / 3
        nns_t.set_method(method)
        knn_step()
        stream.synchronize()
        fp32_peak_b = 148 * FP32_NOFMA_LANES * pk["sm_max_mhz"] * 1e6 / 1e12
        ach_b = 8.0 * N_NODES * N_QUERIES / (brute_ms * 1e-3) / 1e12
        brute = {"value": world * N_QUERIES / (brute_ms * 1e-3), "unit": "queries/s", "ms_per_step": brute_ms, "steps": 3,
                 "identical_to_grid_search": bool(torch.equal(b_idx, d_idx) and torch.equal(b_d2.view(torch.int32), d_d2.view(torch.int32))),
                 "roofline": {"bound": "fp32", "kernel": "knn_brute_kernel", "achieved": ach_b, "peak": fp32_peak_b, "unit": "TFLOP/s",
                              "frac": ach_b / fp32_peak_b, "flops_per_pair": 8, "pairs_per_step": float(N_NODES) * N_QUERIES,
                              "note": "8 non-FMA FP32 operations per (query, node) pair is the algorithmic work of the exact distance "
                                      "(SURVEY 8(d)); the kernel spends most pairs on a 3-FFMA2 conservative filter (two pairs per issue "
                                      "slot) and the exact arithmetic on the survivors, so the fraction is against the non-FMA peak of "
                                      "148 SM x 128 lanes x sm_max_mhz"}}
    l1 = L.bgr_launch_count()                       # launches per step, counted on one isolated step
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

    def knn_e2e_idx_step():
        nns_t.knn(qh_np, K_NN, indices=ih_np, want_dists2=False)

    e2e_ms = max_over_ranks(timed_steps(torch, stream, knn_e2e_step, args.steps, args.warmup, flush, barrier, 0.1, run_ahead=False))
    assert np.array_equal(ih_np, d_idx.cpu().numpy()), "e2e and device-resident results differ"
    e2e_idx_ms = max_over_ranks(timed_steps(torch, stream, knn_e2e_idx_step, args.steps, args.warmup, flush, barrier, 0.1, run_ahead=False))
    # the three legs of the e2e step alone (every rank at once): what the host link allows at this N
    h2d_ms = max_over_ranks(copy_only_ms(torch, stream, [(d_q, qh)], 20, barrier))
    d2h_ms = max_over_ranks(copy_only_ms(torch, stream, [(ih, d_idx), (dh, d_d2)], 20, barrier))

    ms_per_step = knn_ms / args.steps
    value = world * N_QUERIES / (ms_per_step * 1e-3)
    pairs_per_step = float(pairs)
    fp32_peak = 148 * FP32_NOFMA_LANES * pk["sm_max_mhz"] * 1e6 / 1e12     # TFLOP/s, non-FMA
    achieved = 8.0 * pairs_per_step / (ms_per_step * 1e-3) / 1e12
    alg_bytes = 12.0 * N_NODES + 12.0 * N_QUERIES + 8.0 * K_NN * N_QUERIES
    kname = "grid_knn_group_kernel" if args.method != "brute" else "knn_brute_kernel"
    d2h_bytes = int(8 * K_NN * N_QUERIES)
    out.update({
        "value": value, "ms_per_step": ms_per_step, "ms_per_step_per_rank": knn_ms_ranks, "brute_scan": brute,
        "config": {"workload": "configs[1]: batched kNN, 1,000,000-node 3-D tree (U[0,25.6)^3, MT19937 seed 2), "
                               "65,536 queries/GPU (seed 3+rank), k=16, eps=0, exact, method=%s" % args.method,
                   "nodes": N_NODES, "queries_per_gpu": N_QUERIES, "k": K_NN,
                   "l2": "flushed between steps (256 MiB fill on the same stream, untimed)",
                   "prewarm": "0.25 s untimed clock ramp before the W warm-up steps; the K timed steps are enqueued "
                              "behind a 1.5 ms spin kernel (no host launch gap inside an event pair)",
                   "index_build_ms": build_ms, "parallelism": "replicas x%d" % world,
                   "per_rank_note": "every rank searches its own query batch (seed 3 + rank); the kernel time depends on the batch by "
                                    "+-4 % (profiles/r2_knn_query_order_and_bounds.txt: seeds 3..6 on one GPU), so the max over ranks "
                                    "at N > 1 sits a few % above the N = 1 figure with no collective in the path"},
        "e2e": {"value": world * N_QUERIES / (e2e_ms / args.steps * 1e-3), "unit": "queries/s",
                "ms_per_step": e2e_ms / args.steps,
                "h2d_bytes_per_step": int(12 * N_QUERIES), "d2h_bytes_per_step": d2h_bytes,
                "legs_alone_ms": {"h2d": h2d_ms, "kernel": ms_per_step, "d2h": d2h_ms,
                                  "d2h_gb_per_s_per_gpu": d2h_bytes / (d2h_ms * 1e-3) / 1e9,
                                  "note": "each leg timed alone with all %d GPUs' copies in flight at once (max over ranks); the "
                                          "chunked call overlaps kernel and D2H, so its floor is about h2d + d2h" % world},
                "indices_only": {"value": world * N_QUERIES / (e2e_idx_ms / args.steps * 1e-3), "ms_per_step": e2e_idx_ms / args.steps,
                                 "d2h_bytes_per_step": d2h_bytes // 2,
                                 "note": "bgr_nns_knn with d2 = NULL: the reference consumes only the indices (MS:1312-1319)"}},
        "gpu_launches": int(per_step_launches * args.steps),
        # top-level figures: HBM roofline of the kernel (the contract's "hbm" bound: algorithmic bytes of
        # SURVEY 8(d) / measured launch time / measured HBM peak).  The node index is L2 resident and the
        # kernel is bound by L2 row reads and the issue pipes, not HBM: those rooflines are in "l2" and "fp32".
        "roofline": {"bound": "hbm", "kernel": kname,
                     "achieved": alg_bytes / (ms_per_step * 1e-3) / 1e9, "peak": pk["hbm_gbs"], "unit": "GB/s",
                     "frac": alg_bytes / (ms_per_step * 1e-3) / 1e9 / pk["hbm_gbs"], "peak_src": pk["src"],
                     "algorithmic_bytes": alg_bytes, "algorithmic_bytes_def": "12 B/node + 12 B/query + 8 B x k per query (SURVEY 8(d))",
                     "traffic": ncu_traffic(kname),
                     "l2": l2_roofline(kname, ms_per_step, c["l2_peak"]["rows_16MiB"], "random 512 B rows, LDG.128, 16 MiB resident buffer"),
                     "fp32": {"achieved": achieved, "peak": fp32_peak, "unit": "TFLOP/s", "frac": achieved / fp32_peak,
                              "peak_src": "148 SM x 128 FP32 lanes x %s sm_max_mhz (non-FMA: the bit-exact contract "
                                          "forbids FMA)" % pk["src"],
                              "pairs_evaluated_per_step": pairs_per_step, "flops_per_pair": 8},
                     "note": "node index L2 resident (DRAM traffic per L2-flushed launch = the index read once); the "
                             "kernel is bound by L2 row reads / dependent loads and instruction issue (ncu: profiles/), see DESIGN.md 3.2"},
    })
    return pts, q_host


def edge_section(c, out):
    """configs[2]: 1 M cylinder edges vs the 512^3 map; sphere state checks on the same map"""
    torch, B, synth, L, pk, dev, stream = c["torch"], c["B"], c["synth"], c["L"], c["pk"], c["dev"], c["stream"]
    args, rank, world, max_over_ranks, barrier, flush = c["args"], c["rank"], c["world"], c["max_over_ranks"], c["barrier"], c["flush"]
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

    edge_e2e_ms = max_over_ranks(timed_steps(torch, stream, edge_e2e_step, args.steps, args.warmup, flush, barrier, 0.1, run_ahead=False))
    assert np.array_equal(hn, d_hit.cpu().numpy()) and np.array_equal(bn, d_band.cpu().numpy()), \
        "edge checks: host-buffer and device-resident results differ"
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
                     "algorithmic_bytes": ebytes, "algorithmic_bytes_def": "28 B/edge + the 16 MiB bitmap once (SURVEY 8(d))",
                     "peak_src": pk["src"], "traffic": ncu_traffic("edges_sweep_kernel"),
                     "l2": l2_roofline("edges_sweep_kernel", sweep_ms, c["l2_peak"]["sectors_32MiB"],
                                       "random 32 B sectors (one 8 B brick word per load), 32 MiB resident buffer (both maps)"),
                     "kernel_ms": kms, "kernel_share_of_step": sweep_ms / max(sum(kms.values()), 1e-9),
                     "note": "both 16 MiB maps are L2 resident: the sweep is bound by instruction issue and the L1/L2 "
                             "probe rate, not HBM (ncu: profiles/); the HBM figure is the contract of SURVEY 8(d)",
                     "work_items_per_s": stats["work_items"] / (sweep_ms * 1e-3)},
    }

    # sphere state checks (isStateValid MS:310-351) on the same map: the edges' first endpoints
    d_shit = torch.empty(N_EDGES, dtype=torch.uint8, device=dev)
    d_sband = torch.empty(N_EDGES, dtype=torch.uint8, device=dev)

    def sphere_step():
        omap.collide_spheres(d_p1, 0.1, hit=d_shit, band=d_sband)

    sphere_ms = max_over_ranks(timed_steps(torch, stream, sphere_step, args.steps, args.warmup, flush, barrier, 0.1)) / args.steps
    shh, sbh = torch.empty(N_EDGES, dtype=torch.uint8).pin_memory(), torch.empty(N_EDGES, dtype=torch.uint8).pin_memory()
    shn, sbn = shh.numpy(), sbh.numpy()

    def sphere_e2e_step():
        omap.collide_spheres(p1n, 0.1, hit=shn, band=sbn)

    sphere_e2e_ms = max_over_ranks(timed_steps(torch, stream, sphere_e2e_step, args.steps, args.warmup, flush, barrier, 0.1, run_ahead=False)) / args.steps
    assert np.array_equal(shn, d_shit.cpu().numpy()) and np.array_equal(sbn, d_sband.cpu().numpy()), \
        "sphere checks: host-buffer and device-resident results differ"
    sbytes = 14.0 * N_EDGES + map_bytes
    out["state_checks"] = {
        "value": world * N_EDGES / (sphere_ms * 1e-3), "unit": "sphere state checks/s", "ms_per_step": sphere_ms,
        "config": {"workload": "1,000,000 spheres/GPU (r=0.1) vs the configs[2] map", "hit_fraction": float(d_shit.float().mean().item()),
                   "in_band": int(d_sband.sum().item())},
        "e2e": {"value": world * N_EDGES / (sphere_e2e_ms * 1e-3), "unit": "sphere state checks/s", "ms_per_step": sphere_e2e_ms,
                "h2d_bytes_per_step": int(12 * N_EDGES), "d2h_bytes_per_step": int(2 * N_EDGES)},
        "roofline": {"bound": "hbm", "kernel": "collide_spheres_lane_kernel", "achieved": sbytes / (sphere_ms * 1e-3) / 1e9,
                     "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": sbytes / (sphere_ms * 1e-3) / 1e9 / pk["hbm_gbs"],
                     "algorithmic_bytes": sbytes, "algorithmic_bytes_def": "14 B/sphere + the 16 MiB bitmap once",
                     "peak_src": pk["src"], "traffic": ncu_traffic("collide_spheres_lane_kernel"),
                     "l2": l2_roofline("collide_spheres_lane_kernel", sphere_ms, c["l2_peak"]["sectors_32MiB"],
                                       "random 32 B sectors (one 8 B brick word per load), 32 MiB resident buffer (both maps)")}}
    return boxes, p1, p2


def sharded_section(c):
    """configs[4](ii): index-range-sharded kNN, 6.25 M nodes per GPU (50 M at N = 8), 65,536 replicated queries,
    k = 16: local search + ONE exchange (NCCL all-gather + merge kernel | NVLink peer-memory merge kernel).
    N = 1 runs the one-shard search without an exchange (the base of the weak-scaling curve).  Every rank count
    checks a 128-query subsample against the oracle's brute force over the same nodes."""
    torch, B, synth, dev, stream = c["torch"], c["B"], c["synth"], c["dev"], c["stream"]
    args, rank, world, max_over_ranks, barrier, flush, dist = c["args"], c["rank"], c["world"], c["max_over_ranks"], c["barrier"], c["flush"], c["dist"]
    from bottleneck_guided_rrt_b200.dist import ShardedNNSearch, shard_range, merge_slabs_host
    lo, hi = shard_range(SHARD_NODES * world, rank, world)
    shard_host = synth.points(5 + 1000 * rank, hi - lo, 102.4)
    shard = torch.from_numpy(shard_host).to(dev)
    sq_host = synth.points(6, N_QUERIES, 102.4)
    sq = torch.from_numpy(sq_host).to(dev)
    sh = {"nodes_total": SHARD_NODES * world, "nodes_per_gpu": SHARD_NODES, "queries": N_QUERIES, "k": K_NN, "unit": "queries/s"}
    results = {}
    for exch in (("nccl", "peer") if world > 1 else ("none",)):
        try:
            snn = ShardedNNSearch(shard, lo, stream=stream.cuda_stream, exchange=exch, max_entries=N_QUERIES * K_NN)
            snn.nns.build_index()

            def shard_step():
                snn.knn(sq, K_NN)

            sh_ms = max_over_ranks(timed_steps(torch, stream, shard_step, args.steps, args.warmup, flush, barrier,
                                               prewarm_iters=20, run_ahead=False))
            # local search and exchange + merge apart: events inside the call, same flush discipline
            loc_ms = exc_ms = 0.0
            for _ in range(args.steps):
                ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
                flush()
                snn.knn(sq, K_NN, events=ev)
                stream.synchronize()
                loc_ms += ev[0].elapsed_time(ev[1]) / args.steps
                exc_ms += ev[1].elapsed_time(ev[2]) / args.steps
            ri, rd = snn.knn(sq, K_NN)
            stream.synchronize()
            results[exch] = (ri.clone(), rd.clone())
            sh[exch] = {"value": N_QUERIES / (sh_ms / args.steps * 1e-3), "ms_per_step": sh_ms / args.steps,
                        "local_ms": max_over_ranks(loc_ms), "exchange_ms": max_over_ranks(exc_ms)}
            if exch == "peer":
                sh[exch]["spin_timeouts"] = snn.peer.status(stream.cuda_stream)
            del snn
        except Exception as ex:      # reported, never hidden: the other exchange still stands
            sh[exch] = {"error": repr(ex)[:300]}
    if len(results) == 2:
        sh["exchanges_identical"] = bool(torch.equal(results["nccl"][0], results["peer"][0]) and
                                         torch.equal(results["nccl"][1], results["peer"][1]))
    # parity against the oracle on the real nodes: every rank brute-forces the subsample over ITS shard (global
    # indices), the per-shard lists are gathered and merged by the numpy statement of the merge rule
    if not args.skip_cpu and results:
        from oracle import oracle as O
        sub = np.arange(0, N_QUERIES, N_QUERIES // 128)[:128]
        oi, od = O.knn_brute(shard_host, sq_host[sub], K_NN, threads=max(1, host_cores() // world))
        oi = np.where(oi >= 0, oi + lo, oi).astype(np.int32)
        if dist is not None:
            gi = [torch.empty((128, K_NN), dtype=torch.int32, device=dev) for _ in range(world)]
            gd = [torch.empty((128, K_NN), dtype=torch.float32, device=dev) for _ in range(world)]
            dist.all_gather(gi, torch.from_numpy(oi).to(dev))
            dist.all_gather(gd, torch.from_numpy(od).to(dev))
            mi, md = merge_slabs_host(np.stack([t.cpu().numpy() for t in gi]), np.stack([t.cpu().numpy() for t in gd]))
        else:
            mi, md = oi, od
        par = {"checked": int(len(sub)), "oracle": "oracle.knn_brute over each rank's 6.25 M nodes + numpy merge by (d2, index)"}
        for exch, (ri, rd) in results.items():
            gi_, gd_ = ri.cpu().numpy()[sub], rd.cpu().numpy()[sub]
            bad = int(((gi_ != mi).any(axis=1) | (gd_.view(np.uint32) != md.view(np.uint32)).any(axis=1)).sum())
            par["mismatches_" + exch] = bad
        par["mismatches"] = int(max_over_ranks(float(sum(v for k_, v in par.items() if k_.startswith("mismatches_")))))
        sh["parity"] = par
    best = max((v for v in (sh.get("peer"), sh.get("nccl"), sh.get("none")) if v and "value" in v), key=lambda v: v["value"], default=None)
    if best:
        sh["value"], sh["ms_per_step"] = best["value"], best["ms_per_step"]
        sh["local_ms"], sh["exchange_ms"] = best["local_ms"], best["exchange_ms"]
        # the local search is the grid kNN over a 6.25 M-node index (100 MB: L2 resident at the edge); its HBM-contract
        # roofline on SURVEY 8(d) bytes per GPU
        alg = 12.0 * SHARD_NODES + 12.0 * N_QUERIES + 8.0 * K_NN * N_QUERIES
        sh["roofline"] = {"bound": "hbm", "kernel": "grid_knn_group_kernel (local search)", "achieved": alg / (best["local_ms"] * 1e-3) / 1e9,
                          "peak": c["pk"]["hbm_gbs"], "unit": "GB/s", "frac": alg / (best["local_ms"] * 1e-3) / 1e9 / c["pk"]["hbm_gbs"],
                          "algorithmic_bytes": alg, "exchange_bytes_per_rank": int(8 * K_NN * N_QUERIES * (world - 1) // max(world, 1))}
    sh["exchange"] = {"nccl": "ONE all_gather_into_tensor of the packed (idx,d2) slabs (%d B per rank) + bgr_knn_merge_strided_dev" % (8 * K_NN * N_QUERIES),
                      "peer": "bgr_peer_merge_dev: one kernel, peer loads of this rank's query slice from all slabs over NVLink, "
                              "merge by (d2,index), peer stores of the merged lists (%d B in + out per rank)" % (8 * K_NN * N_QUERIES * (world - 1) // world)}
    return sh


def door_milestones(occ, n_map, n_walls=3):
    """the 3 milestone points = the door centres of the 3 walls (what the CNN is trained to predict)"""
    ms = []
    for w in range(n_walls):
        x0 = (w + 1) * n_map // (n_walls + 1)
        zz, yy = np.nonzero(occ[:, :, x0] == 0)
        ms.append([(x0 + 0.5) * 0.05, (yy.mean() + 0.5) * 0.05, (zz.mean() + 0.5) * 0.05])
    return np.asarray(ms, dtype=np.float32)


QUERY_OCCUPANCY = 0.065   # configs[4](i) maps: 3 walls (4.7 % of the cells) with one 0.4 m door each + 1.8 % clutter
QUERY_DOOR_CELLS = 8      # (see planner_bench for why these differ from the configs[3] map)
QUERIES_PER_GPU = 512
QUERY_MAX_NODES = 60_000


def planner_bench(c):
    """Tree growth of the batched planner (bgr_planner_*): configs[0] = RRT* with 10 % goal bias on a
    64^3 map to 5,000 nodes, k=10 (batch 1 = the reference's sequential loop, and batch 256);
    configs[3] = bottleneck-guided RRT* (3 milestones) on a 256^3 narrow-passage map to 200,000 nodes,
    batch 4096, K=16, edge checks + rewiring; configs[4](i) = 512 independent start/goal queries per GPU on a
    256^3 map (one distinct map per GPU), each run to the goal."""
    torch, B, synth, stream = c["torch"], c["B"], c["synth"], c["stream"]
    rank, world, max_over_ranks, barrier = c["rank"], c["world"], c["max_over_ranks"], c["barrier"]
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
                "in_band": st["in_band"], "map": "%d^3 @0.05 m, %d leaf boxes" % (n_map, len(boxes)),
                "steps_in_persistent_kernel": pl.persist_stats()["steps"], "us_per_step": 1e6 * el / max(st["batches"], 1)}

    g64 = (3.0, 3.0, 3.0)
    res["config0_rrtstar_64_batch1"] = grow(64, 1, 3, 6, B.PLAN_MODE_RS, 10, 1, B.PLAN_QUIRKS, 5000, None, g64)
    res["config0_rrtstar_64_batch256"] = grow(64, 1, 3, 6, B.PLAN_MODE_RS, 10, 256, 0, 5000, None, g64)
    occ3 = synth.make_map(256, 5, n_walls=3, wall_thick=4, door_cells=5)
    ms = door_milestones(occ3, 256)
    # wall clock of ~860 steps: sensitive to whatever else the host is doing, so the growth is
    # repeated (same seed, same tree) and every run is listed; the headline is the fastest
    runs3 = [grow(256, 5, 3, 5, B.PLAN_MODE_MS, 16, 4096, B.PLAN_EDGE_CHECK | B.PLAN_REWIRE, 200_000, ms, (12.6, 12.6, 12.6))
             for _ in range(3)]
    best3 = min(runs3, key=lambda r: r["seconds"])
    best3["seconds_all_runs"] = [r["seconds"] for r in runs3]
    res["config3_bottleneck_256_batch4096"] = best3

    # configs[4](i): QUERIES_PER_GPU independent start/goal planning queries per GPU, each run to the goal (or to
    # QUERY_MAX_NODES nodes), on one 256^3 map per GPU (map seed 5 + rank: 8 distinct maps at N = 8).
    # Measured on this planner (scripts/try_queries2.py, 32 random start/goal pairs, profiles/r2_queries_*.txt):
    #   * the configs[3] map (0.25 m doors, 10 % clutter): 9 of 32 reach the goal.  Two causes: with 10 % clutter the
    #     free space of a 0.1 m sphere is ~30 % of the volume, at the percolation threshold, so many pairs are not
    #     connected at all; and a 0.25 m door leaves the 0.1 m sphere a +-0.025 m corridor, which a point-validity
    #     RRT with 0.15 m steps (the reference's extension rule) enters only from a ~6 degree cone in front of it;
    #   * 0.4 m doors, 1.8 % clutter, no BGR_PLAN_DEDUP: 6 of 32 — every batch appends hundreds of copies of the goal /
    #     milestone candidates (the reference's sequential loop piles them up too) and the node budget is spent on piles;
    #   * the same with BGR_PLAN_DEDUP: 29-30 of 32.
    # So the queries run on 0.4 m doors / 1.8 % clutter with BGR_PLAN_DEDUP on.
    occq = synth.make_map(256, 5 + rank, n_walls=3, wall_thick=4, door_cells=QUERY_DOOR_CELLS, occupancy=QUERY_OCCUPANCY)
    msq = door_milestones(occq, 256)
    boxesq = synth.grid_to_boxes(occq, 0.05, (0, 0, 0), 3)
    nconc = 16
    qstreams = [torch.cuda.Stream() for _ in range(nconc)]
    qmaps = []
    for qs in qstreams:
        qm = B.OccupancyMap(0.05, (0, 0, 0), (256,) * 3, stream=qs.cuda_stream)
        qm.set_boxes(boxesq)
        qmaps.append(qm)
    cand = synth.points(77 + rank, 4 * QUERIES_PER_GPU + 2000, 12.8)
    chit, _ = qmaps[0].collide_spheres(cand, 0.1)
    free = cand[np.asarray(chit) == 0]
    assert len(free) >= 2 * QUERIES_PER_GPU
    table = synth.points(9 + rank, 200_000, 12.8)

    # one planner per worker (map handle + stream each), re-armed for every query (bgr_planner_reset): a worker takes
    # the next (start, goal, seed) as soon as its query ends (bgr_planner_run_queue) — no query waits for the slowest
    # of a group, no buffer is allocated or freed while queries run
    P0 = B.PlannerParams(tuple(free[0]), tuple(free[1]), mode=B.PLAN_MODE_MS, k=16, batch=1024,
                         flags=B.PLAN_EDGE_CHECK | B.PLAN_REWIRE | B.PLAN_DEDUP, seed=100)
    workers = [B.Planner(P0, qmaps[j], table, msq, stream=qstreams[j].cuda_stream) for j in range(nconc)]
    starts, goals = free[0:2 * QUERIES_PER_GPU:2], free[1:2 * QUERIES_PER_GPU:2]
    seeds = np.arange(QUERIES_PER_GPU, dtype=np.uint32) + np.uint32(100 + rank * QUERIES_PER_GPU)

    def run_queries(q0, q1, nworkers):
        r_, n_ = B.Planner.run_queue(workers[:nworkers], starts[q0:q1], goals[q0:q1], seeds[q0:q1], QUERY_MAX_NODES)
        torch.cuda.synchronize()
        return [bool(v) for v in r_], [int(v) for v in n_]

    run_queries(0, 2 * nconc, nconc)                        # warm-up: allocations, first launches
    barrier()
    t0 = time.perf_counter()
    reached, nodes = run_queries(0, QUERIES_PER_GPU, nconc)
    el = max_over_ranks(time.perf_counter() - t0)
    barrier()
    t0 = time.perf_counter()
    run_queries(0, 2 * nconc, 1)
    el_seq = max_over_ranks(time.perf_counter() - t0)
    ok_nodes = [n for n, r in zip(nodes, reached) if r]
    n_ok = int(max_over_ranks(float(sum(reached))))        # max over ranks would overstate: report rank 0 and say so
    res["config4_independent_queries_256"] = {
        "queries": world * QUERIES_PER_GPU, "queries_per_gpu": QUERIES_PER_GPU, "seconds": el,
        "queries_per_s": world * QUERIES_PER_GPU / el,
        "reached_goal_rank0": int(sum(reached)), "success_rate_rank0": float(np.mean(reached)), "reached_goal_max_over_ranks": n_ok,
        "successful_queries_per_s": world * float(sum(reached)) / el,
        "median_nodes_to_goal_rank0": int(np.median(ok_nodes)) if ok_nodes else None,
        "median_nodes_all_rank0": int(np.median(nodes)),
        "one_after_the_other": {"queries": 2 * nconc, "seconds": el_seq, "queries_per_s": world * 2 * nconc / el_seq},
        "batch": 1024, "k": 16, "max_nodes": QUERY_MAX_NODES, "concurrent_queries": nconc,
        "map": "256^3 @0.05 m, 3 walls with one %.2f m door each, %.1f%% occupied, %d leaf boxes, map seed 5 + rank"
               % (0.05 * QUERY_DOOR_CELLS, 100 * occq.mean(), len(boxesq)),
        "workload": "configs[4](i): %d independent (start, goal) pairs per GPU (random free points, sphere r = 0.1), planner seeds "
                    "100 + 512 rank + q, MS sampler with the 3 door-centre milestones, edge checks + rewiring + BGR_PLAN_DEDUP; "
                    "bgr_planner_run_queue, %d worker planners (one host thread, stream and map handle each), re-armed per query" % (QUERIES_PER_GPU, nconc)}
    res["note"] = ("wall clock of bgr_planner_run; batches of <= 8 candidates without edge checks run whole steps inside one "
                   "launch (plan_persist_kernel, steps_in_persistent_kernel), the others are host-driven (~30 launches and one "
                   "sync per batch); nodes_per_s is the "
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


def cpu_planner_baseline(budget_s):
    """the oracle's restatement of the planner loop (oracle/src/planner.cpp, rrt_star_milestones.cpp:1160-1406 /
    rrt_star.cpp:1140-1300) timed on the host: configs[0] in full (batch 1 = the reference's own sequential
    loop, literal loop over every leaf box in the state check), and a bounded number of batches of the
    configs[3] shape (bucket-grid broadphase in front of the same MPR: the literal loop over 795 k boxes
    would take hours)."""
    import bottleneck_guided_rrt_b200 as B
    from bottleneck_guided_rrt_b200 import synth
    from oracle import oracle as O
    cores = host_cores()
    out = {}
    occ = synth.make_map(64, 1, n_walls=3, wall_thick=2, door_cells=6)
    boxes = synth.grid_to_boxes(occ, 0.05, (0, 0, 0), 3)
    table = synth.points(9, 200_000, 3.2)
    for batch, flags, key in ((1, B.PLAN_QUIRKS, "config0_rrtstar_64_batch1"), (256, 0, "config0_rrtstar_64_batch256")):
        P = B.PlannerParams((0.2, 0.2, 0.2), (3.0, 3.0, 3.0), mode=B.PLAN_MODE_RS, k=10, batch=batch, flags=flags | B.PLAN_NO_GOAL, seed=5489)
        op = O.Planner(P, boxes, table)
        t0 = time.perf_counter()
        while len(op) < 5000 and time.perf_counter() - t0 < budget_s:
            op.run(5000, 64 if batch == 1 else 1)
        el = time.perf_counter() - t0
        out[key] = {"value": (len(op) - 1) / el, "unit": "nodes/s", "nodes": len(op), "seconds": el, "cores": 1 if batch == 1 else cores,
                    "kind": "port", "sample": "configs[0] %s: %d of 5000 nodes, %d leaf boxes, k=10, batch %d; state check = the "
                    "reference's loop over every leaf box + MPR (oracle mode 1)" % ("in full" if len(op) >= 5000 else "bounded by the time budget",
                                                                                     len(op), len(boxes), batch)}
    occ3 = synth.make_map(256, 5, n_walls=3, wall_thick=4, door_cells=5)
    boxes3 = synth.grid_to_boxes(occ3, 0.05, (0, 0, 0), 3)
    ms = door_milestones(occ3, 256)
    table3 = synth.points(9, 200_000, 12.8)
    P = B.PlannerParams((0.2, 0.2, 0.2), (12.6, 12.6, 12.6), mode=B.PLAN_MODE_MS, k=16, batch=4096,
                        flags=B.PLAN_EDGE_CHECK | B.PLAN_REWIRE | B.PLAN_NO_GOAL, seed=5489)
    op = O.Planner(P, boxes3, table3, ms, collision_mode=2)
    op.step()                                               # builds the bucket grid (untimed, as the GPU map upload is)
    n0, t0, nb = len(op), time.perf_counter(), 0
    while time.perf_counter() - t0 < budget_s and len(op) < 200_000:
        op.step()
        nb += 1
    el = time.perf_counter() - t0
    out["config3_bottleneck_256_batch4096"] = {
        "value": (len(op) - n0) / el, "unit": "nodes/s", "nodes": len(op), "batches": nb, "seconds": el, "cores": cores, "kind": "port",
        "sample": "configs[3] shape, the first %d batches after the first (%d of 200,000 nodes; the per-batch cost grows with "
                  "the tree, so this flatters the CPU): brute-force Nearest / kNN over the tree, bucket-grid broadphase + MPR "
                  "collision (oracle mode 2), OpenMP over candidates" % (nb, len(op))}
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
           "config": {"workload": "configs[1]: batched kNN, 1,000,000-node 3-D tree (U[0,25.6)^3, MT19937 seed 2), "
                                  "65,536 queries/GPU (seed 3+rank), k=16, eps=0, exact, method=%s" % args.method,
                      "cpu_arm": "libnabo 1.0.7 algorithm restated in oracle/ (kind: port), OpenMP over queries, %d threads" % cores},
           "cpu_baseline": cb,
           "e2e": {"value": value, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    if not args.skip_edges:
        occ = synth.make_map(MAP_N, 4, n_walls=8, wall_thick=4, door_cells=5)
        boxes = synth.grid_to_boxes(occ, 0.05, (0, 0, 0), 3)
        p1, p2 = synth.edges(4, N_EDGES, occ)
        out["edge_checks"] = cpu_edge_baseline(boxes, p1, p2, budget_s=15.0)
    if not args.skip_planner:
        out["planner"] = {"cpu_baseline": cpu_planner_baseline(budget_s=12.0)}
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
    ap.add_argument("--skip-brute", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
