#!/usr/bin/env python
"""Benchmark of the prioritized sample + gather + TD/priority-update loop.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one pass of the hot path over one batch of B=512 records on a
2M-record shard of Atari-shaped (4 x 84 x 84 uint8) synthetic records:
K3 proportional search -> K4 frame gather -> K5 double-DQN TD/loss/|delta|^alpha
-> K2 in-place priority write-back.  Prints ONE JSON line.
Under torchrun every rank owns a 2M shard and the global batch is 512 x N (weak scaling).
"""
import argparse
import collections
import json
import os
import subprocess
import sys
import threading
import time


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


if "reference" in sys.argv:
    # The CPU arm uses every host core it can, also under torchrun (which exports OMP_NUM_THREADS=1 to its
    # workers): the OpenMP / BLAS pools read these variables when numpy and torch are first imported.
    for _v in ("OMP_NUM_THREADS", "MKL_NUM_THREADS", "OPENBLAS_NUM_THREADS"):
        os.environ[_v] = str(host_threads())

import numpy as np   # noqa: E402

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

F_BYTES = 7056          # 84 x 84 uint8
ALPHA = 0.6
GAMMA_N = 0.99 ** 3
NCU_DRAM_BYTES_PER_LAUNCH = 26_140_000      # sample_gather_kernel, B=512, C=4 (profiles/r02_ncu_full_sample_gather.csv)
NCU_TRAFFIC_SOURCE = ("profiles/r02_ncu_full_sample_gather.csv (ncu --set full, B=512, C=4, three launches: "
                      "dram__bytes_read.sum 25.92 MB + dram__bytes_write.sum 0.20-0.26 MB per launch)")
NCU_DRAM_BYTES_SHARDED = 26_230_000         # sample_gather_sharded_kernel, 512 rows, group of one (ncu is single-GPU only)
NCU_TRAFFIC_SOURCE_SHARDED = ("profiles/r02_ncu_full_sample_gather_sharded_world1.csv (ncu --set full of the same kernel run as "
                              "a group of one on one GPU, 512 rows, C=4: dram__bytes_read.sum 25.93 MB + dram__bytes_write.sum "
                              "0.22-0.41 MB per launch; a rank of an N-GPU run does the same per-rank work)")
METRIC = "prioritized sample+update transitions/s (2M cap, B=512)"
DTYPE = "u8 frames / u64 fixed-point tree / f32 TD"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--capacity", type=int, default=1 << 21)
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--frames-per-state", type=int, default=4)
    ap.add_argument("--actions", type=int, default=4)
    ap.add_argument("--lanes", type=int, default=32)
    ap.add_argument("--gather-variant", type=int, default=int(os.environ.get("APEX_GATHER_VARIANT", "1")))
    ap.add_argument("--unfused", action="store_true", help="separate K3 and K4 launches instead of the fused kernel")
    ap.add_argument("--no-pdl", action="store_true", help="plain stream-ordered launches (no programmatic dependent launch)")
    ap.add_argument("--no-graph", action="store_true", help="launch every step from Python instead of a CUDA graph")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="budget of the CPU-baseline leg")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-learner-loop", action="store_true", help="skip the short learner-loop (configs[2]) extra key")
    ap.add_argument("--min-ms", type=float, default=50.0,
                    help="repeat the K-step block until at least this much time has been measured; the median block is reported")
    ap.add_argument("--nccl-exchange", action="store_true",
                    help="multi-GPU: all-gather the shard totals with NCCL + separate routing/search/gather kernels "
                         "instead of the peer-memory exchange + fused sharded kernel")
    ap.add_argument("--no-push", action="store_true",
                    help="multi-GPU: exchange the totals with one shard_exchange launch per step instead of "
                         "publishing them from the TD/update kernel")
    ap.add_argument("--ring-fed", action="store_true",
                    help="--workload learner (sharded): the actors' steps arrive through shared-memory staging rings")
    ap.add_argument("--sharded-loop", action="store_true",
                    help="--workload learner at N=1: run the eager sharded-learner loop (the N>1 code path) on one GPU")
    ap.add_argument("--workload", default="replay", choices=["replay", "learner"],
                    help="replay = BASELINE configs[1] microbench (default, the metric's workload); "
                         "learner = configs[2]: the same loop with the duelling network fwd/bwd + RMSprop in it")
    return ap.parse_args()


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the GPU is under load."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
        rows = [r for r in self.rows if len(r) >= 6 and r[0].isdigit()]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in rows for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        sm = sorted(int(r[0]) for r in rows)
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": int(rows[0][1]), "reasons": reasons, "samples": len(rows)}


def synth_chunk(torch, gen, m, C, nA, dev):
    """Synthetic Atari-shaped records: incompressible frames, clipped rewards, 1% terminals."""
    frames = torch.randint(0, 256, (m, F_BYTES), dtype=torch.uint8, device=dev, generator=gen)
    A = torch.randint(0, nA, (m,), dtype=torch.int32, device=dev, generator=gen)
    R = torch.randn(m, device=dev, generator=gen).clamp_(-1, 1)
    G = torch.where(torch.rand(m, device=dev, generator=gen) < 0.01, 0.0, GAMMA_N).float()
    prio = torch.randn(m, device=dev, generator=gen).abs_() + 1e-3
    return frames, A, R, G, prio


def fill_shard(torch, apex, mem, args, dev, seed):
    gen = torch.Generator(device=dev).manual_seed(seed)
    lane_cap = args.capacity // args.lanes
    chunk = min(8192, lane_cap - (args.frames_per_state - 1 + 3))
    for lane in range(args.lanes):
        done = 0
        while done < lane_cap:
            m = min(chunk, lane_cap - done)
            age = (done + np.arange(m)) % 1000                     # episodes of 1000 steps
            rel = torch.as_tensor(apex.stream_offsets(args.frames_per_state, np.full(m, 3), age)).to(dev)
            frames, A, R, G, prio = synth_chunk(torch, gen, m, args.frames_per_state, args.actions, dev)
            mem.insert(lane, frames, rel, A, R, G, prio=prio)
            done += m
    torch.cuda.synchronize()


def r_literal_quote():
    """R-literal: the unmodified reference replay.py cannot run at 2M records (O(N^2) per update) and
    its source is not on the GPU box; the committed build-container measurement + N^2 fit is quoted, labelled as such."""
    try:
        with open(os.path.join(ROOT, "profiles", "r02_r_literal_fit.json")) as f:
            d = json.load(f)
        return {"source": "profiles/r02_r_literal_fit.json (tools/r_literal_fit.py, build container, 1 Python thread)",
                "measured_transitions_per_s": {str(r["N"]): r["transitions_per_s"] for r in d["measured"]},
                "extrapolated_transitions_per_s_at_2M_B512": d["extrapolated"]["2097152"]["transitions_per_s_B512"],
                "label": "not timed in this run; the 2M figure is EXTRAPOLATED from the measured N^2 law"}
    except Exception:
        return None


def config_dict(args, world):
    """The workload description both arms print (a function of the command line and the rank count only)."""
    N, B, C = args.capacity, args.batch, args.frames_per_state
    fused = not args.unfused
    peer_path = world > 1 and not args.nccl_exchange
    graph = (not args.no_graph) and (world == 1 or peer_path)
    return {"workload": "replay microbench: %d-record shard per GPU, batch %d per GPU, C=%d 84x84 uint8, "
                        "n=3, alpha=%.1f; sample(K3)+gather(K4)+TD/loss(K5)+priority update(K2) per step" % (N, B, C, ALPHA),
            "capacity_per_gpu": N, "batch_per_gpu": B, "frames_per_state": C, "lanes": args.lanes,
            "launch": "cuda graph of 4 steps" if graph else "python loop",
            "kernels_per_step": ("sample_gather_kernel + td_small_kernel" if fused else
                                 "tree_sample_kernel + gather kernel (variant %d) + td_small_kernel" % args.gather_variant)
            if world == 1 else ("sample_gather_sharded_kernel + td_small_kernel" if peer_path else
                                "nccl all-gather + shard_targets + tree_sample + gather + td kernels"),
            "pdl": not args.no_pdl,
            "l2": "frame store %.1f GB >> 126 MB L2 (random gathers miss); the %.1f MB tree is L2-resident by design"
                  % (N * F_BYTES / 1e9, N * 8 * 33 / 32 / 1e6),
            "parallelism": ("1 shard per GPU; every rank's TD/update kernel pushes its new shard total into all peers' "
                            "memory over NVLink, the next fused routing+search+gather kernel waits for them (no NCCL on "
                            "the step path; outputs hold the whole global batch: a rank receives every draw it owns)"
                            if peer_path else "1 shard per GPU, NCCL all-gather of shard totals")
            if world > 1 else "single shard"}


def block_stats(ms_list):
    v = sorted(ms_list)
    return v[len(v) // 2], v[0], v[-1]


def run_ours(args):
    import torch
    import torch.distributed as dist
    import apex_b200
    from apex_b200 import _C

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    _C.call("apexb200_set_gather_variant", args.gather_variant)
    _C.call("apexb200_set_pdl", 0 if args.no_pdl else 1)
    fused = not args.unfused
    B, C, nA, N = args.batch, args.frames_per_state, args.actions, args.capacity
    Bg = B * world
    mem = apex_b200.ReplayMemory(N, {"priority_exponent": ALPHA, "layout": "stream", "lanes": args.lanes, "n_step": 3,
                                     "state_shape": (C, 84, 84), "seed": 4321})
    fill_shard(torch, apex_b200, mem, args, dev, 1234 + rank)
    clocks = ClockSampler(local)
    clocks.start()

    gen = torch.Generator(device=dev).manual_seed(99 + rank)
    P = 4                                                         # pool of Q-value inputs, cycled
    rows = B if world == 1 else Bg                               # sharded outputs hold the whole global batch
    qpool = [[torch.randn(rows, nA, device=dev, generator=gen) for _ in range(3)] for _ in range(P)]
    td_out = tuple(torch.empty(rows, device=dev) for _ in range(3)) + (torch.empty(1, device=dev),)
    sampler = apex_b200.ShardedSampler(mem, rank=rank, world_size=world) if world > 1 else None
    peer_path = world > 1 and not args.nccl_exchange
    max_rows = Bg                                                 # no quota: a rank receives every draw it owns
    if peer_path:
        sampler.connect_peers()
        sampler.push_on_update(not args.no_push)                  # K5+K2 publishes the new total to the peers itself
        sampler.exchange()                                        # first round (after the fill)
    ugen = torch.Generator(device=dev).manual_seed(4321)          # same draws on every rank
    u_table = ((torch.arange(Bg, device=dev, dtype=torch.float64) +
                torch.rand(64, Bg, device=dev, dtype=torch.float64, generator=ugen)) / Bg) if world > 1 else None
    step_no = [0]
    graph_steps = None
    shard_out = mem._outputs(Bg, False) if world > 1 else None

    def step(k):
        q = qpool[k % P]
        if world == 1:
            b = mem.sample_tensors(B, fused=fused)                # K3 (in-kernel stratified draws) + K4
            mem.td_update(b.idx, q[0], q[1], q[2], b.A, b.R, b.Gamma, out=td_out)      # K5 + K2
        elif peer_path:
            if args.no_push:
                sampler.exchange()
            b, _pos, cnt = sampler.sample_fused(Bg, max_rows)     # waits for the peers' totals, K8+K3+K4 fused
            mem.td_update(b.idx, q[0], q[1], q[2], b.A, b.R, b.Gamma, out=td_out, count_dev=cnt, mean_divisor=float(Bg))
        else:
            b, _pos, cnt = sampler.sample(u_table[(step_no[0] if graph_steps is None else k) % 64], out=shard_out)
            mem.td_update(b.idx, q[0], q[1], q[2], b.A, b.R, b.Gamma, out=td_out, count_dev=cnt, mean_divisor=float(Bg))
            step_no[0] += 1

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(values):
        if world == 1:
            return [float(v) for v in values]
        t = torch.tensor(list(values), device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(v) for v in t.tolist()]

    use_graph = (not args.no_graph) and (world == 1 or peer_path)   # the NCCL path stays a python loop
    W = max(args.warmup, 3)
    for k in range(W):
        step(k)
    torch.cuda.synchronize()
    graph = None
    if use_graph:
        graph_steps = P
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side):
                for k in range(P):
                    step(k)
        torch.cuda.current_stream().wait_stream(side)
        for _ in range(3):
            graph.replay()
        torch.cuda.synchronize()

    # ---- parity round at the bench size, before anything is timed (the oracle is the checker here)
    parity = parity_round(torch, dist, apex_b200, mem, sampler, args, dev, rank, world, peer_path)

    # ---- device-timed region: blocks of EXACTLY K steps, each bracketed by barrier + synchronize on both sides and
    # timed with CUDA events; the block is repeated until >= 50 ms of device time has been measured (a single
    # block of the driver's K=20 is 0.3 ms); value = the MEDIAN block, max over ranks per block.
    K = args.steps
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def timed_block():
        barrier()
        e0.record()
        if graph is not None:
            for _ in range(K // P):
                graph.replay()
            for k in range(K % P):
                step(k)
        else:
            for k in range(K):
                step(k)
        e1.record()
        barrier()
        return e0.elapsed_time(e1)

    pilot = max_over_ranks([timed_block()])[0]
    n_blocks = int(min(400, max(3, -(-args.min_ms // max(pilot, 1e-3)))))
    if peer_path:
        sampler.peer_wait_stats(reset=True)
    blocks = max_over_ranks([timed_block() for _ in range(n_blocks)])
    peer_wait_us = max_over_ranks([sampler.peer_wait_stats(reset=True)[0]])[0] if peer_path else None
    ms, ms_min, ms_max = block_stats(blocks)
    launches_per_step = (2 if fused else 3) if world == 1 else (2 if peer_path else 5 + (1 if Bg > 1024 else 0))
    value = K * Bg / (ms * 1e-3)
    timing = {"blocks": n_blocks, "steps_per_block": K, "block_ms_median": round(ms, 5), "block_ms_min": round(ms_min, 5),
              "block_ms_max": round(ms_max, 5), "device_ms_total": round(sum(blocks), 3),
              "how": "each block = exactly K steps between barrier+synchronize, CUDA events, max over ranks; median block reported"}

    # ---- roofline of the dominant kernel (fused K3+K4, single-GPU or sharded form; K4 alone with --unfused).
    # Launch duration: CUDA events around replays of a CUDA graph of R back-to-back launches of that kernel
    # (fresh draws every launch, R x 28.9 MB of distinct source frames per replay >> L2), / launches.
    peak, peak_src = peaks()
    bytes_per_rec = 4 * C * F_BYTES + 32
    roofline = None
    multi = None

    def graph_of(fn, reps_in=16, replays=25):
        sd = torch.cuda.Stream()
        sd.wait_stream(torch.cuda.current_stream())
        g = torch.cuda.CUDAGraph()
        with torch.cuda.stream(sd):
            with torch.cuda.graph(g, stream=sd):
                for r in range(reps_in):
                    fn(r)
        torch.cuda.current_stream().wait_stream(sd)
        for _ in range(3):
            g.replay()
        barrier()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record()
        for _ in range(replays):
            g.replay()
        a1.record()
        barrier()
        return a0.elapsed_time(a1) * 1e3 / (reps_in * replays)

    if peer_path:
        sampler.push_on_update(False)
        counts = []
        for _ in range(8):
            _b, _p, c_ = sampler.sample_fused(Bg, max_rows)
            counts.append(int(c_.item()))
        rows_owned = sum(counts) / len(counts)
        x_us = graph_of(lambda r: sampler.exchange())
        s_us = max_over_ranks([graph_of(lambda r: sampler.sample_fused(Bg, max_rows))])[0]
        sampler.push_on_update(not args.no_push)
        sampler.exchange()
        rows_all = max_over_ranks([rows_owned])[0]
        ach = bytes_per_rec * rows_owned / (s_us * 1e-6) / 1e9
        roofline = {"kernel": "sample_gather_sharded_kernel", "bound": "hbm", "achieved": round(ach, 1), "peak": peak,
                    "unit": "GB/s", "frac": round(ach / peak, 4),
                    "traffic": NCU_DRAM_BYTES_SHARDED if (B == 512 and C == 4) else None,
                    "traffic_source": NCU_TRAFFIC_SOURCE_SHARDED,
                    "dram_frac": round(NCU_DRAM_BYTES_SHARDED / (s_us * 1e-6) / 1e9 / peak, 4) if (B == 512 and C == 4) else None,
                    "algorithmic_bytes_per_launch": int(bytes_per_rec * rows_owned), "rows_owned_mean_rank0": rows_owned,
                    "rows_owned_max_over_ranks": rows_all, "launch_us_mean": round(s_us, 3),
                    "launches_timed": 16 * 25, "peak_source": peak_src,
                    "timing": "rank 0's rows x SURVEY 8(d) bytes / launch duration (max over ranks) from CUDA events around 25 "
                              "replays of a CUDA graph of 16 back-to-back launches",
                    "share_of_step": round(s_us / (ms / K * 1e3), 3)}
        multi = {"shard_exchange_us": round(x_us, 3), "sample_gather_sharded_us": round(s_us, 3),
                 "peer_wait_us": round(peer_wait_us, 3),
                 "peer_wait_note": "mean time per step, inside the timed region, that the sharded sampling kernel spent polling for "
                                   "the slowest peer's total (pushed over NVLink by that peer's TD/update kernel); max over ranks"}
    if world == 1:
        o = mem._outputs(B, False)

        def dominant(r=0):
            if fused:
                mem.sample_tensors(B)
            else:
                mem.gather(o["idx"], fsel=o["fsel"], out=o)
        R_, reps = 16, 25
        dominant()
        torch.cuda.synchronize()

        def with_k3(r):
            if not fused:
                mem.sample_stratified(B, idx_out=o["idx"], fsel_out=o["fsel"])
            dominant()
        g_mean_us = graph_of(with_k3, R_, reps)
        if not fused:                                             # subtract the interleaved K3 launches
            g_mean_us -= graph_of(lambda r: mem.sample_stratified(B, idx_out=o["idx"], fsel_out=o["fsel"]), R_, reps)
        # the same launches with the OUTPUT batch rotating through R_ buffer sets (16 x 28.9 MB = 462 MB > L2):
        # every launch's 28.9 MB of writes then has to evict a previous batch to HBM instead of overwriting an
        # L2-resident one
        rot_us = None
        if fused:
            key = (B, False)
            sets = [dict(o)] + [{k: (v.clone() if v is not None else None) for k, v in o.items()} for _ in range(R_ - 1)]

            def rotating(r):
                mem._out_cache[key] = sets[r % R_]
                mem.sample_tensors(B)
            rot_us = graph_of(rotating, R_, reps)
            mem._out_cache[key] = sets[0]
            del sets

        n_r = min(K, 400)
        ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(n_r)]
        big = torch.zeros(1 << 30, dtype=torch.uint8, device=dev)
        torch.cuda.synchronize()
        for _ in range(40):
            big.add_(1)                                           # ~13 ms of queued GPU work: the CPU runs ahead
        for k in range(n_r):
            q = qpool[k % P]
            if fused:
                ev[k][0].record()
                mem.sample_tensors(B)
            else:
                mem.sample_stratified(B, idx_out=o["idx"], fsel_out=o["fsel"])
                ev[k][0].record()
                mem.gather(o["idx"], fsel=o["fsel"], out=o)
            ev[k][1].record()
            mem.td_update(o["idx"], q[0], q[1], q[2], o["A"], o["R"], o["Gamma"], out=td_out)
            ev[k][2].record()
        torch.cuda.synchronize()
        g_us = sorted(e[0].elapsed_time(e[1]) * 1e3 for e in ev)
        t_us = sorted(e[1].elapsed_time(e[2]) * 1e3 for e in ev)
        del big
        ach = bytes_per_rec * B / (g_mean_us * 1e-6) / 1e9
        traffic = NCU_DRAM_BYTES_PER_LAUNCH if (fused and B == 512 and C == 4) else None
        kname = "sample_gather_kernel" if fused else ("gather_tma_kernel" if args.gather_variant == 1 else "gather_vec_kernel")
        roofline = {"kernel": kname, "bound": "hbm", "achieved": round(ach, 1), "peak": peak, "unit": "GB/s",
                    "frac": round(ach / peak, 4), "traffic": traffic, "traffic_source": NCU_TRAFFIC_SOURCE,
                    "dram_frac": round(traffic / (g_mean_us * 1e-6) / 1e9 / peak, 4) if traffic else None,
                    "dram_frac_note": "traffic / launch duration / peak: the bytes DRAM itself moves per launch (the output batch "
                                      "stays in L2 and S_t/S_tpn share one frame in eight); `frac` is the SURVEY 8(d) accounting "
                                      "of algorithmic bytes",
                    "algorithmic_bytes_per_launch": bytes_per_rec * B,
                    "launch_us_mean": round(g_mean_us, 3), "launches_timed": R_ * reps, "peak_source": peak_src,
                    "rotating_outputs": None if rot_us is None else {
                        "launch_us_mean": round(rot_us, 3), "achieved": round(bytes_per_rec * B / (rot_us * 1e-6) / 1e9, 1),
                        "frac": round(bytes_per_rec * B / (rot_us * 1e-6) / 1e9 / peak, 4),
                        "note": "same launches, output batch rotating through %d buffer sets (%.0f MB > 126 MB L2): the "
                                "writes reach HBM" % (R_, R_ * 2 * B * C * F_BYTES / 1e6)},
                    "timing": "cuda events around %d replays of a CUDA graph of %d back-to-back launches "
                              "(includes one inter-launch gap per launch); outputs reuse one L2-resident batch buffer"
                              % (reps, R_),
                    "share_of_step": round(g_mean_us / (ms / K * 1e3), 3),
                    "in_step_event_pairs": {"launch_us_mean": round(sum(g_us) / len(g_us), 3),
                                            "launch_us_min": round(g_us[0], 3), "launches_timed": len(g_us),
                                            "td_small_kernel_us_mean": round(sum(t_us) / len(t_us), 3),
                                            "note": "event pair per launch inside an eager pass of the full step; "
                                                    "inflated by event overhead"},
                    "note": "the launch is a chain of latencies (tree descent, frame-slot row, one HBM round trip and one L2 "
                            "write per frame), not a saturated memory system; see profiles/"}

    # ---- end to end through the host-facing C-ABI step: inputs start in pinned HOST memory every step, loss /
    # priorities / indices (/ draw positions, row count) are read back on the host every step.  Blocks of K
    # steps of wall-clock time, repeated until >= 50 ms, median block, max over ranks.
    e2e = None
    if world == 1 or peer_path:
        rng = np.random.default_rng(4321 + rank)
        if world == 1:
            pipe = apex_b200.HostStepPipeline(mem, B, nA)
            q_host = rng.standard_normal((P, 3, B, nA)).astype(np.float32)
            u_host = (np.arange(B) + rng.random((64, B))) / B     # host-side stratified draws

            def submit(k):
                return pipe.submit(u_host[k % 64], q_host[k % P])  # fills the pinned inputs, one C call
        else:
            pipe = apex_b200.ShardedHostStepPipeline(sampler, Bg, nA, max_rows=max_rows)
            # the host writes fresh Q rows every step for the rows this rank can plausibly own (twice its fair share;
            # the count that comes back is checked against it below); the H2D copy and the kernels cover all max_rows
            fill_rows = min(max_rows, 2 * B)
            q_host = rng.standard_normal((P, 3, fill_rows, nA)).astype(np.float32)
            pipe.submit(rng.standard_normal((3, max_rows, nA)).astype(np.float32))     # every row of every slot valid once
            pipe.submit(rng.standard_normal((3, max_rows, nA)).astype(np.float32))
            pipe.result(pipe.submit(rng.standard_normal((3, max_rows, nA)).astype(np.float32)))

            def submit(k):
                return pipe.submit(q_host[k % P], rows=fill_rows)
        state = {"loss": 0.0, "rows": 0, "rows_max": 0}

        def take(ticket):
            r = pipe.result(ticket)
            state["loss"] += r[0]
            if world > 1:
                state["rows"] += r[4]
                state["rows_max"] = max(state["rows_max"], r[4])

        def e2e_run(n_blocks):
            """(n_blocks + 1) x K steps through the pipeline without draining it in between; a wall-clock stamp is
            taken whenever another K results have been read on the host.  The first K steps are the ramp; the
            differences between later stamps are steady-state blocks of exactly K steps."""
            barrier()
            pending, stamps, taken = collections.deque(), [], 0
            for k in range((n_blocks + 1) * K):
                pending.append(submit(k))
                if len(pending) == pipe.DEPTH:
                    take(pending.popleft())                       # host reads step k-2 while steps k-1, k are queued
                    taken += 1
                    if taken % K == 0:
                        stamps.append(time.perf_counter())
            while pending:
                take(pending.popleft())
                taken += 1
                if taken % K == 0:
                    stamps.append(time.perf_counter())
            barrier()
            return [(b - a) * 1e3 for a, b in zip(stamps, stamps[1:])]
        pilot_e = max_over_ranks([block_stats(e2e_run(2))[0]])[0]
        n_eb = int(min(2000, max(3, -(-args.min_ms // max(pilot_e, 1e-3)))))
        state.update(loss=0.0, rows=0, rows_max=0)
        eb = max_over_ranks(e2e_run(n_eb))
        em, emin, emax = block_stats(eb)
        n_taken = (n_eb + 1) * K
        e2e = {"value": round(K * Bg / (em * 1e-3), 1), "unit": "transitions/s", "h2d_bytes_per_step": pipe.in_bytes,
               "d2h_bytes_per_step": pipe.out_words * 4, "steps": K, "ms_per_step": round(em / K, 5),
               "blocks": n_eb, "block_ms_median": round(em, 4), "block_ms_min": round(emin, 4), "block_ms_max": round(emax, 4),
               "api": ("apex_b200.HostStepPipeline.submit(u_host, q_host)/result() -> C ABI apexb200_host_step: pinned H2D of "
                       "draws+Q on a copy stream, fused sample+gather, TD+update, D2H of loss/priorities/indices on a copy "
                       "stream, ordered by events" if world == 1 else
                       "apex_b200.ShardedHostStepPipeline.submit(q_host)/result() -> C ABI apexb200_host_step_sharded on every rank: "
                       "pinned H2D of Q rows for the whole global batch (%d rows: no quota), fused sharded routing+search+gather "
                       "(draws generated in-kernel, identical on every rank), TD+update+total push, D2H of loss/priorities/"
                       "indices/draw positions/row count" % max_rows) +
                      "; every step's results are read on the host, %d steps behind (%d buffer sets); wall clock, bytes are "
                      "per rank" % (pipe.DEPTH - 1, pipe.DEPTH),
               "mean_loss": round(state["loss"] / n_taken, 6),
               "how": "one continuous run of (blocks + 1) x K pipelined steps; wall-clock stamp every K results read on "
                      "the host; first K steps (ramp) discarded; median block, max over ranks per block"}
        if world > 1:
            e2e["host_rows_filled_per_step"] = fill_rows
            e2e["rows_within_filled"] = bool(state["rows_max"] <= fill_rows)
            e2e["rows_owned_mean_rank0"] = round(state["rows"] / n_taken, 2)
            e2e["rows_owned_max_rank0"] = state["rows_max"]
    clk = clocks.stop()
    mem.check_status()

    learner_loop = None
    if rank == 0 and world == 1 and fused and not args.no_learner_loop:
        learner_loop = learner_loop_short(torch, apex_b200, mem, args, dev)

    cpu, cpu_c = None, None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu, cpu_c = cpu_baseline_from_device(torch, mem, args, dev)

    if rank == 0:
        ref = lambda a, b: round(a / b, 2) if (a and b) else None           # noqa: E731
        line = {
            "metric": METRIC, "value": round(value, 1),
            "unit": "transitions/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": round(ms / K, 6), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": DTYPE, "data": "synthetic", "config": config_dict(args, world),
            "gpu_launches": K * launches_per_step, "timing": timing, "parity": parity, "clocks": clk, "roofline": roofline,
            "multi_gpu_kernels": multi, "e2e": e2e, "cpu_baseline": cpu, "cpu_baseline_c_openmp": cpu_c,
            "vs_cpu_c_openmp": None if not cpu_c else {"e2e_ratio": ref(e2e and e2e["value"], cpu_c["value"]),
                                                      "value_ratio": ref(value, cpu_c["value"]),
                                                      "note": "ours / the hand-written C+OpenMP port on this box's cores "
                                                              "(a stricter baseline than the reference's numpy/torch path)"},
            "learner_loop": learner_loop, "cpu_baseline_r_literal": r_literal_quote(),
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def parity_round(torch, dist, apex, mem, sampler, args, dev, rank, world, peer_path):
    """One sample at the bench size checked against the oracle: record indices bit-exact against the cumulative
    search over the (concatenated) leaves for the same draws, every gathered byte equal to the frame store."""
    from apex_b200 import _C
    from oracle import replay_oracle as ro
    B, C, N = args.batch, args.frames_per_state, args.capacity
    Bg = B * world
    store = mem.buffer(_C.BUF_FRAMES).view(-1, F_BYTES)
    fidx = mem.buffer(_C.BUF_FIDX).view(N, 2 * C)
    leaves = mem.buffer(_C.BUF_LEAVES).clone()

    def bytes_ok(b, n):
        sel = b.fsel[:n].long()
        ok = torch.equal(b.fsel[:n], fidx.index_select(0, b.idx[:n].long()))
        ok &= torch.equal(b.S_t[:n].reshape(n * C, F_BYTES), store.index_select(0, sel[:, :C].reshape(-1)))
        ok &= torch.equal(b.S_tpn[:n].reshape(n * C, F_BYTES), store.index_select(0, sel[:, C:].reshape(-1)))
        for name, which in (("A", _C.BUF_ACTION), ("R", _C.BUF_REWARD), ("Gamma", _C.BUF_GAMMA)):
            ok &= torch.equal(getattr(b, name)[:n], mem.buffer(which).index_select(0, b.idx[:n].long()))
        return bool(ok)

    if world == 1:
        u_out = torch.empty(B, dtype=torch.float64, device=dev)
        b = mem.sample_tensors(B, u_out=u_out)
        want = ro.tree_sample(leaves.cpu().numpy().view(np.uint64), u_out.cpu().numpy())
        return {"indices_bit_exact": bool((b.idx.cpu().numpy() == want).all()), "bytes_exact": bytes_ok(b, B), "world": 1,
                "draws": B, "records": N, "checked_against": "oracle.replay_oracle.tree_sample (searchsorted(cumsum(leaves), "
                "target, 'right')) on the device's leaves and the kernel's own draws; every gathered byte vs the frame store"}
    if not peer_path:
        return None
    u_out = torch.zeros(Bg, dtype=torch.float64, device=dev)
    b, pos, cnt = sampler.sample_fused(Bg, u_out=u_out)
    n = int(cnt.item())
    u_glob = torch.zeros(Bg, dtype=torch.float64, device=dev)
    owner = torch.zeros(Bg, dtype=torch.int64, device=dev)
    u_glob[pos[:n].long()] = u_out[:n]
    owner[pos[:n].long()] = 1
    dist.all_reduce(u_glob)
    dist.all_reduce(owner)
    every = [torch.empty_like(leaves) for _ in range(world)]
    dist.all_gather(every, leaves)
    whole = ro.tree_sample(torch.cat(every).cpu().numpy().view(np.uint64), u_glob.cpu().numpy())
    mine = np.nonzero(whole // N == rank)[0]
    idx_ok = bool((owner == 1).all().item()) and n == len(mine) and bool((pos.cpu().numpy()[:n] == mine).all()) and \
        bool((b.idx.cpu().numpy()[:n] == whole[mine] % N).all())
    flags = torch.tensor([1 if idx_ok else 0, 1 if bytes_ok(b, n) else 0], device=dev)
    dist.all_reduce(flags, op=dist.ReduceOp.MIN)
    cmin, cmax = torch.tensor([n], device=dev), torch.tensor([n], device=dev)
    dist.all_reduce(cmin, op=dist.ReduceOp.MIN)
    dist.all_reduce(cmax, op=dist.ReduceOp.MAX)
    return {"indices_bit_exact": bool(flags[0].item()), "bytes_exact": bool(flags[1].item()), "world": world, "draws": Bg,
            "records": N * world, "rows_per_rank_min": int(cmin.item()), "rows_per_rank_max": int(cmax.item()),
            "checked_against": "oracle.replay_oracle.tree_sample over the concatenated leaves of all shards for the global draw "
                               "vector reassembled from every rank's rows (each draw owned exactly once); every gathered byte "
                               "vs the rank's frame store; min over ranks"}


def learner_loop_short(torch, apex, mem, args, dev):
    """BASELINE configs[2] in short (<= ~2 s): the learner loop on the same shard, whole step replayed from one CUDA
    graph — fused sample+gather writing the network's float32 input directly, three forwards, backward, RMSprop,
    fused TD/priority update.  Reported as an extra key; not the metric's workload."""
    try:
        B, C, nA = args.batch, args.frames_per_state, args.actions
        lp = {"remove_old_xp_freq": 100, "q_target_sync_freq": 2500, "min_replay_mem_size": 20000, "replay_sample_size": B,
              "load_saved_state": False}
        out = {}
        for fmt in ("f32", "u8"):
            lp["batch_format"] = fmt
            L = apex.Learner({"state_shape": [C, 84, 84], "action_dim": nA}, dict(lp), {}, mem)
            L.capture_step()
            for _ in range(5):
                L.step_graphed()
            torch.cuda.synchronize()
            n = 60
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(n):
                loss = L.step_graphed()
            e1.record()
            torch.cuda.synchronize()
            out[fmt] = (e0.elapsed_time(e1) / n, float(loss))
            del L
        mem.check_status()
        ms = out["f32"][0]
        return {"ms_per_step": round(ms, 4), "transitions_per_s": round(B / (ms * 1e-3), 1), "steps": 60,
                "ms_per_step_u8_gather_plus_float_cast": round(out["u8"][0], 4), "final_loss": out["f32"][1],
                "workload": "Ape-X learner loop (BASELINE configs[2]): 2M-record replay, batch %d, n=3 double-DQN, duelling net "
                            "fp32 (PyTorch/cuDNN), RMSprop; one CUDA graph per step; f32 = gather writes the float32 "
                            "[B,C,84,84] network input itself, u8 = uint8 gather + two .float() passes" % B}
    except Exception as e:                                        # an extra key must never take the headline line down
        return {"error": "%s: %s" % (type(e).__name__, e)}


class NumpyReplay:
    """R-vec of BASELINE.md §3: the reference's own algorithm and libraries (numpy / torch-CPU),
    vectorised over the batch — np.random.choice's cumsum + right-sided search (replay.py:24,52),
    batch assembly (replay.py:54-56, learner.py:35-38), learner.py:43-52 arithmetic, dict-update
    write-back (replay.py:40).  torch.index_select does the frame copies on all host threads;
    numpy's cumsum / searchsorted are single-threaded by construction."""

    def __init__(self, leaves, frames, fidx, A, R, G, C, alpha):
        import torch
        self.torch = torch
        self.leaves, self.fidx, self.A, self.R, self.G, self.C, self.alpha = leaves, fidx, A, R, G, C, alpha
        self.frames = torch.from_numpy(frames)
        self.threads = torch.get_num_threads()

    def step(self, u, q1, q2, q3):
        from oracle import replay_oracle as ro, td_oracle as to
        c = np.cumsum(self.leaves, dtype=np.uint64)
        idx = np.searchsorted(c, ro.sample_target(u, int(c[-1])), side="right")
        f = self.torch.from_numpy(self.fidx[idx].astype(np.int64))
        S_t = self.torch.index_select(self.frames, 0, f[:, :self.C].reshape(-1))
        S_tpn = self.torch.index_select(self.frames, 0, f[:, self.C:].reshape(-1))
        o = to.td_forward(q1, q2, q3, self.A[idx], self.R[idx], self.G[idx], alpha=self.alpha)
        self.leaves[idx] = ro.priority_to_leaf(o["prio"])
        return idx, S_t, S_tpn, o


def time_cpu(obj, B, nA, seconds, max_steps, rng):
    from oracle import replay_oracle as ro
    q = [rng.standard_normal((B, nA)).astype(np.float32) for _ in range(3)]
    for _ in range(2):
        obj.step(ro.stratified_uniforms(B, rng), *q)
    t0, n = time.perf_counter(), 0
    while True:
        obj.step(ro.stratified_uniforms(B, rng), *q)
        n += 1
        dt = time.perf_counter() - t0
        if dt > seconds or n >= max_steps:
            return n, dt


def cpu_baselines(leaves, frames, fidx, A, R, G, args, n_records, only_c=False):
    """-> (cpu_baseline dict [numpy/torch port], extra dict [C/OpenMP port])."""
    from oracle.c_oracle import CpuReplay
    C, B, nA = args.frames_per_state, args.batch, args.actions
    rng = np.random.default_rng(7)
    base = None
    if not only_c:
        ref = NumpyReplay(leaves.copy(), frames, fidx, A, R, G, C, ALPHA)
        n, dt = time_cpu(ref, B, nA, args.cpu_seconds * 0.6, 5000, rng)
        base = {"value": round(n * B / dt, 1), "unit": "transitions/s", "cores": ref.threads, "kind": "port",
                "sample": "%d steps of B=%d on a %d-record host-resident shard (%.1f s): numpy/torch-CPU restatement of the "
                          "reference algorithm (u64 cumsum + right-sided searchsorted per np.random.choice, index_select "
                          "frame copies on %d torch threads, learner.py TD math in numpy, fancy-index write-back)"
                          % (n, B, n_records, dt, ref.threads),
                "ms_per_step": round(dt / n * 1e3, 4)}
    cpu = CpuReplay(leaves.copy(), frames, fidx, A, R, G, C, F_BYTES, ALPHA)
    n2, dt2 = time_cpu(cpu, B, nA, args.cpu_seconds * 0.4, 5000, rng)
    extra = {"value": round(n2 * B / dt2, 1), "unit": "transitions/s", "cores": cpu.threads, "kind": "port (optimised)",
             "sample": "%d steps (%.1f s) of the same step hand-written in C with OpenMP on %d threads "
                       "(oracle/apex_oracle.c): a stricter baseline than the reference's numpy path" % (n2, dt2, cpu.threads),
             "ms_per_step": round(dt2 / n2 * 1e3, 4)}
    return base, extra


def cpu_baseline_from_device(torch, mem, args, dev):
    """CPU baselines on the host cores, on the SAME shard contents (copied back from HBM)."""
    from apex_b200 import _C
    from oracle import replay_oracle as ro
    try:
        import psutil
        avail = psutil.virtual_memory().available
    except Exception:
        avail = 0
    N, C, B = args.capacity, args.frames_per_state, args.batch
    need = N * F_BYTES * 1.15
    if avail and avail < need + (8 << 30):
        return {"value": None, "unit": "transitions/s", "cores": os.cpu_count(), "kind": "port",
                "sample": "skipped: host RAM %.0f GB < %.0f GB needed for the frame store" % (avail / 1e9, need / 1e9)}, None
    # numpy-allocated host copy, as in the reference arm (the numpy port runs up to 2x slower on a frame store that
    # torch's CPU allocator handed out)
    frames = np.empty((N, F_BYTES), np.uint8)
    torch.from_numpy(frames).copy_(mem.buffer(_C.BUF_FRAMES).view(-1, F_BYTES))
    leaves = mem.buffer(_C.BUF_LEAVES).cpu().numpy().view(np.uint64).copy()
    fidx = mem.buffer(_C.BUF_FIDX).cpu().numpy().reshape(N, 2 * C)
    A, R, G = (mem.buffer(w).cpu().numpy() for w in (_C.BUF_ACTION, _C.BUF_REWARD, _C.BUF_GAMMA))
    # parity at full size: GPU and CPU draw the same records from the same leaves for the same uniforms
    u = ro.stratified_uniforms(B, np.random.default_rng(5))
    gpu_idx = mem.sample_indices(torch.as_tensor(u).to(dev)).cpu().numpy()
    parity = bool((gpu_idx == ro.tree_sample(leaves, u)).all())
    base, extra = cpu_baselines(leaves, frames, fidx, A, R, G, args, N)
    base["index_parity_vs_gpu_at_full_size"] = parity
    return base, extra


def run_reference(args):
    """CPU arm: the reference's algorithm for the path (oracle port on the host cores); rank 0 only.
    Same workload description as the GPU arm.  At N > 1 the one host stands in for all N shards: the
    global batch of N x B draws is taken over N x capacity records (or as many as host RAM holds)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import torch
    from oracle import replay_oracle as ro
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    torch.set_num_threads(host_threads())
    N, C, B, nA = args.capacity, args.frames_per_state, args.batch, args.actions
    Bg = B * world
    try:
        import psutil
        avail = psutil.virtual_memory().available
    except Exception:
        avail = 1 << 62
    n_cpu = N * world
    while n_cpu * F_BYTES * 1.1 + (4 << 30) > avail and n_cpu > 1 << 16:
        n_cpu //= 2
    rng = np.random.default_rng(1234)
    block = rng.integers(0, 256, (1 << 12, F_BYTES), dtype=np.uint8)
    frames = np.empty((n_cpu, F_BYTES), np.uint8)
    for lo in range(0, n_cpu, 1 << 12):
        hi = min(n_cpu, lo + (1 << 12))
        frames[lo:hi] = block[: hi - lo]
        frames[lo:hi, 0] = (lo >> 12) & 255
    lane_cap = n_cpu // args.lanes
    pos = np.arange(n_cpu) % lane_cap
    rel = np.concatenate([np.arange(-3 - C + 1, -2), np.arange(-C + 1, 1)])[None, :]
    fidx = ((np.arange(n_cpu) // lane_cap * lane_cap)[:, None] + (pos[:, None] + rel) % lane_cap).astype(np.int32)
    A = rng.integers(0, nA, n_cpu).astype(np.int32)
    R = np.clip(rng.standard_normal(n_cpu), -1, 1).astype(np.float32)
    G = np.where(rng.random(n_cpu) < 0.01, 0, GAMMA_N).astype(np.float32)
    leaves = ro.priority_to_leaf(ro.pow_priority(np.abs(rng.standard_normal(n_cpu)) + 1e-3, ALPHA))
    ref = NumpyReplay(leaves.copy(), frames, fidx, A, R, G, C, ALPHA)
    q = [[rng.standard_normal((Bg, nA)).astype(np.float32) for _ in range(3)] for _ in range(4)]
    W, K = max(args.warmup, 3), args.steps
    for k in range(W):
        ref.step(ro.stratified_uniforms(Bg, rng), *q[k % 4])
    us = [ro.stratified_uniforms(Bg, rng) for _ in range(min(K, 256))]

    def block_of_k():
        t0 = time.perf_counter()
        for k in range(K):
            ref.step(us[k % len(us)], *q[k % 4])
        return (time.perf_counter() - t0) * 1e3
    blocks = [block_of_k()]
    while sum(blocks) < args.min_ms or len(blocks) < 3:
        if sum(blocks) > 120e3:                                   # bounded: at most ~2 minutes of CPU blocks
            break
        blocks.append(block_of_k())
    ms, ms_min, ms_max = block_stats(blocks)
    value = round(K * Bg / (ms * 1e-3), 1)
    sample = ("each step = one full step of %d draws (B=%d x %d shard(s)) over %d host-resident records (%s); numpy/torch-CPU "
              "restatement of the reference algorithm (u64 cumsum + right-sided searchsorted per np.random.choice, "
              "torch.index_select frame copies on %d threads, learner.py TD math, fancy-index write-back); %d block(s) of %d "
              "steps, median block; the literal reference is O(N^2) per update and cannot run at this size (BASELINE.md §2)"
              % (Bg, B, world, n_cpu, "all shards" if n_cpu == N * world else "as many as host RAM holds, of %d" % (N * world),
                 ref.threads, len(blocks), K))
    args.cpu_seconds = min(args.cpu_seconds, 10.0)
    args_c = argparse.Namespace(**vars(args))
    args_c.batch = Bg
    _, extra = cpu_baselines(leaves, frames, fidx, A, R, G, args_c, n_cpu, only_c=True)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value,
        "unit": "transitions/s", "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": round(ms / K, 5),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": DTYPE,
        "data": "synthetic", "config": config_dict(args, world),
        "timing": {"blocks": len(blocks), "steps_per_block": K, "block_ms_median": round(ms, 3), "block_ms_min": round(ms_min, 3),
                   "block_ms_max": round(ms_max, 3), "how": "wall clock around blocks of K steps, median block"},
        "cpu_baseline": {"value": value, "unit": "transitions/s", "cores": ref.threads, "kind": "port", "sample": sample,
                         "host_records": n_cpu, "draws_per_step": Bg},
        "cpu_baseline_c_openmp": extra, "cpu_baseline_r_literal": r_literal_quote(),
        "e2e": {"value": value, "unit": "transitions/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def run_learner(args):
    """BASELINE configs[2]: Ape-X learner loop, 2M-record replay, batch 512, n=3 double-DQN, 1 GPU; under
    torchrun (N > 1) configs[3]: the replay sharded over N GPUs (2M records each), data-parallel learner with
    a gradient all-reduce, and 256/8 = 32 synthetic actors per GPU inserting through K7 + K6 every step.
    Not the headline metric's workload; prints its own JSON line."""
    import torch
    import apex_b200
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1 or args.sharded_loop:
        return run_learner_sharded(args, world)
    torch.cuda.set_device(0)
    dev = torch.device("cuda:0")
    B, C, nA, N = args.batch, args.frames_per_state, args.actions, args.capacity
    mem = apex_b200.ReplayMemory(N, {"priority_exponent": ALPHA, "layout": "stream", "lanes": args.lanes, "n_step": 3,
                                     "state_shape": (C, 84, 84), "seed": 4321})
    fill_shard(torch, apex_b200, mem, args, dev, 1234)
    lp = {"remove_old_xp_freq": 100, "q_target_sync_freq": 2500, "min_replay_mem_size": 20000, "replay_sample_size": B,
          "load_saved_state": False}
    L = apex_b200.Learner({"state_shape": [C, 84, 84], "action_dim": nA}, lp, {}, mem)
    for _ in range(max(args.warmup, 3)):
        L.step_fast()
    torch.cuda.synchronize()
    K = min(args.steps, 300)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        loss = L.step_fast()
    e1.record()
    torch.cuda.synchronize()
    ms_eager = e0.elapsed_time(e1)
    L.capture_step()                                             # whole step (model included) as one CUDA graph
    for _ in range(3):
        L.step_graphed()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(K):
        loss = L.step_graphed()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    # the data path alone inside the same loop
    d0, d1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    q = [torch.randn(B, nA, device=dev) for _ in range(3)]
    d0.record()
    for _ in range(K):
        b = mem.sample_tensors(B)
        mem.td_update(b.idx, q[0], q[1], q[2], b.A, b.R, b.Gamma)
    d1.record()
    torch.cuda.synchronize()
    mem.check_status()
    print(json.dumps({"metric": "Ape-X learner loop transitions/s (2M cap, B=512, duelling net fp32, 1 GPU)",
                      "value": round(K * B / (ms * 1e-3), 1), "unit": "transitions/s", "n_gpus": 1, "steps": K,
                      "ms_per_step": round(ms / K, 4), "ms_per_step_eager": round(ms_eager / K, 4),
                      "data_path_ms_per_step_eager": round(d0.elapsed_time(d1) / K, 4),
                      "final_loss": float(loss), "higher_is_better": True, "data": "synthetic",
                      "config": {"workload": "learner loop: sample+gather -> 3 forwards + backward + RMSprop (PyTorch/cuDNN fp32) -> "
                                             "fused TD/priority update; whole step replayed from one CUDA graph",
                                 "model": "duelling conv net 3.33M params (unchanged)"}}))


def run_learner_sharded(args, world):
    import torch
    import torch.distributed as dist
    import apex_b200
    rank, local = int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce_(t, op=None):
        if world > 1:
            dist.all_reduce(t, op=op if op is not None else dist.ReduceOp.SUM)
        return t
    B, C, nA, N = args.batch, args.frames_per_state, args.actions, args.capacity
    mem = apex_b200.ReplayMemory(N, {"priority_exponent": ALPHA, "layout": "stream", "lanes": args.lanes, "n_step": 3,
                                     "state_shape": (C, 84, 84), "seed": 4321})
    fill_shard(torch, apex_b200, mem, args, dev, 1234 + rank)
    lp = {"remove_old_xp_freq": 100, "q_target_sync_freq": 2500, "min_replay_mem_size": 20000, "replay_sample_size": B,
          "load_saved_state": False}
    L = apex_b200.ShardedLearner({"state_shape": [C, 84, 84], "action_dim": nA}, lp, {}, mem)
    L.setup()
    # 256 actors over the box = 256 / world per GPU, one ring lane each, T env steps per learner step
    E, T, every = min(max(256 // world, 1), args.lanes), 8, 4        # chunks of 8 env steps every 4th learner step
    rings, ring_ingest = [], None
    if args.ring_fed:
        # the actors' transport of SURVEY 8(f).1: one shared-memory staging ring per actor (written here by this process
        # the way an actor process writes it: plain memcpys, no CUDA), drained by RingIngest.poll
        names = ["apexring_bench_%d_%d_%d" % (os.getpid(), rank, e) for e in range(E)]
        rings = [apex_b200.ActorStagingRing(nm, 4, T, F_BYTES, nA, create=True) for nm in names]
        ring_ingest = apex_b200.RingIngest(mem, rings, n_step=3, gamma=0.99)
        rrng = np.random.default_rng(5 + rank)
        payload = (rrng.integers(0, 256, (T, F_BYTES), dtype=np.uint8), rrng.integers(0, nA, T).astype(np.int32),
                   np.clip(rrng.standard_normal(T), -1, 1).astype(np.float32), np.zeros(T, bool),
                   rrng.standard_normal((T, nA)).astype(np.float32))
    ingest = None if args.ring_fed else apex_b200.VectorActorIngest(mem, E, n_step=3, gamma=0.99, lanes=list(range(E)))
    gen = torch.Generator(device=dev).manual_seed(77 + rank)
    pool = [(torch.randint(0, 256, (E, T, F_BYTES), dtype=torch.uint8, device=dev, generator=gen),
             torch.randint(0, nA, (E, T), dtype=torch.int32, device=dev, generator=gen),
             torch.randn(E, T, device=dev, generator=gen).clamp_(-1, 1),
             np.zeros((E, T), bool), torch.randn(E, T, nA, device=dev, generator=gen)) for _ in range(4)]

    def step(k):
        if k % every == 0:
            if ring_ingest is not None:
                for r_ in rings:                                  # every actor has one more chunk of T steps ready
                    r_.push(*payload)
                ring_ingest.poll()                                # one pack kernel + the three ingest launches for all rings
            else:
                ingest.push(*pool[(k // every) % 4])              # n-step + initial priority + insert into E lanes (device inputs)
            L.after_insert()                                      # totals changed at different rates: one exchange launch
        return L.step_sharded()

    for k in range(max(args.warmup, 3)):
        step(k)
    K = min(args.steps, 200)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(K):
        loss = step(k)
    e1.record()
    barrier()
    ms = reduce_(torch.tensor([e0.elapsed_time(e1)], device=dev), dist.ReduceOp.MAX)
    total = reduce_(loss.clone().reshape(1))
    mem.check_status()
    csum = torch.stack([p.double().sum() for p in L.Q.parameters()]).sum().reshape(1)
    both = [torch.zeros_like(csum) for _ in range(world)]
    if world > 1:
        dist.all_gather(both, csum)
    if ring_ingest is not None:
        ring_ingest.flush()
        torch.cuda.synchronize()
        for r_ in rings:
            r_.close(unlink=True)
    if rank == 0:
        ms = float(ms)
        print(json.dumps({"metric": "Ape-X sharded learner loop transitions/s (2M cap per GPU, B=512 per GPU, duelling net fp32)",
                          "value": round(K * B * world / (ms * 1e-3), 1), "unit": "transitions/s", "n_gpus": world, "steps": K,
                          "ms_per_step": round(ms / K, 4), "final_global_loss": float(total), "scaling": "weak",
                          "replicas_equal": len({float(x) for x in both}) == 1,
                          "inserted_per_step_per_gpu": E * T / every, "higher_is_better": True, "data": "synthetic",
                          "actor_transport": ("%d shared-memory staging rings per GPU -> RingIngest.poll (apexb200_ingest_push_blobs)" % E)
                          if args.ring_fed else "device tensors -> VectorActorIngest.push (apexb200_ingest_push)",
                          "config": {"workload": "sharded learner loop: %d actors x %d env steps inserted per GPU every 4th step (K7+K6) -> "
                                                 "shard exchange -> fused sharded sample+gather (peer memory) -> 3 forwards + backward "
                                                 "(cuDNN fp32) -> flat gradient all-reduce (NCCL) -> RMSprop -> fused TD/priority "
                                                 "update with total push; eager launches" % (E, T),
                                     "model": "duelling conv net 3.33M params (unchanged)"}}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    elif a.workload == "learner":
        run_learner(a)
    else:
        run_ours(a)
