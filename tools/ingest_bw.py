#!/usr/bin/env python
"""Actor -> replay ingest throughput (replaces main.py:21-25 + actor.py:29-70,127-143,181-187).

(a) VectorActorIngest.push = apexb200_ingest_push: n-step assembly + initial priority + ring insert for E lock-stepped
    actors, all state on the device; inputs on the device / in host arrays; eager calls and a CUDA-graph replay.
(b) RingIngest.poll = apexb200_ingest_push_blobs: 256 shared-memory staging rings (filled here by the same process the
    way actor processes fill them: plain memcpys, no CUDA), every poll moves all ready chunks with one pack kernel that
    reads the page-locked rings in place.
Prints one JSON line per configuration."""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import apex_b200  # noqa: E402

torch.cuda.set_device(0)
dev = torch.device("cuda:0")
C, F, nA, n = 4, 7056, 4, 3


def vector_push():
    for E, T in ((32, 8), (256, 8), (256, 64)):
        lane_cap = 8192
        mem = apex_b200.ReplayMemory(E * lane_cap, {"priority_exponent": 0.6, "layout": "stream", "lanes": E, "n_step": n,
                                                    "state_shape": (C, 84, 84)})
        ing = apex_b200.VectorActorIngest(mem, E, n_step=n, gamma=0.99, num_actions=nA)
        g = torch.Generator(device=dev).manual_seed(1)
        dev_in = (torch.randint(0, 256, (E, T, F), dtype=torch.uint8, device=dev, generator=g),
                  torch.randint(0, nA, (E, T), dtype=torch.int32, device=dev, generator=g),
                  torch.randn(E, T, device=dev, generator=g),
                  (torch.rand(E, T, device=dev, generator=g) < 0.01).to(torch.uint8),
                  torch.randn(E, T, nA, device=dev, generator=g))
        host_in = tuple(x.cpu().numpy() for x in dev_in)
        for name, args, reps in (("device inputs", dev_in, 200), ("host inputs (pageable numpy)", host_in, 20)):
            for _ in range(3):
                ing.push(*args)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(reps):
                ing.push(*args)
            host_dt = (time.perf_counter() - t0) / reps
            torch.cuda.synchronize()
            dt = (time.perf_counter() - t0) / reps
            print(json.dumps({"path": "apexb200_ingest_push", "actors": E, "env_steps_per_push": T, "inputs": name,
                              "ms_per_push": round(dt * 1e3, 4), "host_ms_per_push": round(host_dt * 1e3, 4),
                              "records_per_s": round(E * T / dt, 0), "frame_GBps": round(E * T * F / dt / 1e9, 2)}))
        # the same push replayed from a CUDA graph (heads and carry advance on the device)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.stream(side):
            with torch.cuda.graph(gr, stream=side):
                ing.push(*dev_in)
        torch.cuda.current_stream().wait_stream(side)
        for _ in range(3):
            gr.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(100):
            gr.replay()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 100
        ing.sync_heads()
        print(json.dumps({"path": "apexb200_ingest_push, CUDA graph replay", "actors": E, "env_steps_per_push": T,
                          "device_ms_per_push": round(ms, 4), "records_per_s": round(E * T / (ms * 1e-3), 0),
                          "frame_GBps_read_plus_write": round(2 * E * T * F / (ms * 1e-3) / 1e9, 1)}))
        mem.check_status()
        del ing, mem
        torch.cuda.empty_cache()


def ring_transport(E=256, T=8, slots=4, rounds=40):
    lane_cap = 4096
    mem = apex_b200.ReplayMemory(E * lane_cap, {"priority_exponent": 0.6, "layout": "stream", "lanes": E, "n_step": n,
                                                "state_shape": (C, 84, 84)})
    names = ["apexring_bw_%d_%d" % (os.getpid(), e) for e in range(E)]
    rings = [apex_b200.ActorStagingRing(nm, slots, T, F, nA, create=True) for nm in names]
    try:
        ingest = apex_b200.RingIngest(mem, rings, n_step=n)
        rng = np.random.default_rng(0)
        payload = (rng.integers(0, 256, (T, F), dtype=np.uint8), rng.integers(0, nA, T).astype(np.int32),
                   rng.standard_normal(T).astype(np.float32), rng.random(T) < 0.01, rng.standard_normal((T, nA)).astype(np.float32))

        def produce():                                    # what 256 actor processes do concurrently; not timed
            for r in rings:
                while r.push(*payload):
                    pass
        produce()
        ingest.poll()
        ingest.flush()
        polls, recs, host_s = 0, 0, 0.0
        torch.cuda.synchronize()
        dev_ms = 0.0
        for _ in range(rounds):
            produce()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0 = time.perf_counter()
            e0.record()
            got = ingest.poll()                           # `slots` chunks per ring: `slots` push_blobs calls
            e1.record()
            host_s += time.perf_counter() - t0
            ingest.flush()
            dev_ms += e0.elapsed_time(e1)
            recs += got
            polls += 1
        wall = max(host_s, dev_ms * 1e-3)
        print(json.dumps({"path": "RingIngest.poll -> apexb200_ingest_push_blobs", "actors": E, "env_steps_per_chunk": T,
                          "chunks_per_ring_per_poll": slots, "records_per_poll": recs // polls,
                          "host_ms_per_poll": round(host_s / polls * 1e3, 4), "device_ms_per_poll": round(dev_ms / polls, 4),
                          "records_per_s": round(recs / wall, 0), "pcie_GBps": round(recs * F / (dev_ms * 1e-3) / 1e9, 2),
                          "note": "rings are filled by this process between polls (not timed); device time from CUDA events "
                                  "around the poll's launches; records/s = records / max(host, device) time"}))
        mem.check_status()
    finally:
        torch.cuda.synchronize()
        for r in rings:
            r.close(unlink=True)


if __name__ == "__main__":
    vector_push()
    ring_transport()
