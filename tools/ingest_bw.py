#!/usr/bin/env python
"""Actor -> replay ingest throughput: VectorActorIngest.push (K7 + initial priority + multi-lane K6) for E lock-stepped
actors, inputs already on the device, and from host arrays.  Prints one JSON line per configuration."""
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
for E, T in ((32, 8), (256, 8), (256, 64)):
    lane_cap = 8192
    mem = apex_b200.ReplayMemory(E * lane_cap, {"priority_exponent": 0.6, "layout": "stream", "lanes": E, "n_step": n,
                                                "state_shape": (C, 84, 84)})
    ing = apex_b200.VectorActorIngest(mem, E, n_step=n, gamma=0.99)
    g = torch.Generator(device=dev).manual_seed(1)
    dev_in = (torch.randint(0, 256, (E, T, F), dtype=torch.uint8, device=dev, generator=g),
              torch.randint(0, nA, (E, T), dtype=torch.int32, device=dev, generator=g),
              torch.randn(E, T, device=dev, generator=g), np.zeros((E, T), bool),
              torch.randn(E, T, nA, device=dev, generator=g))
    host_in = tuple(x.cpu().numpy() if isinstance(x, torch.Tensor) else x for x in dev_in)
    for name, args, reps in (("device inputs", dev_in, 50), ("host inputs (pageable numpy)", host_in, 10)):
        for _ in range(3):
            ing.push(*args)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            ing.push(*args)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / reps
        print(json.dumps({"actors": E, "env_steps_per_push": T, "inputs": name, "ms_per_push": round(dt * 1e3, 3),
                          "records_per_s": round(E * T / dt, 0), "frame_GBps": round(E * T * F / dt / 1e9, 2)}))
    mem.check_status()
    del ing, mem
    torch.cuda.empty_cache()
