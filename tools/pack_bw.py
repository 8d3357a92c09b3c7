#!/usr/bin/env python
"""Throughput and ratio of the lossless checkpoint packing (apexb200_frames_pack*) on one GPU:
65 536 frames of 84x84 uint8, an Atari-like stream (each frame = its predecessor with a 16x16 sprite moved) and
incompressible random frames.  One JSON line per case."""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import apex_b200  # noqa: E402
from apex_b200 import _C  # noqa: E402

torch.cuda.set_device(0)
dev = torch.device("cuda:0")
N, F = 1 << 16, 7056
mem = apex_b200.ReplayMemory(N, {"priority_exponent": 0.6, "layout": "stream", "lanes": 1, "n_step": 3, "state_shape": (4, 84, 84)})
mem._ensure()
store = mem.buffer(_C.BUF_FRAMES).view(N, 84, 84)
g = torch.Generator(device=dev).manual_seed(0)
for kind in ("atari-like (16x16 sprite moves by a pixel per frame over a fixed background)", "random bytes"):
    if kind.startswith("random"):
        store.copy_(torch.randint(0, 256, (N, 84, 84), dtype=torch.uint8, device=dev, generator=g))
    else:
        bg = torch.randint(0, 256, (84, 84), dtype=torch.uint8, device=dev, generator=g)
        store.copy_(bg.expand(N, 84, 84))
        t = torch.arange(N, device=dev)
        x, y = (t % 68), ((t // 68) % 68)
        for dy in range(16):
            for dx in range(16):
                store[t, y + dy, x + dx] = 255
    orig = store.clone()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    sizes, offsets, packed = mem.pack_frames(0, N)
    torch.cuda.synchronize()
    t_pack = time.perf_counter() - t0
    store.zero_()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    mem.unpack_frames(offsets, packed, 0)
    torch.cuda.synchronize()
    t_unpack = time.perf_counter() - t0
    assert torch.equal(store, orig)
    print(json.dumps({"frames": N, "frame_bytes": F, "data": kind, "raw_MB": round(N * F / 1e6, 1),
                      "packed_MB": round(packed.numel() / 1e6, 2), "ratio": round(N * F / packed.numel(), 2),
                      "pack_ms": round(t_pack * 1e3, 2), "pack_GBps_of_raw": round(N * F / t_pack / 1e9, 1),
                      "unpack_ms": round(t_unpack * 1e3, 2), "unpack_GBps_of_raw": round(N * F / t_unpack / 1e9, 1),
                      "round_trip_exact": True}))
