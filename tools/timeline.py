#!/usr/bin/env python
"""Where the time of one fused sample+gather launch goes: runs the instrumented kernel
(apexb200_set_gather_variant(3)) inside the step loop and prints, over all warps of the last launch, the
quantiles of each phase boundary (SM clock, us at 1.965 GHz, relative to the warp's first instruction).
With a library built with -DAPEX_TD_TIMELINE (APEXB200_LIB=...) it also prints the kernel boundaries around the
TD/update kernel from global-timer stamps."""
import argparse
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import apex_b200  # noqa: E402
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--capacity", type=int, default=1 << 21)
    ap.add_argument("--mhz", type=float, default=1965.0)
    a = ap.parse_args()
    args = argparse.Namespace(capacity=a.capacity, lanes=32, frames_per_state=4, actions=4, batch=a.batch)
    torch.cuda.set_device(0)
    dev = torch.device("cuda:0")
    mem = apex_b200.ReplayMemory(args.capacity, {"priority_exponent": 0.6, "layout": "stream", "lanes": 32, "n_step": 3,
                                                 "state_shape": (4, 84, 84), "seed": 1})
    bench.fill_shard(torch, apex_b200, mem, args, dev, 1234)
    B = a.batch
    q = [torch.randn(B, 4, device=dev) for _ in range(3)]
    out = tuple(torch.empty(B, device=dev) for _ in range(3)) + (torch.empty(1, device=dev),)
    apex_b200._C.call("apexb200_set_gather_variant", 3)

    def step():
        b = mem.sample_tensors(B)
        mem.td_update(b.idx, q[0], q[1], q[2], b.A, b.R, b.Gamma, out=out)
    modes = ("step loop, ends with sample", "step loop, ends with TD/update", "back-to-back sample launches")
    for mode in modes:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        g = torch.cuda.CUDAGraph()
        with torch.cuda.stream(side):
            with torch.cuda.graph(g, stream=side):
                for _ in range(8):
                    if mode.startswith("step"):
                        step()
                    else:
                        mem.sample_tensors(B)
                if mode.endswith("sample"):
                    mem.sample_tensors(B)
        torch.cuda.current_stream().wait_stream(side)
        for _ in range(5):
            g.replay()
        torch.cuda.synchronize()
        units = 2 * B
        buf = (C.c_uint64 * (units * 20))()
        apex_b200._C.call("apexb200_debug_timeline", buf, units)
        d = np.frombuffer(buf, dtype=np.uint64).reshape(units, 20).astype(np.int64)
        td = (C.c_uint64 * 8)()
        apex_b200._C.call("apexb200_debug_td_timeline", td)
        td = [int(x) for x in td]
        names = ["dependency wait done", "search done", "loads issued (first lane)", "first frame landed", "last frame landed",
                 "all stores issued", "staging free (reads done)", "writes complete"]
        print("== %s: %d warps, %d SMs" % (mode, units, len(set(d[:, 10].tolist()))))
        for j, nme in enumerate(names):
            v = d[:, 1 + j] / a.mhz
            print("  %-28s min %6.2f  p10 %6.2f  median %6.2f  p90 %6.2f  max %6.2f us" %
                  (nme, v.min(), np.percentile(v, 10), np.median(v), np.percentile(v, 90), v.max()))
        lv = ["draw formed", "top level(s) searched"] + ["level %d searched" % k for k in range(1, 6)]
        for j, nme in enumerate(lv):
            v = d[:, 12 + j] / a.mhz
            if v.max() > 0:
                print("    . %-26s min %6.2f  p10 %6.2f  median %6.2f  p90 %6.2f  max %6.2f us" %
                      (nme, v.min(), np.percentile(v, 10), np.median(v), np.percentile(v, 90), v.max()))
        print("    . frame slots guessed wrong: %d of %d" % (int((d[:, 11] >> 32).sum()), units * 4))
        g0, g9 = d[:, 0], d[:, 9]
        print("  global timer: entry spread %.2f us, first entry -> last exit %.2f us" %
              ((g0.max() - g0.min()) / 1e3, (g9.max() - g0.min()) / 1e3))
        if td[0] and mode.startswith("step"):
            print("  TD kernel body (thread 0, us after its dependency wait): rows evaluated %.2f, leaves written %.2f, ancestor "
                  "updates issued %.2f, loss reduced / end %.2f (last thread %.2f)" %
                  tuple((td[k] - td[1]) / 1e3 for k in (3, 4, 5, 2, 6)))
            wait_done = g0 + (d[:, 1] / a.mhz * 1e3).astype(np.int64)      # absolute time the sample warps were released
            if mode.endswith("sample"):      # ... TD(8) -> sample(9): stamps of the TD kernel that precedes this sample
                print("  TD kernel before this sample: entry -> wait done %.2f us, wait done -> thread 0 end %.2f us, "
                      "thread 0 end -> sample warps released %.2f us (median)" %
                      ((td[1] - td[0]) / 1e3, (td[2] - td[1]) / 1e3, (np.median(wait_done) - td[2]) / 1e3))
            else:                            # ... sample(8) -> TD(8): the TD kernel that follows this sample
                print("  TD kernel after this sample: last sample exit -> TD wait done %.2f us, TD wait done -> thread 0 end %.2f us" %
                      ((td[1] - g9.max()) / 1e3, (td[2] - td[1]) / 1e3))
    mem.check_status()


if __name__ == "__main__":
    main()
