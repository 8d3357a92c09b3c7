#!/usr/bin/env python
"""Weight publication: the reference's Manager-dict pickle (learner.py:74 + actor.py:189-191) against
the weight board (K9).  Prints one JSON line.  Run on the GPU box: python tools/publish_bench.py"""
import json
import multiprocessing as mp
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import apex_b200                                   # noqa: E402
from apex_b200.qnet import DuellingQNet            # noqa: E402


def reference_path(net_cpu, reps):
    """shared_state["Q_state_dict"] = Q.state_dict()  /  Q.load_state_dict(shared_state["Q_state_dict"])."""
    mgr = mp.Manager()
    shared = mgr.dict()
    actor_net = DuellingQNet((1, 84, 84), 4)
    put, get = [], []
    for _ in range(reps):
        t0 = time.perf_counter()
        shared["Q_state_dict"] = net_cpu.state_dict()
        t1 = time.perf_counter()
        actor_net.load_state_dict(shared["Q_state_dict"])
        t2 = time.perf_counter()
        put.append(t1 - t0)
        get.append(t2 - t1)
    mgr.shutdown()
    return float(np.median(put)), float(np.median(get))


def main():
    torch.manual_seed(0)
    dev = torch.device("cuda")
    net = DuellingQNet((1, 84, 84), 4).to(dev)
    ref_put, ref_get = reference_path(DuellingQNet((1, 84, 84), 4), 10)
    # the learner in the reference also pays a device->host state_dict copy when the net is on a GPU
    t0 = time.perf_counter()
    for _ in range(10):
        sd = {k: v.detach().cpu() for k, v in net.state_dict().items()}
    d2h_sd = (time.perf_counter() - t0) / 10
    name = "apexboard_bench_%d" % os.getpid()
    pub = apex_b200.WeightPublisher(net, name=name, create=True)
    out = {"flat_bytes": pub.flat_bytes, "tensors": len(pub.layout)}
    try:
        for _ in range(5):
            pub.snapshot()
        torch.cuda.synchronize()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        reps = 200
        ev[0].record()
        for _ in range(reps):
            pub.snapshot()
        ev[1].record()
        torch.cuda.synchronize()
        pack_us = ev[0].elapsed_time(ev[1]) * 1e3 / reps
        out["pack_kernel_us_back_to_back"] = round(pack_us, 2)
        out["pack_GBps"] = round(2 * pub.flat_bytes / pack_us / 1e3, 1)
        lat, call = [], []
        for _ in range(30):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            v = pub.publish()
            t1 = time.perf_counter()
            pub.wait(v)
            lat.append(time.perf_counter() - t0)
            call.append(t1 - t0)
        out["publish_call_us"] = round(float(np.median(call)) * 1e6, 1)          # what the learner thread pays
        out["publish_to_visible_us"] = round(float(np.median(lat)) * 1e6, 1)
        cpu_actor = DuellingQNet((1, 84, 84), 4)
        sub = apex_b200.WeightSubscriber(cpu_actor, name)
        polls = []
        for _ in range(10):
            pub.wait(pub.publish())
            t0 = time.perf_counter()
            assert sub.poll()
            polls.append(time.perf_counter() - t0)
        out["cpu_actor_poll_us"] = round(float(np.median(polls)) * 1e6, 1)
        t0 = time.perf_counter()
        for _ in range(1000):
            sub.poll()
        out["cpu_actor_poll_unchanged_us"] = round((time.perf_counter() - t0) * 1e3, 3)
        gpu_actor = DuellingQNet((1, 84, 84), 4).to(dev)
        dsub = apex_b200.DeviceWeightSubscriber(gpu_actor, name)
        polls = []
        for _ in range(10):
            pub.wait(pub.publish())
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            assert dsub.poll()
            torch.cuda.synchronize()
            polls.append(time.perf_counter() - t0)
        out["gpu_actor_poll_us"] = round(float(np.median(polls)) * 1e6, 1)
        sub.close()
        dsub.close()
    finally:
        pub.close(unlink=True)
    out["reference_put_ms"] = round(ref_put * 1e3, 2)
    out["reference_get_load_ms"] = round(ref_get * 1e3, 2)
    out["reference_state_dict_d2h_ms"] = round(d2h_sd * 1e3, 2)
    # batched policy: E actors in one forward + K10 against E batch-1 CPU forwards (actor.py:160-161)
    E = 256
    pol = apex_b200.BatchedActorPolicy(net, {"epsilon": 0.4, "alpha": 7, "num_actors": E}, range(E))
    obs = torch.randint(0, 256, (E, 1, 84, 84), dtype=torch.uint8, device=dev)
    for _ in range(5):
        pol.act(obs)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(50):
        a, q = pol.act(obs)
    a.cpu()
    out["batched_policy_256_actors_ms"] = round((time.perf_counter() - t0) / 50 * 1e3, 3)
    cpu_net = DuellingQNet((1, 84, 84), 4)
    torch.set_num_threads(1)
    x = obs[:1].cpu().float()
    with torch.no_grad():
        for _ in range(3):
            cpu_net(x)
        t0 = time.perf_counter()
        for _ in range(20):
            cpu_net(x)[2].squeeze().numpy()
    out["reference_policy_one_actor_ms"] = round((time.perf_counter() - t0) / 20 * 1e3, 3)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
