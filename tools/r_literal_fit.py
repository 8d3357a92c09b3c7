#!/usr/bin/env python
"""R-literal of BASELINE.md §3: the UNMODIFIED /root/reference/replay.py timed in-process (single Python thread — it
cannot use more), sample(B) + set_priorities(B keys) at N in {1k, 2k, 4k, 8k}, with the fitted c*N^2 law
extrapolated to 100k / 2M records.  The reference tree only exists in the build container (not on the GPU box), so
this tool is run there; its output is committed as profiles/r02_r_literal_fit.json and quoted, labelled as such,
beside the bench lines.  Usage: PYTHONDONTWRITEBYTECODE=1 python tools/r_literal_fit.py [/root/reference]"""
import json
import os
import sys
import time

import numpy as np

ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
sys.dont_write_bytecode = True
sys.path.insert(0, ref)
import replay as ref_replay  # noqa: E402

B = 32                                   # parameters.json:23
rows = []
for N in (1000, 2000, 4000, 8000):
    rng = np.random.default_rng(N)
    mem = ref_replay.ReplayMemory(10 ** 9, {"priority_exponent": 0.6})
    state = np.zeros((1, 84, 84))        # shared reference: no copy cost (flatters the reference)
    xp = [ref_replay.N_Step_Transition(state, 0, 0.0, 0.97, None, state, None, str(k)) for k in range(N)]
    mem.memory.extend(xp)
    mem.priorities.update({str(k): np.float64(abs(rng.standard_normal()) + 1e-3) for k in range(N)})
    t0 = time.perf_counter()
    mem.update_sample_probabilities()
    t_first = time.perf_counter() - t0
    t0 = time.perf_counter()
    batch = mem.sample(B)
    t_s = time.perf_counter() - t0
    new = {x.key: np.float64(abs(rng.standard_normal()) + 1e-3) for x in batch}
    t0 = time.perf_counter()
    mem.set_priorities(new)
    t_u = time.perf_counter() - t0
    rows.append({"N": N, "sample_s": round(t_s, 4), "set_priorities_s": round(t_u, 4),
                 "transitions_per_s": round(B / (t_s + t_u), 3)})
    print(rows[-1], flush=True)
Ns = np.array([r["N"] for r in rows], float)
c = float(np.sum(np.array([r["set_priorities_s"] for r in rows]) * Ns ** 2) / np.sum(Ns ** 4))     # least squares, t = c N^2
cs = float(np.sum(np.array([r["sample_s"] for r in rows]) * Ns) / np.sum(Ns ** 2))                  # t = cs N
ext = {str(N): {"set_priorities_s": c * N * N, "sample_s": cs * N,
                "transitions_per_s_B32": B / (c * N * N + cs * N),
                "transitions_per_s_B512": 512 / (c * N * N + 16 * cs * N)} for N in (100_000, 2_097_152)}
out = {"what": "unmodified /root/reference/replay.py, in-process, one Python thread, B=32 (replay.py:18-57)", "host": os.uname().nodename,
       "cpu_count": os.cpu_count(), "measured": rows, "fit": {"set_priorities_s_per_N2": c, "sample_s_per_N": cs},
       "extrapolated": ext, "label": "build-container measurement; 100k / 2M rows are EXTRAPOLATED from the N^2 fit"}
print(json.dumps(out))
with open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "r02_r_literal_fit.json"), "w") as f:
    json.dump(out, f, indent=1)
