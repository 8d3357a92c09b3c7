"""ctypes wrapper of oracle/apex_oracle.c (TEST INFRASTRUCTURE / timed CPU baseline)."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(_HERE, "_build", "liboracle.so")
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            import subprocess
            subprocess.check_call(["bash", os.path.join(_HERE, "build_c.sh")])
        _lib = C.CDLL(LIB)
        _lib.oracle_step.restype = C.c_float
        _lib.oracle_td.restype = C.c_float
        _lib.oracle_max_threads.restype = C.c_int
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class CpuReplay:
    """Host-resident replay shard driven by the C oracle: the CPU baseline of bench.py."""

    def __init__(self, leaves, frames, fidx, A, R, G, C_, F, alpha, threads=0):
        self.leaves = np.ascontiguousarray(leaves, np.uint64)
        self.cum = np.empty_like(self.leaves)
        self.frames, self.fidx = frames, np.ascontiguousarray(fidx, np.int32)
        self.A, self.R, self.G = (np.ascontiguousarray(x, dt) for x, dt in ((A, np.int32), (R, np.float32), (G, np.float32)))
        self.C, self.F, self.alpha = int(C_), int(F), float(alpha)
        if threads:
            lib().oracle_set_threads(int(threads))
        self.threads = lib().oracle_max_threads()
        self._out = {}

    def _outputs(self, B):
        o = self._out.get(B)
        if o is None:
            o = dict(idx=np.empty(B, np.int32), S_t=np.empty((B, self.C, self.F), np.uint8),
                     S_tpn=np.empty((B, self.C, self.F), np.uint8), A=np.empty(B, np.int32), R=np.empty(B, np.float32),
                     G=np.empty(B, np.float32), delta=np.empty(B, np.float32), prio=np.empty(B, np.float32),
                     loss=np.empty(B, np.float32))
            self._out[B] = o
        return o

    def step(self, u, q_on_t, q_on_tpn, q_tg_tpn):
        B, nA = q_on_t.shape
        o = self._outputs(B)
        u = np.ascontiguousarray(u, np.float64)
        mean = lib().oracle_step(_p(self.leaves), _p(self.cum), C.c_int64(self.leaves.size), _p(u), C.c_int(B),
                                 _p(self.frames), _p(self.fidx), _p(self.A), _p(self.R), _p(self.G), C.c_int(self.C),
                                 C.c_int(self.F), _p(q_on_t), _p(q_on_tpn), _p(q_tg_tpn), C.c_int(nA),
                                 C.c_float(self.alpha), _p(o["idx"]), _p(o["S_t"]), _p(o["S_tpn"]), _p(o["A"]),
                                 _p(o["R"]), _p(o["G"]), _p(o["delta"]), _p(o["prio"]), _p(o["loss"]))
        o["loss_mean"] = float(mean)
        return o
