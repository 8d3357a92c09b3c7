"""CPU tests: the C-ABI library loads and exports every symbol include/apexb200.h
declares, the ctypes table covers them all, and the host-side mirror classes behave
like the reference classes (no CUDA calls here)."""
import ctypes
import json
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "apexb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(apexb200_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    import apex_b200
    names = _declared()
    assert len(names) >= 30
    lib = ctypes.CDLL(apex_b200._C.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), "libapexb200.so does not export " + n
    assert sorted(apex_b200._C.SIGNATURES) == names, "ctypes table and header disagree"
    assert apex_b200._C.lib().apexb200_version() == 100
    assert apex_b200._C.lib().apexb200_strerror(-4) == b"size beyond supported limit"


def test_argument_validation_without_a_gpu():
    """Pure host-side checks of the C ABI return codes (no kernel is launched)."""
    import apex_b200
    _C = apex_b200._C
    lib = _C.lib()
    n = ctypes.c_int64()
    assert lib.apexb200_td_workspace_bytes(4096, ctypes.byref(n)) == 0 and n.value == 512
    assert lib.apexb200_td_workspace_bytes(0, ctypes.byref(n)) == _C.EINVAL
    assert lib.apexb200_set_gather_variant(7) == _C.EINVAL
    assert b"unknown variant 7" in lib.apexb200_last_error()
    h = ctypes.c_void_p()
    cfg = _C.ReplayConfig(100, 3, 1, 0, 1, 7056, 0, 0.6, 0)
    assert lib.apexb200_replay_create(ctypes.byref(h), ctypes.byref(cfg)) == _C.EINVAL
    assert lib.apexb200_sample(None, None, 4, None, None, None, None, None) == _C.EINVAL
    # entry points added for ingest, weight publication and the batched actor policy
    assert lib.apexb200_replay_insert_lanes(None, None, 1, 1, None, None, None, None, None, None, None, None, None) == _C.EINVAL
    p = ctypes.c_void_p()
    assert lib.apexb200_params_create(0, None, 0, 0, ctypes.byref(p)) == _C.EINVAL and not p.value
    span = (_C.ParamSpan * 1)()
    span[0].tensor, span[0].flat_offset, span[0].bytes = 4096, 128, 64                   # offset not a multiple of 256
    assert lib.apexb200_params_create(0, ctypes.cast(span, ctypes.c_void_p), 1, 4096, ctypes.byref(p)) == _C.EINVAL
    assert b"256-byte aligned" in lib.apexb200_last_error()
    span[0].flat_offset, span[0].bytes = 0, 8192                                         # beyond flat_bytes
    assert lib.apexb200_params_create(0, ctypes.cast(span, ctypes.c_void_p), 1, 4096, ctypes.byref(p)) == _C.EINVAL
    assert lib.apexb200_params_destroy(None) == 0
    assert lib.apexb200_params_publish(None, None, None, 1, None) == _C.EINVAL
    assert lib.apexb200_params_load(None, None, 0, None) == _C.EINVAL
    assert lib.apexb200_actor_policy(None, None, None, None, None, 0, 0, 4, 4, None, None, None, None) == _C.EINVAL
    assert lib.apexb200_host_register(None, 0) == _C.EINVAL and lib.apexb200_host_unregister(None) == _C.EINVAL
    assert _C.HOST_STEP_SLOTS == 4
    # round-2 entry points: sharded groups, host-facing sharded step, device-side ingest, checkpoint packing
    g = ctypes.c_void_p()
    assert lib.apexb200_ingest_create(None, 1, None, 3, 0.99, 4, 0, ctypes.byref(g)) == _C.EINVAL and not g.value
    assert lib.apexb200_ingest_destroy(None) == 0
    assert lib.apexb200_ingest_push(None, None, 1, 1, None, None, None, None, None, None) == _C.EINVAL
    assert lib.apexb200_ingest_push_blobs(None, None, None, 1, 1, 0, 0, 0, 0, None) == _C.EINVAL
    assert lib.apexb200_ingest_sync_heads(None, None) == _C.EINVAL
    assert lib.apexb200_host_register_mapped(None, 0, ctypes.byref(g)) == _C.EINVAL
    assert lib.apexb200_peer_connect_local(None, 0, 1, None) == _C.EINVAL
    assert lib.apexb200_shard_push(None, None) == _C.EINVAL
    assert lib.apexb200_peer_wait_stats(None, None, 0, None) == _C.EINVAL
    assert lib.apexb200_host_step_sharded(None, 0, 0, 512, 512, None, 4, 0.0, 0, 1.0, None, None, None, None, None, None, None,
                                          None) == _C.EINVAL
    assert lib.apexb200_frames_pack_sizes(None, 0, 1, None, None) == _C.EINVAL
    assert lib.apexb200_frames_pack(None, 0, 1, None, None, None) == _C.EINVAL
    assert lib.apexb200_frames_unpack(None, 0, 1, None, None, None) == _C.EINVAL
    assert lib.apexb200_debug_timeline(None, 4) == _C.EINVAL
    assert (_C.FMT_U8, _C.FMT_F32) == (0, 1)


def test_product_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "ape-x-dqn_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".sh")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.replace("checker", ""), f + " mentions the oracle"


def test_replay_constructor_does_not_touch_cuda():
    import apex_b200
    mem = apex_b200.ReplayMemory(100000, {"priority_exponent": 0.6, "importance_sampling_exponent": 0.4})
    assert mem.size() == 0 and mem.alpha == 0.6 and mem._h is None
    mem.remove_to_fit()
    mem.update_sample_probabilities()
    mem.set_priorities({"a": 1.0})                     # unknown key: remembered (dict.update adds it, replay.py:40)
    assert mem._keys.pending == {"a": 1.0}
    import torch
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError, match="CUDA"):
            mem.sample(4)


def test_key_index_semantics():
    from apex_b200 import KeyIndex
    k = KeyIndex()
    k.assign("10", 0)
    k.assign("11", 1)
    assert k.split({"10": 2.0, "zz": 3.0}) == ([0], [2.0]) and k.pending == {"zz": 3.0}
    k.assign("12", 0)                                  # ring overwrote slot 0: key "10" is gone
    assert "10" not in k.slot_of and k.key_of[0] == "12"
    k.assign("11", 5)                                  # key collision: newest record owns the key
    assert k.slot_of["11"] == 5 and 1 not in k.key_of


def test_stream_offsets():
    from apex_b200 import stream_offsets
    rel = stream_offsets(4, [3, 3, 1], [100, 2, 0])
    assert rel[0].tolist() == [-6, -5, -4, -3, -3, -2, -1, 0]
    assert rel[1].tolist() == [-2, -2, -2, -2, -2, -2, -1, 0]      # clamped at the episode start
    assert rel[2].tolist() == [0] * 8


# ---- a10: ExperienceBuffer against the vectors generated from the reference class
def _feed(buf, case, Transition):
    for s in case["feed"]:
        buf.add(Transition(s["S"], s["A"], s["R"], s["Gamma"], s["q"]))
        if s["done"]:
            buf.construct_nstep_transition(Transition(1000 + s["S"], s["A"], 0, 0.99, s["q"]))


def test_experience_buffer_reference_semantics(golden_dir):
    from apex_b200 import ExperienceBuffer, Transition
    for case in json.load(open(os.path.join(golden_dir, "nstep_literal.json"))):
        buf = ExperienceBuffer(case["spec"]["n"], case["spec"]["actor_id"])
        _feed(buf, case, Transition)
        assert buf.size == len(case["emitted"])
        got = buf.get(buf.size)
        for g, w in zip(got, case["emitted"]):
            assert (g.S_t, g.A_t, g.S_tpn, g.key) == (w["S_t"], w["A_t"], w["S_tpn"], w["key"])
            assert g.R_ttpB == pytest.approx(w["R"], rel=1e-12) and g.Gamma_ttpB == pytest.approx(w["Gamma"], rel=1e-12)
        assert buf.size == 0 and buf.B in range(0, case["spec"]["n"] + 1)


def test_experience_buffer_sliding_semantics_matches_nstep_oracle():
    from apex_b200 import ExperienceBuffer, Transition
    from oracle.nstep_oracle import nstep_intended
    rng = np.random.default_rng(0)
    T, n = 40, 3
    r = rng.standard_normal(T)
    d = rng.random(T) < 0.15
    d[-1] = True
    buf = ExperienceBuffer(n, 2, semantics="sliding")
    for t in range(T):
        buf.add(Transition(t, t % 4, float(r[t]), 0.99, None))
        if d[t]:
            buf.construct_nstep_transition(Transition(t + 1, 0, 0, 0.99, None))
    recs = buf.get(buf.size)
    R, G, span = nstep_intended(r[None], d[None], n, 0.99, np.float64)
    assert [x.S_t for x in recs] == list(range(T))                  # one record per step
    np.testing.assert_allclose([x.R_ttpB for x in recs], R[0], rtol=1e-12)
    np.testing.assert_allclose([x.Gamma_ttpB for x in recs], G[0], rtol=1e-12)
    assert [x.S_tpn - x.S_t for x in recs] == span[0].tolist()
    with pytest.raises(AssertionError):
        buf.get(1)                                                  # actor.py:74


def test_learner_constructs_on_cpu_like_the_reference():
    import apex_b200
    shared = {}
    lp = {"remove_old_xp_freq": 100, "q_target_sync_freq": 2500, "min_replay_mem_size": 20000,
          "replay_sample_size": 32, "load_saved_state": False}
    L = apex_b200.Learner({"state_shape": [1, 84, 84], "action_dim": 4}, lp, shared, object())
    assert set(shared) == {"Q_state_dict"} and L.device is None and L.num_q_updates == 0
    assert sum(p.numel() for p in L.Q.parameters()) == 3321541
    g = L.optimizer.param_groups[0]
    assert (g["lr"], g["weight_decay"], g["eps"]) == (0.00025 / 4, 0.95, 1.5e-7)   # learner.py:26


def test_dropin_modules_and_manager_proxy():
    """main.py:18,40-42: ReplayMemory registered with a BaseManager and built in the server process;
    main.py:6-8: classes imported as replay.ReplayMemory / learner.Learner / actor.Actor."""
    import subprocess
    import sys
    code = r'''
import sys
sys.path.insert(0, %r); sys.path.insert(0, %r)
from multiprocessing.managers import BaseManager
from replay import ReplayMemory
from learner import Learner
from actor import Actor, ExperienceBuffer
BaseManager.register("Memory", ReplayMemory)
if __name__ == "__main__":
    m = BaseManager(); m.start()
    mem = m.Memory(100000, {"soft_capacity": 100000, "priority_exponent": 0.6, "importance_sampling_exponent": 0.4})
    assert mem.size() == 0
    mem.remove_to_fit(); mem.update_sample_probabilities(); mem.set_priorities({"10": 1.0})
    exposed = sorted(x for x in dir(mem) if not x.startswith("_"))
    for name in ("add", "remove_to_fit", "sample", "set_priorities", "size", "update_sample_probabilities"):
        assert name in exposed, name
    m.shutdown()
    print("proxy ok")
''' % (ROOT, os.path.join(ROOT, "ape-x-dqn_b200", "dropin"))
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0 and "proxy ok" in out.stdout, out.stderr[-2000:]


def _ring_producer(name, slots, T, F, nA, chunks, seed):
    import sys
    sys.path.insert(0, ROOT)
    import time
    import numpy as np
    import apex_b200
    ring = apex_b200.ActorStagingRing(name, slots, T, F, nA, create=False)
    rng = np.random.default_rng(seed)
    for c in range(chunks):
        payload = (rng.integers(0, 256, (T, F), dtype=np.uint8), rng.integers(0, nA, T).astype(np.int32),
                   rng.standard_normal(T).astype(np.float32), rng.random(T) < 0.1, rng.standard_normal((T, nA)).astype(np.float32))
        while not ring.push(*payload):
            time.sleep(0.001)
    ring.close()


def test_actor_staging_ring_cross_process():
    """Actor -> replay transport in shared host memory: a producer process writes chunks,
    this process reads exactly what was written, in order, with back-pressure when the ring is full."""
    import multiprocessing as mp
    import apex_b200
    slots, T, F, nA, chunks, seed = 3, 8, 64, 4, 11, 123
    name = "apexring_test_%d" % os.getpid()
    ring = apex_b200.ActorStagingRing(name, slots, T, F, nA, create=True)
    try:
        p = mp.get_context("spawn").Process(target=_ring_producer, args=(name, slots, T, F, nA, chunks, seed))
        p.start()
        rng = np.random.default_rng(seed)
        got = 0
        import time
        t0 = time.time()
        while got < chunks and time.time() - t0 < 60:
            chunk = ring.pop()
            if chunk is None:
                time.sleep(0.001)
                continue
            f, a, r, d, q = chunk
            np.testing.assert_array_equal(f, rng.integers(0, 256, (T, F), dtype=np.uint8))
            np.testing.assert_array_equal(a, rng.integers(0, nA, T).astype(np.int32))
            np.testing.assert_array_equal(r, rng.standard_normal(T).astype(np.float32))
            np.testing.assert_array_equal(d.astype(bool), rng.random(T) < 0.1)
            np.testing.assert_array_equal(q, rng.standard_normal((T, nA)).astype(np.float32))
            assert int(ring.ctrl[0]) - int(ring.ctrl[1]) <= slots
            ring.release()
            got += 1
        p.join(timeout=30)
        assert got == chunks and p.exitcode == 0 and ring.pop() is None
    finally:
        ring.close(unlink=True)


def test_header_is_plain_c(tmp_path):
    """include/apexb200.h is the drop-in boundary: it must compile as C99 (cgo / JNI / any FFI generator) and
    as C++, and a C program must link against the library with nothing but the header."""
    import shutil
    import subprocess
    if shutil.which("gcc") is None:
        pytest.skip("gcc not available")
    hdr = os.path.join(ROOT, "include", "apexb200.h")
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only", "-x", "c", hdr])
    subprocess.check_call(["g++", "-std=c++17", "-Wall", "-Werror", "-fsyntax-only", "-x", "c++", hdr])
    src = tmp_path / "probe.c"
    src.write_text('#include "apexb200.h"\n#include <stdio.h>\n'
                   'int main(void) { int64_t n = 0; int rc = apexb200_td_workspace_bytes(4096, &n);\n'
                   '  printf("%d %d %lld %s\\n", apexb200_version(), rc, (long long)n, apexb200_strerror(APEXB200_ERANGE));\n'
                   '  return apexb200_sample(0, 0, 4, 0, 0, 0, 0, 0) == APEXB200_EINVAL ? 0 : 1; }\n')
    import apex_b200
    lib = apex_b200._C.LIB_PATH
    exe = tmp_path / "probe"
    subprocess.check_call(["gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe), lib,
                           "-Wl,-rpath," + os.path.dirname(lib)])
    out = subprocess.run([str(exe)], capture_output=True, text=True, timeout=60)
    assert out.returncode == 0 and out.stdout.split()[:3] == ["100", "0", "512"], (out.stdout, out.stderr)
