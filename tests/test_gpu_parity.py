"""GPU parity tests (run on the B200 box with -m gpu): every CUDA entry point, called
through the C ABI (ctypes), against the CPU oracle on the same seeded inputs.

Bars: record indices bit-exact, gathered batches byte-exact, tree nodes bit-exact,
delta / priority / loss within 1e-5 relative (fp32).
"""
import ctypes as C
import json
import os

import numpy as np
import pytest

torch = pytest.importorskip("torch")

from oracle import nstep_oracle as no            # noqa: E402
from oracle import replay_oracle as ro           # noqa: E402
from oracle import td_oracle as to               # noqa: E402

pytestmark = pytest.mark.gpu

RTOL = 1e-5


@pytest.fixture(scope="module")
def apex():
    import apex_b200
    apex_b200._C.lib()
    torch.cuda.set_device(0)
    return apex_b200


def dev(x, dtype=None):
    t = torch.as_tensor(np.ascontiguousarray(x))
    if dtype is not None:
        t = t.to(dtype)
    return t.cuda()


def u64(t):
    return t.cpu().numpy().view(np.uint64) if t.dtype == torch.int64 else t.cpu().numpy()


def make_mem(apex, N, C=1, F=16, layout="materialised", lanes=1, n=3, nA=0, alpha=0.6):
    mem = apex.ReplayMemory(N, {"priority_exponent": alpha, "layout": layout, "lanes": lanes, "n_step": n,
                                "state_shape": (C, 1, F), "num_actions": nA})
    mem._ensure()
    return mem


def load_leaves(apex, mem, leaves):
    t = dev(np.asarray(leaves, np.uint64).view(np.int64))
    apex._C.call("apexb200_tree_load_leaves", mem._h, apex._C.ptr(t), 0, int(t.numel()), apex._C.stream_ptr())
    mem.update_sample_probabilities()          # K1


def read_levels(apex, mem):
    nl, sizes = C.c_int32(), (C.c_int64 * 8)()
    apex._C.call("apexb200_tree_geometry", mem._h, C.byref(nl), sizes)
    out = []
    for l in range(nl.value):
        t = torch.empty(sizes[l], dtype=torch.int64, device="cuda")
        apex._C.call("apexb200_tree_read", mem._h, l, 0, sizes[l], apex._C.ptr(t), apex._C.stream_ptr())
        out.append(u64(t))
    return out


# ------------------------------------------------------------------ K1 / K2 tree
@pytest.mark.parametrize("N", [1, 31, 32, 33, 1000, 1 << 15, (1 << 20) + 17])
def test_tree_build_bit_exact(apex, N):
    rng = np.random.default_rng(N)
    leaves = rng.integers(0, ro.LEAF_MAX, N, dtype=np.uint64)
    leaves[rng.random(N) < 0.3] = 0
    mem = make_mem(apex, N)
    load_leaves(apex, mem, leaves)
    got, want = read_levels(apex, mem), ro.build_levels(leaves)
    assert [len(x) for x in got] == ro.level_sizes(N)
    for g, w in zip(got, want):
        np.testing.assert_array_equal(g, w)
    assert mem.total() == int(leaves.sum(dtype=np.uint64))


@pytest.mark.parametrize("N,count", [(1000, 7), (1000, 1024), (1 << 16, 1025), (1 << 16, 4096), (1 << 21, 512)])
def test_tree_update_last_writer_wins_and_matches_rebuild(apex, N, count):
    rng = np.random.default_rng(count)
    leaves = rng.integers(0, 1 << 30, N, dtype=np.uint64)
    mem = make_mem(apex, N)
    load_leaves(apex, mem, leaves)
    ring_leaves = leaves.copy()
    for rep in range(3):
        idx = rng.integers(0, N, count).astype(np.int32)
        idx[: count // 3] = idx[count // 3: 2 * (count // 3)][: count // 3]      # force duplicates
        vals = rng.integers(0, 1 << 34, count, dtype=np.uint64)
        mem.set_leaves(dev(idx), dev(vals.view(np.int64)))
        for i, v in zip(idx.tolist(), vals.tolist()):
            ring_leaves[i] = v                                                    # sequential = last writer wins
        got = read_levels(apex, mem)
        for g, w in zip(got, ro.build_levels(ring_leaves)):
            np.testing.assert_array_equal(g, w)
    mem.check_status()


def test_tree_update_priority_conversion(apex):
    rng = np.random.default_rng(2)
    N, count = 5000, 700
    mem = make_mem(apex, N, alpha=0.6)
    idx = rng.permutation(N)[:count].astype(np.int32)
    td = (np.abs(rng.standard_normal(count)) * 3).astype(np.float32)
    td[:5] = [0.0, 1.0, 1e-8, 1e7, 2.5]
    mem.update_priorities(dev(idx), dev(td))
    got = u64(mem.buffer(apex._C.BUF_LEAVES))[idx].astype(np.float64)
    want = ro.priority_to_leaf(ro.pow_priority(td, 0.6)).astype(np.float64)
    np.testing.assert_allclose(got, want, rtol=RTOL, atol=2)
    assert got[0] == 0 and got[1] == 1 << 24 and got[3] == ro.LEAF_MAX
    # the leaf is an exact function of the fp32 priority the kernel computed
    levels = read_levels(apex, mem)
    np.testing.assert_array_equal(levels[-1].sum(dtype=np.uint64), levels[0].sum(dtype=np.uint64))


# ---------------------------------------------------------------- K3 descent
@pytest.mark.parametrize("N", [1, 5, 32, 33, 1024, 1000, 40000, 1 << 20, (1 << 21) + 3])
@pytest.mark.parametrize("strat", [False, True])
def test_sample_indices_bit_exact(apex, N, strat):
    rng = np.random.default_rng(N + strat)
    pri = np.abs(rng.standard_normal(N)) + 1e-3
    leaves = ro.priority_to_leaf(ro.pow_priority(pri, 0.6))
    leaves[rng.random(N) < 0.2] = 0
    if leaves.sum() == 0:
        leaves[0] = 5
    mem = make_mem(apex, N)
    load_leaves(apex, mem, leaves)
    B = 2048
    u = ro.stratified_uniforms(B, rng) if strat else rng.random(B)
    u[:4] = [0.0, 1.0 - 2 ** -53, 0.5, 2 ** -60]
    leaf_out = torch.empty(B, dtype=torch.int64, device="cuda")
    total_out = torch.empty(1, dtype=torch.int64, device="cuda")
    idx = mem.sample_indices(dev(u), leaf_out=leaf_out, total_out=total_out).cpu().numpy()
    want = ro.tree_sample(leaves, u)
    np.testing.assert_array_equal(idx, want)
    np.testing.assert_array_equal(u64(leaf_out), leaves[want])
    assert int(u64(total_out)[0]) == int(leaves.sum(dtype=np.uint64))
    assert (leaves[idx] > 0).all()
    mem.check_status()


def test_sample_matches_reference_choice_law(apex, golden_dir):
    """Same priorities, same uniform draws as np.random.choice consumed (replay.py:52)."""
    z = np.load(os.path.join(golden_dir, "choice_law.npz"))
    pri = z["priorities"].astype(np.float32)
    mem = make_mem(apex, len(pri), alpha=float(z["alpha"]))
    mem.update_priorities(dev(np.arange(len(pri), dtype=np.int32)), dev(pri))
    idx = mem.sample_indices(dev(z["u"])).cpu().numpy()
    assert int((idx != z["picks"]).sum()) == 0


def test_stratified_draws_generated_in_kernel(apex):
    rng = np.random.default_rng(21)
    N, B = 50000, 512
    leaves = ro.priority_to_leaf(ro.pow_priority(np.abs(rng.standard_normal(N)) + 1e-3, 0.6))
    mem = make_mem(apex, N)
    mem.seed = 0xC0FFEE
    load_leaves(apex, mem, leaves)
    epoch = 0                                                # tree-changing kernels run on this handle so far
    for call in range(4):                                    # host call index separates draws between updates
        u_out = torch.empty(B, dtype=torch.float64, device="cuda")
        if call % 2:
            idx = mem.sample_tensors(B, u_out=u_out).idx.cpu().numpy()         # fused K3+K4
        else:
            idx = mem.sample_stratified(B, u_out=u_out).cpu().numpy()
        u = u_out.cpu().numpy()
        np.testing.assert_array_equal(u, ro.kernel_stratified_uniforms(0xC0FFEE, epoch, call, B))
        assert ((u >= np.arange(B) / B) & (u < (np.arange(B) + 1) / B)).all()
        np.testing.assert_array_equal(idx, ro.tree_sample(leaves, u))
        if call == 1:                                        # a priority update advances the device epoch
            mem.update_priorities(dev(idx[:7].astype(np.int32)), dev(np.ones(7, np.float32)))
            leaves[idx[:7]] = 1 << 24
            epoch += 1
    mem.check_status()


@pytest.mark.parametrize("pdl", [1, 0])
def test_step_loop_in_cuda_graph_equals_eager(apex, pdl):
    """sample+gather -> TD+update chained on one stream, eager vs replayed from a CUDA graph:
    same in-kernel draws (device call counter), same indices, same tree afterwards."""
    apex._C.call("apexb200_set_pdl", pdl)
    try:
        N, C_, F, B, nA, steps = 1 << 15, 4, 7056, 256, 4, 6

        def build():
            rng = np.random.default_rng(77)
            mem = make_mem(apex, N, C=C_, F=F, layout="stream", lanes=4, n=3)
            for lane in range(4):
                m = N // 4 - 8
                mem.insert(lane, rng.integers(0, 256, (m, F), dtype=np.uint8), apex.stream_offsets(C_, np.full(m, 3), np.arange(m)),
                           rng.integers(0, nA, m).astype(np.int32), rng.standard_normal(m).astype(np.float32),
                           np.full(m, 0.97, np.float32), prio=(np.abs(rng.standard_normal(m)) + 1e-3).astype(np.float32))
            return mem
        torch.manual_seed(5)
        q = [[torch.randn(B, nA, device="cuda") for _ in range(3)] for _ in range(steps)]
        out = tuple(torch.empty(B, device="cuda") for _ in range(3)) + (torch.empty(1, device="cuda"),)

        def run(mem, k, log):
            b = mem.sample_tensors(B)
            mem.td_update(b.idx, q[k][0], q[k][1], q[k][2], b.A, b.R, b.Gamma, out=out)
            log[k].copy_(b.idx)
        eager, graphed = build(), build()
        log_e = torch.zeros(steps, B, dtype=torch.int32, device="cuda")
        log_g = torch.zeros_like(log_e)
        for k in range(steps):
            run(eager, k, log_e)
        run(graphed, 0, log_g)                               # warm-up outside capture (allocations)
        graphed2 = build()
        graphed2.sample_tensors(B); graphed2.workspace(B)    # allocate outputs before capture
        # rewind: a fresh handle so that the draw counter starts at zero again
        graphed2 = build()
        graphed2._outputs(B, False); graphed2.workspace(B)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        g = torch.cuda.CUDAGraph()
        with torch.cuda.stream(side):
            with torch.cuda.graph(g, stream=side):
                for k in range(steps // 2):
                    run(graphed2, k, log_g)
        torch.cuda.current_stream().wait_stream(side)
        g.replay()
        torch.cuda.synchronize()
        # capture does not execute: the replay is steps 0..2 of the sequence
        np.testing.assert_array_equal(log_g.cpu().numpy()[: steps // 2], log_e.cpu().numpy()[: steps // 2])
        ref = build()
        for k in range(steps // 2):
            run(ref, k, torch.zeros_like(log_e))
        for a, b_ in zip(read_levels(apex, graphed2), read_levels(apex, ref)):
            np.testing.assert_array_equal(a, b_)
        graphed2.check_status()
    finally:
        apex._C.call("apexb200_set_pdl", 1)


@pytest.mark.parametrize("N", [3000, 2048, 50000, 40])       # 2048 / 50000: the two top levels are searched as one
def test_sample_targets_and_device_count(apex, N):
    rng = np.random.default_rng(9)
    leaves = rng.integers(1, 1 << 28, N, dtype=np.uint64)
    mem = make_mem(apex, N)
    load_leaves(apex, mem, leaves)
    total = int(leaves.sum(dtype=np.uint64))
    targets = rng.integers(0, total, 500, dtype=np.uint64)
    idx = torch.full((500,), -7, dtype=torch.int32, device="cuda")
    cnt = dev(np.array([321], np.int32))
    apex._C.call("apexb200_sample_targets", mem._h, apex._C.ptr(dev(targets.view(np.int64))), apex._C.ptr(cnt), 500,
                 apex._C.ptr(idx), None, None, apex._C.stream_ptr())
    got = idx.cpu().numpy()
    np.testing.assert_array_equal(got[:321], ro.descend(leaves, targets[:321]))
    assert (got[321:] == -7).all()


def test_empty_tree_sets_status(apex):
    mem = make_mem(apex, 100)
    idx = mem.sample_indices(dev(np.array([0.3, 0.6])))
    assert (idx.cpu().numpy() == -1).all()
    with pytest.raises(apex._C.ApexError, match="empty tree"):
        mem.check_status()
    mem.check_status()            # cleared


# ------------------------------------------------------- K6 insert + K4 gather
@pytest.mark.parametrize("variant", [0, 1])
@pytest.mark.parametrize("C_,F", [(1, 7056), (4, 7056), (4, 64)])
def test_stream_insert_and_gather_byte_exact(apex, variant, C_, F):
    rng = np.random.default_rng(10 * C_ + variant)
    n, lanes, N, nA = 3, 4, 4 * 96, 4
    W = C_ - 1 + n
    mem = make_mem(apex, N, C=C_, F=F, layout="stream", lanes=lanes, n=n, nA=nA)
    ring = ro.RingOracle(N, C_, F, lanes=lanes, fpr=1, lookback=W, alpha=0.6, num_actions=nA)
    apex._C.call("apexb200_set_gather_variant", variant)
    try:
        age0 = np.zeros(lanes, np.int64)
        for it in range(14):                                   # 14 * 20 = 280 records per lane of 96: wraps twice
            for lane in range(lanes):
                m = 20 if it % 3 else 11
                frames = rng.integers(0, 256, (m, F), dtype=np.uint8)
                age = age0[lane] + np.arange(m)
                reset = rng.integers(0, m)
                age[reset:] -= age[reset]                      # an episode boundary inside the chunk
                age0[lane] = age[-1] + 1
                span = rng.integers(1, n + 1, m)
                rel = apex.stream_offsets(C_, span, age)
                A = rng.integers(0, nA, m).astype(np.int32)
                R = rng.standard_normal(m).astype(np.float32)
                G = np.where(rng.random(m) < 0.2, 0, 0.97).astype(np.float32)
                q_t, q_tpn = rng.standard_normal((m, nA)).astype(np.float32), rng.standard_normal((m, nA)).astype(np.float32)
                if it % 2:
                    prio = (np.abs(rng.standard_normal(m)) + 0.01).astype(np.float32)
                    s1 = mem.insert(lane, frames, rel, A, R, G, prio=prio, q_t=q_t, q_tpn=q_tpn)
                    s2 = ring.insert(lane, frames, rel, A, R, G, prio=prio, q_t=q_t, q_tpn=q_tpn)
                else:                                          # priorities from the actor's q-values (actor.py:138-142)
                    s1 = mem.insert(lane, frames, rel, A, R, G, q_t=q_t, q_tpn=q_tpn)
                    s2 = ring.insert(lane, frames, rel, A, R, G, q_t=q_t, q_tpn=q_tpn)
                np.testing.assert_array_equal(s1, s2)
        assert mem.size() == ring.size()
        np.testing.assert_array_equal(mem.buffer(apex._C.BUF_FRAMES).cpu().numpy().reshape(-1, F), ring.frames)
        np.testing.assert_array_equal(mem.buffer(apex._C.BUF_FIDX).cpu().numpy().reshape(N, 2 * C_), ring.fidx)
        leaves = u64(mem.buffer(apex._C.BUF_LEAVES))
        np.testing.assert_allclose(leaves.astype(np.float64), ring.leaves.astype(np.float64), rtol=1e-5, atol=4)
        assert ((leaves == 0) == (ring.leaves == 0)).all()     # dead zone behind each head
        for g, w in zip(read_levels(apex, mem), ro.build_levels(leaves)):
            np.testing.assert_array_equal(g, w)
        ring.leaves[:] = leaves
        for B, fused in ((1, True), (33, False), (512, True), (512, False), (700, True)):
            u = rng.random(B)
            b = mem.sample_tensors(B, u=dev(u), with_q=True, fused=fused)
            want = ring.sample(u)
            np.testing.assert_array_equal(b.idx.cpu().numpy(), want)
            g = ring.gather(want)
            np.testing.assert_array_equal(b.fsel.cpu().numpy(), ring.fidx[want])
            np.testing.assert_array_equal(b.S_t.cpu().numpy(), g["S_t"])
            np.testing.assert_array_equal(b.S_tpn.cpu().numpy(), g["S_tpn"])
            np.testing.assert_array_equal(b.A.cpu().numpy(), g["A"])
            np.testing.assert_array_equal(b.R.cpu().numpy(), g["R"])
            np.testing.assert_array_equal(b.Gamma.cpu().numpy(), g["G"])
            np.testing.assert_array_equal(mem._last_q[0].cpu().numpy(), g["q_t"])
            np.testing.assert_array_equal(mem._last_q[1].cpu().numpy(), g["q_tpn"])
            # without the pre-resolved frame slots (one more dependent lookup) the bytes are the same
            b2 = mem.gather(b.idx.clone(), fsel=None, out={k: (torch.empty_like(v) if v is not None else None)
                                                           for k, v in mem._outputs(B, True).items()})
            np.testing.assert_array_equal(b2.S_t.cpu().numpy(), g["S_t"])
            np.testing.assert_array_equal(b2.S_tpn.cpu().numpy(), g["S_tpn"])
        mem.check_status()
    finally:
        apex._C.call("apexb200_set_gather_variant", 0)


@pytest.mark.parametrize("C_,F,B", [(4, 7056, 96), (1, 7056, 33), (3, 208, 700), (8, 64, 17)])
def test_float32_batches_equal_widened_uint8_batches(apex, C_, F, B):
    """apexb200_sample_gather with APEXB200_FMT_F32 (learner.py:37-38: the model is handed float32 states):
    for the same draws the float32 batch equals the uint8 batch converted by .float(), element for element,
    and the uint8 batch equals the oracle's gather byte for byte; same for the sharded kernel (group of one)."""
    rng = np.random.default_rng(C_ * 1000 + B)
    N, lanes = 4096, 2
    mem = make_mem(apex, N, C=C_, F=F, layout="stream", lanes=lanes, n=3)
    ring = ro.RingOracle(N, C_, F, lanes=lanes, fpr=1, lookback=C_ - 1 + 3, alpha=0.6)
    for lane in range(lanes):
        m = N // lanes - (C_ + 2)
        args = (lane, rng.integers(0, 256, (m, F), dtype=np.uint8), apex.stream_offsets(C_, np.full(m, 3), np.arange(m) % 211),
                rng.integers(0, 4, m).astype(np.int32), rng.standard_normal(m).astype(np.float32), np.full(m, 0.97, np.float32))
        prio = (np.abs(rng.standard_normal(m)) + 1e-3).astype(np.float32)
        mem.insert(*args, prio=prio)
        ring.insert(*args, prio=prio)
    ring.leaves[:] = u64(mem.buffer(apex._C.BUF_LEAVES))
    u = rng.random(B)
    want = ring.gather(ring.sample(u))
    b8 = mem.sample_tensors(B, u=dev(u))
    s8, n8, idx8 = b8.S_t.clone(), b8.S_tpn.clone(), b8.idx.clone()
    np.testing.assert_array_equal(s8.cpu().numpy(), want["S_t"])
    np.testing.assert_array_equal(n8.cpu().numpy(), want["S_tpn"])
    bf = mem.sample_tensors(B, u=dev(u), f32=True)
    assert bf.S_t.dtype == torch.float32 and bf.S_t.shape == (B, C_, F)
    assert torch.equal(bf.idx, idx8)
    assert torch.equal(bf.S_t, s8.float()) and torch.equal(bf.S_tpn, n8.float())
    np.testing.assert_array_equal(bf.R.cpu().numpy(), want["R"])
    # sharded kernel, group of one: same rows in float32 and uint8 for the same in-kernel draws
    sampler = apex.ShardedSampler(mem, rank=0, world_size=1)
    sampler.connect_peers()
    sampler.exchange()
    from apex_b200 import _C
    o8, of = mem._outputs(B, False), mem._outputs(B, False, True)
    for o, fmt in ((o8, _C.FMT_U8), (of, _C.FMT_F32)):
        cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
        u_out = torch.zeros(B, dtype=torch.float64, device="cuda")
        _C.call("apexb200_sample_gather_sharded", mem._h, 77, B, B, _C.ptr(o["idx"]), _C.ptr(o["fsel"]), None, _C.ptr(u_out),
                _C.ptr(cnt), _C.ptr(o["S_t"]), _C.ptr(o["S_tpn"]), _C.ptr(o["A"]), _C.ptr(o["R"]), _C.ptr(o["Gamma"]), None, None,
                fmt, _C.stream_ptr())
        assert int(cnt.item()) == B
        w = ring.gather(ring.sample(u_out.cpu().numpy()))
        got = o["S_t"].cpu().numpy()
        np.testing.assert_array_equal(got, w["S_t"].astype(got.dtype))
        got = o["S_tpn"].cpu().numpy()
        np.testing.assert_array_equal(got, w["S_tpn"].astype(got.dtype))
    mem.check_status()


@pytest.mark.parametrize("seed", range(12))
def test_random_configurations(apex, seed):
    """Ragged shapes: random frames-per-state (odd, 1, the fused kernels' limit 8 and beyond it), frame
    sizes, lane counts, layouts and batch sizes; insert -> sample -> gather -> fused update against the oracle."""
    rng = np.random.default_rng(1000 + seed)
    C_ = int(rng.choice([1, 2, 3, 4, 5, 8, 9]))
    F = int(rng.choice([16, 48, 112, 7056])) if C_ < 8 else int(rng.choice([16, 48]))
    stream = bool(rng.integers(2))
    lanes = int(rng.choice([1, 3, 8]))
    n = int(rng.integers(1, 5))
    nA = int(rng.choice([0, 2, 18]))
    lane_cap = int(rng.integers(40, 90))
    N = lanes * lane_cap
    fpr, W = (1, C_ - 1 + n) if stream else (2 * C_, 0)
    mem = make_mem(apex, N, C=C_, F=F, layout="stream" if stream else "materialised", lanes=lanes, n=n, nA=nA)
    ring = ro.RingOracle(N, C_, F, lanes=lanes, fpr=fpr, lookback=W, alpha=0.6, num_actions=nA)
    for it in range(int(rng.integers(3, 9))):
        lane = int(rng.integers(lanes))
        m = int(rng.integers(1, lane_cap - W))
        frames = rng.integers(0, 256, (m * fpr, F), dtype=np.uint8)
        if stream:
            rel = apex.stream_offsets(C_, rng.integers(1, n + 1, m), rng.integers(0, 2 * W + 2, m))
        else:
            rel = np.tile(np.arange(-(2 * C_ - 1), 1, dtype=np.int32), (m, 1))
        A = rng.integers(0, max(nA, 1), m).astype(np.int32)
        R, G = rng.standard_normal(m).astype(np.float32), rng.random(m).astype(np.float32)
        prio = (np.abs(rng.standard_normal(m)) * (rng.random(m) > 0.2)).astype(np.float32)      # some zero priorities
        q = (rng.standard_normal((m, nA)).astype(np.float32), rng.standard_normal((m, nA)).astype(np.float32)) if nA else (None, None)
        s1 = mem.insert(lane, frames, rel, A, R, G, prio=prio, q_t=q[0], q_tpn=q[1])
        s2 = ring.insert(lane, frames, rel, A, R, G, prio=prio, q_t=q[0], q_tpn=q[1])
        np.testing.assert_array_equal(s1, s2)
    leaves = u64(mem.buffer(apex._C.BUF_LEAVES))
    ring.leaves[:] = leaves
    if leaves.sum() == 0:
        return
    for B in (1, int(rng.integers(2, 700))):
        u = rng.random(B)
        b = mem.sample_tensors(B, u=dev(u), with_q=bool(nA))
        want = ring.sample(u)
        np.testing.assert_array_equal(b.idx.cpu().numpy(), want)
        g = ring.gather(want)
        np.testing.assert_array_equal(b.S_t.cpu().numpy(), g["S_t"])
        np.testing.assert_array_equal(b.S_tpn.cpu().numpy(), g["S_tpn"])
        np.testing.assert_array_equal(b.A.cpu().numpy(), g["A"])
        np.testing.assert_array_equal(b.Gamma.cpu().numpy(), g["G"])
        if nA:
            np.testing.assert_array_equal(mem._last_q[1].cpu().numpy(), g["q_tpn"])
        nq = max(nA, 3)
        qs = [rng.standard_normal((B, nq)).astype(np.float32) for _ in range(3)]
        acts = dev((g["A"] % nq).astype(np.int32))
        _, prio_d, _, _ = mem.td_update(b.idx, dev(qs[0]), dev(qs[1]), dev(qs[2]), acts, b.R, b.Gamma)
        ring.set_leaves(want, ro.priority_to_leaf(prio_d.cpu().numpy()))
        for gl, w in zip(read_levels(apex, mem), ro.build_levels(ring.leaves)):
            np.testing.assert_array_equal(gl, w)
    mem.check_status()


def test_compat_api_matches_reference_semantics(apex, golden_dir):
    """add / sample / set_priorities / size with reference-shaped records (replay.py:32-83)."""
    z = np.load(os.path.join(golden_dir, "choice_law.npz"))
    K = 600
    pri = z["priorities"][:K]
    NT = apex.N_Step_Transition
    mem = apex.ReplayMemory(K, {"priority_exponent": float(z["alpha"])})
    rng = np.random.default_rng(4)
    S = rng.integers(0, 256, (K + 1, 1, 84, 84)).astype(np.float64)            # reference stores float64 (Q20)
    recs = [NT(S[i], i % 4, float(i), 0.97, None, S[i + 1], None, "k%d" % i) for i in range(K)]
    for c in range(0, K, 5):                                                   # parameters.json:15 batches of 5
        chunk = recs[c:c + 5]
        mem.add({r.key: np.float64(pri[c + j]) for j, r in enumerate(chunk)}, chunk)
    assert mem.size() == K
    np.random.seed(99)
    u = np.random.random_sample(256)
    want = ro.reference_law_sample(pri, float(z["alpha"]), u)                  # fp64 np.random.choice law
    np.random.seed(99)
    got = mem.sample(256)
    assert [r.key for r in got] == ["k%d" % i for i in want]
    for r, i in zip(got, want):
        assert r.S_t.dtype == np.uint8 and r.S_t.shape == (1, 84, 84)
        np.testing.assert_array_equal(r.S_t, S[i].astype(np.uint8))
        np.testing.assert_array_equal(r.S_tpn, S[i + 1].astype(np.uint8))
        assert (r.A_t, r.R_ttpB) == (i % 4, float(i)) and r.Gamma_ttpB == pytest.approx(0.97)
    # overwrite priorities: only k7 and k9 keep mass
    mem.set_priorities({"k%d" % i: 0.0 for i in range(K)})
    mem.set_priorities({"k7": 1.0, "k9": 3.0, "k7 ": 2.0})                     # unknown key is remembered, not an error
    keys = {r.key for r in mem.sample(64)}
    assert keys == {"k7", "k9"}
    # ring eviction: 5 more records overwrite k0..k4 and their priorities
    more = [NT(S[i], 0, 0.0, 0.0, None, S[i], None, "new%d" % i) for i in range(5)]
    mem.add({r.key: 1e6 for r in more}, more)
    assert mem.size() == K
    assert {r.key for r in mem.sample(64)} <= {"k7", "k9"} | {r.key for r in more}
    mem.remove_to_fit()
    mem.update_sample_probabilities()
    assert mem.size() == K


def test_compat_add_sample_equals_reference_on_equal_priorities(apex, golden_dir):
    """End to end against the reference's own ReplayMemory.add + sample (replay_literal.json,
    'equal_priority_sample'): with equal priorities its literal (uniform) law and the intended one
    coincide, so the same seed must return the same records in the same order."""
    with open(os.path.join(golden_dir, "replay_literal.json")) as f:
        g = json.load(f)["equal_priority_sample"]
    NT = apex.N_Step_Transition
    mem = apex.ReplayMemory(100, {"priority_exponent": 0.6})
    frame = lambda v: np.full((1, 84, 84), v % 256, np.uint8)           # the state id, as an image
    recs = [NT(frame(i), i % 4, float(i) / 7, 0.97, None, frame(i + 1), None, "e%d" % i) for i in range(g["n"])]
    for c in range(0, g["n"], 5):
        mem.add({r.key: np.float64(2.5) for r in recs[c:c + 5]}, recs[c:c + 5])
    np.random.seed(g["seed"])
    got = mem.sample(len(g["keys"]))
    assert [r.key for r in got] == g["keys"]
    assert [r.A_t for r in got] == g["A"]
    np.testing.assert_allclose([r.R_ttpB for r in got], g["R"], rtol=1e-6)
    assert [int(r.S_t[0, 0, 0]) for r in got] == g["S_t"] and [int(r.S_tpn[0, 0, 0]) for r in got] == g["S_tpn"]


def test_checkpoint_round_trip(apex):
    rng = np.random.default_rng(12)
    mem = make_mem(apex, 256, C=2, F=32, layout="stream", lanes=2, n=2)
    for lane in range(2):
        m = 100
        mem.insert(lane, rng.integers(0, 256, (m, 32), dtype=np.uint8), apex.stream_offsets(2, np.full(m, 2), np.arange(m)),
                   np.zeros(m, np.int32), np.ones(m, np.float32), np.ones(m, np.float32), prio=np.ones(m, np.float32))
    sd = mem.state_dict()
    mem2 = apex.ReplayMemory(256, {"priority_exponent": 0.6, "layout": "stream", "lanes": 2, "n_step": 2})
    mem2.load_state_dict(sd)
    u = dev(rng.random(64))
    a, b = mem.sample_tensors(64, u=u), mem2.sample_tensors(64, u=u)
    assert torch.equal(a.idx, b.idx) and torch.equal(a.S_t, b.S_t) and torch.equal(a.S_tpn, b.S_tpn)
    assert mem2.size() == mem.size() and mem2.total() == mem.total()


def _sprite_frames(rng, count, F, change=40):
    """Atari-like stream: every frame equals its predecessor except for a short run of bytes."""
    fr = np.zeros((count, F), np.uint8)
    fr[0] = rng.integers(0, 256, F)
    for j in range(1, count):
        fr[j] = fr[j - 1]
        if j % 17:                                                       # some frames do not change at all
            k = int(rng.integers(0, max(1, F - change)))
            fr[j, k:k + change] = rng.integers(0, 256, min(change, F - k))
    return fr


@pytest.mark.parametrize("F,count,first", [(7056, 200, 0), (64, 65, 3), (16, 1, 0), (528, 130, 64), (7056, 64, 128)])
def test_frame_packing_equals_oracle_and_round_trips(apex, F, count, first):
    """apexb200_frames_pack_sizes / _pack / _unpack ((f).4, lossless checkpoint packing) against oracle/pack_oracle.py:
    sizes and packed bytes identical, unpack restores the store byte for byte (also over random, incompressible
    frames)."""
    from oracle import pack_oracle as po
    rng = np.random.default_rng(F + count)
    N = 512
    mem = make_mem(apex, N, C=1, F=F, layout="stream", lanes=1, n=1)
    store = mem.buffer(apex._C.BUF_FRAMES).view(N, F)
    for kind in ("sprites", "random"):
        fr = _sprite_frames(rng, N, F) if kind == "sprites" else rng.integers(0, 256, (N, F), dtype=np.uint8)
        store.copy_(dev(fr))
        sizes, offsets, packed = mem.pack_frames(first, count)
        w_sizes, w_off, w_packed = po.pack_frames(fr[first:first + count])
        np.testing.assert_array_equal(sizes.cpu().numpy().view(np.uint32), w_sizes)
        np.testing.assert_array_equal(offsets.cpu().numpy().astype(np.uint64), w_off)
        np.testing.assert_array_equal(packed.cpu().numpy(), w_packed)
        if kind == "sprites" and F >= 528 and count >= 64:
            assert packed.numel() * 4 < count * F                        # a sprite stream shrinks well below a quarter
        store[first:first + count].zero_()
        mem.unpack_frames(offsets, packed, first)
        assert torch.equal(store, dev(fr))
    mem.check_status()


def test_checkpoint_round_trip_with_packed_frames(apex):
    rng = np.random.default_rng(13)
    C_, F, lanes, cap = 2, 528, 2, 512
    mem = make_mem(apex, cap, C=C_, F=F, layout="stream", lanes=lanes, n=2)
    for lane in range(lanes):
        m = 200
        mem.insert(lane, _sprite_frames(rng, m, F), apex.stream_offsets(C_, np.full(m, 2), np.arange(m)),
                   np.zeros(m, np.int32), np.ones(m, np.float32), np.ones(m, np.float32),
                   prio=(np.abs(rng.standard_normal(m)) + 0.1).astype(np.float32))
    sd = mem.state_dict(pack_frames=True)
    assert sd["frames"] is None and sum(p.numel() for _, _, p in sd["frames_packed"]) * 3 < cap * F
    mem2 = apex.ReplayMemory(cap, {"priority_exponent": 0.6, "layout": "stream", "lanes": lanes, "n_step": 2})
    mem2.load_state_dict(sd)
    assert torch.equal(mem.buffer(apex._C.BUF_FRAMES), mem2.buffer(apex._C.BUF_FRAMES))
    u = dev(rng.random(64))
    a, b = mem.sample_tensors(64, u=u), mem2.sample_tensors(64, u=u)
    assert torch.equal(a.idx, b.idx) and torch.equal(a.S_t, b.S_t) and torch.equal(a.S_tpn, b.S_tpn)
    assert mem2.size() == mem.size() and mem2.total() == mem.total()


# ------------------------------------------------------------------ K5 TD / loss
@pytest.mark.parametrize("B", [1, 32, 512, 1024, 1025, 4096, 10000])
@pytest.mark.parametrize("kind", [to.LOSS_HALF_SQ, to.LOSS_HUBER])
@pytest.mark.parametrize("nA", [4, 18])
def test_td_forward_backward(apex, B, kind, nA):
    rng = np.random.default_rng(B + kind + nA)
    q = [rng.standard_normal((B, nA)).astype(np.float32) * 2 for _ in range(3)]
    q[1][::7, 1] = q[1][::7].max(axis=1) + 0.0                 # ties: first maximum must win
    A = rng.integers(0, nA, B).astype(np.int32)
    R = np.clip(rng.standard_normal(B), -1, 1).astype(np.float32)
    G = np.where(rng.random(B) < 0.05, 0, 0.970299).astype(np.float32)
    o = to.td_forward(q[0], q[1], q[2], A, R, G, alpha=0.6, loss_kind=kind, kappa=1.0)
    delta, prio, loss, mean = apex.ops.td_forward(dev(q[0]), dev(q[1]), dev(q[2]), dev(A), dev(R), dev(G), 0.6, 0.0, kind, 1.0)
    np.testing.assert_allclose(delta.cpu().numpy(), o["delta"], rtol=RTOL, atol=1e-6)
    np.testing.assert_allclose(loss.cpu().numpy(), o["loss"], rtol=RTOL, atol=1e-7)
    np.testing.assert_allclose(prio.cpu().numpy(), o["prio"], rtol=RTOL, atol=1e-7)
    np.testing.assert_allclose(mean.item(), o["loss_mean"], rtol=RTOL)
    gq = apex.ops.td_backward(delta, dev(A), dev(np.array([1.5], np.float32)), nA, kind, 1.0)
    np.testing.assert_allclose(gq.cpu().numpy(), to.td_backward(o["delta"], A, nA, 1.5, kind, 1.0), rtol=RTOL, atol=1e-9)
    # determinism of the fixed-order mean
    again = apex.ops.td_forward(dev(q[0]), dev(q[1]), dev(q[2]), dev(A), dev(R), dev(G), 0.6, 0.0, kind, 1.0)[3]
    assert again.item() == mean.item()


def test_td_kernel_matches_reference_learner_function(apex, golden_dir):
    """K5 against the numbers the reference's Learner.compute_loss_and_priorities produced
    (tests/golden/td_reference.npz, generated by executing learner.py:29-52)."""
    z = np.load(os.path.join(golden_dir, "td_reference.npz"))
    for c in range(int(z["cases"])):
        q = [dev(z["c%d_%s" % (c, k)].astype(np.float32)) for k in ("q_on_t", "q_on_tpn", "q_tg_tpn")]
        delta, prio, loss, mean = apex.ops.td_forward(q[0], q[1], q[2], dev(z["c%d_A" % c]), dev(z["c%d_R" % c].astype(np.float32)),
                                                      dev(z["c%d_Gamma" % c].astype(np.float32)), 0.6)
        np.testing.assert_allclose(delta.abs().cpu().numpy(), z["c%d_abs_td" % c], rtol=RTOL, atol=1e-6)
        np.testing.assert_allclose(mean.item(), z["c%d_loss_mean" % c], rtol=RTOL)
        np.testing.assert_allclose(prio.cpu().numpy(), z["c%d_abs_td" % c] ** 0.6, rtol=1e-4, atol=1e-6)


def test_importance_weights_and_weighted_loss(apex):
    rng = np.random.default_rng(61)
    N, B, nA, beta = 20000, 512, 4, 0.4
    leaves = ro.priority_to_leaf(ro.pow_priority(np.abs(rng.standard_normal(N)) + 1e-3, 0.6))
    mem = make_mem(apex, N)
    load_leaves(apex, mem, leaves)
    u = rng.random(B)
    leaf_out = torch.empty(B, dtype=torch.int64, device="cuda")
    total_out = torch.empty(1, dtype=torch.int64, device="cuda")
    idx = mem.sample_indices(dev(u), leaf_out=leaf_out, total_out=total_out).cpu().numpy()
    w = apex.ops.is_weights(leaf_out, total_out, N, beta)
    want_w = to.is_weights(leaves[idx], int(leaves.sum(dtype=np.uint64)), N, beta)
    np.testing.assert_allclose(w.cpu().numpy(), want_w, rtol=1e-5)
    assert w.max().item() == 1.0
    q = [rng.standard_normal((B, nA)).astype(np.float32) for _ in range(3)]
    A = rng.integers(0, nA, B).astype(np.int32)
    R, G = rng.standard_normal(B).astype(np.float32), np.full(B, 0.97, np.float32)
    t1 = torch.tensor(q[0], device="cuda", requires_grad=True)
    loss, delta, prio, per = apex.td_loss(t1, dev(q[1]), dev(q[2]), dev(A), dev(R), dev(G), alpha=0.6, weights=w)
    loss.backward()
    o = to.td_forward(q[0], q[1], q[2], A, R, G, alpha=0.6, weights=want_w)
    np.testing.assert_allclose(per.cpu().numpy(), o["loss"], rtol=RTOL, atol=1e-7)
    np.testing.assert_allclose(loss.item(), o["loss_mean"], rtol=RTOL)
    np.testing.assert_allclose(prio.cpu().numpy(), o["prio"], rtol=RTOL, atol=1e-7)       # priorities are unweighted
    np.testing.assert_allclose(t1.grad.cpu().numpy(), to.td_backward(o["delta"], A, nA, weights=want_w), rtol=RTOL, atol=1e-9)


def test_td_loss_autograd_matches_torch(apex):
    torch.manual_seed(0)
    B, nA = 300, 6
    q1 = torch.randn(B, nA, device="cuda", requires_grad=True)
    q2, q3 = torch.randn(B, nA, device="cuda"), torch.randn(B, nA, device="cuda")
    A = torch.randint(0, nA, (B,), device="cuda", dtype=torch.int32)
    R, G = torch.randn(B, device="cuda"), torch.full((B,), 0.97, device="cuda")
    loss, delta, prio, _ = apex.td_loss(q1, q2, q3, A, R, G, alpha=0.6)
    (loss * 2).backward()
    r1 = q1.detach().clone().requires_grad_(True)
    with torch.no_grad():                                                          # learner.py:43-45
        tgt = R + G * q3.gather(1, torch.argmax(q2, 1).view(-1, 1)).squeeze()
    qsa = r1.gather(1, A.long().view(-1, 1)).squeeze()                             # learner.py:46
    ref = (0.5 * (tgt - qsa) ** 2).mean()                                          # learner.py:48,52
    (ref * 2).backward()
    torch.testing.assert_close(loss, ref, rtol=RTOL, atol=1e-7)
    torch.testing.assert_close(q1.grad, r1.grad, rtol=RTOL, atol=1e-9)
    torch.testing.assert_close(prio, (tgt - qsa).abs().pow(0.6), rtol=RTOL, atol=1e-7)


@pytest.mark.parametrize("cluster", [1, 0])
@pytest.mark.parametrize("B", [64, 100, 512, 1024, 3000])
def test_td_update_writes_priorities_in_place(apex, B, cluster):
    apex._C.call("apexb200_set_td_cluster", cluster)
    try:
        _td_update_case(apex, B, sorted_idx=False)
        _td_update_case(apex, B, sorted_idx=True)
    finally:
        apex._C.call("apexb200_set_td_cluster", 0)


def test_out_of_range_action_is_clamped_consistently_and_reported(apex):
    """An action outside [0, num_actions) is clamped the same way by the forward and the backward kernel (loss and
    gradient refer to the same column) and the handle-bound call raises the status bit."""
    rng = np.random.default_rng(12)
    B, nA, N = 64, 4, 1024
    mem = make_mem(apex, N)
    load_leaves(apex, mem, rng.integers(1, 1 << 20, N, dtype=np.uint64))
    q = [dev(rng.standard_normal((B, nA)).astype(np.float32)) for _ in range(3)]
    A = rng.integers(0, nA, B).astype(np.int32)
    A[5], A[9] = 7, -2                                                            # clamped to nA-1 and 0
    R, G = dev(rng.standard_normal(B).astype(np.float32)), dev(np.full(B, 0.9, np.float32))
    idx = dev(np.sort(rng.integers(0, N, B)).astype(np.int32))
    delta, prio, loss, mean = mem.td_update(idx, q[0], q[1], q[2], dev(A), R, G)
    o = to.td_forward(*(x.cpu().numpy() for x in q), np.clip(A, 0, nA - 1), R.cpu().numpy(), G.cpu().numpy(), alpha=0.6)
    np.testing.assert_allclose(delta.cpu().numpy(), o["delta"], rtol=RTOL, atol=1e-6)
    g = apex.ops.td_backward(delta, dev(A), torch.ones(1, device="cuda"), nA).cpu().numpy()
    np.testing.assert_allclose(g, to.td_backward(o["delta"], np.clip(A, 0, nA - 1), nA), rtol=RTOL, atol=1e-7)
    with pytest.raises(apex._C.ApexError, match="action"):
        mem.check_status()


def _td_update_case(apex, B, sorted_idx):
    rng = np.random.default_rng(B)
    N, nA = 1 << 16, 4
    mem = make_mem(apex, N)
    leaves = rng.integers(1, 1 << 26, N, dtype=np.uint64)
    load_leaves(apex, mem, leaves)
    idx = rng.integers(0, N, B).astype(np.int32)
    idx[1::9] = idx[0]                                                             # duplicates: last row wins
    if sorted_idx:
        idx = np.sort(idx)                                                         # stratified draws: adjacent duplicates
    q = [rng.standard_normal((B, nA)).astype(np.float32) for _ in range(3)]
    A = rng.integers(0, nA, B).astype(np.int32)
    R, G = rng.standard_normal(B).astype(np.float32), np.full(B, 0.97, np.float32)
    delta, prio, loss, mean = mem.td_update(dev(idx), dev(q[0]), dev(q[1]), dev(q[2]), dev(A), dev(R), dev(G))
    o = to.td_forward(q[0], q[1], q[2], A, R, G, alpha=0.6)
    np.testing.assert_allclose(prio.cpu().numpy(), o["prio"], rtol=RTOL, atol=1e-7)
    np.testing.assert_allclose(mean.item(), o["loss_mean"], rtol=RTOL)
    new = ro.priority_to_leaf(prio.cpu().numpy())                                  # exact function of the kernel's fp32 prio
    for i, v in zip(idx.tolist(), new.tolist()):
        leaves[i] = v
    for g, w in zip(read_levels(apex, mem), ro.build_levels(leaves)):
        np.testing.assert_array_equal(g, w)
    mem.check_status()


# ---------------------------------------------------------------- K7 / a11 / K8
def test_nstep_kernel(apex):
    rng = np.random.default_rng(3)
    for (E, T, n) in [(1, 6, 3), (7, 50, 3), (64, 129, 5), (3, 10, 1)]:
        r = rng.standard_normal((E, T)).astype(np.float32)
        d = (rng.random((E, T)) < 0.1)
        R, G, span = apex.ops.nstep_returns(dev(r), dev(d.astype(np.uint8)), n, 0.99)
        wR, wG, wspan = no.nstep_intended(r, d, n, 0.99)
        np.testing.assert_array_equal(span.cpu().numpy(), wspan)
        np.testing.assert_allclose(R.cpu().numpy(), wR, rtol=RTOL, atol=1e-7)
        np.testing.assert_allclose(G.cpu().numpy(), wG, rtol=RTOL, atol=0)


def test_initial_priority_against_reference_vectors(apex, golden_dir):
    with open(os.path.join(golden_dir, "initial_priority.json")) as f:
        cases = json.load(f)
    for c in cases:
        got = apex.ops.initial_priority(dev(np.array(c["q_t"], np.float32)), dev(np.array(c["q_tpn"], np.float32)),
                                        dev(np.array(c["A"], np.int32)), dev(np.array(c["R"], np.float32)),
                                        dev(np.array(c["Gamma"], np.float32))).cpu().numpy()
        np.testing.assert_allclose(got, c["per_row_abs_td"], rtol=RTOL, atol=1e-6)


def test_shard_routing_equals_single_tree(apex):
    """§8(e) on one GPU: G shard handles + K8 routing == one tree over the concatenated leaves."""
    rng = np.random.default_rng(8)
    G_, n_per, B = 4, 5000, 1024
    shards = [rng.integers(0, 1 << 28, n_per, dtype=np.uint64) for _ in range(G_)]
    shards[2][:] = 0                                                               # an empty shard owns no draw
    mems = []
    for s in shards:
        m = make_mem(apex, n_per)
        load_leaves(apex, m, s)
        mems.append(m)
    totals = torch.stack([torch.tensor(m.total() - (1 << 64) if m.total() >= 1 << 63 else m.total()) for m in mems]).cuda()
    u = rng.random(B)
    whole = ro.tree_sample(np.concatenate(shards), u)
    t, shard, local_t = ro.shard_split([int(s.sum(dtype=np.uint64)) for s in shards], u)
    got = np.full(B, -1, np.int64)
    for g in range(G_):
        lt, lp, lc = apex.ops.shard_targets(totals, g, dev(u))
        cnt = int(lc.item())
        assert cnt == int((shard == g).sum())
        np.testing.assert_array_equal(lp.cpu().numpy()[:cnt], np.nonzero(shard == g)[0])
        np.testing.assert_array_equal(u64(lt)[:cnt], local_t[shard == g])
        idx = torch.full((B,), -1, dtype=torch.int32, device="cuda")
        apex._C.call("apexb200_sample_targets", mems[g]._h, apex._C.ptr(lt), apex._C.ptr(lc), B, apex._C.ptr(idx), None,
                     None, apex._C.stream_ptr())
        got[lp.cpu().numpy()[:cnt]] = g * n_per + idx.cpu().numpy()[:cnt]
    np.testing.assert_array_equal(got, whole)


@pytest.mark.parametrize("N", [1 << 13, 1 << 11, 1 << 16])   # 2048 / 65536: merged top levels in the sharded kernel
def test_peer_exchange_and_fused_sharded_sample_world1(apex, N):
    """The peer-memory exchange and the fused routing+search+gather kernel with a group of one
    (the multi-rank behaviour is checked by tests/multi_gpu_check.py under torchrun)."""
    rng = np.random.default_rng(15)
    C_, F, B = 2, 256, 300
    mem = make_mem(apex, N, C=C_, F=F, layout="stream", lanes=2, n=3)
    ring = ro.RingOracle(N, C_, F, lanes=2, fpr=1, lookback=C_ - 1 + 3, alpha=0.6)
    for lane in range(2):
        m = N // 2 - 8
        args = (lane, rng.integers(0, 256, (m, F), dtype=np.uint8), apex.stream_offsets(C_, np.full(m, 3), np.arange(m)),
                rng.integers(0, 4, m).astype(np.int32), rng.standard_normal(m).astype(np.float32), np.full(m, 0.97, np.float32))
        prio = (np.abs(rng.standard_normal(m)) + 1e-3).astype(np.float32)
        mem.insert(*args, prio=prio)
        ring.insert(*args, prio=prio)
    ring.leaves[:] = u64(mem.buffer(apex._C.BUF_LEAVES))
    sampler = apex.ShardedSampler(mem, rank=0, world_size=1)
    sampler.connect_peers()
    totals = torch.zeros(1, dtype=torch.int64, device="cuda")
    for rnd in range(1, 4):
        sampler.exchange(totals)
        assert int(u64(totals)[0]) == ring.total()
        for max_rows in (B, B + 37):
            u_out = torch.full((max_rows,), -1.0, dtype=torch.float64, device="cuda")
            b, pos, cnt = sampler.sample_fused(B, max_rows, u_out=u_out)
            assert int(cnt.item()) == B
            np.testing.assert_array_equal(pos.cpu().numpy()[:B], np.arange(B))
            u = u_out.cpu().numpy()[:B]
            assert ((u >= np.arange(B) / B) & (u <= (np.arange(B) + 1) / B)).all()
            want = ring.sample(u)
            np.testing.assert_array_equal(b.idx.cpu().numpy()[:B], want)
            g = ring.gather(want)
            np.testing.assert_array_equal(b.S_t.cpu().numpy()[:B], g["S_t"])
            np.testing.assert_array_equal(b.S_tpn.cpu().numpy()[:B], g["S_tpn"])
            np.testing.assert_array_equal(b.R.cpu().numpy()[:B], g["R"])
        # default row capacity = the global batch: nothing can be dropped
        b, pos, cnt = sampler.sample_fused(B)
        assert int(cnt.item()) == B and b.S_t.shape[0] == B
        # output arrays smaller than the batch are a caller-side limit: reported, not silent
        b, pos, cnt = sampler.sample_fused(B, B - 50)
        assert int(cnt.item()) == B - 50
        with pytest.raises(apex._C.ApexError, match="quota"):
            mem.check_status()
        mem.update_priorities(dev(np.arange(5, dtype=np.int32)), dev(np.full(5, 2.0 * rnd, np.float32)))
        ring.set_priorities(np.arange(5), np.full(5, 2.0 * rnd, np.float32))
        ring.leaves[:5] = u64(mem.buffer(apex._C.BUF_LEAVES))[:5]
    mem.check_status()


@pytest.mark.parametrize("skew", [1.0, 8.0, 200.0])
def test_sharded_group_on_one_gpu_returns_every_draw_whatever_the_skew(apex, skew):
    """§8(e) with four shards driven in lock step from this process (apexb200_peer_connect_local): the fused
    sharded kernel + the push from the TD/update kernel.  Shard 0 carries `skew` times the priority mass of
    shards 1 and 2 and shard 3 is empty, so shard 0 owns far more than its fair share of the global batch
    (several waves inside one launch, several passes of the single-CTA TD kernel).  Over all shards the rows
    must be exactly the records one tree over the concatenated leaves returns for the same draws (indices
    bit-exact, batches byte-exact, no draw lost or duplicated), the loss parts must add up to the global mean,
    and every tree level must equal a rebuild after the updates."""
    rng = np.random.default_rng(int(skew) + 77)
    _run_sharded_group(apex, rng, G_=4, n=4096, C_=2, F=256, lanes=2, Bg=1024, scales=[skew ** (1 / 0.6), 1.0, 1.0, 0.0],
                       steps=3, expect_heavy_first=skew >= 8)


@pytest.mark.parametrize("seed", range(6))
def test_sharded_group_random_configurations(apex, seed):
    """The same checks over random group sizes, shard sizes, frame shapes, batch sizes and priority splits
    (including shards without any mass)."""
    rng = np.random.default_rng(1000 + seed)
    G_ = int(rng.choice([2, 3, 5, 7]))
    scales = [float(x) for x in rng.choice([0.0, 0.3, 1.0, 1.0, 4.0, 60.0], G_)]
    if not any(scales):
        scales[int(rng.integers(G_))] = 1.0
    _run_sharded_group(apex, rng, G_=G_, n=int(rng.choice([512, 2048, 4096])), C_=int(rng.choice([1, 2, 3])),
                       F=int(rng.choice([64, 208])), lanes=int(rng.choice([1, 2, 4])), Bg=int(rng.choice([64, 300, 1000])),
                       scales=scales, steps=2, expect_heavy_first=False)


def _run_sharded_group(apex, rng, G_, n, C_, F, lanes, Bg, scales, steps, expect_heavy_first):
    nA = 4
    W = C_ - 1 + 3
    mems, rings, samplers = [], [], []
    for g in range(G_):
        mem = make_mem(apex, n, C=C_, F=F, layout="stream", lanes=lanes, n=3)
        mem.seed = 1234                                                        # identical draws on every shard
        ring = ro.RingOracle(n, C_, F, lanes=lanes, fpr=1, lookback=W, alpha=0.6)
        for lane in range(lanes):
            m = n // lanes - W
            args = (lane, rng.integers(0, 256, (m, F), dtype=np.uint8),
                    apex.stream_offsets(C_, np.full(m, 3), np.minimum(np.arange(m), 300)),
                    rng.integers(0, nA, m).astype(np.int32), rng.standard_normal(m).astype(np.float32),
                    np.where(rng.random(m) < 0.05, 0, 0.97).astype(np.float32))
            prio = ((np.abs(rng.standard_normal(m)) + 1e-2) * scales[g]).astype(np.float32)
            mem.insert(*args, prio=prio)
            ring.insert(*args, prio=prio)
        ring.leaves[:] = u64(mem.buffer(apex._C.BUF_LEAVES))
        mems.append(mem)
        rings.append(ring)
        samplers.append(apex.ShardedSampler(mem, rank=g, world_size=G_))
    for s_ in samplers:
        s_.connect_local(samplers)
        s_.push_on_update(True)
    for s_ in samplers:
        s_.push()                                                              # round 1, nobody waits
    rnd = 1
    for step in range(steps):
        concat = np.concatenate([r.leaves for r in rings])
        seen = np.zeros(Bg, np.int64)
        u_glob = np.zeros(Bg)
        got_idx = np.full(Bg, -1, np.int64)
        outs = []
        for g in range(G_):                                                    # all samples first ...
            u_out = torch.full((Bg,), -1.0, dtype=torch.float64, device="cuda")
            b, pos, cnt = samplers[g].sample_fused(Bg, u_out=u_out)
            c = int(cnt.item())
            pos_h = pos.cpu().numpy()[:c]
            seen[pos_h] += 1
            u_glob[pos_h] = u_out.cpu().numpy()[:c]
            got_idx[pos_h] = g * n + b.idx.cpu().numpy()[:c]
            assert (np.diff(pos_h) == 1).all() or c <= 1                       # a rank owns one contiguous run of draws
            outs.append((b, pos_h, c, {k: getattr(b, k)[:c].clone() for k in ("idx", "S_t", "S_tpn", "A", "R", "Gamma")}))
        assert (seen == 1).all(), "every global draw must be owned by exactly one shard"
        np.testing.assert_array_equal(u_glob, ro.kernel_stratified_uniforms(1234, rnd, step, Bg))
        whole = ro.tree_sample(concat, u_glob)
        np.testing.assert_array_equal(got_idx, whole)
        for g in range(G_):
            if scales[g] == 0.0:
                assert outs[g][2] == 0                                         # a shard without mass owns nothing
        if expect_heavy_first:
            assert outs[0][2] > Bg // 2                                        # far beyond one wave of the fair share
        qrng = np.random.default_rng(100 + step)
        qg = [qrng.standard_normal((Bg, nA)).astype(np.float32) for _ in range(3)]
        loss_sum, A_g, R_g, G_g = 0.0, np.zeros(Bg, np.int32), np.zeros(Bg, np.float32), np.zeros(Bg, np.float32)
        for g in range(G_):                                                    # ... then all updates (they push round + 1)
            b, pos_h, c, rows = outs[g]
            want = rings[g].gather(rows["idx"].cpu().numpy())
            for k, wk in (("S_t", "S_t"), ("S_tpn", "S_tpn"), ("A", "A"), ("R", "R"), ("Gamma", "G")):
                np.testing.assert_array_equal(rows[k].cpu().numpy(), want[wk])
            A_g[pos_h], R_g[pos_h], G_g[pos_h] = want["A"], want["R"], want["G"]
            q = [dev(np.concatenate([x[pos_h], np.zeros((Bg - c, nA), np.float32)])) for x in qg]
            cnt = torch.tensor([c], dtype=torch.int32, device="cuda")
            delta, prio, loss, mean = mems[g].td_update(b.idx, q[0], q[1], q[2], b.A, b.R, b.Gamma, count_dev=cnt,
                                                        mean_divisor=float(Bg))
            loss_sum += float(mean.item())
            if c:
                o = to.td_forward(qg[0][pos_h], qg[1][pos_h], qg[2][pos_h], want["A"], want["R"], want["G"], alpha=0.6)
                np.testing.assert_allclose(prio.cpu().numpy()[:c], o["prio"], rtol=RTOL, atol=1e-7)
                np.testing.assert_allclose(delta.cpu().numpy()[:c], o["delta"], rtol=RTOL, atol=1e-6)
            rings[g].set_leaves(rows["idx"].cpu().numpy(), ro.priority_to_leaf(prio.cpu().numpy()[:c]))
            for gl, w in zip(read_levels(apex, mems[g]), ro.build_levels(rings[g].leaves)):
                np.testing.assert_array_equal(gl, w)
        o = to.td_forward(qg[0], qg[1], qg[2], A_g, R_g, G_g, alpha=0.6)
        np.testing.assert_allclose(loss_sum, o["loss_mean"], rtol=1e-5)
        rnd += 1
    for m_ in mems:
        m_.check_status()                                                      # no quota / timeout / index bit anywhere


@pytest.mark.parametrize("sorted_idx", [True, False])
def test_td_update_with_more_live_rows_than_threads(apex, sorted_idx):
    """Sharded call shape of K5+K2: arrays sized for a whole global batch, the live count on the device and
    several times the rank's fair share -> the single-CTA kernel takes several passes (duplicates across
    pass boundaries included); rows beyond the live count are ignored."""
    rng = np.random.default_rng(5 + sorted_idx)
    N, nA, rows, live, Bg = 1 << 15, 4, 3000, 2477, 256
    mem = make_mem(apex, N)
    leaves = rng.integers(1, 1 << 26, N, dtype=np.uint64)
    load_leaves(apex, mem, leaves)
    idx = rng.choice(rng.integers(0, N, 600), rows).astype(np.int32)               # ~4 rows per record: runs cross the
    idx[310:330] = idx[330]                                                        # pass boundaries (320 threads per pass)
    if sorted_idx:
        idx[:live] = np.sort(idx[:live])
    q = [rng.standard_normal((rows, nA)).astype(np.float32) for _ in range(3)]
    A = rng.integers(0, nA, rows).astype(np.int32)
    R, G = rng.standard_normal(rows).astype(np.float32), np.full(rows, 0.97, np.float32)
    cnt = torch.tensor([live], dtype=torch.int32, device="cuda")
    out = tuple(torch.full((k,), -7.0, device="cuda") for k in (rows, rows, rows, 1))
    delta, prio, loss, mean = mem.td_update(dev(idx), dev(q[0]), dev(q[1]), dev(q[2]), dev(A), dev(R), dev(G), out=out,
                                            count_dev=cnt, mean_divisor=float(Bg))
    o = to.td_forward(q[0][:live], q[1][:live], q[2][:live], A[:live], R[:live], G[:live], alpha=0.6)
    np.testing.assert_allclose(prio.cpu().numpy()[:live], o["prio"], rtol=RTOL, atol=1e-7)
    np.testing.assert_allclose(delta.cpu().numpy()[:live], o["delta"], rtol=RTOL, atol=1e-6)
    np.testing.assert_allclose(mean.item(), float(np.sum(o["loss"].astype(np.float64)) / Bg), rtol=RTOL)
    assert (prio.cpu().numpy()[live:] == -7.0).all()                               # dead rows untouched
    new = ro.priority_to_leaf(prio.cpu().numpy()[:live])
    for i, v in zip(idx[:live].tolist(), new.tolist()):
        leaves[i] = v
    for g, w in zip(read_levels(apex, mem), ro.build_levels(leaves)):
        np.testing.assert_array_equal(g, w)
    mem.check_status()


@pytest.mark.parametrize("use_graph,depth", [(False, 1), (False, 2), (True, 2), (False, 3), (False, 4), (True, 3)])
def test_host_step_pipeline(apex, use_graph, depth):
    """apexb200_host_step / HostStepPipeline: host draws + Q in, host loss / priorities / indices out,
    equal to the oracle's sample -> TD -> write-back on the same numbers, step after step."""
    rng = np.random.default_rng(41)
    N, C_, F, B, nA = 1 << 13, 2, 256, 96, 5
    mem = make_mem(apex, N, C=C_, F=F, layout="stream", lanes=2, n=3)
    ring = ro.RingOracle(N, C_, F, lanes=2, fpr=1, lookback=C_ - 1 + 3, alpha=0.6)
    for lane in range(2):
        m = N // 2 - 8
        args = (lane, rng.integers(0, 256, (m, F), dtype=np.uint8), apex.stream_offsets(C_, np.full(m, 3), np.arange(m)),
                rng.integers(0, nA, m).astype(np.int32), rng.standard_normal(m).astype(np.float32), np.full(m, 0.97, np.float32))
        prio = (np.abs(rng.standard_normal(m)) + 1e-3).astype(np.float32)
        mem.insert(*args, prio=prio)
        ring.insert(*args, prio=prio)
    ring.leaves[:] = u64(mem.buffer(apex._C.BUF_LEAVES))
    pipe = apex.HostStepPipeline(mem, B, nA, use_graph=use_graph, depth=depth)
    with pytest.raises(ValueError):
        apex.HostStepPipeline(mem, B, nA, depth=apex._C.HOST_STEP_SLOTS + 1)
    tickets, inputs, steps, checked = [], [], 11, 0

    def check(k):
        t, (u0, q0) = tickets[k], inputs[k]
        loss_mean, prio, idx = pipe.result(t)
        want = ring.sample(u0)
        np.testing.assert_array_equal(idx, want)
        g = ring.gather(want)
        o = to.td_forward(q0[0], q0[1], q0[2], g["A"], g["R"], g["G"], alpha=0.6)
        np.testing.assert_allclose(prio, o["prio"], rtol=RTOL, atol=1e-7)
        assert loss_mean == pytest.approx(o["loss_mean"], rel=RTOL)
        ring.set_leaves(want, ro.priority_to_leaf(prio.copy()))

    for k in range(steps):
        u = ro.stratified_uniforms(B, rng)
        q = rng.standard_normal((3, B, nA)).astype(np.float32)
        tickets.append(pipe.submit(u, q))
        inputs.append((u, q))
        if k >= depth - 1:                                           # read depth-1 steps behind, as the bench does
            check(checked)
            checked += 1
    while checked < steps:
        check(checked)
        checked += 1
    mem.check_status()


@pytest.mark.parametrize("depth,max_rows", [(1, None), (3, None), (3, 160)])
def test_sharded_host_step_pipeline_world1(apex, depth, max_rows):
    """apexb200_host_step_sharded / ShardedHostStepPipeline with a group of one: host Q rows in; loss part, priorities,
    record indices, global draw positions and the row count out — equal, step after step, to the oracle's
    sample -> TD -> write-back on the draws the kernel generates (restated by kernel_stratified_uniforms).
    max_rows larger than the batch: rows beyond the count are ignored."""
    rng = np.random.default_rng(43)
    N, C_, F, B, nA = 1 << 13, 2, 256, 96, 4
    mem = make_mem(apex, N, C=C_, F=F, layout="stream", lanes=2, n=3)
    mem.seed = 4242
    ring = ro.RingOracle(N, C_, F, lanes=2, fpr=1, lookback=C_ - 1 + 3, alpha=0.6)
    for lane in range(2):
        m = N // 2 - 8
        args = (lane, rng.integers(0, 256, (m, F), dtype=np.uint8), apex.stream_offsets(C_, np.full(m, 3), np.arange(m) % 97),
                rng.integers(0, nA, m).astype(np.int32), rng.standard_normal(m).astype(np.float32), np.full(m, 0.97, np.float32))
        prio = (np.abs(rng.standard_normal(m)) + 1e-3).astype(np.float32)
        mem.insert(*args, prio=prio)
        ring.insert(*args, prio=prio)
    ring.leaves[:] = u64(mem.buffer(apex._C.BUF_LEAVES))
    sampler = apex.ShardedSampler(mem, rank=0, world_size=1)
    sampler.connect_peers()
    sampler.push_on_update(True)
    sampler.exchange()                                                   # round 1
    pipe = apex.ShardedHostStepPipeline(sampler, B, nA, max_rows=max_rows, depth=depth)
    M = pipe.M
    tickets, inputs, steps, checked = [], [], 9, 0

    def check(k):
        loss_part, prio, idx, pos, n = pipe.result(tickets[k])
        q0 = inputs[k]
        assert n == B and (pos == np.arange(B)).all()
        u = ro.kernel_stratified_uniforms(4242, 1 + k, k, B)             # exchange round advances with every update
        want = ring.sample(u)
        np.testing.assert_array_equal(idx, want)
        g = ring.gather(want)
        o = to.td_forward(q0[0][:B], q0[1][:B], q0[2][:B], g["A"], g["R"], g["G"], alpha=0.6)
        np.testing.assert_allclose(prio, o["prio"], rtol=RTOL, atol=1e-7)
        assert loss_part == pytest.approx(o["loss_mean"], rel=RTOL)
        ring.set_leaves(want, ro.priority_to_leaf(prio.copy()))

    for k in range(steps):
        q = rng.standard_normal((3, M, nA)).astype(np.float32)
        tickets.append(pipe.submit(q))
        inputs.append(q)
        if k >= depth - 1:
            check(checked)
            checked += 1
    while checked < steps:
        check(checked)
        checked += 1
    assert mem.total() == ring.total()
    mem.check_status()


def test_vector_actor_ingest_matches_sliding_buffer(apex):
    """K7 + K6 ingest of E lock-stepped environments == the host ExperienceBuffer(semantics="sliding")
    fed the same steps, record by record (a10/a11/a3 on the device, frames stored once)."""
    rng = np.random.default_rng(52)
    E, n, C_, F, nA, lane_cap, gamma = 3, 3, 4, 16, 4, 256, 0.99
    mem = apex.ReplayMemory(E * lane_cap, {"priority_exponent": 0.6, "layout": "stream", "lanes": E, "n_step": n,
                                           "state_shape": (C_, 1, F), "num_actions": 0})
    ing = apex.VectorActorIngest(mem, E, gamma=gamma)
    steps = 150
    frames = np.zeros((E, steps, F), np.uint8)
    frames[:, :, 0] = np.arange(E)[:, None]
    frames[:, :, 1] = np.arange(steps)[None, :] % 256                   # frame (e, tau) names itself
    acts = rng.integers(0, nA, (E, steps)).astype(np.int32)
    rews = rng.standard_normal((E, steps)).astype(np.float32)
    dones = rng.random((E, steps)) < 0.06
    qs = rng.standard_normal((E, steps, nA)).astype(np.float32)
    for lo, hi in ((0, 40), (40, 41), (41, 97), (97, 150)):                # ragged chunks, including a 1-step chunk
        ing.push(frames[:, lo:hi], acts[:, lo:hi], rews[:, lo:hi], dones[:, lo:hi], qs[:, lo:hi])
    A_d, R_d, G_d = (mem.buffer(w).cpu().numpy() for w in (apex._C.BUF_ACTION, apex._C.BUF_REWARD, apex._C.BUF_GAMMA))
    leaves = u64(mem.buffer(apex._C.BUF_LEAVES))
    fidx = mem.buffer(apex._C.BUF_FIDX).cpu().numpy().reshape(-1, 2 * C_)
    store = mem.buffer(apex._C.BUF_FRAMES).cpu().numpy().reshape(-1, F)
    for e in range(E):
        buf = apex.ExperienceBuffer(n, e, semantics="sliding", gamma=gamma)
        for t in range(steps):
            buf.add(apex.Transition(t, int(acts[e, t]), float(rews[e, t]), gamma, qs[e, t]))
            if dones[e, t]:
                buf.construct_nstep_transition(apex.Transition(t + 1, 0, 0, gamma, qs[e, min(t + 1, steps - 1)]))
        recs = {r.S_t: r for r in buf.get(buf.size)}                       # keyed by step index t
        ep_start = 0
        starts = np.zeros(steps, np.int64)
        for t in range(steps):
            starts[t] = ep_start
            if dones[e, t]:
                ep_start = t + 1
        assert (leaves[e * lane_cap: e * lane_cap + n] == 0).all()         # no record on the first n frames
        for tau in range(n, steps):
            t = tau - n
            slot = e * lane_cap + tau
            r = recs[t]                                                    # sliding buffer emitted step t (window complete by tau)
            assert A_d[slot] == r.A_t
            assert R_d[slot] == pytest.approx(r.R_ttpB, rel=RTOL, abs=1e-6)
            assert G_d[slot] == pytest.approx(r.Gamma_ttpB, rel=RTOL)
            want_p = abs(r.R_ttpB + r.Gamma_ttpB * float(qs[e, tau].max()) - float(qs[e, t, r.A_t]))
            got_p = leaves[slot] / float(1 << 24)
            assert got_p == pytest.approx(want_p ** 0.6, rel=1e-4, abs=2e-7)
            # S_t = the C frames ending at t, repeated at the episode start; each frame names (env, time)
            want_t = [max(t - b, starts[t]) for b in range(C_ - 1, -1, -1)]
            got = store[fidx[slot, :C_]]
            assert (got[:, 0] == e).all() and got[:, 1].tolist() == [x % 256 for x in want_t]
            if r.Gamma_ttpB != 0.0:                                        # S_tpn only matters when we bootstrap
                want_n = [max(tau - b, starts[tau]) for b in range(C_ - 1, -1, -1)]
                got = store[fidx[slot, C_:]]
                assert (got[:, 0] == e).all() and got[:, 1].tolist() == [x % 256 for x in want_n]
                assert starts[tau] == starts[t]
    mem.check_status()


@pytest.mark.parametrize("E,lane_cap,nA,store_q", [(5, 64, 4, False), (2, 40, 6, True), (300, 32, 3, False)])
def test_device_ingest_equals_ring_oracle_fed_with_nstep_oracle(apex, E, lane_cap, nA, store_q):
    """apexb200_ingest_push (n-step assembly, episode ages, initial priority, ring insert, all state on the device)
    against the oracles directly: nstep_oracle.nstep_intended over the whole step history gives R / Gamma of every
    record, stream offsets + RingOracle.insert give frame slots, eviction and the dead zone.  Ragged chunks, chunks
    shorter than n, pushes for a subset of the environments, several ring wrap-arounds; every tree level equals a
    rebuild at the end."""
    rng = np.random.default_rng(E * 7 + lane_cap)
    n, C_, F, gamma, steps = 3, 3, 32, 0.99, 150
    W = C_ - 1 + n
    mem = apex.ReplayMemory(E * lane_cap, {"priority_exponent": 0.6, "layout": "stream", "lanes": E, "n_step": n,
                                           "state_shape": (C_, 1, F), "num_actions": nA if store_q else 0})
    lanes = [int(x) for x in rng.permutation(E)]
    ing = apex.VectorActorIngest(mem, E, gamma=gamma, lanes=lanes)
    ring = ro.RingOracle(E * lane_cap, C_, F, lanes=E, fpr=1, lookback=W, alpha=0.6, num_actions=nA if store_q else 0)
    frames = rng.integers(0, 256, (E, steps, F), dtype=np.uint8)
    acts = rng.integers(0, nA, (E, steps)).astype(np.int32)
    rews = rng.standard_normal((E, steps)).astype(np.float32)
    dones = rng.random((E, steps)) < 0.08
    qs = rng.standard_normal((E, steps, nA)).astype(np.float32)
    R_all, G_all, span = no.nstep_intended(rews, dones, n, gamma)
    age = np.zeros((E, steps), np.int64)                                   # frames since the episode start, per frame
    for t in range(1, steps):
        age[:, t] = np.where(dones[:, t - 1], 0, age[:, t - 1] + 1)
    back = np.arange(C_ - 1, -1, -1)
    cuts = [0, 2, 3, 11, 12, 30, 31, 33, 60, 61, 90, 120, 150] if lane_cap >= 40 else list(range(0, 151, 25))
    for a_, b_ in zip(cuts, cuts[1:]):
        # every other chunk goes to a subset of the environments first, then to the rest (same steps, two calls)
        groups = [list(range(E))] if (a_ // 3) % 2 else [list(range(0, E, 2)), list(range(1, E, 2))]
        for envs in [g for g in groups if g]:
            sl = np.array(envs)
            ing.push(frames[sl, a_:b_], acts[sl, a_:b_], rews[sl, a_:b_], dones[sl, a_:b_], qs[sl, a_:b_],
                     envs=None if len(envs) == E else envs)
            for e in envs:
                tau = np.arange(a_, b_)
                t = tau - n
                ok = t >= 0
                tc = np.maximum(t, 0)
                A = np.where(ok, acts[e, tc], 0).astype(np.int32)
                R = np.where(ok, R_all[e, tc], 0).astype(np.float32)
                G = np.where(ok, G_all[e, tc], 0).astype(np.float32)
                assert (span[e, tc][ok] > 0).all()                         # the window of step t is complete at tau
                rel = np.concatenate([-(n + np.minimum(back[None, :], np.where(ok, age[e, tc], 0)[:, None])),
                                      -np.minimum(back[None, :], age[e, tau][:, None])], 1).astype(np.int32)
                prio = np.abs(R + G * qs[e, tau].max(1) - qs[e, tc, A]) * ok
                ring.insert(lanes[e], frames[e, a_:b_], rel, A, R, G, prio=prio.astype(np.float32),
                            q_t=np.where(ok[:, None], qs[e, tc], 0) if store_q else None, q_tpn=qs[e, tau] if store_q else None)
    assert mem.size() == ring.size()
    written = np.zeros(E * lane_cap, bool)
    for e in range(E):
        written[lanes[e] * lane_cap + np.arange(min(steps, lane_cap))] = True
    get = lambda w: mem.buffer(w).cpu().numpy()                            # noqa: E731
    np.testing.assert_array_equal(get(apex._C.BUF_FRAMES).reshape(-1, F)[written], ring.frames[written])
    np.testing.assert_array_equal(get(apex._C.BUF_FIDX).reshape(-1, 2 * C_)[written], ring.fidx[written])
    np.testing.assert_array_equal(get(apex._C.BUF_ACTION)[written], ring.A[written])
    np.testing.assert_allclose(get(apex._C.BUF_REWARD)[written], ring.R[written], rtol=RTOL, atol=1e-6)
    np.testing.assert_allclose(get(apex._C.BUF_GAMMA)[written], ring.G[written], rtol=RTOL)
    if store_q:
        np.testing.assert_array_equal(get(apex._C.BUF_QT).reshape(-1, nA)[written], ring.q_t[written])
        np.testing.assert_array_equal(get(apex._C.BUF_QTPN).reshape(-1, nA)[written], ring.q_tpn[written])
    leaves = u64(mem.buffer(apex._C.BUF_LEAVES))
    assert ((leaves == 0) == (ring.leaves == 0)).all()                     # same live / dead records
    np.testing.assert_allclose(leaves / float(1 << 24), ring.leaves / float(1 << 24), rtol=2e-4, atol=2e-6)
    for l, (got, want) in enumerate(zip(read_levels(apex, mem), ro.build_levels(leaves))):
        np.testing.assert_array_equal(got, want, err_msg="tree level %d" % l)
    mem.check_status()


@pytest.mark.parametrize("E,lane_cap,T,rounds", [(3, 64, 10, 9), (70, 32, 5, 4), (1, 128, 40, 5)])
def test_insert_lanes_equals_per_lane_inserts(apex, E, lane_cap, T, rounds):
    """apexb200_replay_insert_lanes (one launch pair for E lanes, 64 lanes per launch) == E single-lane inserts,
    through several ring wrap-arounds, with given priorities and with priorities computed from stored Q rows."""
    rng = np.random.default_rng(E * 1000 + T)
    C_, F, nA, n = 2, 32, 3, 3
    params = {"priority_exponent": 0.6, "layout": "stream", "lanes": E, "n_step": n, "state_shape": (C_, 1, F),
              "num_actions": nA}
    a, b = apex.ReplayMemory(E * lane_cap, params), apex.ReplayMemory(E * lane_cap, params)
    lanes = list(rng.permutation(E))                                       # any order, each lane once
    for r in range(rounds):
        frames = rng.integers(0, 256, (E, T, F), dtype=np.uint8)
        rel = np.stack([apex.stream_offsets(C_, np.full(T, n), np.arange(T) + r * T) for _ in range(E)])
        A = rng.integers(0, nA, (E, T)).astype(np.int32)
        R = rng.standard_normal((E, T)).astype(np.float32)
        G = np.where(rng.random((E, T)) < 0.1, 0.0, 0.97).astype(np.float32)
        q_t, q_tpn = (rng.standard_normal((E, T, nA)).astype(np.float32) for _ in range(2))
        prio = None if r % 2 else (np.abs(rng.standard_normal((E, T))) + 1e-3).astype(np.float32)
        a.insert_lanes(lanes, frames, rel, A, R, G, prio=prio, q_t=q_t, q_tpn=q_tpn)
        for k, lane in enumerate(lanes):
            b.insert(lane, frames[k], rel[k], A[k], R[k], G[k], prio=None if prio is None else prio[k], q_t=q_t[k], q_tpn=q_tpn[k])
        assert a.size() == b.size() and a.total() == b.total()
    written = min(rounds * T, lane_cap)
    for which in (apex._C.BUF_FRAMES, apex._C.BUF_FIDX, apex._C.BUF_ACTION, apex._C.BUF_REWARD, apex._C.BUF_GAMMA,
                  apex._C.BUF_QT, apex._C.BUF_QTPN):
        x, y = (m_.buffer(which).view(E, lane_cap, -1)[:, :written] for m_ in (a, b))
        assert torch.equal(x, y), which
    assert torch.equal(a.buffer(apex._C.BUF_LEAVES), b.buffer(apex._C.BUF_LEAVES))
    # the tree above the leaves is what a rebuild gives
    top_a = a.total()
    apex._C.call("apexb200_tree_build", a._h, apex._C.stream_ptr())
    assert a.total() == top_a
    with pytest.raises(apex._C.ApexError):
        a.insert_lanes([0, 0][:min(2, E)] if E > 1 else [5], np.zeros((min(2, E), 1, F), np.uint8), np.zeros((min(2, E), 1, 2 * C_), np.int32),
                       np.zeros((min(2, E), 1), np.int32), np.zeros((min(2, E), 1), np.float32), np.zeros((min(2, E), 1), np.float32),
                       prio=np.ones((min(2, E), 1), np.float32))           # a lane listed twice / out of range
    a.check_status()
    b.check_status()


@pytest.mark.parametrize("lane_cap,lanes", [(40, 1), (44, 8), (1032, 1), (96, 3), (33, 2)])
def test_insert_across_ring_wrap_keeps_every_tree_level_exact(apex, lane_cap, lanes):
    """Inserts whose slots wrap around the ring inside one warp of K6 (lane capacity not a multiple of 32):
    after every insert each inner level of the sum tree must equal a rebuild from the leaves."""
    rng = np.random.default_rng(lane_cap * 131 + lanes)
    C_, F, n = 1, 16, 1
    mem = make_mem(apex, lane_cap * lanes, C=C_, F=F, layout="stream", lanes=lanes, n=n)
    look = C_ - 1 + n
    sizes = [5, 7, 13, 29, 31, 32, lane_cap - look, 3, 17, 11, 23, 19, 9, 30, 6, 27]
    for r, m in enumerate(sizes * 3):
        m = min(m, lane_cap - look)
        lane = r % lanes
        mem.insert(lane, rng.integers(0, 256, (m, F), dtype=np.uint8), apex.stream_offsets(C_, np.full(m, n), np.arange(m)),
                   rng.integers(0, 4, m).astype(np.int32), rng.standard_normal(m).astype(np.float32),
                   np.full(m, 0.99, np.float32), prio=(np.abs(rng.standard_normal(m)) * 10 + 1e-3).astype(np.float32))
        levels = read_levels(apex, mem)
        for l, (got, want) in enumerate(zip(levels, ro.build_levels(levels[0]))):
            np.testing.assert_array_equal(got, want, err_msg="level %d after insert %d (%d records)" % (l, r, m))
    mem.check_status()


def _ring_actor(name, slots, T, F, nA, chunks, seed):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import time
    import apex_b200
    ring = apex_b200.ActorStagingRing(name, slots, T, F, nA, create=False)
    rng = np.random.default_rng(seed)
    for c in range(chunks):
        payload = (rng.integers(0, 256, (T, F), dtype=np.uint8), rng.integers(0, nA, T).astype(np.int32),
                   rng.standard_normal(T).astype(np.float32), rng.random(T) < 0.1, rng.standard_normal((T, nA)).astype(np.float32))
        while not ring.push(*payload):
            time.sleep(0.001)
    ring.close()


def test_ring_ingest_from_actor_processes(apex):
    """Two actor processes (no CUDA) stream step chunks through shared-memory rings; the replay's
    process drains them into the device ring.  Result == feeding the same chunks directly."""
    import multiprocessing as mp
    import time
    E, T, F, nA, C_, n, chunks, slots, lane_cap = 2, 16, 64, 4, 2, 3, 9, 3, 512
    params = {"priority_exponent": 0.6, "layout": "stream", "lanes": E, "n_step": n, "state_shape": (C_, 1, F)}
    mem = apex.ReplayMemory(E * lane_cap, params)
    ref = apex.ReplayMemory(E * lane_cap, params)
    names = ["apexring_gpu_%d_%d" % (os.getpid(), e) for e in range(E)]
    rings = [apex.ActorStagingRing(nm, slots, T, F, nA, create=True) for nm in names]
    try:
        procs = [mp.get_context("spawn").Process(target=_ring_actor, args=(names[e], slots, T, F, nA, chunks, 900 + e))
                 for e in range(E)]
        for p in procs:
            p.start()
        ingest = apex.RingIngest(mem, rings, n_step=n)
        assert all(r._registered for r in rings)                       # page-locked for async H2D
        want, t0 = E * chunks * T, time.time()
        got = 0
        while got < want and time.time() - t0 < 120:
            got += ingest.poll()
            time.sleep(0.002)
        ingest.flush()                                                     # every slot handed back to its actor
        for p in procs:
            p.join(timeout=30)
        assert got == want and all(p.exitcode == 0 for p in procs)
        assert all(int(r.ctrl[0]) == int(r.ctrl[1]) == chunks for r in rings)
        direct = [apex.VectorActorIngest(ref, 1, n_step=n, lanes=[e]) for e in range(E)]
        for e in range(E):
            rng = np.random.default_rng(900 + e)
            for c in range(chunks):
                f, a, r = rng.integers(0, 256, (T, F), dtype=np.uint8), rng.integers(0, nA, T).astype(np.int32), rng.standard_normal(T).astype(np.float32)
                d, q = rng.random(T) < 0.1, rng.standard_normal((T, nA)).astype(np.float32)
                direct[e].push(f[None], a[None], r[None], d[None], q[None])
        written = chunks * T                                               # records per lane; the rest of the ring was never written
        for which in (apex._C.BUF_FRAMES, apex._C.BUF_FIDX, apex._C.BUF_ACTION, apex._C.BUF_REWARD, apex._C.BUF_GAMMA):
            a_, b_ = (m_.buffer(which).view(E, lane_cap, -1)[:, :written] for m_ in (mem, ref))
            assert torch.equal(a_, b_), which
        assert torch.equal(mem.buffer(apex._C.BUF_LEAVES), ref.buffer(apex._C.BUF_LEAVES))
        assert mem.total() == ref.total() and mem.size() == ref.size() == want
        mem.check_status()
    finally:
        torch.cuda.synchronize()
        for r in rings:
            r.close(unlink=True)


# --------------------------------------------------------------- error behaviour
def test_argument_errors(apex):
    _C = apex._C
    h = C.c_void_p()
    bad = _C.ReplayConfig(100, 3, 1, 0, 1, 7056, 0, 0.6, 0)
    assert _C.lib().apexb200_replay_create(C.byref(h), C.byref(bad)) == _C.EINVAL          # capacity % lanes
    bad = _C.ReplayConfig((1 << 24) + 32, 1, 1, 0, 1, 16, 0, 0.6, 0)
    assert _C.lib().apexb200_replay_create(C.byref(h), C.byref(bad)) == _C.ERANGE
    bad = _C.ReplayConfig(100, 1, 1, 0, 1, 7050, 0, 0.6, 0)
    assert _C.lib().apexb200_replay_create(C.byref(h), C.byref(bad)) == _C.EINVAL          # frame_bytes % 16
    mem = make_mem(apex, 64, layout="stream", lanes=1, n=3, C=4)
    with pytest.raises(_C.ApexError, match="exceeds lane capacity"):
        mem.insert(0, np.zeros((60, 16), np.uint8), np.zeros((60, 8), np.int32), np.zeros(60, np.int32),
                   np.zeros(60, np.float32), np.zeros(60, np.float32), prio=np.ones(60, np.float32))
    with pytest.raises(_C.ApexError):
        apex.ops.td_forward(torch.zeros(2000, 4).cuda(), torch.zeros(2000, 4).cuda(), torch.zeros(2000, 4).cuda(),
                            torch.zeros(2000, dtype=torch.int32).cuda(), torch.zeros(2000).cuda(), torch.zeros(2000).cuda(),
                            0.6, workspace=torch.empty(256, dtype=torch.uint8, device="cuda"))   # workspace too small


# -------------------------------------------- full-size properties (BASELINE sizes)
def test_full_size_properties_2m(apex):
    """2M-record shard, B=512: sortedness of indices for sorted draws, gather round trip of
    self-identifying frames, tree == rebuild after many fused updates (size-independent checks)."""
    N, C_, F, B, nA = 1 << 21, 4, 7056, 512, 4
    mem = apex.ReplayMemory(N, {"priority_exponent": 0.6, "layout": "stream", "lanes": 32, "n_step": 3,
                                "state_shape": (C_, 84, 84)})
    mem._ensure()
    g = torch.Generator(device="cuda").manual_seed(1234)
    lane_cap = N // 32
    chunk = 8192
    rel_host = apex.stream_offsets(C_, np.full(chunk, 3), np.minimum(np.arange(chunk) + 100, 1000)).astype(np.int64)
    rel = dev(rel_host.astype(np.int32))
    for lane in range(32):
        for c in range(lane_cap // chunk):
            frames = torch.randint(0, 256, (chunk, F), dtype=torch.uint8, device="cuda", generator=g)
            base = lane * lane_cap + c * chunk
            tag = (base + torch.arange(chunk, device="cuda", dtype=torch.int64))
            frames[:, :8] = tag.view(-1, 1).bitwise_right_shift(torch.arange(0, 64, 8, device="cuda")).bitwise_and(255).to(torch.uint8)
            A = torch.randint(0, nA, (chunk,), dtype=torch.int32, device="cuda", generator=g)
            R = torch.randn(chunk, device="cuda", generator=g).clamp(-1, 1)
            G = torch.full((chunk,), 0.970299, device="cuda")
            prio = torch.randn(chunk, device="cuda", generator=g).abs() + 1e-3
            mem.insert(lane, frames, rel, A, R, G, prio=prio)
    assert mem.size() == N
    store = mem.buffer(apex._C.BUF_FRAMES).view(-1, F)
    rng = np.random.default_rng(4321)
    leaves = u64(mem.buffer(apex._C.BUF_LEAVES))
    for step in range(20):
        u = ro.stratified_uniforms(B, rng)
        b = mem.sample_tensors(B, u=dev(u))
        idx = b.idx.cpu().numpy()
        assert (np.diff(idx) >= 0).all()                                           # sorted draws -> sorted indices
        if step % 5 == 0:
            np.testing.assert_array_equal(idx, ro.tree_sample(leaves, u))          # and bit-exact vs the oracle
        fsel = b.fsel.cpu().numpy().astype(np.int64)
        tags = b.S_tpn.cpu().numpy()[:, :, :8].astype(np.uint64)
        got = sum(tags[:, :, k] << np.uint64(8 * k) for k in range(8))
        np.testing.assert_array_equal(got, fsel[:, C_:].astype(np.uint64))         # each gathered frame names its slot
        np.testing.assert_array_equal(fsel[:, -1], idx)                            # newest S_tpn frame = the record's own
        # the frame slots of every sampled record, recomputed on the host from the insert's offsets ...
        p = idx % lane_cap
        want_f = (idx // lane_cap * lane_cap)[:, None] + (p[:, None] + rel_host[p % chunk]) % lane_cap
        np.testing.assert_array_equal(fsel, want_f)
        # ... and every byte of both gathered states against the frame store itself (2 x B x C x 7056 bytes)
        sel = b.fsel.long()
        assert torch.equal(b.S_t.view(B * C_, F), store.index_select(0, sel[:, :C_].reshape(-1)))
        assert torch.equal(b.S_tpn.view(B * C_, F), store.index_select(0, sel[:, C_:].reshape(-1)))
        for name, which in (("A", apex._C.BUF_ACTION), ("R", apex._C.BUF_REWARD), ("Gamma", apex._C.BUF_GAMMA)):
            assert torch.equal(getattr(b, name), mem.buffer(which).index_select(0, b.idx.long()))
        q = [torch.randn(B, nA, device="cuda", generator=g) for _ in range(3)]
        _, prio, _, _ = mem.td_update(b.idx, q[0], q[1], q[2], b.A, b.R, b.Gamma)
        new = ro.priority_to_leaf(prio.cpu().numpy())
        for i, v in zip(idx.tolist(), new.tolist()):
            leaves[i] = v
    for gl, w in zip(read_levels(apex, mem), ro.build_levels(leaves)):
        np.testing.assert_array_equal(gl, w)
    mem.check_status()


# ------------------------------------------------------------ a8/a9: Learner mirror
def _small_learner(apex, mem, B=32, C_=1):
    lp = {"remove_old_xp_freq": 100, "q_target_sync_freq": 3, "min_replay_mem_size": 10, "replay_sample_size": B,
          "load_saved_state": False}
    torch.manual_seed(0)
    return apex.Learner({"state_shape": [C_, 84, 84], "action_dim": 4}, lp, {}, mem)


def test_learner_compat_and_fast_tier(apex):
    """learner.py:29-61 through the mirror class: compat records in -> (loss, {key: |delta|}); fast tier
    sample -> model -> fused TD + in-place priority update -> RMSprop step."""
    rng = np.random.default_rng(31)
    NT = apex.N_Step_Transition
    K = 200
    mem = apex.ReplayMemory(K, {"priority_exponent": 0.6})
    S = rng.integers(0, 256, (K + 1, 1, 84, 84)).astype(np.uint8)
    recs = [NT(S[i], int(rng.integers(4)), float(rng.standard_normal()), 0.97, None, S[i + 1], None, "1%d" % i) for i in range(K)]
    mem.add({r.key: 1.0 + i % 3 for i, r in enumerate(recs)}, recs)
    L = _small_learner(apex, mem)
    np.random.seed(3)
    batch = mem.sample(32)
    loss, prios = L.compute_loss_and_priorities(batch)                      # learner.py:70
    assert loss.requires_grad and set(prios) == {r.key for r in batch}
    # the same numbers from plain torch on the same networks (learner.py:43-52 restated in fp32)
    St = torch.from_numpy(np.stack([r.S_t for r in batch])).cuda().float()
    Sn = torch.from_numpy(np.stack([r.S_tpn for r in batch])).cuda().float()
    A = torch.tensor([r.A_t for r in batch], device="cuda")
    R = torch.tensor([r.R_ttpB for r in batch], device="cuda", dtype=torch.float32)
    G = torch.tensor([r.Gamma_ttpB for r in batch], device="cuda", dtype=torch.float32)
    with torch.no_grad():
        tgt = R + G * L.Q_double(Sn)[2].gather(1, torch.argmax(L.Q(Sn)[2], 1).view(-1, 1)).squeeze()
    qsa = L.Q(St)[2].gather(1, A.view(-1, 1)).squeeze()
    ref = (0.5 * (tgt - qsa) ** 2).mean()
    torch.testing.assert_close(loss, ref, rtol=1e-4, atol=1e-6)
    want = (tgt - qsa).abs().detach().cpu().numpy()
    for j, r in enumerate(batch):                                           # per-record |delta|, last dup wins
        if [x.key for x in batch].index(r.key) == j and [x.key for x in batch].count(r.key) == 1:
            assert prios[r.key] == pytest.approx(want[j], rel=1e-4, abs=1e-6)
    g_ref = torch.autograd.grad(ref, list(L.Q.parameters()), allow_unused=True)
    L.optimizer.zero_grad()
    loss.backward()
    for p, g in zip(L.Q.parameters(), g_ref):                               # cuDNN backward is not bit-reproducible
        torch.testing.assert_close(p.grad, g, rtol=2e-2, atol=1e-3)
    L.update_Q(L.compute_loss_and_priorities(batch)[0])                     # learner.py:73
    mem.set_priorities(prios)                                               # learner.py:76
    assert L.num_q_updates == 1
    # fast tier: the batch never leaves HBM, priorities are written in place
    before = mem.total()
    losses = [float(L.step_fast()) for _ in range(5)]
    assert L.num_q_updates == 6 and all(np.isfinite(losses)) and mem.total() != before
    sd_q, sd_t = L.Q.state_dict(), L.Q_double.state_dict()
    assert all(torch.equal(sd_q[k], sd_t[k]) for k in sd_q)                 # target sync at update 6 (freq 3)
    # the same step captured as one CUDA graph (model included): replays keep learning and keep updating priorities
    del loss, ref, qsa, g_ref
    w_before = [p.detach().clone() for p in L.Q.parameters()]
    t_before = mem.total()
    g_losses = [float(L.step_graphed()) for _ in range(4)]
    assert all(np.isfinite(g_losses)) and mem.total() != t_before
    assert any((a != b).any().item() for a, b in zip(w_before, L.Q.parameters()))
    mem.check_status()


def test_max_capacity_16m_records(apex):
    """The largest shard the ABI allows (2^24 records, 118 GB of frames in HBM): same size-independent
    properties as the 2M test — sorted draws give sorted indices, gathered frames name their slots,
    the tree total equals the sum of the leaves after fused updates."""
    import gc
    gc.collect()
    torch.cuda.empty_cache()
    free, _ = torch.cuda.mem_get_info()
    N, C_, F, B, nA, lanes = 1 << 24, 4, 7056, 512, 4, 64
    if free < N * F + (6 << 30):
        pytest.skip("not enough free HBM for a 16M-record shard (%d GB free)" % (free >> 30))
    mem = apex.ReplayMemory(N, {"priority_exponent": 0.6, "layout": "stream", "lanes": lanes, "n_step": 3,
                                "state_shape": (C_, 84, 84)})
    mem._ensure()
    g = torch.Generator(device="cuda").manual_seed(7)
    lane_cap, chunk = N // lanes, 16384
    rel_host = apex.stream_offsets(C_, np.full(chunk, 3), np.full(chunk, 100)).astype(np.int64)
    rel = dev(rel_host.astype(np.int32))
    block = torch.randint(0, 256, (chunk, F), dtype=torch.uint8, device="cuda", generator=g)
    shifts = torch.arange(0, 64, 8, device="cuda")
    for lane in range(lanes):
        for c in range(lane_cap // chunk):
            base = lane * lane_cap + c * chunk
            tag = base + torch.arange(chunk, device="cuda", dtype=torch.int64)
            block[:, :8] = tag.view(-1, 1).bitwise_right_shift(shifts).bitwise_and(255).to(torch.uint8)
            mem.insert(lane, block, rel, torch.zeros(chunk, dtype=torch.int32, device="cuda"),
                       torch.zeros(chunk, device="cuda"), torch.full((chunk,), 0.97, device="cuda"),
                       prio=torch.rand(chunk, device="cuda", generator=g) + 1e-3)
    assert mem.size() == N
    store = mem.buffer(apex._C.BUF_FRAMES).view(-1, F)
    for step in range(4):
        u_out = torch.empty(B, dtype=torch.float64, device="cuda")
        b = mem.sample_tensors(B, u_out=u_out)
        idx = b.idx.cpu().numpy()
        assert (np.diff(idx) >= 0).all() and idx.min() >= 0 and idx.max() < N
        fsel = b.fsel.cpu().numpy().astype(np.uint64)
        tags = b.S_tpn.cpu().numpy()[:, :, :8].astype(np.uint64)
        got = sum(tags[:, :, k] << np.uint64(8 * k) for k in range(8))
        np.testing.assert_array_equal(got, fsel[:, C_:])
        np.testing.assert_array_equal(fsel[:, -1], idx.astype(np.uint64))
        p = idx.astype(np.int64) % lane_cap                                  # frame slots recomputed on the host
        want_f = (idx.astype(np.int64) // lane_cap * lane_cap)[:, None] + (p[:, None] + rel_host[p % chunk]) % lane_cap
        np.testing.assert_array_equal(fsel.astype(np.int64), want_f)
        sel = b.fsel.long()                                                  # every gathered byte against the store
        assert torch.equal(b.S_t.view(B * C_, F), store.index_select(0, sel[:, :C_].reshape(-1)))
        assert torch.equal(b.S_tpn.view(B * C_, F), store.index_select(0, sel[:, C_:].reshape(-1)))
        q = [torch.randn(B, nA, device="cuda", generator=g) for _ in range(3)]
        mem.td_update(b.idx, q[0], q[1], q[2], b.A, b.R, b.Gamma)
    leaves = mem.buffer(apex._C.BUF_LEAVES)
    assert mem.total() == int(leaves.sum().item())                      # u64 total < 2^63 here
    # the last draws, replayed on the host from the same leaves (before the last update) would need a
    # 128 MB cumsum; the index law itself is covered bit-exactly at 2M and below
    mem.check_status()
    mem.close()
