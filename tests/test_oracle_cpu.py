"""CPU tests: the oracle against the golden vectors produced by the reference
(tests/golden/make_golden.py) and against its own invariants."""
import json
import os

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from oracle import nstep_oracle as no
from oracle import replay_oracle as ro
from oracle import td_oracle as to


def _load(golden_dir, name):
    with open(os.path.join(golden_dir, name)) as f:
        return json.load(f)


# ---------------------------------------------------------------- n-step (a10)
def test_literal_nstep_matches_reference(golden_dir):
    for case in _load(golden_dir, "nstep_literal.json"):
        sp = case["spec"]
        buf = no.LiteralExperienceBuffer(sp["n"], sp["actor_id"])
        for s in case["feed"]:
            buf.add(no.Step(s["S"], s["A"], s["R"], s["Gamma"], s["q"]))
            if s["done"]:
                buf.flush(no.Step(1000 + s["S"], s["A"], 0, 0.99, s["q"]))
        assert len(buf.ready) == len(case["emitted"])
        for got, want in zip(buf.ready, case["emitted"]):
            assert (got.S_t, got.A_t, got.S_tpn, got.key) == (want["S_t"], want["A_t"], want["S_tpn"], want["key"])
            assert got.R_ttpB == pytest.approx(want["R"], rel=1e-12)
            assert got.Gamma_ttpB == pytest.approx(want["Gamma"], rel=1e-12)
            assert list(got.qS_t) == want["qS_t"] and list(got.qS_tpn) == want["qS_tpn"]


def test_survey_known_answers(golden_dir):
    case = _load(golden_dir, "nstep_literal.json")[0]        # ExperienceBuffer(3, 1), r = 10^t
    e = case["emitted"]
    assert len(e) == 2 and e[0]["key"] == "10" and e[1]["key"] == "11"
    assert e[0]["R"] == pytest.approx(118.81) and e[0]["Gamma"] == pytest.approx(0.9801)
    assert (e[0]["S_t"], e[0]["S_tpn"], e[1]["S_t"], e[1]["S_tpn"]) == (0, 3, 4, 7)
    # intended value for the same rewards: r0 + g r1 + g^2 r2 = 108.91, Gamma = g^3
    R, G, span = no.nstep_intended(np.array([[1, 10, 100, 1000]], np.float64), np.zeros((1, 4)), 3, 0.99, np.float64)
    assert R[0, 0] == pytest.approx(108.91) and G[0, 0] == pytest.approx(0.970299) and span[0, 0] == 3


def test_nstep_intended_terminal_and_tail():
    r = np.array([[1, 2, 3, 4, 5, 6]], np.float32)
    d = np.array([[0, 0, 1, 0, 0, 0]], bool)
    R, G, span = no.nstep_intended(r, d, 3, 0.5)
    assert R[0].tolist() == [1 + 1 + 0.75, 2 + 1.5, 3, 4 + 2.5 + 1.5, 0, 0]
    assert G[0].tolist() == [0, 0, 0, 0.125, 0, 0]
    assert span[0].tolist() == [3, 2, 1, 3, 0, 0]


def test_initial_priority_matches_reference(golden_dir):
    for case in _load(golden_dir, "initial_priority.json"):
        got = no.initial_priority(case["q_t"], case["q_tpn"], case["A"], case["R"], case["Gamma"])
        np.testing.assert_allclose(got, case["per_row_abs_td"], rtol=1e-12, atol=1e-12)
        # the literal dict maps every key to the LAST row's |TD|
        assert all(v == pytest.approx(case["per_row_abs_td"][-1]) for v in case["literal"].values())
    first = _load(golden_dir, "initial_priority.json")[0]
    np.testing.assert_allclose(first["per_row_abs_td"], [3.9005, 2.0199], rtol=1e-6)


# ------------------------------------------------------------- sampling (a5/a6)
def test_choice_law_matches_numpy_choice(golden_dir):
    z = np.load(os.path.join(golden_dir, "choice_law.npz"))
    got = ro.reference_law_sample(z["priorities"], float(z["alpha"]), z["u"])
    np.testing.assert_array_equal(got, z["picks"])


def test_choice_mechanism_through_reference_sample(golden_dir):
    g = _load(golden_dir, "replay_literal.json")
    ls = g["literal_sample"]
    cdf = np.cumsum(ls["probs"])
    cdf /= cdf[-1]
    np.testing.assert_array_equal(np.searchsorted(cdf, ls["u"], side="right"), ls["picked"])
    assert g["literal_probs_1_2_4"] == pytest.approx([1 / 3] * 3)          # Q1: literal = uniform
    assert (g["size_after_add8"], g["size_after_trim"], g["priorities_after_trim"]) == (8, 5, 8)
    assert len(g["sample6_after_trim_keys"]) < 6                           # Q5/Q6


def test_fixed_point_tree_agrees_with_fp64_law(golden_dir):
    """SURVEY H1 regression: u64 fixed-point leaves reproduce the fp64 choice law."""
    z = np.load(os.path.join(golden_dir, "choice_law.npz"))
    leaves = ro.priority_to_leaf(ro.pow_priority(z["priorities"], float(z["alpha"])))
    got = ro.tree_sample(leaves, z["u"])
    assert int((got != z["picks"]).sum()) == 0
    rng = np.random.default_rng(0)
    pri = np.abs(rng.standard_normal(1 << 18)) + 1e-3
    u = rng.random(50_000)
    leaves = ro.priority_to_leaf(ro.pow_priority(pri, 0.6))
    a = ro.tree_sample(leaves, u)
    b = ro.reference_law_sample(pri.astype(np.float32).astype(np.float64), 0.6, u)
    # fp32 pow + 2^-24 quantisation move a cdf boundary across a draw only rarely
    assert (a != b).mean() < 2e-3 and np.abs(a - b).max() <= 1


@settings(max_examples=60, deadline=None)
@given(st.lists(st.integers(0, 1 << 30), min_size=1, max_size=300), st.lists(st.floats(0, 1, exclude_max=True), min_size=1, max_size=20))
def test_descent_equals_level_walk(leaves, us):
    leaves = np.array(leaves, np.uint64)
    if leaves.sum() == 0:
        leaves[0] = 1
    u = np.array(us)
    idx = ro.tree_sample(leaves, u)
    assert (leaves[idx] > 0).all()                                         # zero leaves never picked
    levels = ro.build_levels(leaves)
    assert int(levels[-1].sum()) == int(leaves.sum())
    # walk the 32-ary tree the way the kernel does and compare
    tgt = ro.sample_target(u, int(leaves.sum()))
    for t, want in zip(tgt.tolist(), idx.tolist()):
        node = 0
        for lvl in reversed(levels):
            grp = lvl[node * 32:(node + 1) * 32]
            c = int(np.searchsorted(np.cumsum(grp, dtype=np.uint64), np.uint64(t), side="right"))
            t -= int(grp[:c].sum())
            node = node * 32 + c
        assert node == want


def test_leaf_format():
    assert ro.priority_to_leaf(np.float32(1.0)) == 1 << 24
    assert ro.priority_to_leaf(np.float32(1e9)) == ro.LEAF_MAX
    assert ro.priority_to_leaf(np.float32(np.nan)) == 0 and ro.priority_to_leaf(np.float32(-1)) == 0
    assert ro.sample_target(np.array([0.0, 0.999999999999]), 10).tolist() == [0, 9]
    assert ro.sample_target(np.array([1.0 - 2 ** -53]), (1 << 60) + 1)[0] <= (1 << 60)


# ------------------------------------------------------------------ ring (a3/a6)
def test_ring_stream_layout_and_dead_zone():
    C, n = 4, 3
    W = C - 1 + n
    ring = ro.RingOracle(capacity=64, frames_per_state=C, frame_bytes=16, lanes=2, fpr=1, lookback=W, alpha=1.0)
    rng = np.random.default_rng(1)
    rel = np.concatenate([np.arange(-n - C + 1, -n + 1), np.arange(-C + 1, 1)])[None, :].astype(np.int32)
    for it in range(9):
        m = 5
        fr = rng.integers(0, 256, (m, 16), dtype=np.uint8)
        slots = ring.insert(it % 2, fr, np.repeat(rel, m, 0), np.arange(m), np.ones(m), np.ones(m), prio=np.ones(m))
        assert ((slots // 32) == it % 2).all()
    assert ring.size() == 45
    # lane 0 got 25 records: positions 25..31 never written; dead zone = next W positions after head
    head0 = 25
    assert (ring.leaves[head0:head0 + W] == 0).all() and (ring.leaves[:head0] == 1 << 24).all()
    b = ring.gather(np.array([10]))
    assert b["S_tpn"].shape == (1, C, 16)
    np.testing.assert_array_equal(b["S_tpn"][0], ring.frames[7:11])
    np.testing.assert_array_equal(b["S_t"][0], ring.frames[4:8])
    # wrap: 8 more inserts of 5 in lane 0 -> 65 records in a 32-ring
    for it in range(8):
        ring.insert(0, rng.integers(0, 256, (5, 16), dtype=np.uint8), np.repeat(rel, 5, 0), np.arange(5), np.ones(5), np.ones(5), prio=np.ones(5))
    head = 65 % 32
    dead = [(head + w) % 32 for w in range(W)]
    assert (ring.leaves[dead] == 0).all()
    live = [p for p in range(32) if p not in dead]
    assert (ring.leaves[live] > 0).all()
    assert (ring.fidx[:32] >= 0).all() and (ring.fidx[:32] < 32).all()


def test_ring_last_writer_wins_and_total():
    ring = ro.RingOracle(32, 1, 16, alpha=1.0)
    ring.insert(0, np.zeros((8, 16), np.uint8), np.zeros((8, 2), np.int32), np.zeros(8), np.zeros(8), np.zeros(8), prio=np.ones(8))
    ring.set_leaves([3, 5, 3], [10, 20, 30])
    assert ring.leaves[3] == 30 and ring.leaves[5] == 20
    assert ring.total() == 6 * (1 << 24) + 50


# --------------------------------------------------------------------- TD (a8)
def test_td_oracle_vs_torch_fp64():
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(3)
    B, A = 257, 6
    q1, q2, q3 = (rng.standard_normal((B, A)).astype(np.float32) for _ in range(3))
    act = rng.integers(0, A, B)
    R = rng.standard_normal(B).astype(np.float32)
    G = np.where(rng.random(B) < 0.1, 0, 0.970299).astype(np.float32)
    for kind in (to.LOSS_HALF_SQ, to.LOSS_HUBER):
        o = to.td_forward(q1, q2, q3, act, R, G, alpha=0.6, loss_kind=kind)
        tq1 = torch.tensor(q1, dtype=torch.float64, requires_grad=True)
        tq2, tq3 = torch.tensor(q2, dtype=torch.float64), torch.tensor(q3, dtype=torch.float64)
        with torch.no_grad():                                            # learner.py:43-45
            g = torch.tensor(R, dtype=torch.float64) + torch.tensor(G, dtype=torch.float64) * \
                tq3.gather(1, torch.argmax(tq2, 1).view(-1, 1)).squeeze()
        qsa = tq1.gather(1, torch.from_numpy(act).reshape(-1, 1)).squeeze()   # learner.py:46
        td = g - qsa
        loss = 0.5 * td ** 2 if kind == to.LOSS_HALF_SQ else torch.nn.functional.huber_loss(qsa, g, reduction="none", delta=1.0)
        loss.mean().backward()
        np.testing.assert_allclose(o["delta"], td.detach().numpy(), rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(o["loss"], loss.detach().numpy(), rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(o["loss_mean"], loss.mean().item(), rtol=1e-5)
        np.testing.assert_allclose(o["prio"], np.abs(td.detach().numpy()) ** 0.6, rtol=1e-5, atol=1e-7)
        np.testing.assert_allclose(to.td_backward(o["delta"], act, A, loss_kind=kind), tq1.grad.numpy(), rtol=1e-5, atol=1e-8)


def test_td_oracle_matches_reference_learner_function(golden_dir):
    """td_reference.npz: outputs of the reference's own Learner.compute_loss_and_priorities
    (learner.py:29-52), run by make_golden.py under an operator shim that restores torch-0.4's
    ndarray coercion.  Per-row |TD| read from the returned dict by rotating the batch."""
    z = np.load(os.path.join(golden_dir, "td_reference.npz"))
    for c in range(int(z["cases"])):
        o = to.td_forward(z["c%d_q_on_t" % c], z["c%d_q_on_tpn" % c], z["c%d_q_tg_tpn" % c], z["c%d_A" % c],
                          z["c%d_R" % c], z["c%d_Gamma" % c], alpha=0.6)
        np.testing.assert_allclose(o["abs_td"], z["c%d_abs_td" % c], rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(o["loss_mean"], z["c%d_loss_mean" % c], rtol=1e-5)


def test_shard_split_equals_single_tree():
    rng = np.random.default_rng(5)
    shards = [rng.integers(0, 1 << 28, 1000).astype(np.uint64) for _ in range(4)]
    u = rng.random(2000)
    whole = ro.tree_sample(np.concatenate(shards), u)
    t, shard, local_t = ro.shard_split([int(s.sum(dtype=np.uint64)) for s in shards], u)
    got = np.array([g * 1000 + ro.descend(shards[g], np.array([lt]))[0] for g, lt in zip(shard, local_t)])
    np.testing.assert_array_equal(got, whole)


def test_c_oracle_matches_numpy_oracle():
    """oracle/apex_oracle.c (the timed CPU baseline) == the numpy oracle, bit for bit on indices/bytes."""
    from oracle.c_oracle import CpuReplay
    rng = np.random.default_rng(17)
    N, C_, F, B, nA = 200_000, 2, 64, 256, 4
    ring = ro.RingOracle(N, C_, F, lanes=4, fpr=1, lookback=C_ - 1 + 3, alpha=0.6)
    for lane in range(4):
        m = 40_000
        rel = np.tile(np.array([[-4, -3, -1, 0]], np.int32), (m, 1))
        ring.insert(lane, rng.integers(0, 256, (m, F), dtype=np.uint8), rel, rng.integers(0, nA, m), rng.standard_normal(m),
                    np.full(m, 0.97), prio=np.abs(rng.standard_normal(m)) + 1e-3)
    cpu = CpuReplay(ring.leaves.copy(), ring.frames, ring.fidx, ring.A, ring.R, ring.G, C_, F, 0.6)
    assert cpu.threads >= 1
    for step in range(4):
        u = ro.stratified_uniforms(B, rng) if step % 2 else rng.random(B)
        q = [rng.standard_normal((B, nA)).astype(np.float32) for _ in range(3)]
        want_idx = ring.sample(u)
        g = ring.gather(want_idx)
        o = cpu.step(u, *q)
        np.testing.assert_array_equal(o["idx"], want_idx)
        np.testing.assert_array_equal(o["S_t"], g["S_t"])
        np.testing.assert_array_equal(o["S_tpn"], g["S_tpn"])
        t = to.td_forward(q[0], q[1], q[2], g["A"], g["R"], g["G"], alpha=0.6)
        np.testing.assert_array_equal(o["delta"], t["delta"])
        np.testing.assert_allclose(o["prio"], t["prio"], rtol=1e-6)
        np.testing.assert_allclose(o["loss_mean"], t["loss_mean"], rtol=1e-6)
        ring.set_leaves(want_idx, ro.priority_to_leaf(o["prio"]))
        np.testing.assert_array_equal(cpu.leaves, ring.leaves)


def test_pack_oracle_round_trip_and_format():
    """oracle/pack_oracle.py ((f).4 lossless frame packing): unpack(pack(x)) == x; a frame equal to its predecessor
    costs only its bitmap; key frames every 64 frames are stored whole; random frames grow by the bitmap only."""
    from oracle import pack_oracle as po
    rng = np.random.default_rng(3)
    for F, count in ((7056, 130), (16, 5), (528, 64)):
        nb = F // 16
        head = ((nb + 31) // 32 * 4 + 15) // 16 * 16
        fr = rng.integers(0, 256, (count, F), dtype=np.uint8)
        fr[3] = fr[2]                                                    # an unchanged frame
        sizes, offsets, packed = po.pack_frames(fr)
        assert (po.unpack_frames(sizes, offsets, packed, F) == fr).all()
        assert sizes[0] == head + F and sizes[3] == head
        if count > 64:
            assert sizes[64] == head + F
        assert (offsets % 16 == 0).all() and packed.size == int(sizes.sum())
