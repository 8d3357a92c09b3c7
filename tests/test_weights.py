"""Weight board (K9) and batched actor policy (K10): SURVEY §8(f).2.

CPU part: the oracle against the reference's own outputs (tests/golden/policy_reference.npz) and the
shared-memory board protocol.  GPU part: the kernels against the oracle and the golden vectors."""
import os
import time

import numpy as np
import pytest
import torch

from oracle import policy_oracle as po


def _golden(golden_dir):
    return np.load(os.path.join(golden_dir, "policy_reference.npz"))


# ------------------------------------------------------------------------------- CPU
def test_policy_oracle_matches_reference(golden_dir):
    g = _golden(golden_dir)
    for c in range(int(g["cases"])):
        q = po.duelling_q_batch1(g["c%d_value" % c], g["c%d_adv" % c])
        np.testing.assert_allclose(q, g["c%d_q_forward" % c], rtol=1e-6, atol=1e-6)
        acts, explored = po.epsilon_greedy(g["c%d_q_policy" % c], float(g["c%d_eps" % c]), g["c%d_u_explore" % c],
                                           g["c%d_u_action" % c])
        np.testing.assert_array_equal(acts, g["c%d_actions" % c])
        assert 0 < explored.sum() < len(acts)                       # both branches of actor.py:122 are exercised


def test_policy_oracle_edges():
    q = np.array([[1, 3, 3, 2], [np.nan, 5, np.nan, 1], [0, 0, 0, 0]], np.float32)
    acts, ex = po.epsilon_greedy(q, [0.5, 0.5, 0.5], [0.5, 0.9, 0.49999], [0.0, 0.0, 0.999999])
    assert acts.tolist() == [1, 0, 3] and ex.tolist() == [False, False, True]   # u == eps is greedy; first max; NaN first
    assert po.actor_epsilons(0.4, 7, 8, [0, 7]).tolist() == pytest.approx([0.4, 0.4 ** 8], rel=1e-6)


def test_param_layout_matches_oracle_layout():
    import apex_b200
    from apex_b200.qnet import DuellingQNet
    sd = DuellingQNet((1, 84, 84), 4).state_dict()
    layout, total = apex_b200.param_layout(sd)
    want, want_total = po.flat_layout([(tuple(t.shape), t.element_size()) for t in sd.values()])
    assert [(o, n) for (_, _, _, o, n) in layout] == want and total == want_total
    assert all(o % 256 == 0 for o, _ in want) and total >= sum(n for _, n in want)


def _board_writer(name, flat_bytes, n_tensors, versions):
    """Stands in for the library's copy stream + host callback: same control-word protocol."""
    from apex_b200.weights import _Board
    b = _Board(name, flat_bytes, n_tensors, create=False)
    for v in range(1, versions + 1):
        b.ctrl[1] = v
        b.slots[v & 1][:] = v % 251                                  # the whole slot carries one value
        b.ctrl[0] = v
        time.sleep(0.002)
    b.close()


def test_board_reads_are_never_torn():
    """Seqlock protocol of the shared-memory board: a reader racing a writer only ever sees whole versions."""
    import multiprocessing as mp
    from apex_b200.weights import _Board
    flat_bytes, versions = 1 << 20, 60
    name = "apexboard_test_%d" % os.getpid()
    board = _Board(name, flat_bytes, 3, create=True)
    try:
        with pytest.raises(ValueError):
            _Board(name, flat_bytes + 256, 3, create=False)          # another model's layout is refused
        p = mp.get_context("spawn").Process(target=_board_writer, args=(name, flat_bytes, 3, versions))
        p.start()
        scratch = np.empty(flat_bytes, np.uint8)
        seen, last, t0 = 0, 0, time.time()
        while last < versions and time.time() - t0 < 60:
            v = board.read_consistent(lambda slot: np.copyto(scratch, slot))
            if v:
                assert v >= last and scratch.min() == scratch.max() == v % 251
                seen += v != last
                last = v
        p.join(timeout=30)
        assert last == versions and seen >= 2 and p.exitcode == 0
    finally:
        board.close(unlink=True)


def test_subscriber_loads_a_published_state_dict_on_cpu():
    """WeightSubscriber (the CPU actor side) needs no CUDA: a board filled by hand loads into a module."""
    import apex_b200
    from apex_b200.weights import _Board
    from apex_b200.qnet import DuellingQNet
    torch.manual_seed(3)
    src, dst = DuellingQNet((1, 84, 84), 4), DuellingQNet((1, 84, 84), 4)
    layout, total = apex_b200.param_layout(src.state_dict())
    name = "apexboard_sub_%d" % os.getpid()
    board = _Board(name, total, len(layout), create=True)
    try:
        sub = apex_b200.WeightSubscriber(dst, name)
        assert sub.poll() is False                                   # nothing published yet
        flat = po.pack_flat([t.numpy() for t in src.state_dict().values()])
        board.ctrl[1] = 1
        board.slots[1][:] = flat
        board.ctrl[0] = 1
        assert sub.poll() is True and sub.poll() is False
        for a, b in zip(src.state_dict().values(), dst.state_dict().values()):
            assert torch.equal(a, b)
        sub.close()
    finally:
        board.close(unlink=True)


# ------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
def test_actor_policy_kernel_matches_reference_golden(golden_dir):
    import apex_b200
    g = _golden(golden_dir)
    dev = torch.device("cuda")
    for c in range(int(g["cases"])):
        E = len(g["c%d_actions" % c])
        eps = torch.full((E,), float(g["c%d_eps" % c]), device=dev)
        # duelling combine: the batched kernel reproduces E batch-1 forwards of the reference model
        _, q, _ = apex_b200.actor_policy(torch.tensor(g["c%d_adv" % c], device=dev), eps,
                                         value=torch.tensor(g["c%d_value" % c], device=dev),
                                         u_explore=torch.zeros(E, dtype=torch.float64, device=dev) + 2.0,
                                         u_action=torch.zeros(E, dtype=torch.float64, device=dev))
        np.testing.assert_allclose(q.cpu().numpy(), g["c%d_q_forward" % c], rtol=1e-6, atol=1e-6)
        np.testing.assert_array_equal(q.cpu().numpy(), po.duelling_q_batch1(g["c%d_value" % c], g["c%d_adv" % c]))
        # epsilon-greedy on the reference's own random stream
        acts, _, ex = apex_b200.actor_policy(torch.tensor(g["c%d_q_policy" % c], device=dev), eps,
                                             u_explore=torch.tensor(g["c%d_u_explore" % c], device=dev),
                                             u_action=torch.tensor(g["c%d_u_action" % c], device=dev), want_explored=True)
        np.testing.assert_array_equal(acts.cpu().numpy(), g["c%d_actions" % c])
        assert 0 < int(ex.sum()) < E


@pytest.mark.gpu
@pytest.mark.parametrize("E,A", [(1, 2), (257, 4), (5000, 18), (64, 64)])
def test_actor_policy_kernel_matches_oracle(E, A):
    import apex_b200
    rng = np.random.default_rng(E * 100 + A)
    dev = torch.device("cuda")
    value = rng.standard_normal(E).astype(np.float32)
    adv = rng.standard_normal((E, A)).astype(np.float32)
    adv[::7] = np.round(adv[::7])                                    # ties
    if E > 3:
        adv[3, A // 2] = np.nan
    eps = rng.random(E).astype(np.float32)
    ue, ua = rng.random(E), rng.random(E)
    ue[::5] = eps[::5].astype(np.float64)                            # u == eps: greedy
    q_want = po.duelling_q_batch1(value, adv)
    acts_want, ex_want = po.epsilon_greedy(q_want, eps, ue, ua)
    acts, q, ex = apex_b200.actor_policy(torch.tensor(adv, device=dev), torch.tensor(eps, device=dev),
                                         value=torch.tensor(value, device=dev), u_explore=torch.tensor(ue, device=dev),
                                         u_action=torch.tensor(ua, device=dev), want_explored=True)
    np.testing.assert_array_equal(q.cpu().numpy(), q_want)
    np.testing.assert_array_equal(acts.cpu().numpy(), acts_want)
    np.testing.assert_array_equal(ex.cpu().numpy().astype(bool), ex_want)
    # in-kernel uniforms: same decisions as the oracle fed with the restated generator
    ue2, ua2 = po.kernel_uniforms(1234, 9, E)
    acts2, _, ex2 = apex_b200.actor_policy(torch.tensor(q_want, device=dev), torch.tensor(eps, device=dev), seed=1234, step=9,
                                           want_explored=True)
    w2, e2 = po.epsilon_greedy(q_want, eps, ue2, ua2)
    np.testing.assert_array_equal(acts2.cpu().numpy(), w2)
    np.testing.assert_array_equal(ex2.cpu().numpy().astype(bool), e2)


@pytest.mark.gpu
def test_actor_policy_argument_checks():
    import apex_b200
    from apex_b200 import _C
    dev = torch.device("cuda")
    q = torch.zeros((4, 65), device=dev)
    with pytest.raises(_C.ApexError):
        apex_b200.actor_policy(q, torch.zeros(4, device=dev))        # A > APEXB200_MAX_ACTIONS
    with pytest.raises(_C.ApexError):
        apex_b200.actor_policy(torch.zeros((4, 4), device=dev), torch.zeros(4, device=dev),
                               u_explore=torch.zeros(4, dtype=torch.float64, device=dev))   # one uniform array only


def _cpu_actor(name, state_shape, A, want_versions, out_q):
    """A reference-style CPU actor process: no CUDA, polls the board, reports a checksum per version."""
    import apex_b200
    from apex_b200.qnet import DuellingQNet
    torch.set_num_threads(1)
    net = DuellingQNet(state_shape, A)
    sub = apex_b200.WeightSubscriber(net, name)
    t0, got = time.time(), 0
    while got < want_versions and time.time() - t0 < 120:
        if sub.poll():
            out_q.put((sub.seen, float(sum(t.double().sum() for t in net.state_dict().values()))))
            got += 1
        else:
            time.sleep(0.001)
    sub.close()


@pytest.mark.gpu
def test_weight_board_learner_to_cpu_and_gpu_actors():
    """learner.py:74 -> actor.py:189-191 through the board: byte-exact state_dicts, versions in order,
    CPU subscriber in another process, device subscriber through the pinned slot and through the snapshot."""
    import multiprocessing as mp
    import apex_b200
    from apex_b200.qnet import DuellingQNet
    dev = torch.device("cuda")
    torch.manual_seed(11)
    shape, A = (1, 84, 84), 4
    learner_q = DuellingQNet(shape, A).to(dev)
    name = "apexboard_gpu_%d" % os.getpid()
    pub = apex_b200.WeightPublisher(learner_q, name=name, create=True)
    ctx = mp.get_context("spawn")
    out_q = ctx.Queue()
    try:
        sums = {}
        v1 = pub.publish()
        pub.wait(v1)
        sums[v1] = float(sum(t.double().sum() for t in learner_q.state_dict().values()))
        # the flat device snapshot is the oracle's packing of the same tensors
        flat_want = po.pack_flat([t.cpu().numpy() for t in learner_q.state_dict().values()])
        np.testing.assert_array_equal(pub.flat().cpu().numpy(), flat_want)
        np.testing.assert_array_equal(pub.board.slots[v1 & 1], flat_want)
        p = ctx.Process(target=_cpu_actor, args=(name, shape, A, 2, out_q))
        p.start()
        gpu_actor_q = DuellingQNet(shape, A).to(dev)
        dsub = apex_b200.DeviceWeightSubscriber(gpu_actor_q, name)
        assert dsub.poll() is True and dsub.poll() is False
        torch.cuda.synchronize()
        for a, b in zip(learner_q.state_dict().values(), gpu_actor_q.state_dict().values()):
            assert torch.equal(a, b)
        got = [out_q.get(timeout=120)]                               # the CPU actor has read version 1
        with torch.no_grad():                                        # an "optimizer step", then publish twice in a row
            for t in learner_q.parameters():
                t.add_(0.25)
        v2 = pub.publish()
        with torch.no_grad():
            for t in learner_q.parameters():
                t.mul_(0.5)                                          # must not leak into snapshot v2
        pub.wait(v2)
        want_v2 = flat_want.copy()
        layout, _ = apex_b200.param_layout(learner_q.state_dict())
        for (_, shp, dt, off, nb), t in zip(layout, learner_q.state_dict().values()):
            want_v2[off:off + nb] = ((t.cpu() / 0.5)).numpy().view(np.uint8).reshape(-1)
        np.testing.assert_array_equal(pub.board.slots[v2 & 1], want_v2)
        sums[v2] = float(sum((t.double() / 0.5).sum() for t in learner_q.state_dict().values()))
        got.append(out_q.get(timeout=120))
        p.join(timeout=30)
        assert p.exitcode == 0 and [v for v, _ in got] == [v1, v2]
        for v, s in got:
            assert s == pytest.approx(sums[v], rel=1e-9)
        # device-to-device path: scatter the publisher's snapshot (v2) straight into the GPU actor's model
        dsub.load_from(pub.flat())
        torch.cuda.synchronize()
        for a, b in zip(learner_q.state_dict().values(), gpu_actor_q.state_dict().values()):
            assert torch.equal(a / 0.5, b)
        dsub.close()
    finally:
        pub.close(unlink=True)


@pytest.mark.gpu
def test_weight_board_odd_tensors():
    """Tensors of odd sizes / dtypes (tails shorter than a vector, > one chunk, 1 byte)."""
    import apex_b200
    dev = torch.device("cuda")

    class Odd(torch.nn.Module):
        def __init__(self):
            super().__init__()
            g = torch.Generator().manual_seed(5)
            self.a = torch.nn.Parameter(torch.randn(3, generator=g))
            self.b = torch.nn.Parameter(torch.randn(100003, generator=g))
            self.register_buffer("c", torch.randint(0, 255, (33,), generator=g, dtype=torch.uint8))
            self.register_buffer("d", torch.randint(0, 1000, (7, 5), generator=g, dtype=torch.int64))
            self.e = torch.nn.Parameter(torch.randn(1, generator=g).to(torch.float16))

    src, dst = Odd().to(dev), Odd().to(dev)
    with torch.no_grad():
        for t in dst.state_dict().values():
            t.zero_()
    name = "apexboard_odd_%d" % os.getpid()
    pub = apex_b200.WeightPublisher(src, name=name, create=True)
    try:
        v = pub.publish()
        pub.wait(v)
        np.testing.assert_array_equal(pub.board.slots[v & 1], po.pack_flat([t.cpu().numpy() for t in src.state_dict().values()]))
        sub = apex_b200.DeviceWeightSubscriber(dst, name)
        assert sub.poll()
        torch.cuda.synchronize()
        for a, b in zip(src.state_dict().values(), dst.state_dict().values()):
            assert torch.equal(a, b)
        # a burst of publications without waiting (more than the library keeps callback slots for):
        # versions complete in order and the last slot holds the last state
        for k in range(20):
            with torch.no_grad():
                src.b.add_(1.0)
            last = pub.publish()
        pub.wait(last)
        assert pub.version == last == v + 20 and int(pub.board.ctrl[1]) == last
        np.testing.assert_array_equal(pub.board.slots[last & 1], po.pack_flat([t.cpu().numpy() for t in src.state_dict().values()]))
        assert sub.poll() and sub.seen == last
        torch.cuda.synchronize()
        assert torch.equal(src.b, dst.b)
        sub.close()
    finally:
        pub.close(unlink=True)


@pytest.mark.gpu
def test_batched_actor_policy_equals_batch1_actors():
    """BatchedActorPolicy for E actors == E reference-style actors at batch 1 (same weights, same uniforms)."""
    import apex_b200
    from apex_b200.qnet import DuellingQNet
    dev = torch.device("cuda")
    torch.manual_seed(21)
    E, A = 16, 4
    net = DuellingQNet((1, 84, 84), A).to(dev)
    params = {"epsilon": 0.4, "alpha": 7, "num_actors": E}
    pol = apex_b200.BatchedActorPolicy(net, params, range(E), seed=5)
    np.testing.assert_allclose(pol.eps.cpu().numpy(), po.actor_epsilons(0.4, 7, E, range(E)), rtol=1e-6)
    obs = torch.randint(0, 256, (E, 1, 84, 84), dtype=torch.uint8, device=dev)
    rng = np.random.default_rng(4)
    ue, ua = rng.random(E), rng.random(E)
    tf32 = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False     # compare fp32 with fp32
    try:
        acts, q = pol.act(obs, u_explore=torch.tensor(ue, device=dev), u_action=torch.tensor(ua, device=dev))
        with torch.no_grad():
            q1 = torch.cat([net(obs[i:i + 1].float())[2] for i in range(E)]).cpu().numpy()  # actor.py:160-161
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = tf32
    np.testing.assert_allclose(q.cpu().numpy(), q1, rtol=1e-3, atol=1e-3)                 # cuDNN picks algorithms per batch size
    want, _ = po.epsilon_greedy(q.cpu().numpy(), pol.eps.cpu().numpy(), ue, ua)
    np.testing.assert_array_equal(acts.cpu().numpy(), want)
    acts2, _ = pol.act(obs)                                                               # in-kernel uniforms, step 1
    ue2, ua2 = po.kernel_uniforms(5, 1, E)
    want2, _ = po.epsilon_greedy(q.cpu().numpy(), pol.eps.cpu().numpy(), ue2, ua2)
    np.testing.assert_array_equal(acts2.cpu().numpy(), want2)


class _CounterEnv:
    """Deterministic toy environment (gym protocol): every frame names (env, episode, step)."""

    def __init__(self, eid, episode_len):
        self.eid, self.L, self.ep, self.k = eid, episode_len, -1, 0
        self.log = []                                    # (frame tag, action, reward, done) per step

    def _obs(self):
        f = np.zeros((4, 4), np.uint8)
        f[0, 0], f[0, 1], f[0, 2] = self.eid, self.ep % 256, self.k % 256
        f[1:] = (self.eid * 31 + self.ep * 7 + self.k) % 251
        return f

    def reset(self):
        self.ep, self.k = self.ep + 1, 0
        return self._obs()

    def step(self, a):
        tag = (self.ep, self.k)
        self.k += 1
        r = float(((a + 1) * (self.k % 5) - 4) / 4.0)
        d = self.k >= self.L
        self.log.append((tag, int(a), r, d))
        return self._obs(), r, d, {}


class _TinyDuelling(torch.nn.Module):
    def __init__(self, C_, A):
        super().__init__()
        self.v = torch.nn.Linear(C_ * 16, 1)
        self.a = torch.nn.Linear(C_ * 16, A)

    def forward(self, x):
        x = x.reshape(x.shape[0], -1) / 255.0
        v, a = self.v(x), self.a(x)
        return v, a, v + (a - a.sum() / a.shape[1])


@pytest.mark.gpu
def test_vector_actor_end_to_end():
    """E lock-stepped actors: what lands in the replay is what the environments and the policy produced
    (frames, stacks with episode-start repeats, actions, sliding n-step returns, Gamma = 0 at terminals,
    initial priorities from the policy's own Q rows), and the actors' weights follow the learner's board."""
    import apex_b200 as apex
    dev = torch.device("cuda")
    torch.manual_seed(8)
    E, C_, A, n, gamma, lane_cap, T = 5, 3, 4, 3, 0.9, 256, 90
    mem = apex.ReplayMemory(E * lane_cap, {"priority_exponent": 0.6, "layout": "stream", "lanes": E, "n_step": n,
                                           "state_shape": (C_, 4, 4), "num_actions": 0})
    learner_q, actor_q = _TinyDuelling(C_, A).to(dev), _TinyDuelling(C_, A).to(dev)
    name = "apexboard_va_%d" % os.getpid()
    pub = apex.WeightPublisher(learner_q, name=name, create=True)
    try:
        pub.wait(pub.publish())
        sub = apex.DeviceWeightSubscriber(actor_q, name)
        envs = [_CounterEnv(e, 7 + 3 * e) for e in range(E)]
        params = {"epsilon": 0.4, "alpha": 7, "num_actors": E, "num_steps": n, "gamma": gamma, "Q_network_sync_freq": 10,
                  "n_step_transition_batch_size": 4}
        va = apex.VectorActor(envs, actor_q, mem, params, weights=sub, seed=3)
        seen_q = []
        act0 = va.policy.act
        va.policy.act = lambda obs, **kw: (lambda r: (seen_q.append((obs.clone(), r[1].clone())), r)[1])(act0(obs, **kw))
        assert va.run(50) == E * 50
        with torch.no_grad():
            for t in learner_q.parameters():
                t.add_(0.125)
        pub.wait(pub.publish())
        assert va.run(T - 50) == E * (T - 50)
        torch.cuda.synchronize()
        assert sub.seen == 2
        for a, b in zip(learner_q.state_dict().values(), actor_q.state_dict().values()):
            assert torch.equal(a, b)
        assert len(va.episode_returns) == sum(T // env.L for env in envs)
        A_d, R_d, G_d = (mem.buffer(w).cpu().numpy() for w in (apex._C.BUF_ACTION, apex._C.BUF_REWARD, apex._C.BUF_GAMMA))
        leaves = mem.buffer(apex._C.BUF_LEAVES).cpu().numpy().view(np.uint64)
        fidx = mem.buffer(apex._C.BUF_FIDX).cpu().numpy().reshape(-1, 2 * C_)
        store = mem.buffer(apex._C.BUF_FRAMES).cpu().numpy().reshape(-1, 16)
        qs = torch.stack([q for _, q in seen_q], 1).cpu().numpy()               # [E, T, A]: Q(s_tau) the policy used
        stacks = torch.stack([o for o, _ in seen_q], 1).cpu().numpy()           # [E, T, C, 4, 4]
        for e, env in enumerate(envs):
            assert len(env.log) == T
            buf = apex.ExperienceBuffer(n, e, semantics="sliding", gamma=gamma)
            for t, (tag, a, r, d) in enumerate(env.log):
                buf.add(apex.Transition(t, a, r, gamma, qs[e, t]))
                if d:
                    buf.construct_nstep_transition(apex.Transition(t + 1, 0, 0, gamma, qs[e, min(t + 1, T - 1)]))
            recs = {r.S_t: r for r in buf.get(buf.size)}
            for tau in range(n, T):
                t, slot = tau - n, e * lane_cap + tau
                r = recs[t]
                assert A_d[slot] == r.A_t == env.log[t][1]
                assert R_d[slot] == pytest.approx(r.R_ttpB, rel=1e-5, abs=1e-6)
                assert G_d[slot] == pytest.approx(r.Gamma_ttpB, rel=1e-5)
                want_p = abs(r.R_ttpB + r.Gamma_ttpB * float(qs[e, tau].max()) - float(qs[e, t, r.A_t]))
                assert leaves[slot] / float(1 << 24) == pytest.approx(want_p ** 0.6, rel=1e-4, abs=2e-7)
                # the state the learner will gather for this record is the state the policy saw at step t
                np.testing.assert_array_equal(store[fidx[slot, :C_]].reshape(C_, 4, 4), stacks[e, t])
                if r.Gamma_ttpB != 0.0:
                    np.testing.assert_array_equal(store[fidx[slot, C_:]].reshape(C_, 4, 4), stacks[e, tau])
        mem.check_status()
        sub.close()
    finally:
        pub.close(unlink=True)


@pytest.mark.gpu
def test_host_register_replaces_a_stale_registration():
    """A shared-memory mapping that went away without cudaHostUnregister leaves its virtual range registered;
    apexb200_host_register of a later mapping at the same address must still succeed (and leave no error behind)."""
    from apex_b200 import _C
    buf = np.zeros(1 << 20, np.uint8)
    addr = buf.ctypes.data & ~4095
    _C.host_register(addr + 4096, 1 << 18)
    _C.host_register(addr + 4096, 1 << 18)               # same range again: replaced, not an error
    _C.host_unregister(addr + 4096)
    with pytest.raises(_C.ApexError):
        _C.host_unregister(addr + 4096)                  # not registered any more
    x = torch.ones(8, device="cuda")                     # the CUDA error state is clean
    assert float(x.sum()) == 8.0


def test_policy_oracle_properties():
    """Size-independent properties of the policy restatement: eps = 0 is always greedy, eps = 1 always explores,
    the duelling combine is centred per row and invariant to what other rows hold (batch-1 semantics)."""
    rng = np.random.default_rng(0)
    for E, A in ((1, 2), (33, 4), (500, 18)):
        value, adv = rng.standard_normal(E).astype(np.float32), rng.standard_normal((E, A)).astype(np.float32)
        q = po.duelling_q_batch1(value, adv)
        np.testing.assert_allclose((q - value[:, None]).mean(1), 0, atol=2e-6)
        other = adv.copy()
        other[1:] += 100.0                                           # other rows change: row 0 must not (Q19)
        np.testing.assert_array_equal(po.duelling_q_batch1(value, other)[0], q[0])
        ue, ua = rng.random(E), rng.random(E)
        greedy, ex0 = po.epsilon_greedy(q, 0.0, ue, ua)
        np.testing.assert_array_equal(greedy, q.argmax(1))
        assert not ex0.any()
        rnd, ex1 = po.epsilon_greedy(q, 1.0, ue, ua)
        assert ex1.all() and (rnd == np.minimum((ua * A).astype(np.int64), A - 1)).all()
        assert rnd.min() >= 0 and rnd.max() < A
    u1, u2 = po.kernel_uniforms(7, 3, 1000)
    assert 0 <= u1.min() and u1.max() < 1 and 0 <= u2.min() and u2.max() < 1 and abs(u1.mean() - 0.5) < 0.05
    assert not np.array_equal(u1, u2) and not np.array_equal(u1, po.kernel_uniforms(7, 4, 1000)[0])
