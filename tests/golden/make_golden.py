#!/usr/bin/env python
"""Generate the golden vectors under tests/golden/ by EXECUTING the reference.

Run in the build container only (it imports /root/reference, which does not
exist on the GPU box):   python tests/golden/make_golden.py

The reference ships no tests or fixtures, so these files are the
only pinned outputs: every value below is produced by the unmodified reference
classes (replay.ReplayMemory, actor.ExperienceBuffer, Actor.compute_priorities)
or by the numpy routine the reference calls (np.random.choice, replay.py:52).
"""
import json
import os
import sys
import types

import numpy as np

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
sys.dont_write_bytecode = True
sys.path.insert(0, REF)
sys.modules.setdefault("gym", types.ModuleType("gym"))      # env.py:1 imports gym; never called here
if not hasattr(np, "int"):
    np.int = int                                            # actor.py:133

import replay as ref_replay          # noqa: E402
import actor as ref_actor            # noqa: E402


def nstep_cases():
    cases = []
    specs = [
        dict(n=3, actor_id=1, steps=9, done_at=None),
        dict(n=3, actor_id=7, steps=1, done_at=0),
        dict(n=1, actor_id=0, steps=6, done_at=None),
        dict(n=5, actor_id=12, steps=20, done_at=13),
        dict(n=3, actor_id=3, steps=14, done_at=4),
        dict(n=2, actor_id=5, steps=11, done_at=10),
    ]
    for sp in specs:
        buf = ref_actor.ExperienceBuffer(sp["n"], sp["actor_id"])
        feed = []
        for t in range(sp["steps"]):
            r = float(10 ** (t % 4)) if sp["n"] == 3 and sp["actor_id"] == 1 else float((t * 7) % 5 - 2) * 0.5
            q = (np.arange(4, dtype=np.float32) + t).tolist()
            feed.append(dict(S=t, A=t % 4, R=r, Gamma=0.99, q=q, done=(sp["done_at"] == t)))
            buf.add(ref_actor.Transition(t, t % 4, r, 0.99, np.array(q, np.float32)))
            if sp["done_at"] == t:     # actor.py:172-174: flush with (next obs, reward 0, stale q)
                buf.construct_nstep_transition(ref_actor.Transition(1000 + t, t % 4, 0, 0.99, np.array(q, np.float32)))
        out = [dict(S_t=x.S_t, A_t=x.A_t, R=float(x.R_ttpB), Gamma=float(x.Gamma_ttpB),
                    qS_t=np.asarray(x.qS_t).tolist(), S_tpn=x.S_tpn, qS_tpn=np.asarray(x.qS_tpn).tolist(),
                    key=x.key) for x in buf.local_nstep_buffer]
        cases.append(dict(spec=sp, feed=feed, emitted=out))
    return cases


def priority_cases():
    rng = np.random.RandomState(7)
    NT = ref_actor.N_Step_Transition
    actor_cls = ref_actor.Actor
    stub = types.SimpleNamespace()                       # compute_priorities uses no instance state
    cases = []
    for case in range(4):
        m = [2, 5, 5, 17][case]
        recs = []
        for i in range(m):
            recs.append(NT(None, int(rng.randint(4)), float(rng.randn()), float(0.99 ** 3 if rng.rand() > .2 else 0.0),
                           rng.randn(4).astype(np.float32), None, rng.randn(4).astype(np.float32), "k%d" % i))
        if case == 0:   # SURVEY §8(c) vector (3)
            recs = [NT(None, 1, 1.0, .9801, np.array([1, 2, 3, 4], np.float32), None, np.array([0, 5, 1, 1], np.float32), "10"),
                    NT(None, 2, 0.0, .9801, np.array([1, 2, 3, 4], np.float32), None, np.array([0, 1, 1, 1], np.float32), "11")]
        literal = actor_cls.compute_priorities(stub, recs)            # every key -> last |TD| (Q12)
        per_row = [float(list(actor_cls.compute_priorities(stub, [r]).values())[0]) for r in recs]
        cases.append(dict(A=[r.A_t for r in recs], R=[r.R_ttpB for r in recs], Gamma=[r.Gamma_ttpB for r in recs],
                          q_t=[r.qS_t.tolist() for r in recs], q_tpn=[r.qS_tpn.tolist() for r in recs],
                          keys=[r.key for r in recs],
                          literal={k: float(v) for k, v in literal.items()}, per_row_abs_td=per_row))
    return cases


def replay_cases():
    NT = ref_replay.N_Step_Transition
    out = {}
    # (4) literal probabilities are uniform; (6) trimmed keys stay sampleable.
    mem = ref_replay.ReplayMemory(5, {"priority_exponent": 0.6})
    mem.set_priorities({"a": np.float64(1), "b": np.float64(2), "c": np.float64(4)})
    out["literal_probs_1_2_4"] = [float(mem.sample_probabilities[k]) for k in "abc"]
    mem = ref_replay.ReplayMemory(5, {"priority_exponent": 0.6})
    recs = [NT(i, i % 4, float(i), 0.97, None, i + 3, None, "k%d" % i) for i in range(8)]
    mem.add({r.key: np.float64(1 + r.S_t) for r in recs}, recs)
    out["size_after_add8"] = mem.size()
    mem.remove_to_fit()
    out["size_after_trim"] = mem.size()
    out["priorities_after_trim"] = len(mem.priorities)
    np.random.seed(11)
    out["sample6_after_trim_keys"] = [r.key for r in mem.sample(6)]
    # choice mechanism through the real ReplayMemory.sample: same stream, same picks
    mem = ref_replay.ReplayMemory(100, {"priority_exponent": 0.6})
    recs = [NT(i, 0, 0.0, 0.97, None, i, None, "k%d" % i) for i in range(37)]
    mem.add({r.key: np.float64(1 + (i % 5)) for i, r in enumerate(recs)}, recs)
    np.random.seed(5)
    picked = [int(r.key[1:]) for r in mem.sample(64)]
    np.random.seed(5)
    u = np.random.random_sample(64)
    out["literal_sample"] = dict(n=37, picked=picked, u=u.tolist(),
                                 probs=[float(v) for v in mem.sample_probabilities.values()])
    # equal priorities: literal (uniform, Q1) and intended (proportional) laws coincide, so the reference's
    # own add + sample can be compared end to end with the drop-in class on the same seed
    mem = ref_replay.ReplayMemory(100, {"priority_exponent": 0.6})
    recs = [NT(i, i % 4, float(i) / 7, 0.97, None, i + 1, None, "e%d" % i) for i in range(50)]
    for c in range(0, 50, 5):
        mem.add({r.key: np.float64(2.5) for r in recs[c:c + 5]}, recs[c:c + 5])
    np.random.seed(2024)
    got = mem.sample(40)
    out["equal_priority_sample"] = dict(n=50, seed=2024, keys=[r.key for r in got], A=[r.A_t for r in got],
                                        R=[r.R_ttpB for r in got], S_t=[r.S_t for r in got], S_tpn=[r.S_tpn for r in got])
    return out


def choice_law():
    """np.random.choice(K, p=P) for intended P = p^alpha / sum, with the uniform stream it consumed."""
    rs = np.random.RandomState(123)
    K, B, alpha = 5000, 4096, 0.6
    pri = np.abs(rs.randn(K)) + 1e-3
    P = pri ** alpha
    P /= P.sum()
    np.random.seed(99)
    picks = np.array([np.random.choice(K, p=P) for _ in range(B)], np.int64)   # replay.py:52 call shape
    np.random.seed(99)
    u = np.random.random_sample(B)
    np.savez_compressed(os.path.join(HERE, "choice_law.npz"), priorities=pri, alpha=alpha, u=u, picks=picks)


def td_cases():
    """Learner.compute_loss_and_priorities (learner.py:29-52) executed for real.

    learner.py:44 multiplies a numpy array by a torch Tensor, which torch 0.4 (the reference's pin)
    coerced and torch 2.x refuses.  The only thing changed to run it is an operator
    shim that restores that coercion (ndarray operand -> torch.from_numpy) for Tensor.__rmul__ /
    __radd__; learner.py itself is imported unmodified.  The three Q tensors the function computes
    are captured with forward hooks; the returned dict gives every key the LAST row's |TD| (Q12), so
    the batch is rotated B times to read every row's |TD| (the duelling head's batch-wide advantage
    sum, Q19, is invariant under row permutation).
    """
    import torch
    import learner as ref_learner

    def coerce(fn):
        def wrapped(self, other):
            if isinstance(other, np.ndarray):
                other = torch.from_numpy(other)
            return fn(self, other)
        return wrapped
    saved = torch.Tensor.__rmul__, torch.Tensor.__radd__
    torch.Tensor.__rmul__, torch.Tensor.__radd__ = coerce(saved[0]), coerce(saved[1])
    out = {}
    try:
        torch.manual_seed(1234)
        rng = np.random.RandomState(5)
        NT = ref_learner.N_Step_Transition
        L = ref_learner.Learner({"state_shape": [1, 84, 84], "action_dim": 4},
                                {"load_saved_state": False, "q_target_sync_freq": 2500}, {}, None)
        for p_, q_ in zip(L.Q_double.parameters(), L.Q.parameters()):       # a target net that differs from the online net
            p_.data.copy_(q_.data + 0.01 * torch.randn_like(q_))
        captured = []
        hooks = [m.register_forward_hook(lambda mod, inp, outp, tag=tag: captured.append((tag, outp[2].detach().numpy().copy())))
                 for tag, m in (("online", L.Q), ("target", L.Q_double))]
        for case, B in enumerate((8, 5)):
            recs = [NT(rng.rand(1, 84, 84) * 4, int(rng.randint(4)), float(rng.randn()),
                       float(0.99 ** 3 if rng.rand() > 0.25 else 0.0), None, rng.rand(1, 84, 84) * 4, None, "k%d" % i)
                    for i in range(B)]
            abs_td = np.zeros(B)
            for rot in range(B):
                batch = recs[rot + 1:] + recs[:rot + 1]                      # row `rot` is last
                del captured[:]
                loss, prios = L.compute_loss_and_priorities(batch)
                assert len(set(float(v) for v in prios.values())) == 1       # Q12: one value for every key
                abs_td[rot] = float(list(prios.values())[0])
                if rot == B - 1:                                             # batch in original order
                    tags = [t for t, _ in captured]
                    assert tags == ["target", "online", "online"], tags     # Q_double(S_tpn), Q(S_tpn), Q(S_t)
                    out["c%d_q_tg_tpn" % case], out["c%d_q_on_tpn" % case], out["c%d_q_on_t" % case] = (c for _, c in captured)
                    out["c%d_loss_mean" % case] = np.float64(loss.item())
            out["c%d_A" % case] = np.array([r.A_t for r in recs], np.int32)
            out["c%d_R" % case] = np.array([r.R_ttpB for r in recs], np.float64)
            out["c%d_Gamma" % case] = np.array([r.Gamma_ttpB for r in recs], np.float64)
            out["c%d_abs_td" % case] = abs_td
        for h_ in hooks:
            h_.remove()
    finally:
        torch.Tensor.__rmul__, torch.Tensor.__radd__ = saved
    out["cases"] = np.int32(2)
    np.savez_compressed(os.path.join(HERE, "td_reference.npz"), **out)


def policy_cases():
    """The actor's policy executed for real: DuellingDQN at batch 1 (actor.py:160-161 ->
    duelling_network.py:21-28) and Actor.epsilon_greedy_Q (actor.py:121-125) on the `random`
    stream it consumes.  A second random.Random with the same seed replays that stream to record
    the uniform behind every explore decision and the action random.choice returned."""
    import random
    import torch
    from duelling_network import DuellingDQN
    out = {}
    for case, (A, E, eps, seed) in enumerate(((4, 24, 0.4, 9), (6, 17, 0.7, 31))):
        torch.manual_seed(100 + case)
        net = DuellingDQN((1, 84, 84), A)
        obs = torch.rand(E, 1, 84, 84) * 255
        with torch.no_grad():
            rows = [net(obs[i:i + 1]) for i in range(E)]                 # batch 1, as every actor does
        value = np.concatenate([r[0].numpy() for r in rows]).reshape(E)
        adv = np.concatenate([r[1].numpy() for r in rows])
        q = np.concatenate([r[2].numpy() for r in rows])
        q_fwd = q.copy()
        q[3, :] = q[3, 1]                                                # a row of ties: np.argmax -> first
        stub = types.SimpleNamespace(epsilon=eps)
        random.seed(seed)
        acts = [int(ref_actor.Actor.epsilon_greedy_Q(stub, q[i])) for i in range(E)]
        replay = random.Random(seed)
        u_explore, u_action = np.zeros(E), np.full(E, 0.5)
        for i in range(E):
            u_explore[i] = replay.random()
            if not (u_explore[i] >= eps):
                picked = replay.choice(list(range(A)))
                assert picked == acts[i]
                u_action[i] = (picked + 0.5) / A
        out.update({"c%d_value" % case: value, "c%d_adv" % case: adv, "c%d_q_forward" % case: q_fwd,
                    "c%d_q_policy" % case: q, "c%d_eps" % case: np.float64(eps),
                    "c%d_u_explore" % case: u_explore, "c%d_u_action" % case: u_action,
                    "c%d_actions" % case: np.array(acts, np.int32)})
    out["cases"] = np.int32(2)
    np.savez_compressed(os.path.join(HERE, "policy_reference.npz"), **out)


def main():
    policy_cases()
    td_cases()
    with open(os.path.join(HERE, "nstep_literal.json"), "w") as f:
        json.dump(nstep_cases(), f, indent=1)
    with open(os.path.join(HERE, "initial_priority.json"), "w") as f:
        json.dump(priority_cases(), f, indent=1)
    with open(os.path.join(HERE, "replay_literal.json"), "w") as f:
        json.dump(replay_cases(), f, indent=1)
    choice_law()
    print("golden vectors written to", HERE)


if __name__ == "__main__":
    main()
