"""CPU oracle: weight-board packing and the batched actor policy (TEST INFRASTRUCTURE).

Restates, for E environments at once, what every reference actor does at batch 1:
    q = value + (advantage - (1/A) * advantage.sum())     /root/reference/duelling_network.py:27
    greedy if random.random() >= epsilon else random.choice(range(A))   /root/reference/actor.py:121-125
    epsilon_i = eps ** (1 + alpha * i / (N - 1))          /root/reference/actor.py:111-114
and the flat layout of a state_dict on the weight board (learner.py:74 / actor.py:189-191 move the
same tensors as a pickle).  PINNED by tests/golden/policy_reference.npz: the reference model run at
batch 1 and Actor.epsilon_greedy_Q run on its own `random` stream (tests/golden/make_golden.py).
"""
from __future__ import annotations

import numpy as np


def duelling_q_batch1(value, adv):
    """fp32, in the reference's operation order; one row = one batch-1 forward."""
    value = np.asarray(value, np.float32).reshape(-1, 1)
    adv = np.asarray(adv, np.float32)
    A = adv.shape[1]
    s = np.zeros(adv.shape[0], np.float32)
    for a in range(A):                                     # sequential fp32 sum, as the kernel does
        s = (s + adv[:, a]).astype(np.float32)
    centre = (np.float32(1.0 / A) * s).astype(np.float32)
    return (value + (adv - centre[:, None])).astype(np.float32)


def epsilon_greedy(q, eps, u_explore, u_action):
    """-> (actions int32 [E], explored bool [E])."""
    q = np.asarray(q, np.float32)
    E, A = q.shape
    eps = np.broadcast_to(np.asarray(eps, np.float32), (E,))
    greedy = np.argmax(q, axis=1)                          # first maximum; NaN counts as largest
    explored = np.asarray(u_explore, np.float64) < eps.astype(np.float64)
    rnd = np.minimum((np.asarray(u_action, np.float64) * A).astype(np.int64), A - 1)
    return np.where(explored, rnd, greedy).astype(np.int32), explored


def kernel_uniforms(seed, step, E):
    """The uniforms apexb200_actor_policy draws itself: splitmix64(seed, step, e, which)."""
    M = (1 << 64) - 1
    out = np.zeros((2, E), np.float64)
    for which in (1, 2):
        for e in range(E):
            z = (seed ^ ((step * 0xD1342543DE82EF95) & M) ^ ((which * 0xA24BAED4963EE407) & M)) & M
            z = (z + (e + 1) * 0x9E3779B97F4A7C15) & M
            z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M
            z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M
            z ^= z >> 31
            out[which - 1, e] = float(z >> 11) * 2.0 ** -53
    return out[0], out[1]


def actor_epsilons(eps, alpha, num_actors, actor_ids):
    return np.array([eps ** (1 + alpha * i / (num_actors - 1)) if num_actors > 1 else eps for i in actor_ids], np.float32)


def flat_layout(shapes_and_itemsizes):
    """[(offset, nbytes)], flat_bytes: tensors in order, each at the next multiple of 256."""
    out, off = [], 0
    for shape, itemsize in shapes_and_itemsizes:
        nbytes = int(np.prod(shape, dtype=np.int64)) * itemsize
        out.append((off, nbytes))
        off += (nbytes + 255) // 256 * 256
    return out, max(off, 256)


def pack_flat(arrays):
    layout, total = flat_layout([(a.shape, a.dtype.itemsize) for a in arrays])
    flat = np.zeros(total, np.uint8)
    for (off, nbytes), a in zip(layout, arrays):
        flat[off:off + nbytes] = np.ascontiguousarray(a).view(np.uint8).reshape(-1)
    return flat
