"""CPU oracle: prioritized replay memory (TEST INFRASTRUCTURE — see oracle/__init__.py).

Restates, in numpy, the *intended* semantics of the reference replay
(/root/reference/replay.py) as fixed by SURVEY.md §2.1 / §8(c):

* sampling law      replay.py:24 (P_i = p_i^alpha / sum) + replay.py:52
                    (np.random.choice == searchsorted(cumsum(P), u, 'right'),
                    i.i.d. with replacement)               -> reference_law_sample
* exact-index law   the same search on exact u64 fixed-point leaves
                                                           -> tree_sample
* priority overwrite replay.py:40 (dict.update: last writer wins)
                                                           -> RingOracle.set_leaves / set_priorities
* FIFO eviction     replay.py:77-80 (oldest first)         -> RingOracle.insert (ring overwrite)
* batch content     replay.py:54-56 (rows in sampled order, fields of
                    N_Step_Transition, replay.py:5)        -> RingOracle.gather

The literal reference (uniform sampling, O(N^2) updates) is only
exercised by tests/golden/make_golden.py, which imports the real file.
"""
from __future__ import annotations

import numpy as np

# ---- fixed-point leaf format (shared contract with the CUDA kernels) -------
FRAC_BITS = 24                      # leaf = floor(priority^alpha * 2^24)
LEAF_ONE = 1 << FRAC_BITS
LEAF_MAX = (1 << 36) - 1            # saturation: priority^alpha < 4096
FANOUT = 32                         # sum-tree fan-out used by the kernels


def pow_priority(p, alpha: float, eps: float = 0.0) -> np.ndarray:
    """|delta| -> priority^alpha in fp32 (replay.py:24 numerator, Q16)."""
    p = np.abs(np.asarray(p, dtype=np.float32))
    if eps > 0.0:
        p = np.maximum(p, np.float32(eps))
    return np.power(p, np.float32(alpha)).astype(np.float32)


def priority_to_leaf(prio_alpha) -> np.ndarray:
    """fp32 priority^alpha -> u64 fixed-point leaf (exact function of the fp32 bits)."""
    x = np.asarray(prio_alpha, dtype=np.float32).astype(np.float64)
    x = np.where(np.isfinite(x) & (x > 0.0), x, 0.0)
    scaled = np.floor(x * float(LEAF_ONE))
    scaled = np.minimum(scaled, float(LEAF_MAX))
    return scaled.astype(np.uint64)


def sample_target(u, total: int) -> np.ndarray:
    """Uniform draw u in [0,1) -> integer target in [0,total).

    target = min(trunc(u * double(total)), total-1); one fp64 multiply, the same
    expression the kernel evaluates (no FMA involved).
    """
    u = np.asarray(u, dtype=np.float64)
    total = int(total)
    if total <= 0:
        raise ValueError("empty tree: total priority is zero")
    t = (u * np.float64(total)).astype(np.uint64)
    return np.minimum(t, np.uint64(total - 1))


def descend(leaves_u64: np.ndarray, targets: np.ndarray) -> np.ndarray:
    """index = #{cumulative sums <= target}  (searchsorted side='right')."""
    c = np.cumsum(np.asarray(leaves_u64, dtype=np.uint64), dtype=np.uint64)
    return np.searchsorted(c, np.asarray(targets, dtype=np.uint64), side="right").astype(np.int64)


def tree_sample(leaves_u64: np.ndarray, u: np.ndarray) -> np.ndarray:
    """Exact-arithmetic cumulative-priority search (the bit-exact contract)."""
    leaves_u64 = np.asarray(leaves_u64, dtype=np.uint64)
    total = int(leaves_u64.sum(dtype=np.uint64))
    return descend(leaves_u64, sample_target(u, total))


def reference_law_sample(priorities: np.ndarray, alpha: float, u: np.ndarray) -> np.ndarray:
    """fp64 restatement of replay.py:24 + np.random.choice (replay.py:52).

    np.random.choice(K, p=P) consumes one random_sample() u and returns
    searchsorted(cumsum(P)/cumsum(P)[-1], u, side='right') (numpy legacy
    RandomState.choice; pinned by tests/golden/choice_law.npz).
    """
    p = np.asarray(priorities, dtype=np.float64) ** float(alpha)
    p = p / p.sum()
    cdf = np.cumsum(p)
    cdf /= cdf[-1]
    return np.searchsorted(cdf, np.asarray(u, dtype=np.float64), side="right").astype(np.int64)


def stratified_uniforms(batch: int, rng: np.random.Generator) -> np.ndarray:
    """u_j = (j + v_j)/B, v ~ U[0,1): the PER stratification."""
    v = rng.random(batch)
    return (np.arange(batch, dtype=np.float64) + v) / float(batch)


def kernel_stratified_uniforms(seed: int, epoch: int, call: int, batch: int) -> np.ndarray:
    """The draws apexb200_sample_stratified generates on the device, restated:
    u_j = (j + v_j)/B with v_j = splitmix64-finaliser(seed, update epoch, call index, j) * 2^-53."""
    M = (1 << 64) - 1
    out = np.empty(batch, np.float64)
    for j in range(batch):
        z = (seed ^ ((epoch * 0xD1342543DE82EF95) & M) ^ ((call * 0xA24BAED4963EE407) & M)) & M
        z = (z + (j + 1) * 0x9E3779B97F4A7C15) & M
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M
        z ^= z >> 31
        out[j] = (np.float64(j) + np.float64(z >> 11) * 2.0 ** -53) / np.float64(batch)
    return out


def level_sizes(n_leaves: int) -> list[int]:
    """Entries per level of the 32-ary tree, leaves first; top level has <= 32."""
    sizes = [int(n_leaves)]
    while sizes[-1] > FANOUT:
        sizes.append((sizes[-1] + FANOUT - 1) // FANOUT)
    return sizes


def build_levels(leaves_u64: np.ndarray) -> list[np.ndarray]:
    """All levels of the 32-ary sum tree (level l+1[i] = sum level l[32i:32i+32])."""
    cur = np.asarray(leaves_u64, dtype=np.uint64)
    levels = [cur.copy()]
    while cur.size > FANOUT:
        pad = (-cur.size) % FANOUT
        cur = np.concatenate([cur, np.zeros(pad, np.uint64)]).reshape(-1, FANOUT).sum(axis=1, dtype=np.uint64)
        levels.append(cur)
    return levels


class RingOracle:
    """Ring-buffer replay with the device layout restated on the host.

    capacity N records split into `lanes` equal rings; each record owns `fpr`
    consecutive frame slots and may reference up to `lookback` older records'
    frames of the same lane through relative offsets (stream layout), or only
    its own frames (materialised layout, fpr = 2C, lookback = 0).
    """

    def __init__(self, capacity, frames_per_state, frame_bytes=7056, lanes=1, fpr=1,
                 lookback=0, alpha=0.6, num_actions=0):
        assert capacity % lanes == 0
        self.N, self.C, self.F = int(capacity), int(frames_per_state), int(frame_bytes)
        self.lanes, self.fpr, self.W = int(lanes), int(fpr), int(lookback)
        self.lane_cap = self.N // self.lanes
        self.alpha = float(alpha)
        self.nA = int(num_actions)
        self.frames = np.zeros((self.N * self.fpr, self.F), np.uint8)
        self.fidx = np.zeros((self.N, 2 * self.C), np.int32)
        self.A = np.zeros(self.N, np.int32)
        self.R = np.zeros(self.N, np.float32)
        self.G = np.zeros(self.N, np.float32)
        self.q_t = np.zeros((self.N, self.nA), np.float32)
        self.q_tpn = np.zeros((self.N, self.nA), np.float32)
        self.leaves = np.zeros(self.N, np.uint64)
        self.lane_count = np.zeros(self.lanes, np.int64)

    # -- a3 / a11 / K6 ------------------------------------------------------
    def insert(self, lane, frames, rel, A, R, G, prio=None, q_t=None, q_tpn=None):
        """replay.py:59-69 (add) with ring eviction (replay.py:77-80 as overwrite).

        frames [m*fpr, F] u8; rel [m, 2C] int32 offsets (<= 0, in frame units,
        relative to the record's own last frame, wrapping inside the lane);
        prio: raw priorities (|TD|, actor.py:138-142) or None -> computed from
        q_t/q_tpn like Actor.compute_priorities.
        """
        m = len(A)
        assert m + self.W <= self.lane_cap
        frames = np.asarray(frames, np.uint8).reshape(m * self.fpr, self.F)
        rel = np.asarray(rel, np.int32).reshape(m, 2 * self.C)
        if prio is None:
            from .nstep_oracle import initial_priority
            prio = initial_priority(q_t, q_tpn, A, R, G)
        leaf = priority_to_leaf(pow_priority(prio, self.alpha))
        head = int(self.lane_count[lane])
        lane_frames = self.lane_cap * self.fpr
        for j in range(m):
            p = (head + j) % self.lane_cap
            s = lane * self.lane_cap + p
            self.frames[s * self.fpr:(s + 1) * self.fpr] = frames[j * self.fpr:(j + 1) * self.fpr]
            last = p * self.fpr + self.fpr - 1
            self.fidx[s] = lane * lane_frames + (last + rel[j]) % lane_frames
            self.A[s], self.R[s], self.G[s] = A[j], R[j], G[j]
            if self.nA and q_t is not None:
                self.q_t[s], self.q_tpn[s] = q_t[j], q_tpn[j]
            self.leaves[s] = leaf[j]
        for w in range(self.W):           # records whose lookback frames were just overwritten
            p = (head + m + w) % self.lane_cap
            self.leaves[lane * self.lane_cap + p] = 0
        self.lane_count[lane] += m
        return lane * self.lane_cap + (head + np.arange(m)) % self.lane_cap

    def size(self):
        return int(np.minimum(self.lane_count, self.lane_cap).sum())

    # -- a4 / K2 --------------------------------------------------------------
    def set_leaves(self, idx, leaves):
        for i, v in zip(np.asarray(idx).tolist(), np.asarray(leaves, np.uint64).tolist()):
            self.leaves[i] = v             # sequential => last writer wins (replay.py:40)

    def set_priorities(self, idx, prio, eps=0.0):
        self.set_leaves(idx, priority_to_leaf(pow_priority(prio, self.alpha, eps)))

    def total(self):
        return int(self.leaves.sum(dtype=np.uint64))

    # -- a6 / K3 + K4 ---------------------------------------------------------
    def sample(self, u):
        return tree_sample(self.leaves, u)

    def gather(self, idx):
        idx = np.asarray(idx, np.int64)
        f = self.fidx[idx]                                  # [B, 2C]
        S_t = self.frames[f[:, :self.C]]                    # [B, C, F]
        S_tpn = self.frames[f[:, self.C:]]
        out = dict(S_t=S_t, S_tpn=S_tpn, A=self.A[idx], R=self.R[idx], G=self.G[idx])
        if self.nA:
            out.update(q_t=self.q_t[idx], q_tpn=self.q_tpn[idx])
        return out


def shard_split(totals, u):
    """§8(e): global targets -> (owning shard, local target) for every draw.

    totals: per-shard tree totals (python ints); identical on every rank after
    the all-gather.  Returns (global_target, shard, local_target).
    """
    totals = [int(t) for t in totals]
    T = sum(totals)
    t = sample_target(u, T)
    bounds = np.cumsum(np.array(totals, dtype=np.uint64), dtype=np.uint64)
    shard = np.searchsorted(bounds, t, side="right").astype(np.int64)
    prefix = np.concatenate([[np.uint64(0)], bounds[:-1]]).astype(np.uint64)
    return t, shard, t - prefix[shard]
