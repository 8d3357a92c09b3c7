"""Actor-side n-step buffer and initial priorities with the reference's interface.

Mirror of /root/reference/actor.py:15-191 (ExperienceBuffer, Actor).  Only the
n-step assembly and the initial-priority computation are on the accelerated
path; environment stepping and the epsilon-greedy policy
are host code kept for interface completeness.

ExperienceBuffer has two semantics:
  * ``semantics="reference"`` (default): emits exactly what the reference class
    emits for the same calls — non-overlapping chunks, re-accumulated returns,
    Gamma = gamma^(B-1), keys str(id)+str(seq) — pinned by
    tests/golden/nstep_literal.json;
  * ``semantics="sliding"``: one record per environment step with
    R = sum_k gamma^k r_{t+k}, Gamma = gamma^m and Gamma = 0 at terminals, the
    arithmetic the K7 kernel (apexb200_nstep) implements for vectorised actors.
"""
from __future__ import annotations

import random
from collections import deque, namedtuple

import numpy as np
import torch.multiprocessing as mp

from .replay import N_Step_Transition

Transition = namedtuple("Transition", ["S", "A", "R", "Gamma", "q"])     # actor.py:11


class ExperienceBuffer(object):
    def __init__(self, n, actor_id, semantics="reference", gamma=0.99):
        if semantics not in ("reference", "sliding"):
            raise ValueError("semantics must be 'reference' or 'sliding'")
        self.capacity = int(n)
        self.id = actor_id
        self.gamma = gamma                      # actor.py:25
        self.semantics = semantics
        self.local_1step_buffer = deque() if semantics == "sliding" else []
        self.local_nstep_buffer = []
        self.n_step_seq_num = 0

    # -- shared -----------------------------------------------------------------
    def _emit(self, first, R, Gamma, nxt):
        key = str(self.id) + str(self.n_step_seq_num)                    # actor.py:47
        self.local_nstep_buffer.append(N_Step_Transition(first.S, first.A, R, Gamma, first.q, nxt.S, nxt.q, key))
        self.n_step_seq_num += 1

    @property
    def idx(self):
        return len(self.local_1step_buffer) - 1                          # actor.py:23,55,64

    @property
    def B(self):
        return len(self.local_1step_buffer)                              # actor.py:79-85

    @property
    def size(self):
        return len(self.local_nstep_buffer)                              # actor.py:87-93

    def get(self, batch_size):
        assert batch_size <= self.size, "Requested n-step transitions batch size is more than available"
        head = self.local_nstep_buffer[:batch_size]
        del self.local_nstep_buffer[:batch_size]
        return head

    # -- reference semantics ------------------------------------------------------
    def update_buffer(self):
        """actor.py:29-43, literally: each pass adds the discounted current R of every later
        entry onto what an entry already holds; the newest entry is left untouched."""
        buf = self.local_1step_buffer
        for i in range(len(buf) - 1):
            total, disc = buf[i].R, 1
            for later in list(buf)[i + 1:]:
                disc *= self.gamma
                total += disc * later.R
            buf[i] = buf[i]._replace(R=total, Gamma=disc)

    def construct_nstep_transition(self, data):
        """actor.py:44-55: close the open chunk with `data` as the bootstrap state."""
        if not self.local_1step_buffer:
            return
        if self.semantics == "sliding":
            self._flush_sliding(data, terminal=True)
            return
        first = self.local_1step_buffer[0]
        self._emit(first, first.R, first.Gamma, data)
        self.local_1step_buffer.clear()

    def add(self, data):
        """actor.py:57-70."""
        if self.semantics == "sliding":
            self._add_sliding(data)
        elif len(self.local_1step_buffer) < self.capacity:
            self.local_1step_buffer.append(data)
            self.update_buffer()
        else:
            self.construct_nstep_transition(data)

    # -- sliding-window semantics -------------------------------------------------
    def _window_return(self, start):
        acc, disc = 0.0, 1.0
        for k in range(start, len(self.local_1step_buffer)):
            acc += disc * self.local_1step_buffer[k].R
            disc *= self.gamma
        return acc, disc

    def _add_sliding(self, data):
        buf = self.local_1step_buffer
        if len(buf) == self.capacity:            # oldest step now has its n successors
            acc, disc = self._window_return(0)
            self._emit(buf[0], acc, disc, data)
            buf.popleft()
        buf.append(data)

    def _flush_sliding(self, data, terminal):
        buf = self.local_1step_buffer
        while buf:
            acc, disc = self._window_return(0)
            self._emit(buf[0], acc, 0.0 if terminal else disc, data)
            buf.popleft()


class Actor(mp.Process):
    """actor.py:96-191.  `env_factory(name)` replaces env.make_local_env when gym is absent."""

    def __init__(self, actor_id, env_conf, shared_state, shared_replay_mem, actor_params, env_factory=None,
                 q_factory=None):
        super(Actor, self).__init__()
        self.actor_id = actor_id
        self.state_shape = tuple(env_conf["state_shape"])
        self.action_dim = env_conf["action_dim"]
        self.env_name = env_conf.get("name")
        self.params = actor_params
        self.shared_state = shared_state
        self.T = self.params["T"]
        self.global_replay_queue = shared_replay_mem
        self._env_factory = env_factory
        self._q_factory = q_factory
        self.local_experience_buffer = ExperienceBuffer(self.params["num_steps"], self.actor_id,
                                                        semantics=self.params.get("nstep_semantics", "reference"))
        eps, N, alpha = self.params["epsilon"], self.params["num_actors"], self.params["alpha"]
        self.epsilon = eps ** (1 + alpha * self.actor_id / (N - 1)) if N > 1 else eps      # actor.py:114 (Q22)
        self.gamma = self.params["gamma"]
        self.policy = self.epsilon_greedy_Q
        self.Q = None
        self.env = None

    def epsilon_greedy_Q(self, qS_t):
        if random.random() >= self.epsilon:                                            # actor.py:121-125
            return int(np.argmax(qS_t))
        return random.randrange(len(qS_t))

    def obs_preproc(self, obs):
        """Observation -> uint8 state_shape.  (The reference's float64 reshape/np.resize pipeline,
        actor.py:117-119, is not reproduced: SURVEY Q20.)"""
        x = np.asarray(obs)
        if x.ndim == 3 and x.shape[-1] == 3:
            x = x @ np.array([0.299, 0.587, 0.114])
        x = x.reshape(x.shape[-2], x.shape[-1]) if x.ndim > 2 else x
        H, W = self.state_shape[1:]
        rows = (np.arange(H) * x.shape[0] // H)
        cols = (np.arange(W) * x.shape[1] // W)
        x = x[rows][:, cols]
        return np.broadcast_to(np.clip(np.rint(x), 0, 255).astype(np.uint8), self.state_shape).copy()

    def compute_priorities(self, n_step_transitions):
        """actor.py:127-143 with per-record |TD|, computed by apexb200_initial_priority."""
        import torch
        from . import ops
        recs = list(n_step_transitions)
        dev = torch.device("cuda")
        q_t = torch.tensor(np.stack([np.asarray(r.qS_t, np.float32) for r in recs]), device=dev)
        q_tpn = torch.tensor(np.stack([np.asarray(r.qS_tpn, np.float32) for r in recs]), device=dev)
        A = torch.tensor([int(r.A_t) for r in recs], dtype=torch.int32, device=dev)
        R = torch.tensor([float(r.R_ttpB) for r in recs], dtype=torch.float32, device=dev)
        G = torch.tensor([float(r.Gamma_ttpB) for r in recs], dtype=torch.float32, device=dev)
        td = ops.initial_priority(q_t, q_tpn, A, R, G).cpu().numpy()
        return {r.key: td[i] for i, r in enumerate(recs)}

    def _setup(self):
        import torch
        # This runs in a forked child (main.py:51-54).  The parent has already used torch's OpenMP
        # pool (building the learner's networks), and a forked copy of that pool deadlocks on its
        # first parallel region; batch-1 inference wants one thread anyway.
        torch.set_num_threads(1)
        if self.Q is None:
            from .learner import make_q_network
            self.Q = (self._q_factory or make_q_network)(self.state_shape, self.action_dim)
            self.Q.load_state_dict(self.shared_state["Q_state_dict"])                   # actor.py:106
        self._weights = None
        if self.params.get("weight_board"):               # K9 board published by the learner (weights.py)
            import time
            from .weights import WeightSubscriber
            deadline = time.time() + float(self.params.get("weight_board_wait_s", 60.0))
            while True:                                   # the learner creates the board inside learn() (its own process)
                try:
                    self._weights = WeightSubscriber(self.Q, self.params["weight_board"])
                    break
                except FileNotFoundError:
                    if time.time() > deadline:
                        raise
                    time.sleep(0.05)
            self._weights.poll()
        if self.env is None:
            if self._env_factory is not None:
                self.env = self._env_factory(self.env_name)
            else:
                from env import make_local_env                                       # reference env.py:3-4
                self.env = make_local_env(self.env_name)

    def run(self):
        """actor.py:146-191."""
        import torch
        self._setup()
        obs = self.obs_preproc(self.env.reset())
        batch = self.params["n_step_transition_batch_size"]
        for t in range(self.T):
            with torch.no_grad():
                qS_t = self.Q(torch.from_numpy(obs).unsqueeze(0).float())[2].squeeze().numpy()
            action = self.policy(qS_t)
            next_obs, reward, done, _ = self.env.step(action)
            self.local_experience_buffer.add(Transition(obs, action, reward, self.gamma, qS_t))
            obs = self.obs_preproc(next_obs)
            if done:
                self.local_experience_buffer.construct_nstep_transition(Transition(obs, action, 0, self.gamma, qS_t))
                obs = self.obs_preproc(self.env.reset())
            if self.local_experience_buffer.size >= batch:
                xp = self.local_experience_buffer.get(batch)
                self.global_replay_queue.put([self.compute_priorities(xp), xp])          # actor.py:181-187
            if t % self.params["Q_network_sync_freq"] == 0:                             # actor.py:189-191
                if self._weights is not None:
                    self._weights.poll()
                else:
                    self.Q.load_state_dict(self.shared_state["Q_state_dict"])
