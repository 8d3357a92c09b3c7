"""E reference actors stepped in lock-step in one process, with the network on the GPU.

Actor.run (actor.py:146-191) per environment is: batch-1 CPU forward -> epsilon-greedy -> env.step
-> ExperienceBuffer.add -> compute_priorities -> queue.put -> periodic load_state_dict.  Here the
E forwards are one batched forward (BatchedActorPolicy, K10), the n-step assembly / initial
priority / replay insert are VectorActorIngest (K7, K6) straight into the device ring, and the
weights arrive through the weight board (K9).  Environment stepping stays host code.

Frames: each environment contributes one new uint8 frame per step (the newest plane of its
observation); the C-frame state the network sees is rebuilt from the last C frames of the episode
(the first frame repeated at an episode start), which is exactly what the stream-layout replay
hands the learner for that record.
"""
from __future__ import annotations

import numpy as np
import torch

from .ingest import VectorActorIngest
from .weights import BatchedActorPolicy


class VectorActor:
    def __init__(self, envs, q_network, mem, actor_params, actor_ids=None, weights=None, lanes=None, chunk_steps=None,
                 frame_of=None, seed=0):
        """envs: E objects with reset() -> obs and step(a) -> (obs, reward, done, info) (the gym
        protocol of actor.py:150, :165).  frame_of(obs) -> uint8 [H, W] newest frame (default:
        the last plane of a uint8 observation).  weights: a DeviceWeightSubscriber or None."""
        mem._ensure()
        self.envs, self.E, self.mem = list(envs), len(envs), mem
        ids = list(range(self.E)) if actor_ids is None else list(actor_ids)
        self.policy = BatchedActorPolicy(q_network, actor_params, ids, seed=seed)
        self.n = int(actor_params.get("num_steps", mem.n_step))
        self.gamma = float(actor_params["gamma"])
        self.ingest = VectorActorIngest(mem, self.E, n_step=self.n, gamma=self.gamma, lanes=lanes)
        self.sync_freq = int(actor_params.get("Q_network_sync_freq", 500))
        self.chunk = int(chunk_steps or actor_params.get("n_step_transition_batch_size", 5))
        self.weights = weights
        self.frame_of = frame_of or (lambda obs: np.asarray(obs, np.uint8).reshape((-1,) + np.asarray(obs).shape[-2:])[-1])
        self.C = mem.C
        self.t = 0
        self._stack = None            # device [E, C, H, W] uint8: what the network sees
        self._pending = []            # per step: (frames [E,F] u8 host, actions, rewards, dones, q device)
        self.episode_returns = []
        self._ret = np.zeros(self.E)

    def _flush(self):
        if not self._pending:
            return 0
        frames = torch.as_tensor(np.stack([p[0] for p in self._pending], 1)).to(self.mem.device, non_blocking=True)
        acts = torch.stack([p[1] for p in self._pending], 1)
        rews = np.stack([p[2] for p in self._pending], 1)
        dones = np.stack([p[3] for p in self._pending], 1)
        q = torch.stack([p[4] for p in self._pending], 1)
        self._pending = []
        return self.ingest.push(frames, acts, rews, dones, q)

    def run(self, T):
        """T lock-steps of all E environments; returns the number of records written to the replay."""
        dev = self.mem.device
        if self._stack is None:
            self._cur = [self.frame_of(env.reset()) for env in self.envs]
            first = torch.as_tensor(np.stack(self._cur)).to(dev)
            self._stack = first[:, None].expand(-1, self.C, *first.shape[1:]).contiguous()   # [E, C, H, W] uint8
        written = 0
        for _ in range(T):
            if self.weights is not None and self.t % self.sync_freq == 0:       # actor.py:189-191
                self.weights.poll()
            actions, q = self.policy.act(self._stack)                           # actor.py:160-163
            a_host = actions.cpu().numpy()
            frames = np.stack([f.reshape(-1) for f in self._cur])
            rewards, dones = np.zeros(self.E, np.float32), np.zeros(self.E, bool)
            for e, env in enumerate(self.envs):                                 # actor.py:165
                obs, r, d, _ = env.step(int(a_host[e]))
                rewards[e], dones[e] = r, d
                self._ret[e] += r
                if d:                                                           # actor.py:172-178
                    self.episode_returns.append(self._ret[e])
                    self._ret[e] = 0.0
                    obs = env.reset()
                self._cur[e] = self.frame_of(obs)
            new = torch.as_tensor(np.stack(self._cur)).to(dev, non_blocking=True)[:, None]   # one copy for all E
            restart = torch.as_tensor(dones).to(dev)[:, None, None, None]
            self._stack = torch.where(restart, new.expand_as(self._stack), torch.cat([self._stack[:, 1:], new], 1))
            self._pending.append((frames, actions, rewards, dones, q))
            self.t += 1
            if len(self._pending) >= self.chunk:                                # actor.py:181-187
                written += self._flush()
        return written + self._flush()


__all__ = ["VectorActor"]
