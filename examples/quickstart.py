#!/usr/bin/env python
"""Smallest end-to-end use of the fast tier on one B200: vectorised actors fill a device-resident
replay, a learner samples / trains / writes priorities back, weights reach the actors through the
shared-memory board.  Synthetic environment (gym is not needed).  Run: python examples/quickstart.py"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import apex_b200                                     # noqa: E402


class NoiseEnv:
    """gym-protocol stand-in: 84x84 uint8 frames, random rewards, episodes of 200 steps."""

    def __init__(self, seed):
        self.rng, self.t = np.random.default_rng(seed), 0

    def reset(self):
        self.t = 0
        return self.rng.integers(0, 256, (84, 84), dtype=np.uint8)

    def step(self, action):
        self.t += 1
        return self.rng.integers(0, 256, (84, 84), dtype=np.uint8), float(self.rng.integers(-1, 2)), self.t % 200 == 0, {}


def main():
    torch.cuda.set_device(0)
    E, C, A, n = 16, 4, 4, 3
    env_conf = {"state_shape": [C, 84, 84], "action_dim": A}
    # replay.py:9-16 -> a device ring: one lane per actor, frames stored once (stream layout)
    replay = apex_b200.ReplayMemory(E * 4096, {"priority_exponent": 0.6, "layout": "stream", "lanes": E, "n_step": n,
                                               "state_shape": (C, 84, 84)})
    learner = apex_b200.Learner(env_conf, {"replay_sample_size": 128, "q_target_sync_freq": 100, "remove_old_xp_freq": 100,
                                           "min_replay_mem_size": 1000, "load_saved_state": False}, {}, replay)
    learner._to_device()
    # learner.py:74 / actor.py:189-191 -> versioned board in shared memory
    board = apex_b200.WeightPublisher(learner.Q, name="apex_quickstart_%d" % os.getpid())
    board.wait(board.publish())
    actor_q = apex_b200.learner.make_q_network(env_conf["state_shape"], A).cuda()
    sub = apex_b200.DeviceWeightSubscriber(actor_q, board.name)
    actors = apex_b200.VectorActor([NoiseEnv(s) for s in range(E)], actor_q, replay,
                                   {"epsilon": 0.4, "alpha": 7, "num_actors": E, "num_steps": n, "gamma": 0.99,
                                    "Q_network_sync_freq": 50, "n_step_transition_batch_size": 8}, weights=sub)
    try:
        for it in range(5):
            written = actors.run(100)                 # actor.py:146-191 for E actors in lock-step (K10, K7, K6)
            losses = [float(learner.step_fast()) for _ in range(20)]     # learner.py:66-76 (K3+K4, model, K5+K2)
            board.publish()                           # picked up by the actors at their next sync
            print("iteration %d: %d records written, replay size %d, mean loss %.4f, weights v%d"
                  % (it, written, replay.size(), float(np.mean(losses)), board.version))
        replay.check_status()
    finally:
        sub.close()
        board.close(unlink=True)


if __name__ == "__main__":
    main()
