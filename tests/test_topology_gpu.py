"""The reference's process topology (main.py:28-62) on top of the drop-in classes: a replay
hosted in a BaseManager server process, a learner process, actor processes feeding a Manager
queue and a pump process calling replay.add — with CUDA living in the server / learner /
actor processes and never in the parent (which forks them, as main.py does).

gym is not installed and /root/reference is not present on the GPU box, so the environment is
a synthetic Atari-shaped one and the ~30 lines of process wiring are restated here."""
import os
import subprocess
import sys
import textwrap

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = textwrap.dedent(r'''
    import sys, time
    sys.path.insert(0, %r); sys.path.insert(0, %r)
    import numpy as np
    import torch.multiprocessing as mp
    from multiprocessing.managers import BaseManager
    from replay import ReplayMemory                    # drop-in modules (ape-x-dqn_b200/dropin)
    from learner import Learner
    from actor import Actor

    class FakeAtari:                                   # stands in for gym.make(name): 210x160x3 uint8 frames
        def __init__(self, name):
            self.rng = np.random.default_rng(abs(hash(name)) %% 1000)
            self.t = 0
        def reset(self):
            self.t = 0
            return self.rng.integers(0, 256, (210, 160, 3), dtype=np.uint8)
        def step(self, action):
            self.t += 1
            return self.rng.integers(0, 256, (210, 160, 3), dtype=np.uint8), float(self.rng.integers(-1, 2)), self.t %% 23 == 0, {}

    BaseManager.register("Memory", ReplayMemory)       # main.py:18

    def pump(shared_mem, replay_mem, stop):            # main.py:21-25
        while not stop.is_set():
            while shared_mem.qsize() or not shared_mem.empty():
                priorities, xp_batch = shared_mem.get()
                replay_mem.add(priorities, xp_batch)
            time.sleep(0.01)

    if __name__ == "__main__":
        env_conf = {"state_shape": [1, 84, 84], "action_dim": 4, "name": "FakeNoFrameskip-v4"}
        actor_params = {"num_actors": 2, "T": 240, "num_steps": 3, "epsilon": 0.4, "alpha": 7, "gamma": 0.99,
                        "n_step_transition_batch_size": 5, "Q_network_sync_freq": 100}
        learner_params = {"remove_old_xp_freq": 10, "q_target_sync_freq": 10, "min_replay_mem_size": 40,
                          "replay_sample_size": 32, "load_saved_state": False}
        replay_params = {"soft_capacity": 2000, "priority_exponent": 0.6, "importance_sampling_exponent": 0.4}
        import os
        board = os.environ.get("APEX_TEST_BOARD")       # weights through the shared-memory board instead of the Manager dict
        if board:
            learner_params["weight_board"] = actor_params["weight_board"] = board
            learner_params["weight_publish_freq"] = 5
            learner_params["weight_board_unlink"] = False   # this script reads the final version, then removes the board
        mp_manager = mp.Manager()                      # main.py:37-42
        shared_state, shared_mem, stop = mp_manager.dict(), mp_manager.Queue(), mp_manager.Event()
        replay_manager = BaseManager(); replay_manager.start()
        replay_mem = replay_manager.Memory(replay_params["soft_capacity"], replay_params)
        learner = Learner(env_conf, learner_params, shared_state, replay_mem)          # main.py:45-47
        w0 = {k: v.clone() for k, v in shared_state["Q_state_dict"].items()}
        learner_proc = mp.Process(target=learner.learn, args=(25,)); learner_proc.start()
        actors = [Actor(i, env_conf, shared_state, shared_mem, actor_params, env_factory=FakeAtari) for i in range(2)]
        for a in actors: a.start()                     # main.py:51-54
        pump_proc = mp.Process(target=pump, args=(shared_mem, replay_mem, stop)); pump_proc.start()
        learner_proc.join(timeout=150)
        for a in actors: a.join(timeout=60)
        stop.set(); pump_proc.join(timeout=30)
        for p in [learner_proc, pump_proc] + actors:
            if p.exitcode is None:
                p.terminate()
        size = replay_mem.size()
        w1 = shared_state["Q_state_dict"]
        changed = any((w0[k] != w1[k]).any().item() for k in w0)
        version = -1
        if board:
            import apex_b200
            shm = apex_b200._C.open_shared_memory(board, False)
            version = int(np.ndarray((2,), np.int64, shm.buf, 0)[0])
            shm.close(); apex_b200._C.unlink_shared_memory(shm)
        print("RESULT learner_exit=%%s actors_exit=%%s size=%%d weights_changed=%%s board_version=%%d" %% (
            learner_proc.exitcode, [a.exitcode for a in actors], size, changed, version))
        replay_manager.shutdown()
''') % (ROOT, os.path.join(ROOT, "ape-x-dqn_b200", "dropin"))


@pytest.mark.parametrize("use_board", [False, True])
def test_reference_topology_runs_on_dropin_classes(tmp_path, use_board):
    script = tmp_path / "topology.py"
    script.write_text(SCRIPT)
    env = dict(os.environ)
    env.pop("APEX_TEST_BOARD", None)
    if use_board:
        env["APEX_TEST_BOARD"] = "apexboard_topo_%d" % os.getpid()
    out = subprocess.run([sys.executable, str(script)], capture_output=True, text=True, timeout=300, env=env)
    line = [l for l in out.stdout.splitlines() if l.startswith("RESULT")]
    assert out.returncode == 0 and line, (out.stdout[-1500:], out.stderr[-3000:])
    assert "learner_exit=0" in line[0] and "actors_exit=[0, 0]" in line[0], line[0]
    size = int(line[0].split("size=")[1].split()[0])
    assert size > 40
    if use_board:       # learner.py:74 went through K9: 25 steps, one publication at start-up + every 5th step
        assert int(line[0].split("board_version=")[1]) >= 5, line[0]
    else:
        assert "weights_changed=True" in line[0], line[0]
