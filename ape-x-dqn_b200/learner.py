"""Learner with the reference's interface over the fused TD/priority kernels.

Mirror of /root/reference/learner.py:10-80.  The duelling conv/MLP forward and
backward stay on PyTorch/cuDNN (the model class is used unchanged when
``duelling_network.DuellingDQN`` is importable, otherwise an equivalent module
from qnet.py); everything between the model outputs and the replay — n-step
double-DQN target, TD error, loss, |delta|^alpha priority and its write-back —
runs in libapexb200 kernels (K5 + K2).
"""
from __future__ import annotations

import time

import numpy as np
import torch

from . import _C, ops
from .replay import N_Step_Transition, ReplayMemory


def make_q_network(state_shape, action_dim):
    try:
        from duelling_network import DuellingDQN          # the reference's model file, unchanged
        return DuellingDQN(state_shape, action_dim)
    except ImportError:
        from .qnet import DuellingQNet
        return DuellingQNet(state_shape, action_dim)


class _TDLoss(torch.autograd.Function):
    """loss_mean = mean_b loss(R + Gamma*Q_tgt(S_tpn)[argmax Q(S_tpn)] - Q(S_t)[A]); grad flows to Q(S_t) only
    (learner.py:43-48: the target is built under no_grad)."""

    @staticmethod
    def forward(ctx, q_on_t, q_on_tpn, q_tg_tpn, A, R, Gamma, replay, idx, alpha, eps, loss_kind, kappa, weights=None,
                count_dev=None, mean_divisor=0.0):
        q_on_t, q_on_tpn, q_tg_tpn = (x.detach().contiguous().float() for x in (q_on_t, q_on_tpn, q_tg_tpn))
        if replay is not None:
            delta, prio, loss, mean = replay.td_update(idx, q_on_t, q_on_tpn, q_tg_tpn, A, R, Gamma, loss_kind, kappa,
                                                       weights=weights, count_dev=count_dev, mean_divisor=mean_divisor)
        else:
            if count_dev is not None:
                raise ValueError("count_dev / mean_divisor need a replay (sharded batches)")
            delta, prio, loss, mean = ops.td_forward(q_on_t, q_on_tpn, q_tg_tpn, A, R, Gamma, alpha, eps, loss_kind, kappa,
                                                     weights=weights)
        ctx.save_for_backward(delta, A)
        ctx.weights, ctx.count_dev = weights, count_dev
        ctx.meta = (q_on_t.shape[1], loss_kind, kappa, float(mean_divisor))
        ctx.mark_non_differentiable(delta, prio, loss)
        return mean.reshape(()), delta, prio, loss

    @staticmethod
    def backward(ctx, g_mean, _gd, _gp, _gl):
        delta, A = ctx.saved_tensors
        nA, loss_kind, kappa, mean_divisor = ctx.meta
        g = g_mean.detach().reshape(1).float().contiguous()
        return (ops.td_backward(delta, A, g, nA, loss_kind, kappa, count_dev=ctx.count_dev, mean_divisor=mean_divisor,
                                weights=ctx.weights),) + (None,) * 14


def td_loss(q_on_t, q_on_tpn, q_tg_tpn, A, R, Gamma, alpha=0.6, eps=0.0, loss_kind=_C.LOSS_HALF_SQ, kappa=1.0,
            replay=None, idx=None, weights=None, count_dev=None, mean_divisor=0.0):
    """Differentiable fused loss.  Returns (loss_mean, delta, prio, loss).  With `replay` and `idx`
    the new priorities are also written into the replay's sum tree in the same call; `weights`
    (f32 [B], e.g. ops.is_weights) multiply the per-record loss and gradient.  Sharded batches:
    count_dev (int32 [1], device) = live rows, mean_divisor = global batch size, so that the sum of
    the ranks' loss_mean (and of their gradients) is the mean over the global batch."""
    return _TDLoss.apply(q_on_t, q_on_tpn, q_tg_tpn, A, R, Gamma, replay, idx, alpha, eps, loss_kind, kappa, weights,
                         count_dev, mean_divisor)


class Learner(object):
    def __init__(self, env_conf, learner_params, shared_state, shared_replay_memory):
        # learner.py:11-27.  Networks are built on the CPU here; they move to the GPU on first use
        # inside the process that runs learn() (main.py:45-47 forks after construction).
        self.state_shape = env_conf["state_shape"]
        action_dim = env_conf["action_dim"]
        self.action_dim = action_dim
        self.params = learner_params
        self.shared_state = shared_state
        self.Q = make_q_network(self.state_shape, action_dim)
        self.Q_double = make_q_network(self.state_shape, action_dim)
        if self.params.get("load_saved_state"):
            try:
                self.Q.load_state_dict(torch.load(self.params["load_saved_state"])["Q_state"])
            except FileNotFoundError:
                print("WARNING: No trained model found. Training from scratch")
        self.shared_state["Q_state_dict"] = self.Q.state_dict()
        self.replay_memory = shared_replay_memory
        self.optimizer = torch.optim.RMSprop(self.Q.parameters(), lr=0.00025 / 4, weight_decay=0.95, eps=1.5e-7)
        self.num_q_updates = 0
        self._graph = None
        self.loss_kind = _C.LOSS_HUBER if self.params.get("loss", "half_sq") == "huber" else _C.LOSS_HALF_SQ
        self.kappa = float(self.params.get("huber_kappa", 1.0))
        self.alpha = float(self.params.get("priority_exponent", getattr(shared_replay_memory, "alpha", 0.6)))
        # learner.py:37-38 hands the model float32 states.  "f32" (default): the gather kernel writes the float32
        # batch itself; "u8": uint8 gather followed by two .float() passes (the round-1 path, kept for comparison).
        self.batch_f32 = self.params.get("batch_format", "f32") != "u8"
        self.device = None

    # ------------------------------------------------------------------ helpers
    def _to_device(self):
        if self.device is None:
            if not torch.cuda.is_available():
                raise RuntimeError("apex_b200.Learner needs a CUDA device (no CPU fallback)")
            self.device = torch.device("cuda", torch.cuda.current_device())
            self.Q.to(self.device)
            self.Q_double.to(self.device)
            # optimizer state is created lazily by torch, after the move
        return self.device

    def _fast(self):
        return isinstance(self.replay_memory, ReplayMemory)

    def _f32(self):
        """Float32 batches straight from the gather when the fused kernel can stage half a record (C <= 8)."""
        m = self.replay_memory
        return bool(self.batch_f32 and m.C <= 8 and m.C * m.F + 256 <= 226 * 1024)

    @staticmethod
    def _states(b, shape):
        if b.S_t.dtype == torch.float32:
            return b.S_t.view(shape), b.S_tpn.view(shape)
        return b.S_t.view(shape).float(), b.S_tpn.view(shape).float()

    def _loss_from_tensors(self, S_t, S_tpn, A, R, Gamma, replay=None, idx=None, count_dev=None, mean_divisor=0.0):
        with torch.no_grad():                                                  # learner.py:43-45
            q_tg_tpn = self.Q_double(S_tpn)[2]
            q_on_tpn = self.Q(S_tpn)[2]
        q_on_t = self.Q(S_t)[2]                                                # learner.py:46
        return td_loss(q_on_t, q_on_tpn, q_tg_tpn, A, R, Gamma, self.alpha, 0.0, self.loss_kind, self.kappa,
                       replay=replay, idx=idx, count_dev=count_dev, mean_divisor=mean_divisor)

    # ------------------------------------------------------------- reference API
    def compute_loss_and_priorities(self, xp_batch):
        """learner.py:29-52: list of records -> (loss.mean(), {key: |delta|})."""
        dev = self._to_device()
        n = N_Step_Transition(*zip(*xp_batch))
        S_t = torch.from_numpy(np.array(n.S_t)).to(dev).float()
        S_tpn = torch.from_numpy(np.array(n.S_tpn)).to(dev).float()
        R = torch.tensor(np.array(n.R_ttpB, dtype=np.float32), device=dev)
        G = torch.tensor(np.array(n.Gamma_ttpB, dtype=np.float32), device=dev)
        A = torch.tensor(np.array(n.A_t, dtype=np.int32), device=dev)
        loss, delta, _prio, _ = self._loss_from_tensors(S_t, S_tpn, A, R, G)
        abs_td = delta.abs().cpu().numpy()
        return loss, {k: abs_td[i] for i, k in enumerate(n.key)}                # per-record |delta|

    def update_Q(self, loss):
        """learner.py:54-61."""
        self.optimizer.zero_grad()
        loss.backward()
        self.optimizer.step()
        self.num_q_updates += 1
        freq = self.params["q_target_sync_freq"]
        literal = bool(self.params.get("literal_target_sync", False))
        if (self.num_q_updates % freq != 0) if literal else (self.num_q_updates % freq == 0):
            self.Q_double.load_state_dict(self.Q.state_dict())

    def step_fast(self, u=None):
        """One learner step with the batch resident in HBM: K3 -> K4 -> model -> K5+K2 -> optimiser."""
        self._to_device()
        B = int(self.params["replay_sample_size"])
        b = self.replay_memory.sample_tensors(B, u=u, f32=self._f32())
        shape = (B,) + tuple(self.state_shape)
        S_t, S_tpn = self._states(b, shape)
        loss, delta, prio, _ = self._loss_from_tensors(S_t, S_tpn, b.A, b.R, b.Gamma, replay=self.replay_memory, idx=b.idx)
        self.update_Q(loss)
        return loss.detach()              # no autograd graph leaves this call (capture_step needs them all gone)

    def _maybe_sync_target(self):
        freq = self.params["q_target_sync_freq"]
        literal = bool(self.params.get("literal_target_sync", False))
        if (self.num_q_updates % freq != 0) if literal else (self.num_q_updates % freq == 0):
            self.Q_double.load_state_dict(self.Q.state_dict())

    def capture_step(self):
        """Capture one fast-tier learner step — K3+K4, the three forwards, backward, RMSprop, K5+K2 — in a
        CUDA graph.  Returns nothing; step_graphed() replays it.  The sampling kernel draws
        fresh stratified numbers on every replay (device-side epoch), the target-network sync stays outside.
        Capturing is not free of side effects: the three warm-up passes before the capture are REAL learner
        steps (weights, priorities and num_q_updates advance, the target network is synchronised when one of
        them falls on q_target_sync_freq), exactly as three step_fast() calls would be.
        Losses returned by earlier compute_loss_and_priorities calls must have been released: a live
        autograd graph pins gradient-accumulation nodes to the stream they were created on."""
        dev = self._to_device()
        for g in self.optimizer.param_groups:
            g["capturable"] = True                       # keeps RMSprop's step counters on the device
        for st in self.optimizer.state.values():         # counters created by earlier eager steps live on the host
            if torch.is_tensor(st.get("step")) and st["step"].device != dev:
                st["step"] = st["step"].to(dev)
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):                            # warm-up: optimizer state, cuDNN plans, output buffers
                self.optimizer.zero_grad(set_to_none=True)
                loss = self._step_body()
                loss.backward()
                self.optimizer.step()
                self.num_q_updates += 1
                self._maybe_sync_target()
                del loss
            side.synchronize()
            self._graph = torch.cuda.CUDAGraph()
            self.optimizer.zero_grad(set_to_none=True)
            with torch.cuda.graph(self._graph, stream=side):
                self._graph_loss = self._step_body()
                self._graph_loss.backward()
                self.optimizer.step()
        torch.cuda.current_stream().wait_stream(side)

    def _step_body(self):
        B = int(self.params["replay_sample_size"])
        b = self.replay_memory.sample_tensors(B, f32=self._f32())
        shape = (B,) + tuple(self.state_shape)
        S_t, S_tpn = self._states(b, shape)
        loss, _delta, _prio, _ = self._loss_from_tensors(S_t, S_tpn, b.A, b.R, b.Gamma, replay=self.replay_memory, idx=b.idx)
        return loss

    def step_graphed(self):
        """Replay the captured step; returns the (device) loss tensor of that step."""
        if self._graph is None:
            self.capture_step()
        self._graph.replay()
        self.num_q_updates += 1
        self._maybe_sync_target()
        return self._graph_loss.detach()

    def learn(self, T):
        """learner.py:63-80."""
        if self.device is None:
            torch.set_num_threads(1)      # forked child (main.py:46): the parent's OpenMP pool is unusable here
        board = None
        if self.params.get("weight_board"):                 # K9: versioned shared-memory board instead of a pickle
            from .weights import WeightPublisher            # (created first: the actors wait for it before they step)
            self._to_device()
            board = WeightPublisher(self.Q, name=self.params["weight_board"], create=True)
            board.publish()
        while self.replay_memory.size() <= self.params["min_replay_mem_size"]:
            time.sleep(1)
        fast = self._fast()
        publish = int(self.params.get("weight_publish_freq", 500 if fast else 1))
        for t in range(T):
            if fast:
                self.step_fast()
            else:
                batch = self.replay_memory.sample(int(self.params["replay_sample_size"]))
                loss, priorities = self.compute_loss_and_priorities(batch)
                self.update_Q(loss)
                self.replay_memory.set_priorities(priorities)
            if t % publish == 0:
                if board is not None:
                    board.publish()
                else:
                    self.shared_state["Q_state_dict"] = {k: v.detach().cpu() for k, v in self.Q.state_dict().items()}
            if t % self.params["remove_old_xp_freq"] == 0:
                self.replay_memory.remove_to_fit()
        if board is not None:
            board.wait(board._issued)
            # attached actors keep their mapping after the unlink; set weight_board_unlink=False to leave
            # the segment for a supervisor to inspect and remove
            board.close(unlink=bool(self.params.get("weight_board_unlink", True)))


class FlatGradients:
    """All parameter gradients as views of one flat buffer, so that the data-parallel reduction is one
    collective call (and zeroing is one memset).  Device-agnostic (tested with gloo on the CPU)."""

    def __init__(self, parameters):
        self.params = [p for p in parameters if p.requires_grad]
        total = sum(p.numel() for p in self.params)
        first = self.params[0]
        self.flat = torch.zeros(total, dtype=first.dtype, device=first.device)
        off = 0
        for p in self.params:
            p.grad = self.flat[off:off + p.numel()].view_as(p)
            off += p.numel()

    def zero(self):
        self.flat.zero_()

    def all_reduce(self, group=None):
        import torch.distributed as dist
        if dist.is_initialized() and dist.get_world_size(group) > 1:
            dist.all_reduce(self.flat, group=group)                      # sum: every rank divided by the global batch
        return self.flat


class ShardedLearner(Learner):
    """One rank of the data-parallel learner over a sharded replay (BASELINE configs[3]; no
    counterpart in the single-process reference).  Every rank owns a replay shard and a replica of
    the networks; per step it takes its share of the global stratified batch (fused K8+K3+K4 over
    peer memory), runs the unchanged model on it, all-reduces the flat gradient (NCCL) and applies
    the identical RMSprop step; K5+K2 write the new priorities into the local shard and push the
    shard's new total to the peers.  loss_mean is divided by the global batch on every rank.

    The rank's share is data dependent (proportional to its shard's priority mass) and is delivered in
    full whatever the skew: the sample outputs hold the whole global batch, and the model runs on exactly
    the rows the rank owns (one 4-byte device->host read of the row count per step; no stale rows ever
    enter the network, so the reference model's batch-coupled duelling mean, duelling_network.py:27,
    couples only live rows).  That coupling is per rank: a sharded step cannot reproduce the mean over the
    GLOBAL batch that a single process would subtract; with a per-row mean
    (qnet.DuellingQNet(batch_coupled_mean=False)) the all-reduced gradient equals the single-process
    gradient exactly, which is what tests/multi_gpu_check.py asserts."""

    def __init__(self, env_conf, learner_params, shared_state, shared_replay_memory, group=None, quota=None):
        super().__init__(env_conf, learner_params, shared_state, shared_replay_memory)
        import torch.distributed as dist
        from .sharded import ShardedSampler
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.sampler = ShardedSampler(shared_replay_memory, group=group, rank=self.rank, world_size=self.world)
        B = int(self.params["replay_sample_size"])
        self.global_batch = B * self.world
        # rows the sample outputs hold: the whole global batch (one rank may own every draw).  `quota` caps it
        # for callers that want bounded buffers; a rank that then owns more raises ST_QUOTA in check_status().
        self.max_rows = int(quota) if quota else self.global_batch
        self.last_rows = 0
        self._ready = False

    def setup(self):
        """Collective: move the networks to the device, make every replica equal to rank 0's
        (K9 snapshot -> broadcast -> K9 scatter), lay the gradients out flat, connect the shards."""
        import torch.distributed as dist
        from .weights import _DeviceBoard
        dev = self._to_device()
        for net in (self.Q, self.Q_double):
            board = _DeviceBoard(net)
            from . import _C as C_
            C_.call("apexb200_params_publish", board._h, None, None, 0, C_.stream_ptr())      # pack
            flat = board.flat()
            if self.world > 1:
                dist.broadcast(flat, src=dist.get_global_rank(self.group, 0) if self.group is not None else 0, group=self.group)
            C_.call("apexb200_params_load", board._h, None, 0, C_.stream_ptr())                # scatter
            torch.cuda.current_stream().synchronize()
            board.destroy()
        self.grads = FlatGradients(self.Q.parameters())
        self.sampler.connect_peers()
        self.sampler.push_on_update(True)
        self.sampler.exchange()
        self._ready = True
        return dev

    def after_insert(self):
        """Collective: inserts changed the shard totals at different rates; re-publish them."""
        self.sampler.exchange()

    def step_sharded(self):
        if not self._ready:
            self.setup()
        batch, _pos, cnt = self.sampler.sample_fused(self.global_batch, self.max_rows, f32=self._f32())
        n = int(cnt.item())                                   # this step's live rows (synchronises on the sample)
        self.last_rows = n
        rows = max(n, 1)                                      # a rank that owns no draw still takes part in the collectives
        shape = (rows,) + tuple(self.state_shape)
        S_t, S_tpn = (x[:rows].view(shape) if x.dtype == torch.float32 else x[:rows].view(shape).float()
                      for x in (batch.S_t, batch.S_tpn))
        loss, _delta, _prio, _ = self._loss_from_tensors(S_t, S_tpn, batch.A[:rows], batch.R[:rows], batch.Gamma[:rows],
                                                         replay=self.replay_memory, idx=batch.idx[:rows], count_dev=cnt,
                                                         mean_divisor=float(self.global_batch))
        self.grads.zero()
        loss.backward()
        self.grads.all_reduce(self.group)
        self.optimizer.step()
        self.num_q_updates += 1
        self._maybe_sync_target()
        return loss.detach()              # this rank's share; the sum over ranks is the global mean loss
