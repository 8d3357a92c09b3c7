"""apex_b200 — B200-native prioritized replay + learner data path for Ape-X DQN.

Host-side mirror of the reference's replay.py / learner.py / actor.py classes
over the C-ABI CUDA library libapexb200.so (include/apexb200.h).  The directory
is called ``ape-x-dqn_b200``; import it as ``apex_b200`` through the loader at
the repository root (``import apex_b200``).
"""
from . import _C                                              # noqa: F401
from .replay import ReplayMemory, N_Step_Transition, Batch, KeyIndex, stream_offsets   # noqa: F401
from .actor import ExperienceBuffer, Actor, Transition        # noqa: F401
from .learner import Learner, ShardedLearner, FlatGradients, td_loss   # noqa: F401
from . import ops                                             # noqa: F401
from .sharded import ShardedSampler                           # noqa: F401
from .pipeline import HostStepPipeline, ShardedHostStepPipeline                        # noqa: F401
from .ingest import VectorActorIngest, ActorStagingRing, RingIngest   # noqa: F401
from .weights import (WeightPublisher, WeightSubscriber, DeviceWeightSubscriber, BatchedActorPolicy,   # noqa: F401
                      actor_policy, actor_epsilons, param_layout)
from .vector_actor import VectorActor                         # noqa: F401

__all__ = ["ReplayMemory", "Learner", "Actor", "ExperienceBuffer", "N_Step_Transition", "Transition", "Batch",
           "KeyIndex", "stream_offsets", "td_loss", "ops", "ShardedSampler", "HostStepPipeline", "ShardedHostStepPipeline", "VectorActorIngest", "ActorStagingRing", "RingIngest",
           "WeightPublisher", "WeightSubscriber", "DeviceWeightSubscriber", "BatchedActorPolicy", "actor_policy",
           "actor_epsilons", "param_layout", "VectorActor", "ShardedLearner", "FlatGradients"]
