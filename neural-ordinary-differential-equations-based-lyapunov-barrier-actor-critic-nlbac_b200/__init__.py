"""nlbac_b200 — B200 (sm_100a) drop-in for the NLBAC neural-ODE hot path.

Host mirror of the reference's call surface over the C ABI in include/nlbac.h:
``odeint`` (torchdiffeq-compatible), ``NeuralODEModel``, ``train_step``.  Import as ``nlbac_b200``
(the repo-root shim) — the directory name carries the reference's full name.
"""
from . import _lib
from ._lib import build
from .odeint import odeint, describe
from .model import NeuralODEModel, train_step
from .mlp import QNetwork, LyaNetwork, BarrierNetwork, GaussianPolicy, weights_init_
from .tail import DynamicsModel, FusedAdam, soft_update, hard_update, constraint_sums, lookahead, make_constraint_desc
from . import constraints
from . import parallel
from .replay import ReplayMemory
from .graphs import GraphedTrainStep
from .agent import GraphedUpdate
from .envs import VecUnicycleEnv, VecSimulatedCarsEnv, VecPvtolEnv

__all__ = ["odeint", "NeuralODEModel", "train_step", "build", "describe", "QNetwork", "LyaNetwork", "BarrierNetwork",
           "GaussianPolicy", "weights_init_", "DynamicsModel", "FusedAdam", "soft_update", "hard_update",
           "constraint_sums", "lookahead", "make_constraint_desc", "constraints", "parallel", "ReplayMemory", "GraphedTrainStep", "GraphedUpdate", "VecUnicycleEnv", "VecSimulatedCarsEnv", "VecPvtolEnv"]
