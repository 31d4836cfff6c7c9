"""eam_rl4co_b200 — B200-native construction-rollout path behind the RL4CO API.

Scope: TSP/CVRP `env.step` / `get_action_mask` / `get_reward`, the
AttentionModel/POMO decoder step, the decode strategies and the rollout loop of
`policy(td, env, decode_type=...)`.  Everything runs in hand-written sm_100a kernels behind
the C ABI of include/rl4co_b200.h; there is no CPU or PyTorch fallback.
"""
from . import _lib
from .tensordict_lite import TensorDict, batchify, unbatchify
from .envs import TSPEnv, CVRPEnv, TSPGenerator, CVRPGenerator, RL4COEnvBase, get_env
from .decoder import AttentionModelDecoder, PrecomputedCache
from .decoding import (DecodingStrategy, Greedy, Sampling, Evaluate, get_decoding_strategy,
                       get_log_likelihood, calculate_entropy)
from .encoder import AttentionModelEncoder
from .policy import AttentionModelPolicy
from . import sharding
from . import train
from . import pomo
from . import data
from .data import DeviceGenerator, load_data, load_npz_to_tensordict, save_tensordict_to_npz
from .pomo import StateAugmentation, best_of_starts, pomo_evaluate
from .graphs import GraphedRollout, capture_rollout


def POMOPolicy(env_name: str = "tsp", **kw) -> AttentionModelPolicy:
    """POMO defaults (rl4co/models/zoo/pomo/model.py:58-62)."""
    d = dict(num_encoder_layers=6, normalization="instance", use_graph_context=False)
    d.update(kw)
    return AttentionModelPolicy(env_name=env_name, **d)

__all__ = [n for n in dir() if not n.startswith("_")]
