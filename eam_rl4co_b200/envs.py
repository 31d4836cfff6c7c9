"""Host-side mirror of the reference's TSP / CVRP environments (RL4COEnvBase API).

Same method names, argument meaning, TensorDict keys/dtypes and error texts as
rl4co/envs/common/base.py:122-198, rl4co/envs/routing/tsp/env.py and
rl4co/envs/routing/cvrp/env.py; the arithmetic of `_step`, `get_action_mask` and
`_get_reward` runs in the sm_100a kernels of librl4co_b200.so (no CPU fallback).
`_reset` only allocates state (tsp/env.py:90-115, cvrp/env.py:102-130) and stays in torch.
"""
import ctypes as C
from typing import Optional

import torch

from . import _lib
from .tensordict_lite import TensorDict

# rl4co/envs/routing/cvrp/generator.py:17-32
CAPACITIES = {10: 20.0, 15: 25.0, 20: 30.0, 30: 33.0, 40: 37.0, 50: 40.0, 60: 43.0, 75: 45.0,
              100: 50.0, 125: 55.0, 150: 60.0, 200: 70.0, 500: 100.0, 1000: 150.0}


class TSPGenerator:
    """tsp/generator.py:53-62 — uniform locations in [min_loc, max_loc)."""

    def __init__(self, num_loc: int = 20, min_loc: float = 0.0, max_loc: float = 1.0, **_):
        self.num_loc, self.min_loc, self.max_loc = num_loc, min_loc, max_loc

    def __call__(self, batch_size, device="cpu"):
        bs = [batch_size] if isinstance(batch_size, int) else list(batch_size)
        locs = torch.distributions.Uniform(self.min_loc, self.max_loc).sample((*bs, self.num_loc, 2))
        return TensorDict({"locs": locs.to(device)}, batch_size=bs)


class CVRPGenerator:
    """cvrp/generator.py:120-147 — depot = first sampled loc, integer demand 1..9 / capacity."""

    def __init__(self, num_loc: int = 20, min_loc: float = 0.0, max_loc: float = 1.0,
                 min_demand: int = 1, max_demand: int = 10, vehicle_capacity: float = 1.0,
                 capacity: float = None, **_):
        self.num_loc, self.min_loc, self.max_loc = num_loc, min_loc, max_loc
        self.min_demand, self.max_demand = min_demand, max_demand
        self.vehicle_capacity = vehicle_capacity
        if capacity is None:
            capacity = CAPACITIES[min(CAPACITIES, key=lambda k: abs(k - num_loc))]
        self.capacity = capacity

    def __call__(self, batch_size, device="cpu"):
        bs = [batch_size] if isinstance(batch_size, int) else list(batch_size)
        U = torch.distributions.Uniform
        locs = U(self.min_loc, self.max_loc).sample((*bs, self.num_loc + 1, 2))
        depot, locs = locs[..., 0, :], locs[..., 1:, :]
        demand = U(float(self.min_demand - 1), float(self.max_demand - 1)).sample((*bs, self.num_loc))
        demand = (demand.int() + 1).float()
        capacity = torch.full((*bs, 1), self.capacity)
        td = {"locs": locs, "depot": depot, "demand": demand / self.capacity, "capacity": capacity}
        return TensorDict({k: v.to(device) for k, v in td.items()}, batch_size=bs)


class RL4COEnvBase:
    """The part of rl4co/envs/common/base.py the rollout path uses."""

    name = "base"

    def __init__(self, generator=None, generator_params: dict = None, check_solution: bool = True,
                 device: str = "cuda", **_):
        self.check_solution = check_solution
        self.device = torch.device(device)
        self.generator = generator if generator is not None else self._generator_cls(
            **(generator_params or {}))

    # base.py:122-134
    def step(self, td):
        td = self._step(td)
        return {"next": td}

    # base.py:136-149
    def reset(self, td: Optional[TensorDict] = None, batch_size=None):
        if batch_size is None:
            batch_size = td.batch_size
        if td is None or td.is_empty():
            td = self.generator(batch_size, device=self.device)
        batch_size = [batch_size] if isinstance(batch_size, int) else list(batch_size)
        out = self._reset(td, batch_size=batch_size)
        out.set("done", torch.zeros(*batch_size, 1, dtype=torch.bool, device=out.device))
        return out

    # base.py:188-198
    def get_reward(self, td, actions, check_solution: Optional[bool] = None):
        check = self.check_solution if check_solution is None else check_solution
        return self._get_reward(td, actions, check)

    def check_solution_validity(self, td, actions):
        self._get_reward(td, actions, True)

    def to(self, device):
        self.device = torch.device(device)
        return self

    def _reward_call(self, env_id, td, actions, check):
        locs = td["locs"].contiguous()
        actions = actions.contiguous()
        rows, T = actions.shape
        n = locs.shape[-2]
        reward = torch.empty(rows, dtype=torch.float32, device=locs.device)
        err = _lib.new_err(locs.device) if check else None
        demand = td["demand"].contiguous() if env_id == _lib.ENV_CVRP else None
        cap = td["vehicle_capacity"].contiguous() if env_id == _lib.ENV_CVRP else None
        rc = _lib.lib().rl4co_tour_reward(
            C.c_int32(env_id), _lib.ptr(locs), C.c_int64(locs.shape[0]), _lib.ptr(actions),
            C.c_int64(actions.stride(0)), C.c_int32(T), _lib.ptr(demand),
            C.c_int64(demand.shape[0] if demand is not None else 1), _lib.ptr(cap), _lib.ptr(reward),
            C.c_int32(1 if check else 0), _lib.ptr(err), C.c_int64(rows), C.c_int32(n), _lib.stream_ptr())
        _lib.check(rc, "rl4co_tour_reward")
        if check:
            _lib.raise_on_err(err)
        return reward


class TSPEnv(RL4COEnvBase):
    name = "tsp"
    _generator_cls = TSPGenerator

    def _reset(self, td, batch_size=None):  # tsp/env.py:90-115
        locs = td["locs"]
        dev, n = locs.device, locs.shape[-2]
        cur = torch.zeros(batch_size, dtype=torch.int64, device=dev)
        return TensorDict({
            "locs": locs, "first_node": cur, "current_node": cur,
            "i": torch.zeros(*batch_size, 1, dtype=torch.int64, device=dev),
            "action_mask": torch.ones(*batch_size, n, dtype=torch.bool, device=dev),
            "reward": torch.zeros(*batch_size, 1, dtype=torch.float32, device=dev),
        }, batch_size=batch_size)

    def _step(self, td):  # tsp/env.py:62-88 -> rl4co_tsp_step
        action = td["action"].contiguous()
        mask = td["action_mask"].contiguous()
        rows, n = mask.shape
        dev = mask.device
        i_in = td["i"].contiguous()
        first_in = td["first_node"].contiguous()
        mask_out = torch.empty_like(mask)
        cur_out = torch.empty_like(action)
        first_out = torch.empty_like(first_in)
        i_out = torch.empty_like(i_in)
        done = torch.empty(rows, dtype=torch.bool, device=dev)
        reward = torch.empty(rows, dtype=torch.bool, device=dev)
        flag = torch.empty(1, dtype=torch.int32, device=dev)
        rc = _lib.lib().rl4co_tsp_step(
            _lib.ptr(action), _lib.ptr(mask), _lib.ptr(i_in), _lib.ptr(first_in), _lib.ptr(mask_out),
            _lib.ptr(cur_out), _lib.ptr(first_out), _lib.ptr(i_out), _lib.ptr(done), _lib.ptr(reward),
            _lib.ptr(flag), C.c_int64(rows), C.c_int32(n), _lib.stream_ptr())
        _lib.check(rc, "rl4co_tsp_step")
        td.update({"first_node": first_out, "current_node": cur_out, "i": i_out,
                   "action_mask": mask_out, "reward": reward, "done": done})
        return td

    def _get_reward(self, td, actions, check=False):  # tsp/env.py:152-168
        return self._reward_call(_lib.ENV_TSP, td, actions, check)

    def get_num_starts(self, td):  # ops.py:120-130
        return td["action_mask"].shape[-1]

    def select_start_nodes(self, td, num_starts):  # ops.py:147-151
        return (torch.arange(num_starts, device=td.device).repeat_interleave(td.batch_size[0])
                % self.generator.num_loc)


class CVRPEnv(RL4COEnvBase):
    name = "cvrp"
    _generator_cls = CVRPGenerator

    def _reset(self, td, batch_size=None):  # cvrp/env.py:102-130
        dev = td["locs"].device
        n = td["locs"].shape[-2] + 1
        out = TensorDict({
            "locs": torch.cat((td["depot"][:, None, :], td["locs"]), -2),
            "demand": td["demand"],
            "current_node": torch.zeros(*batch_size, 1, dtype=torch.long, device=dev),
            "used_capacity": torch.zeros(*batch_size, 1, device=dev),
            "vehicle_capacity": torch.full((*batch_size, 1), self.generator.vehicle_capacity, device=dev),
            "visited": torch.zeros(*batch_size, n, dtype=torch.uint8, device=dev),
        }, batch_size=batch_size)
        out.set("action_mask", self.get_action_mask(out))
        return out

    @staticmethod
    def get_action_mask(td):  # cvrp/env.py:132-144 -> rl4co_cvrp_action_mask
        visited = td["visited"].contiguous()
        rows, n = visited.shape
        demand = td["demand"].contiguous()
        mask = torch.empty(rows, n, dtype=torch.bool, device=visited.device)
        rc = _lib.lib().rl4co_cvrp_action_mask(
            _lib.ptr(demand), C.c_int64(demand.shape[0]), _lib.ptr(td["used_capacity"].contiguous()),
            _lib.ptr(td["vehicle_capacity"].contiguous()), _lib.ptr(visited),
            _lib.ptr(td["current_node"].contiguous()), _lib.ptr(mask), C.c_int64(rows), C.c_int32(n),
            _lib.stream_ptr())
        _lib.check(rc, "rl4co_cvrp_action_mask")
        return mask

    def _step(self, td):  # cvrp/env.py:68-100 -> rl4co_cvrp_step_mask
        action = td["action"].contiguous()
        visited = td["visited"].contiguous()
        rows, n = visited.shape
        dev = visited.device
        demand = td["demand"].contiguous()
        used_out = torch.empty(rows, 1, dtype=torch.float32, device=dev)
        visited_out = torch.empty_like(visited)
        cur_out = torch.empty(rows, 1, dtype=torch.int64, device=dev)
        done = torch.empty(rows, dtype=torch.bool, device=dev)
        reward = torch.empty(rows, dtype=torch.bool, device=dev)
        mask = torch.empty(rows, n, dtype=torch.bool, device=dev)
        rc = _lib.lib().rl4co_cvrp_step_mask(
            _lib.ptr(action), _lib.ptr(demand), C.c_int64(demand.shape[0]),
            _lib.ptr(td["used_capacity"].contiguous()), _lib.ptr(td["vehicle_capacity"].contiguous()),
            _lib.ptr(visited), _lib.ptr(used_out), _lib.ptr(visited_out), _lib.ptr(cur_out),
            _lib.ptr(done), _lib.ptr(reward), _lib.ptr(mask), C.c_int64(rows), C.c_int32(n),
            _lib.stream_ptr())
        _lib.check(rc, "rl4co_cvrp_step_mask")
        td.update({"current_node": cur_out, "used_capacity": used_out, "visited": visited_out,
                   "reward": reward, "done": done, "action_mask": mask})
        return td

    def _get_reward(self, td, actions, check=False):  # cvrp/env.py:146-185
        return self._reward_call(_lib.ENV_CVRP, td, actions, check)

    def get_num_starts(self, td):  # ops.py:127-128
        return td["action_mask"].shape[-1] - 1

    def select_start_nodes(self, td, num_starts):  # ops.py:155-161
        return (torch.arange(num_starts, device=td.device).repeat_interleave(td.batch_size[0])
                % self.generator.num_loc + 1)


def get_env(name: str, **kw):
    return {"tsp": TSPEnv, "cvrp": CVRPEnv}[name](**kw)
