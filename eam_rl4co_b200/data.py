"""Instance data either side of the rollout path.

* On-disk format: `.npz` with `locs [M,N,2] f32`; VRP adds `depot [M,2]`, `demand [M,N]` (integers 1..9 stored as f32)
  and `capacity [M]` (rl4co/data/generate_data.py:39-82, rl4co/data/utils.py:11-31); the CVRP loader divides the demand
  by the capacity (cvrp/env.py:187-194).  Host side, numpy - same keys, dtypes and np.random call order as the
  reference, so a seeded file is reproduced.
* On-device generation: `DeviceGenerator` fills a shard's instances with `rl4co_generate_instances` (counter hash keyed
  by the global instance index), so multi-GPU training never touches the host and shards agree with the whole batch.
"""
import ctypes as C
import os

import numpy as np
import torch

from . import _lib
from .tensordict_lite import TensorDict

CAPACITIES = {10: 20.0, 15: 25.0, 20: 30.0, 30: 33.0, 40: 37.0, 50: 40.0, 60: 43.0, 75: 45.0, 100: 50.0, 125: 55.0,
              150: 60.0, 200: 70.0, 500: 100.0, 1000: 150.0}


def generate_tsp_data(dataset_size: int, tsp_size: int) -> dict:
    """generate_data.py:39-42."""
    return {"locs": np.random.uniform(size=(dataset_size, tsp_size, 2)).astype(np.float32)}


def generate_vrp_data(dataset_size: int, vrp_size: int, capacities: dict = None) -> dict:
    """generate_data.py:45-82 (Kool et al. capacities; demand uniform integer 1..9)."""
    caps = dict(CAPACITIES)
    if capacities is not None:
        caps.update({k: v for k, v in capacities.items() if k in caps})
    return {
        "depot": np.random.uniform(size=(dataset_size, 2)).astype(np.float32),
        "locs": np.random.uniform(size=(dataset_size, vrp_size, 2)).astype(np.float32),
        "demand": np.random.randint(1, 10, size=(dataset_size, vrp_size)).astype(np.float32),
        "capacity": np.full(dataset_size, caps[vrp_size]).astype(np.float32),
    }


def check_extension(filename: str, extension: str = ".npz") -> str:
    """utils.py:34-38."""
    return filename if os.path.splitext(filename)[1] == extension else filename + extension


def save_tensordict_to_npz(td, filename: str, compress: bool = False):
    """utils.py:23-31."""
    x = {k: (v.detach().cpu().numpy() if torch.is_tensor(v) else np.asarray(v)) for k, v in td.items()}
    (np.savez_compressed if compress else np.savez)(filename, **x)


def load_npz_to_tensordict(filename: str) -> TensorDict:
    """utils.py:11-20: every array of the file, batch size = leading dimension of the first one."""
    x = {k: torch.from_numpy(v) for k, v in dict(np.load(filename)).items()}
    return TensorDict(x, batch_size=[next(iter(x.values())).shape[0]])


def load_data(env_name: str, fpath: str) -> TensorDict:
    """RL4COEnvBase.load_data; CVRP normalises the demand by the capacity (cvrp/env.py:187-194)."""
    td = load_npz_to_tensordict(fpath)
    if env_name == "cvrp":
        td.set("demand", td["demand"] / td["capacity"][:, None])
    return td


class DeviceGenerator:
    """Generator with the reference's call signature (`generator(batch_size)` -> TensorDict, envs/common/utils.py) whose
    values are produced on the GPU.  `instance_offset` is the global index of the shard's first instance."""

    def __init__(self, env_name: str = "tsp", num_loc: int = 20, capacity: float = None, seed: int = 0,
                 device: str = "cuda"):
        if env_name not in ("tsp", "cvrp"):
            raise NotImplementedError(f"env '{env_name}' is outside the accelerated path (tsp, cvrp)")
        self.env_name, self.num_loc, self.seed, self.device = env_name, num_loc, seed, torch.device(device)
        if capacity is None:
            capacity = CAPACITIES[min(CAPACITIES, key=lambda k: abs(k - num_loc))]
        self.capacity = float(capacity)

    def __call__(self, batch_size, instance_offset: int = 0, seed: int = None) -> TensorDict:
        B = batch_size if isinstance(batch_size, int) else int(batch_size[0])
        n, dev = self.num_loc, self.device
        cvrp = self.env_name == "cvrp"
        locs = torch.empty(B, n, 2, dtype=torch.float32, device=dev)
        depot = torch.empty(B, 2, dtype=torch.float32, device=dev) if cvrp else None
        demand = torch.empty(B, n, dtype=torch.float32, device=dev) if cvrp else None
        L = _lib.lib()
        _lib.check(L.rl4co_generate_instances(
            _lib.ENV_CVRP if cvrp else _lib.ENV_TSP, C.c_uint32(self.seed if seed is None else seed),
            C.c_int64(instance_offset), C.c_int64(B), n, C.c_float(self.capacity), _lib.ptr(locs), _lib.ptr(depot),
            _lib.ptr(demand), _lib.stream_ptr()), "rl4co_generate_instances")
        td = {"locs": locs}
        if cvrp:
            td.update(depot=depot, demand=demand, capacity=torch.full((B, 1), self.capacity, device=dev))
        return TensorDict(td, batch_size=[B])
