"""TEST INFRASTRUCTURE ONLY - numpy restatement of the on-device instance generator (csrc/data_kernels.cu): the same
counter hash as oracle/rollout_oracle.py:_mix32, the input spec of tsp/generator.py:53-62 and
cvrp/generator.py:120-147.  Never imported by the product path."""
import numpy as np

M32 = np.uint64(0xFFFFFFFF)


def _mix32(x):
    x = x.astype(np.uint64) & M32
    x ^= x >> np.uint64(16)
    x = (x * np.uint64(0x7FEB352D)) & M32
    x ^= x >> np.uint64(15)
    x = (x * np.uint64(0x846CA68B)) & M32
    x ^= x >> np.uint64(16)
    return x


def generate(env_name: str, seed: int, inst_offset: int, B: int, n_loc: int, capacity: float = 1.0) -> dict:
    cvrp = env_name == "cvrp"
    n_xy = 2 * (n_loc + (1 if cvrp else 0))
    per = n_xy + (n_loc if cvrp else 0)
    b = (np.arange(B, dtype=np.uint64) + np.uint64(inst_offset)) & M32
    key = _mix32(np.uint64(seed) ^ ((b * np.uint64(0x9E3779B1)) & M32) ^ np.uint64(0x5BD1E995))
    e = np.arange(per, dtype=np.uint64)
    h = _mix32(key[:, None] ^ ((e[None, :] * np.uint64(0x85EBCA77)) & M32))
    u = ((h >> np.uint64(8)).astype(np.float32) * np.float32(5.9604644775390625e-08)).astype(np.float32)
    out = {}
    if cvrp:
        out["depot"] = u[:, :2].copy()
        out["locs"] = u[:, 2:n_xy].reshape(B, n_loc, 2).copy()
        d = (u[:, n_xy:] * np.float32(9.0)).astype(np.int32) + 1
        out["demand"] = (d.astype(np.float32) / np.float32(capacity)).astype(np.float32)
    else:
        out["locs"] = u.reshape(B, n_loc, 2).copy()
    return out
