"""CUDA-graph replay of `policy(td, env, decode_type=...)` for a fixed batch shape.

One rollout is ~40 launches (env.reset fills, 5 encoder kernels per layer, cache precompute, operand pack, the
persistent rollout kernel).  At the reference's multi-GPU shard sizes (config c4 over 8 GPUs: 128 instances per GPU,
rl4co/utils/trainer.py:72-89,119) the device work is ~1.5 ms and the launch overhead of the Python call shows; a
captured graph replays the same launches with one driver call.  Nothing about the arithmetic changes: the graph holds
the very kernels the eager call launches.

    roll = capture_rollout(policy, env, locs_example, decode_type="multistart_greedy")
    out = roll(locs)            # locs: host (pinned) or device [B, N, 2]; out: static device tensors, valid until the next call
    roll.raise_on_error()       # reads the sticky error word (one 12-byte D2H)

Limits: TSP (CVRP's `actions.shape[1]` is the batch-max tour length, a host-side read - constructive/base.py:236), fixed
`B`, `N`, decode type and decoding kwargs; counter-noise sampling replays the captured seed unless `noise_seed_tensor`
semantics are added, so sampling callers should keep the eager call.
"""
import torch

from . import _lib
from .tensordict_lite import TensorDict


class GraphedRollout:
    def __init__(self, policy, env, locs_example: torch.Tensor, decode_type: str = "multistart_greedy", warmup: int = 3,
                 **kwargs):
        if env.name != "tsp":
            raise NotImplementedError("capture_rollout covers TSP; CVRP rollouts read their batch-max length on the host")
        if "sampling" in decode_type and kwargs.get("noise") is None:
            raise NotImplementedError("a captured sampling rollout would replay one noise seed; use the eager call")
        dev = next(policy.parameters()).device
        self.policy, self.env = policy, env
        self.static_in = locs_example.to(dev, copy=True).contiguous()
        B = self.static_in.shape[0]
        kw = dict(decode_type=decode_type, check=False, **kwargs)

        def run():
            td = env.reset(TensorDict({"locs": self.static_in}, batch_size=[B]))
            with torch.no_grad():
                out = policy(td, env, **kw)
            return out, policy.last_status

        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(max(warmup, 1)):          # packs weight tiles, sets kernel attributes, warms the allocator
                run()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.out, self.status = run()

    def __call__(self, locs: torch.Tensor):
        self.static_in.copy_(locs, non_blocking=True)
        self.graph.replay()
        return self.out

    def raise_on_error(self):
        _lib.raise_on_err(self.status[1:])


def capture_rollout(policy, env, locs_example, decode_type="multistart_greedy", **kwargs) -> GraphedRollout:
    return GraphedRollout(policy, env, locs_example, decode_type, **kwargs)
