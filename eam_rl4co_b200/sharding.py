"""Multi-GPU host logic of the rollout path: one process per GPU, instances sharded, no data-path collective.

SURVEY.md §8(e): every instance (with all its S starts) is independent during the rollout, so the
batch axis is split contiguously over the ranks (as Lightning DDP shards the dataset batch in the
reference, rl4co/utils/trainer.py:86-89,119) and each rank runs its own persistent kernel.  The only
communication is *after* the rollout: an optional gather of results for reporting, the max-over-ranks of
device timings, and - for training callers - the flat-buffer all-reduce of gradients / logged means
(rl4co/models/rl/common/base.py:233-240).  Everything here is plain torch.distributed (NCCL on GPUs,
gloo in the CPU tests) and works un-initialised as a single rank.
"""
from typing import Callable, Dict, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def world_info(group=None) -> Tuple[int, int]:
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def shard_bounds(num_instances: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, balanced split of the instance axis: ranks differ by at most one instance."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    return num_instances * rank // world, num_instances * (rank + 1) // world


def shard_instances(inst: Dict[str, torch.Tensor], rank: int, world: int):
    """inst: dict of [B, ...] tensors (or a TensorDict) -> (this rank's slice, global index of its first instance).
    The offset keys the counter noise (`instance_offset=`) so sharded and unsharded sampling agree."""
    B = next(iter(inst.values())).shape[0]
    lo, hi = shard_bounds(B, rank, world)
    return {k: v[lo:hi] for k, v in inst.items()}, lo


def _unbatchify(x: torch.Tensor, S: int) -> torch.Tensor:
    """start-major rows [S*B, ...] -> [B, S, ...] (rl4co/utils/ops.py:35-56)."""
    if S <= 1:
        return x.unsqueeze(1)
    return x.view(S, x.shape[0] // S, *x.shape[1:]).transpose(0, 1)


def _batchify_rows(x: torch.Tensor) -> torch.Tensor:
    """[B, S, ...] -> start-major rows [S*B, ...]."""
    return x.transpose(0, 1).reshape(x.shape[0] * x.shape[1], *x.shape[2:])


def _all_gather_uneven(x: torch.Tensor, group=None) -> Sequence[torch.Tensor]:
    """all_gather of tensors whose first two dims (instances, and for actions the trailing T) may differ."""
    rank, world = world_info(group)
    if world == 1:
        return [x]
    shape = torch.tensor(list(x.shape) + [0] * (4 - x.dim()), dtype=torch.int64, device=x.device)
    shapes = [torch.zeros_like(shape) for _ in range(world)]
    dist.all_gather(shapes, shape, group=group)
    shapes = [s.tolist()[:x.dim()] for s in shapes]
    mx = [max(s[d] for s in shapes) for d in range(x.dim())]
    pad = torch.zeros(mx, dtype=x.dtype, device=x.device)
    pad[tuple(slice(0, n) for n in x.shape)] = x
    bufs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(bufs, pad.contiguous(), group=group)
    return [b[tuple(slice(0, n) for n in s)] for b, s in zip(bufs, shapes)]


def gather_rollout(out: Dict[str, torch.Tensor], num_starts: int, group=None,
                   keys: Sequence[str] = ("reward", "log_likelihood", "actions")) -> Dict[str, torch.Tensor]:
    """Assemble the unsharded result from per-rank rollouts (outside any timed region).

    Rows are start-major *within* a shard (row = s*B_k + b); the global order is start-major over the
    concatenated instances.  CVRP shards can stop at different T (each is its own batch-max, SURVEY.md
    §7.3-5): actions are right-padded with the depot (0), which is what finished rows emit."""
    S = max(int(num_starts), 1)
    res = {}
    for k in keys:
        if k not in out:
            continue
        parts = _all_gather_uneven(_unbatchify(out[k], S).contiguous(), group)
        if parts[0].dim() >= 3:     # [B_k, S, T]: pad T
            T = max(p.shape[2] for p in parts)
            parts = [torch.nn.functional.pad(p, (0, T - p.shape[2])) for p in parts]
        res[k] = _batchify_rows(torch.cat(list(parts), 0)) if S > 1 else torch.cat(list(parts), 0).squeeze(1)
    return res


def sharded_rollout(rollout_fn: Callable[..., Dict[str, torch.Tensor]], inst: Dict[str, torch.Tensor],
                    num_starts: int = 0, gather: bool = False, group=None, **kw) -> Dict[str, torch.Tensor]:
    """Run `rollout_fn(shard, instance_offset=lo, **kw)` on this rank's instances; optionally gather."""
    rank, world = world_info(group)
    shard, lo = shard_instances(inst, rank, world)
    out = rollout_fn(shard, instance_offset=lo, **kw)
    return gather_rollout(out, num_starts, group) if gather else out


def max_over_ranks(values: Sequence[float], device=None, group=None) -> Sequence[float]:
    """Device timings are reported as the MAX over ranks (bench.py contract)."""
    _, world = world_info(group)
    if world == 1:
        return list(values)
    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return t.tolist()


def allreduce_mean_(tensors: Sequence[torch.Tensor], extras: Optional[torch.Tensor] = None, group=None):
    """ONE collective per training step: gradients (or any tensors) flattened into a single fp32 buffer with
    the logged scalars `extras` appended, all-reduced (sum) and divided by the world size - the DDP mean
    semantics of the reference (trainer.py:86-89) plus `log_dict(sync_dist=True)` (rl/common/base.py:233-240).
    Results are written back in place; returns the reduced extras."""
    _, world = world_info(group)
    if world == 1:
        return extras
    flat = [t.reshape(-1).to(torch.float32) for t in tensors]
    if extras is not None:
        flat.append(extras.reshape(-1).to(torch.float32))
    buf = torch.cat(flat)
    dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group)
    buf /= world
    o = 0
    for t in tensors:
        n = t.numel()
        t.copy_(buf[o:o + n].view_as(t))
        o += n
    return buf[o:].view_as(extras) if extras is not None else None
