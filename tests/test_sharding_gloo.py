"""CPU, world_size 2 over gloo: the multi-GPU host logic (instance sharding, noise keyed by the global instance
index, result gather with CVRP's ragged T, max-over-ranks, flat-buffer mean all-reduce).  The oracle stands in
for the CUDA rollout (no GPU here); the GPU twin of the invariance claim is
tests/test_rollout_gpu.py::test_sharding_is_invariant."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import eam_rl4co_b200 as P
from eam_rl4co_b200 import sharding
from oracle import rollout_oracle as O


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _problem(env_name, B, N):
    torch.manual_seed(1234)
    inst = O.generate_tsp(B, N) if env_name == "tsp" else O.generate_cvrp(B, N)
    torch.manual_seed(4321)
    pol = P.POMOPolicy(env_name).eval()
    w = {k: v.detach().clone() for k, v in pol.decoder.state_dict().items()}
    return inst, pol, w


def _encode(pol, env_name, inst):
    td = dict(inst)
    if env_name == "cvrp":      # generator layout -> env layout (depot prepended, cvrp/env.py:112)
        td = {"locs": torch.cat((inst["depot"][:, None, :], inst["locs"]), 1), "demand": inst["demand"]}
    with torch.inference_mode():
        h, _ = pol.encoder(td)
    return h


def _worker(rank, world, port, env_name, B, N, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(2)
        inst, pol, w = _problem(env_name, B, N)

        def rollout_fn(shard, instance_offset=0):
            with torch.inference_mode():
                out = O.rollout(env_name, shard, _encode(pol, env_name, shard), w, "multistart_sampling",
                                use_graph_context=False, noise_seed=11, instance_offset=instance_offset)
            return {k: out[k] for k in ("reward", "log_likelihood", "actions")}

        lo, hi = sharding.shard_bounds(B, rank, world)
        got = sharding.sharded_rollout(rollout_fn, inst, num_starts=N, gather=True)
        t = sharding.max_over_ranks([float(rank + 1), 10.0 - rank])
        g = [torch.full((3,), float(rank)), torch.full((2, 2), 2.0 * rank)]
        extras = sharding.allreduce_mean_(g, torch.tensor([float(hi - lo), 1.0]))
        if rank == 0:
            ret.update(got={k: v.clone() for k, v in got.items()}, t=t, g=[x.clone() for x in g],
                       extras=extras.clone(), bounds=(lo, hi))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("env_name,B,N", [("tsp", 5, 10), ("cvrp", 5, 10)])
def test_sharded_rollout_equals_unsharded(env_name, B, N):
    world = 2
    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(world, _free_port(), env_name, B, N, ret), nprocs=world, join=True)
    inst, pol, w = _problem(env_name, B, N)
    with torch.inference_mode():
        ref = O.rollout(env_name, inst, _encode(pol, env_name, inst), w, "multistart_sampling",
                        use_graph_context=False, noise_seed=11)
    got = ret["got"]
    T = ref["actions"].shape[1]
    assert got["actions"].shape[0] == B * N and got["actions"].shape[1] <= T
    a = torch.nn.functional.pad(got["actions"], (0, T - got["actions"].shape[1]))
    assert torch.equal(a, ref["actions"])                       # uneven shards (2 + 3), ragged T, start-major order
    assert torch.allclose(got["reward"], ref["reward"], rtol=1e-6, atol=1e-6)
    assert torch.allclose(got["log_likelihood"], ref["log_likelihood"], rtol=1e-5, atol=1e-5)
    assert ret["bounds"] == (0, 2)
    assert ret["t"] == [2.0, 10.0]                              # max over ranks
    assert torch.equal(ret["g"][0], torch.full((3,), 0.5)) and torch.equal(ret["g"][1], torch.full((2, 2), 1.0))
    assert torch.equal(ret["extras"], torch.tensor([2.5, 1.0]))  # mean of the per-rank instance counts


def test_shard_bounds_cover_and_balance():
    for B in (1, 7, 128, 1024):
        for world in (1, 2, 4, 8):
            spans = [sharding.shard_bounds(B, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == B
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharding.shard_bounds(8, 2, 2)


def test_single_rank_is_a_no_op():
    x = {"reward": torch.arange(6.0), "actions": torch.arange(12).view(6, 2)}
    out = sharding.gather_rollout(x, num_starts=3)
    assert torch.equal(out["reward"], x["reward"]) and torch.equal(out["actions"], x["actions"])
    assert sharding.max_over_ranks([1.5]) == [1.5]
