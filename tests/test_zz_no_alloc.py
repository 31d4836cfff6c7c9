"""GPU: the allocation-free rollout entry point `rl4co_rollout_noalloc` (include/rl4co_b200.h) against the existing
`rl4co_rollout`, called through the C ABI, on two calls in a row (the static scratch keeps the previous call's
contents).  No counterpart in the reference.  Collected last on
purpose (file name)."""
import ctypes as C

import pytest
import torch

import eam_rl4co_b200 as P
from eam_rl4co_b200 import _lib

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _case(env_name, N, B, seed):
    torch.manual_seed(seed)
    inst = (P.TSPGenerator if env_name == "tsp" else P.CVRPGenerator)(num_loc=N)(B)
    inst = {k: v for k, v in inst.items() if k != "capacity"}
    torch.manual_seed(4321)
    pol = P.POMOPolicy(env_name).eval().to(DEV)
    env = P.get_env(env_name, generator_params=dict(num_loc=N), device=DEV)
    td = env.reset(P.TensorDict({k: v.to(DEV) for k, v in inst.items()}, batch_size=[B]))
    with torch.no_grad():
        h, _ = pol.encoder(td)
        cache = pol.decoder._precompute_cache(h, num_starts=N)      # tables + caller-owned operand image
    return td, cache


def _call(fn, env_name, td, cache, mode, with_image):
    """One rollout through the C ABI; every buffer is allocated HERE, before the call."""
    B, n = td["action_mask"].shape
    S = n if env_name == "tsp" else n - 1
    rows, T = B * S, (n if env_name == "tsp" else 2 * n)
    out = dict(actions=torch.zeros(rows, T, dtype=torch.int64, device=DEV),
               logp=torch.zeros(rows, T, device=DEV), reward=torch.empty(rows, device=DEV),
               ll=torch.empty(rows, device=DEV), lengths=torch.empty(rows, dtype=torch.int32, device=DEV),
               meta=torch.zeros(3, dtype=torch.int32, device=DEV))
    locs = td["locs"].contiguous()
    cvrp = env_name == "cvrp"
    demand = td["demand"].contiguous() if cvrp else None
    cap = td["vehicle_capacity"].contiguous() if cvrp else None
    cs = cache.cstruct()
    if not with_image:
        cs.tc_img = None                                          # the no-alloc entry must not need it
    args = _lib.RolloutArgs(
        env=_lib.ENV_CVRP if cvrp else _lib.ENV_TSP, mode=mode, num_starts=S, max_steps=T, tanh_clipping=10.0,
        temperature=1.0, seed=77, inst_offset=5, locs=locs.data_ptr(),
        demand=demand.data_ptr() if cvrp else None, vehicle_capacity=cap.data_ptr() if cvrp else None,
        gumbel=None, actions_in=None, actions_in_stride=0, actions_in_T=0, actions=out["actions"].data_ptr(),
        logp=out["logp"].data_ptr(), logp_all=None, reward=out["reward"].data_ptr(),
        log_likelihood=out["ll"].data_ptr(), lengths=out["lengths"].data_ptr(), max_len=out["meta"].data_ptr(),
        err=out["meta"].data_ptr() + 4, impl=_lib.IMPL_TC, phase_cycles=None, start_mod=0)
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    rc = fn(C.byref(cs), C.byref(args), _lib.stream_ptr())
    _lib.check(rc, "rollout")
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before                # nothing went through torch's allocator
    assert out["meta"][1].item() == 0, "error word set"
    return out


@pytest.mark.parametrize("env_name,mode", [("tsp", _lib.MODE_GREEDY), ("cvrp", _lib.MODE_SAMPLING)])
def test_noalloc_rollout_matches_rollout_twice(env_name, mode):
    L = _lib.lib()
    # two calls in a row on DIFFERENT shapes and instances: the second must not see the first one's image
    for N, B, seed in ((50, 7, 1), (20, 12, 2)):
        assert B <= L.rl4co_noalloc_max_instances(N + (env_name == "cvrp"))
        td, cache = _case(env_name, N, B, seed)
        ref = _call(L.rl4co_rollout, env_name, td, cache, mode, with_image=True)
        got = _call(L.rl4co_rollout_noalloc, env_name, td, cache, mode, with_image=False)
        assert torch.equal(got["actions"], ref["actions"])
        assert torch.equal(got["lengths"], ref["lengths"]) and got["meta"][0].item() == ref["meta"][0].item()
        # same kernels on the same operand bytes: the tolerance of the parity tests (1e-5 relative) is not even needed
        rel = lambda a, b: ((a - b).abs() / b.abs().clamp_min(1.0)).max().item()
        assert rel(got["reward"], ref["reward"]) <= 1e-5 and rel(got["ll"], ref["ll"]) <= 1e-5
        assert rel(got["logp"], ref["logp"]) <= 1e-5


def test_noalloc_rollout_rejects_oversize_and_unsupported_shapes():
    L = _lib.lib()
    N = 100
    mx = int(L.rl4co_noalloc_max_instances(N))
    assert mx == (8 << 20) // (6 * 112 * 256)
    z = torch.zeros(8, device=DEV)
    p = z.data_ptr()
    cs = _lib.Cache(p, p, p, p, p, p, None, None, mx + 1, N, 128, 8)
    args = _lib.RolloutArgs(env=_lib.ENV_TSP, mode=_lib.MODE_GREEDY, num_starts=N, max_steps=N, tanh_clipping=10.0,
                            temperature=1.0, locs=p, actions=p, logp=p, reward=p, log_likelihood=p, lengths=p,
                            max_len=p, err=p)
    rc = L.rl4co_rollout_noalloc(C.byref(cs), C.byref(args), None)
    assert rc == -1 and b"exceed the static scratch" in L.rl4co_last_error()
    cs.B = 1
    args.num_starts = 1                                           # single-start rows have no operand image
    rc = L.rl4co_rollout_noalloc(C.byref(cs), C.byref(args), None)
    assert rc == -1 and b"tensor-path shapes only" in L.rl4co_last_error()
