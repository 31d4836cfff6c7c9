"""Instance data (eam_rl4co_b200/data.py).  CPU: the npz format against the reference's own
generate_data / utils / CVRPEnv.load_data where the checkout is present.  GPU: the on-device generator bit-exactly
against oracle/data_oracle.py, its distribution, determinism and shard invariance."""
import numpy as np
import pytest
import torch

import eam_rl4co_b200 as P
from eam_rl4co_b200 import data
from oracle import data_oracle as DO
from oracle.ref_loader import load_reference, reference_available

DEV = "cuda"


def test_npz_round_trip_and_cvrp_normalisation(tmp_path):
    np.random.seed(3)
    d = data.generate_vrp_data(7, 20)
    assert d["locs"].shape == (7, 20, 2) and d["depot"].shape == (7, 2) and d["capacity"].shape == (7,)
    assert all(v.dtype == np.float32 for v in d.values())
    assert set(np.unique(d["demand"])) <= set(range(1, 10)) and float(d["capacity"][0]) == 30.0
    f = data.check_extension(str(tmp_path / "vrp20"))
    data.save_tensordict_to_npz({k: torch.from_numpy(v) for k, v in d.items()}, f)
    td = data.load_data("cvrp", f)
    assert td.batch_size[0] == 7
    assert torch.equal(td["locs"], torch.from_numpy(d["locs"]))
    assert torch.equal(td["demand"], torch.from_numpy(d["demand"] / d["capacity"][:, None]))
    t = data.generate_tsp_data(4, 9)
    assert t["locs"].shape == (4, 9, 2) and t["locs"].dtype == np.float32 and set(t) == {"locs"}


@pytest.mark.skipif(not reference_available(), reason="reference checkout not present")
def test_npz_format_matches_reference(tmp_path):
    ref = load_reference()
    import rl4co.data.generate_data as G
    import rl4co.data.utils as U
    for mine, theirs, args in ((data.generate_tsp_data, G.generate_tsp_data, (5, 10)),
                               (data.generate_vrp_data, G.generate_vrp_data, (5, 50))):
        np.random.seed(11)
        a = mine(*args)
        np.random.seed(11)
        b = theirs(*args)
        assert a.keys() == b.keys()
        for k in a:
            assert a[k].dtype == b[k].dtype and np.array_equal(a[k], b[k]), k
    f = str(tmp_path / "x.npz")
    np.random.seed(12)
    d = G.generate_vrp_data(6, 20)
    np.savez(f, **d)
    mine, theirs = data.load_npz_to_tensordict(f), U.load_npz_to_tensordict(f)
    assert sorted(mine.keys()) == sorted(theirs.keys()) and mine.batch_size[0] == theirs.batch_size[0]
    for k in theirs.keys():
        assert torch.equal(mine[k], torch.as_tensor(theirs[k])), k
    td_ref = ref.CVRPEnv.load_data(f)
    assert torch.equal(data.load_data("cvrp", f)["demand"], torch.as_tensor(td_ref["demand"]))


@pytest.mark.gpu
@pytest.mark.parametrize("env_name,n,B,off", [("tsp", 100, 64, 0), ("cvrp", 20, 33, 5), ("cvrp", 100, 16, 1 << 20)])
def test_device_generator_bit_exact_and_sharded(env_name, n, B, off):
    gen = P.DeviceGenerator(env_name, num_loc=n, seed=42, device=DEV)
    td = gen(B, instance_offset=off)
    exp = DO.generate(env_name, 42, off, B, n, gen.capacity)
    for k, v in exp.items():
        assert torch.equal(td[k].cpu(), torch.from_numpy(v)), k
    whole = gen(B + 3, instance_offset=off - 3 if off >= 3 else 0)
    if off >= 3:
        assert torch.equal(whole["locs"][3:], td["locs"])           # a shard equals its slice of the whole
    assert torch.equal(gen(B, instance_offset=off)["locs"], td["locs"])
    assert not torch.equal(gen(B, instance_offset=off, seed=43)["locs"], td["locs"])
    if env_name == "cvrp":
        assert td["capacity"].shape == (B, 1) and td["depot"].shape == (B, 2)
        d = td["demand"] * gen.capacity
        assert bool(((d.round() - d).abs() < 1e-4).all()) and d.min() >= 0.999 and d.max() <= 9.001


@pytest.mark.gpu
def test_device_generator_distribution_and_rollout():
    gen = P.DeviceGenerator("cvrp", num_loc=50, seed=7, device=DEV)
    td = gen(4096)
    x = td["locs"]
    assert 0.0 <= x.min().item() and x.max().item() < 1.0
    assert abs(x.mean().item() - 0.5) < 5e-3 and abs(x.var().item() - 1 / 12) < 2e-3
    counts = torch.bincount((td["demand"] * gen.capacity).round().long().flatten(), minlength=10)[1:].float()
    assert (counts / counts.sum() - 1 / 9).abs().max().item() < 5e-3
    # feeds the path: env.reset -> policy
    env = P.CVRPEnv(generator_params=dict(num_loc=50), device=DEV)
    pol = P.POMOPolicy("cvrp").eval().to(DEV)
    small = gen(8)
    out = pol(env.reset(small), env, decode_type="multistart_greedy")
    assert out["reward"].shape[0] == 8 * 50 and torch.isfinite(out["reward"]).all()
