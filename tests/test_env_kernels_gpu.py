"""GPU: env step / mask / reward kernels, called through the C ABI via the host mirror,
bit-exact against the oracle on random feasible action sequences."""
import pytest
import torch

import eam_rl4co_b200 as P
from oracle import rollout_oracle as O
from helpers import inst_to_env_td

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.mark.parametrize("N,B,S", [(20, 64, 1), (50, 33, 1), (100, 8, 100), (7, 5, 7), (128, 3, 1)])
def test_tsp_step_bit_exact(N, B, S):
    torch.manual_seed(N * 1000 + B)
    inst = O.generate_tsp(B, N)
    env, td = inst_to_env_td(P, "tsp", inst, DEV)
    ref = O.tsp_reset(inst["locs"])
    if S > 1:
        td = P.batchify(td, S)
        ref = {k: O.batchify(v, S) for k, v in ref.items()}
    acts = []
    for step in range(N):
        a = torch.multinomial(ref["action_mask"].float(), 1).squeeze(1)
        acts.append(a)
        ref["action"] = a
        ref = O.tsp_step(ref)
        td.set("action", a.to(DEV))
        td = env.step(td)["next"]
        for k in ("action_mask", "current_node", "first_node", "i", "done", "reward"):
            assert td[k].dtype == ref[k].dtype and td[k].shape == ref[k].shape, k
            assert torch.equal(td[k].cpu(), ref[k]), (k, step)
    assert bool(td["done"].all())
    actions = torch.stack(acts, 1)
    r = env.get_reward(td, actions.to(DEV))
    r_ref = O.tsp_reward(ref["locs"], actions)
    assert torch.allclose(r.cpu(), r_ref, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("N,B,S", [(20, 64, 1), (50, 17, 1), (100, 6, 100), (5, 9, 5), (127, 3, 1)])
def test_cvrp_step_mask_bit_exact(N, B, S):
    torch.manual_seed(N * 1000 + B + 1)
    inst = O.generate_cvrp(B, N)
    env, td = inst_to_env_td(P, "cvrp", inst, DEV)
    ref = O.cvrp_reset(inst["locs"], inst["depot"], inst["demand"])
    assert torch.equal(td["action_mask"].cpu(), ref["action_mask"])
    if S > 1:
        td = P.batchify(td, S)
        ref = {k: O.batchify(v, S) for k, v in ref.items()}
    acts = []
    step = 0
    while not ref["done"].all():
        a = torch.multinomial(ref["action_mask"].float(), 1).squeeze(1)
        acts.append(a)
        ref["action"] = a
        ref = O.cvrp_step(ref)
        td.set("action", a.to(DEV))
        td = env.step(td)["next"]
        for k in ("action_mask", "current_node", "used_capacity", "visited", "done", "reward"):
            assert td[k].dtype == ref[k].dtype and td[k].shape == ref[k].shape, k
            assert torch.equal(td[k].cpu(), ref[k]), (k, step)     # used_capacity: same two fp32 ops
        step += 1
    actions = torch.stack(acts, 1)
    r = env.get_reward(td, actions.to(DEV))      # includes check_solution_validity
    assert torch.allclose(r.cpu(), O.cvrp_reward(ref["locs"], actions), rtol=1e-5, atol=1e-5)
    assert torch.equal(env.get_action_mask(td).cpu(), O.cvrp_action_mask(ref))


def test_validity_checks_raise_reference_messages():
    torch.manual_seed(0)
    inst = O.generate_tsp(4, 10)
    env, td = inst_to_env_td(P, "tsp", inst, DEV)
    good = torch.stack([torch.randperm(10) for _ in range(4)]).to(DEV)
    env.get_reward(td, good)
    bad = good.clone()
    bad[2, 3] = bad[2, 4]
    with pytest.raises(AssertionError, match="Invalid tour"):
        env.get_reward(td, bad)
    env.get_reward(td, bad, check_solution=False)
    inst = O.generate_cvrp(2, 6)
    inst["demand"] = torch.full((2, 6), 0.4)
    env, td = inst_to_env_td(P, "cvrp", inst, DEV)
    ok = torch.tensor([[1, 2, 0, 3, 4, 0, 5, 6, 0]] * 2, device=DEV)
    env.get_reward(td, ok)
    with pytest.raises(AssertionError, match="Used more than capacity"):
        env.get_reward(td, torch.tensor([[1, 2, 3, 0, 4, 0, 5, 6, 0]] * 2, device=DEV))
    with pytest.raises(AssertionError, match="Invalid tour"):
        env.get_reward(td, torch.tensor([[1, 2, 0, 3, 3, 0, 5, 6, 0]] * 2, device=DEV))


def test_empty_batch_and_size_limits():
    env = P.TSPEnv(generator_params=dict(num_loc=10), device=DEV)
    td = P.TensorDict({"locs": torch.zeros(0, 10, 2, device=DEV)}, batch_size=[0])
    td = env.reset(td)
    td.set("action", torch.zeros(0, dtype=torch.int64, device=DEV))
    td = env.step(td)["next"]
    assert td["action_mask"].shape == (0, 10)
    # env kernels have no node limit (only the decoder kernels do)
    inst = O.generate_tsp(2, 300)
    env, td = inst_to_env_td(P, "tsp", inst, DEV)
    td.set("action", torch.tensor([299, 5], device=DEV))
    td = env.step(td)["next"]
    assert int(td["action_mask"].sum()) == 2 * 299
