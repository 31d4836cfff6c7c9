"""GPU: dihedral-8 augmentation and the best-of-starts / best-of-augmentations kernel (eam_rl4co_b200/pomo.py,
csrc/reduce_kernels.cu) against oracle/pomo_oracle.py (pinned to the reference in
tests/test_pomo_oracle_vs_reference.py).  Integer / index results are compared bit-exactly."""
import pytest
import torch

import eam_rl4co_b200 as P
from oracle import pomo_oracle as PO
from oracle import rollout_oracle as O
from helpers import inst_to_env_td

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.mark.parametrize("B,A,S,T", [(5, 8, 20, 20), (3, 0, 7, 9), (130, 8, 100, 100), (2, 8, 50, 61), (1, 8, 2, 3)])
def test_best_of_starts_kernel_matches_oracle(B, A, S, T):
    torch.manual_seed(B * 7 + S)
    rows = B * max(A, 1) * S
    reward = torch.randn(rows).round(decimals=1)              # many exact ties: first-maximum rule
    actions = torch.randint(0, 101, (rows, T))
    exp = PO.pomo_eval_select(reward, actions, A, S)
    got = P.best_of_starts({"reward": reward.to(DEV), "actions": actions.to(DEV)}, A, S)
    for k, v in exp.items():
        assert torch.equal(got[k].cpu(), v), k
    assert torch.equal(got["actions"].cpu(), actions)         # row order untouched


def test_state_augmentation_matches_oracle():
    torch.manual_seed(3)
    inst = O.generate_cvrp(6, 20)
    env, td = inst_to_env_td(P, "cvrp", inst, DEV)
    aug = P.StateAugmentation(8)(td)
    exp = PO.augment_state({"locs": td["locs"].cpu(), "demand": td["demand"].cpu()})
    assert torch.equal(aug["locs"].cpu(), exp["locs"]) and torch.equal(aug["demand"].cpu(), exp["demand"])
    assert aug.batch_size[0] == 48


@pytest.mark.parametrize("env_name,N,B", [("tsp", 20, 4), ("cvrp", 20, 3), ("tsp", 100, 2)])
def test_pomo_evaluate_end_to_end(env_name, N, B):
    torch.manual_seed(5)
    inst = O.generate_tsp(B, N) if env_name == "tsp" else O.generate_cvrp(B, N)
    torch.manual_seed(6)
    pol = P.POMOPolicy(env_name).eval().to(DEV)
    env, td = inst_to_env_td(P, env_name, inst, DEV)
    out = P.pomo_evaluate(pol, env, td, num_augment=8)
    S = out["reward"].shape[0] // (8 * B)
    exp = PO.pomo_eval_select(out["reward"].cpu(), out["actions"].cpu(), 8, S)
    for k, v in exp.items():
        assert torch.equal(out[k].cpu(), v), k
    # the selected tour reproduces the selected reward on the ORIGINAL instance (augmentations are isometries)
    best = out["best_aug_actions"]
    r = env.get_reward(td, best)
    assert torch.allclose(r, out["max_aug_reward"], rtol=1e-5, atol=1e-5)
    assert bool((out["max_aug_reward"] >= out["max_reward"][:, 0] - 1e-6).all())
