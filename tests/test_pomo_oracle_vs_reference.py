"""CPU: oracle/pomo_oracle.py pinned against the reference's own functions (run unmodified through
oracle/ref_loader.py): rl4co/data/transforms.py (dihedral-8, StateAugmentation) and the selection lines of
pomo/model.py:103-145 re-enacted with the reference's rl4co/utils/ops.py.  Skipped where the reference is absent."""
import pytest
import torch

from oracle import pomo_oracle as PO
from oracle.ref_loader import load_reference, reference_available

pytestmark = pytest.mark.skipif(not reference_available(), reason="reference checkout not present")


def test_dihedral8_and_state_augmentation_bit_identical():
    ref = load_reference()
    import rl4co.data.transforms as T
    torch.manual_seed(0)
    xy = torch.rand(5, 11, 2)
    assert torch.equal(PO.dihedral_8_augmentation(xy), T.dihedral_8_augmentation(xy))
    td = ref.TensorDict({"locs": xy, "demand": torch.rand(5, 10)}, batch_size=[5])
    aug = T.StateAugmentation(num_augment=8, augment_fn="dihedral8")(td)
    mine = PO.augment_state({"locs": xy, "demand": td["demand"]})
    assert torch.equal(mine["locs"], aug["locs"]) and torch.equal(mine["demand"], aug["demand"])


@pytest.mark.parametrize("n_aug,n_start", [(8, 7), (0, 7), (8, 1), (8, 0)])
def test_eval_selection_matches_reference_ops(n_aug, n_start):
    ref = load_reference()
    ops = ref.ops
    torch.manual_seed(1)
    B, T = 4, 9
    rows = B * max(n_aug, 1) * max(n_start, 1)
    reward = torch.randn(rows).round(decimals=1)            # ties on purpose
    actions = torch.randint(0, 50, (rows, T))
    got = PO.pomo_eval_select(reward, actions, n_aug, n_start)
    # pomo/model.py:103-145 with the reference's own helpers
    r = ops.unbatchify(reward, (n_aug, n_start))
    exp = {}
    if n_start > 1:
        max_reward, max_idxs = r.max(dim=-1)
        exp["max_reward"] = max_reward
        acts = ops.unbatchify(actions, (n_aug, n_start))
        exp["best_multistart_actions"] = ops.gather_by_index(acts, max_idxs, dim=max_idxs.dim())
    if n_aug > 1:
        r_ = exp["max_reward"] if n_start > 1 else r
        max_aug_reward, idx = r_.max(dim=1)
        exp["max_aug_reward"] = max_aug_reward
        acts_ = exp["best_multistart_actions"] if n_start > 1 else ops.unbatchify(actions, (n_aug, n_start))
        exp["best_aug_actions"] = ops.gather_by_index(acts_, idx)
    assert got.keys() == exp.keys()
    for k in exp:
        assert torch.equal(got[k], exp[k]), k
