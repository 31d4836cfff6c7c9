"""POMO evaluation helpers around the rollout: dihedral-8 state augmentation before the
policy call and the best-of-starts / best-of-augmentations reduction after it (rl4co/data/transforms.py:16-48,107-153,
rl4co/models/zoo/pomo/model.py:115-145).  The reduction is one kernel (`rl4co_best_of_starts`); names follow the
reference's out-dict keys."""
import ctypes as C

import torch

from . import _lib
from .tensordict_lite import batchify


def dihedral_8_augmentation(xy: torch.Tensor) -> torch.Tensor:
    """transforms.py:16-39: [B, N, 2] -> [8B, N, 2], augmentation-major."""
    x, y = xy[..., 0:1], xy[..., 1:2]
    zs = ((x, y), (1 - x, y), (x, 1 - y), (1 - x, 1 - y), (y, x), (1 - y, x), (y, 1 - x), (1 - y, 1 - x))
    return torch.cat([torch.cat(z, dim=-1) for z in zs], dim=0)


class StateAugmentation:
    """transforms.py:107-153 with augment_fn='dihedral8' (POMO's evaluation default, pomo/model.py:70-76);
    first_aug_identity=True, normalize=False."""

    def __init__(self, num_augment: int = 8, augment_fn: str = "dihedral8", feats=None):
        if augment_fn != "dihedral8":
            raise NotImplementedError("only the dihedral-8 augmentation of POMO is provided")
        assert num_augment == 8, "When using the `dihedral8` augmentation function, then num_augment must be 8"
        self.num_augment, self.feats = num_augment, (["locs"] if feats is None else list(feats))

    def __call__(self, td):
        td_aug = batchify(td, self.num_augment)
        for f in self.feats:
            td_aug.set(f, dihedral_8_augmentation(td[f]))
        return td_aug


def best_of_starts(out: dict, num_augment: int, num_starts: int) -> dict:
    """The validation / test branch of POMO.shared_step (pomo/model.py:115-145) on a policy out-dict whose rows are
    start-major over the augmented batch.  Adds `max_reward` [B(,A)], `best_multistart_actions`, and with
    augmentation `max_aug_reward` [B], `best_aug_actions` [B,T].  (The reference also replaces out["actions"] by its
    [B,A,S,T] regrouping; that copy is what the kernel avoids - out["actions"] stays in row order.)"""
    reward, actions = out["reward"], out.get("actions")
    A, S = max(int(num_augment), 1), int(num_starts)
    if S <= 1:
        raise NotImplementedError("best_of_starts is the multistart reduction (num_starts > 1)")
    rows = reward.shape[0]
    B = rows // (A * S)
    assert B * A * S == rows, "reward rows must be num_starts * num_augment * batch"
    dev = reward.device
    T = actions.shape[1] if actions is not None else 0
    max_reward = torch.empty(B, A, dtype=torch.float32, device=dev)
    best_start = torch.empty(B, A, dtype=torch.int64, device=dev)
    ms_actions = torch.empty(B, A, T, dtype=torch.int64, device=dev) if actions is not None else None
    aug = num_augment > 1
    max_aug = torch.empty(B, dtype=torch.float32, device=dev) if aug else None
    best_aug = torch.empty(B, dtype=torch.int64, device=dev) if aug else None
    aug_actions = torch.empty(B, T, dtype=torch.int64, device=dev) if (aug and actions is not None) else None
    L = _lib.lib()
    _lib.check(L.rl4co_best_of_starts(
        _lib.ptr(reward.contiguous()), _lib.ptr(actions.contiguous() if actions is not None else None), C.c_int64(B),
        A, S, T, _lib.ptr(max_reward), _lib.ptr(best_start), _lib.ptr(ms_actions), _lib.ptr(max_aug),
        _lib.ptr(best_aug), _lib.ptr(aug_actions), _lib.stream_ptr()), "rl4co_best_of_starts")
    squeeze = (lambda x: x) if num_augment > 0 else (lambda x: x.squeeze(1))    # n_aug = 0: no augmentation axis
    res = dict(out)
    res["max_reward"] = squeeze(max_reward)
    res["best_start"] = squeeze(best_start)
    if actions is not None:
        res["best_multistart_actions"] = squeeze(ms_actions)
    if aug:
        res["max_aug_reward"] = max_aug
        if actions is not None:
            res["best_aug_actions"] = aug_actions
    return res


@torch.no_grad()
def pomo_evaluate(policy, env, td, num_augment: int = 8, num_starts=None, decode_type: str = "multistart_greedy",
                  **kw) -> dict:
    """POMO.shared_step for phase val / test: augment, one fused multistart rollout, reduce."""
    B = td.batch_size[0]
    if num_augment > 1:
        td = StateAugmentation(num_augment)(td)
    skw = {} if num_starts is None else {"num_starts": num_starts}
    out = policy(td, env, decode_type=decode_type, **skw, **kw)
    S = out["reward"].shape[0] // (B * max(num_augment, 1))
    return best_of_starts(out, num_augment, S)
