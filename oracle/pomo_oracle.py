"""TEST INFRASTRUCTURE ONLY - CPU restatement of the POMO evaluation reduction that follows the rollout
(SURVEY.md §8f rank 2).  Plain torch, each function cites the reference lines it follows; pinned against the
reference's own functions in tests/test_pomo_oracle_vs_reference.py.  Never imported by the product path."""
import torch


def dihedral_8_augmentation(xy: torch.Tensor) -> torch.Tensor:
    """rl4co/data/transforms.py:16-39: the 8 symmetries of the unit square, concatenated along the batch axis
    (augmentation-major: rows [a*B, (a+1)*B) hold transform a of all B instances)."""
    x, y = xy[..., 0:1], xy[..., 1:2]
    zs = [(x, y), (1 - x, y), (x, 1 - y), (1 - x, 1 - y), (y, x), (1 - y, x), (y, 1 - x), (1 - y, 1 - x)]
    return torch.cat([torch.cat(z, dim=-1) for z in zs], dim=0)


def augment_state(td: dict, num_augment: int = 8, feats=("locs",)) -> dict:
    """StateAugmentation.__call__ with augment_fn='dihedral8', first_aug_identity=True, normalize=False
    (transforms.py:107-153, :42-48): batchify every entry x8, replace `feats` by the 8 transforms of the original."""
    assert num_augment == 8
    out = {k: v.repeat(num_augment, *([1] * (v.dim() - 1))) for k, v in td.items()}   # ops.py:13-32 batchify
    for f in feats:
        out[f] = dihedral_8_augmentation(td[f])
    return out


def unbatchify(x: torch.Tensor, shape) -> torch.Tensor:
    """ops.py:35-56 for int or tuple shapes (zero entries are skipped, as in the reference)."""
    shape = [shape] if isinstance(shape, int) else list(shape)
    for s in reversed(shape):
        if s > 0:
            x = x.view(s, x.shape[0] // s, *x.shape[1:]).permute(1, 0, *range(2, x.dim() + 1))
    return x


def gather_by_index(src, idx, dim=1):
    """ops.py:59-71."""
    shape = list(src.shape)
    shape[dim] = -1
    idx = idx.view(idx.shape + (1,) * (src.dim() - idx.dim())).expand(shape)
    g = src.gather(dim, idx)
    return g.squeeze(dim) if idx.size(dim) == 1 else g


def pomo_eval_select(reward: torch.Tensor, actions: torch.Tensor, n_aug: int, n_start: int) -> dict:
    """pomo/model.py:103-145, validation / test branch: reward [S*A*B] and actions [S*A*B, T] of a multistart rollout
    over the augmented batch -> best-of-starts per (instance, augmentation), then best-of-augmentations."""
    out = {}
    r = unbatchify(reward, (n_aug, n_start))
    if n_start > 1:
        max_reward, max_idxs = r.max(dim=-1)
        out["max_reward"] = max_reward
        acts = unbatchify(actions, (n_aug, n_start))
        out["best_multistart_actions"] = gather_by_index(acts, max_idxs, dim=max_idxs.dim())
    if n_aug > 1:
        r_ = out["max_reward"] if n_start > 1 else r
        max_aug_reward, idx = r_.max(dim=1)
        out["max_aug_reward"] = max_aug_reward
        acts_ = out["best_multistart_actions"] if n_start > 1 else unbatchify(actions, (n_aug, n_start))
        out["best_aug_actions"] = gather_by_index(acts_, idx)
    return out
