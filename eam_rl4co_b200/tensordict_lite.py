"""TensorDict container for the rollout API.

The reference keeps env state in `tensordict.TensorDict` (rl4co/envs/common/base.py:19).
If the real package is importable it is used unchanged; otherwise this minimal
dict-of-tensors with a leading batch shape provides the operations the rollout API
touches (get/set/update/clone/to/index/keys), so code written against the reference's
`td["action_mask"]`, `td.set("action", a)`, `env.step(td)["next"]` idiom runs as is.
"""
import torch

try:  # pragma: no cover - not installed in this image
    from tensordict import TensorDict as _RealTensorDict  # type: ignore

    if getattr(_RealTensorDict, "__module__", "").startswith("oracle"):
        raise ImportError
    TensorDict = _RealTensorDict
    HAVE_TENSORDICT = True
except Exception:  # noqa: BLE001
    HAVE_TENSORDICT = False

    class TensorDict:  # type: ignore[no-redef]
        def __init__(self, source=None, batch_size=None, device=None):
            self._d = dict(source or {})
            if batch_size is None:
                batch_size = []
            if isinstance(batch_size, int):
                batch_size = [batch_size]
            self.batch_size = torch.Size(batch_size)
            if device is not None:
                self._d = {k: v.to(device) for k, v in self._d.items()}

        def __getitem__(self, k):
            if isinstance(k, str):
                return self._d[k]
            probe = torch.empty(self.batch_size, device="meta")[k]
            return TensorDict({kk: v[k] for kk, v in self._d.items()}, batch_size=probe.shape)

        def __setitem__(self, k, v):
            self._d[k] = v

        def __contains__(self, k):
            return k in self._d

        def get(self, k, default=None):
            return self._d.get(k, default)

        def set(self, k, v, inplace=False):
            self._d[k] = v
            return self

        def update(self, other, **kw):
            for k, v in other.items():
                self._d[k] = v
            return self

        def keys(self, *a, **k):
            return self._d.keys()

        def items(self):
            return self._d.items()

        def values(self):
            return self._d.values()

        def is_empty(self):
            return len(self._d) == 0

        @property
        def shape(self):
            return self.batch_size

        def dim(self):
            return len(self.batch_size)

        @property
        def device(self):
            for v in self._d.values():
                return v.device
            return torch.device("cpu")

        def clone(self, recurse=True):
            return TensorDict({k: (v.clone() if recurse else v) for k, v in self._d.items()},
                              batch_size=self.batch_size)

        def to(self, device, **kw):
            return TensorDict({k: v.to(device, **kw) for k, v in self._d.items()},
                              batch_size=self.batch_size)

        def cuda(self):
            return self.to("cuda")

        def cpu(self):
            return self.to("cpu")

        def exclude(self, *keys):
            return TensorDict({k: v for k, v in self._d.items() if k not in keys},
                              batch_size=self.batch_size)

        def __repr__(self):
            f = ", ".join(f"{k}: {tuple(v.shape)}" for k, v in self._d.items())
            return f"TensorDict({f}; batch_size={tuple(self.batch_size)})"


def batchify(x, repeats: int):
    """rl4co/utils/ops.py:13-32 — start-major repeat (row = s*B + b); tensors or TensorDicts."""
    if repeats <= 0:
        return x
    if isinstance(x, torch.Tensor):
        s = x.shape
        return x.expand(repeats, *s).contiguous().view(s[0] * repeats, *s[1:])
    bs = x.batch_size
    return TensorDict({k: batchify(v, repeats) for k, v in x.items()},
                      batch_size=[bs[0] * repeats, *bs[1:]])


def unbatchify(x, repeats: int):
    """rl4co/utils/ops.py:35-56 — [S*B, ...] -> [B, S, ...]."""
    if repeats <= 0:
        return x
    if isinstance(x, torch.Tensor):
        s = x.shape
        return x.view(repeats, s[0] // repeats, *s[1:]).permute(1, 0, *range(2, len(s) + 1))
    bs = x.batch_size
    return TensorDict({k: unbatchify(v, repeats) for k, v in x.items()},
                      batch_size=[bs[0] // repeats, repeats, *bs[1:]])
