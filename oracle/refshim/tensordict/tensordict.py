"""TEST INFRASTRUCTURE ONLY — dict-of-tensors with leading batch dims (see __init__)."""
import torch


class TensorDict:
    def __init__(self, source=None, batch_size=None, device=None, **kw):
        self._d = {}
        if batch_size is None:
            batch_size = []
        if isinstance(batch_size, int):
            batch_size = [batch_size]
        self._bs = torch.Size(batch_size)
        self._device = torch.device(device) if device is not None else None
        for k, v in (source or {}).items():
            self.set(k, v)

    # ---- mapping protocol -------------------------------------------------
    def set(self, key, value, inplace=False):
        if isinstance(value, dict):
            value = TensorDict(value, batch_size=self._bs)
        elif not isinstance(value, (torch.Tensor, TensorDict)):
            value = torch.as_tensor(value)
        if self._device is not None and isinstance(value, torch.Tensor):
            value = value.to(self._device)
        self._d[key] = value
        return self

    def get(self, key, default=None):
        return self._d.get(key, default)

    def __getitem__(self, idx):
        if isinstance(idx, str):
            return self._d[idx]
        nb = len(self._bs)
        probe = torch.empty(self._bs, device="meta")[idx]
        out = TensorDict({}, batch_size=probe.shape, device=self._device)
        for k, v in self._d.items():
            out._d[k] = v[idx]
        return out

    def __setitem__(self, key, value):
        if isinstance(key, str):
            self.set(key, value)
        else:
            for k, v in value.items():
                self._d[k][key] = v

    def __contains__(self, key):
        return key in self._d

    def __len__(self):
        return self._bs[0] if len(self._bs) else 0

    def __iter__(self):
        raise TypeError("stub TensorDict is not iterable")

    def keys(self, *a, **k):
        return self._d.keys()

    def values(self):
        return self._d.values()

    def items(self):
        return self._d.items()

    def update(self, other, **kw):
        for k, v in (other.items() if hasattr(other, "items") else other):
            self.set(k, v)
        return self

    def exclude(self, *keys):
        out = TensorDict({}, batch_size=self._bs, device=self._device)
        for k, v in self._d.items():
            if k not in keys:
                out._d[k] = v
        return out

    def select(self, *keys):
        out = TensorDict({}, batch_size=self._bs, device=self._device)
        for k in keys:
            out._d[k] = self._d[k]
        return out

    def is_empty(self):
        return len(self._d) == 0

    # ---- shape / device ---------------------------------------------------
    @property
    def batch_size(self):
        return self._bs

    @batch_size.setter
    def batch_size(self, v):
        self._bs = torch.Size(v)

    @property
    def shape(self):
        return self._bs

    def size(self, dim=None):
        return self._bs if dim is None else self._bs[dim]

    def dim(self):
        return len(self._bs)

    ndim = property(dim)

    @property
    def device(self):
        if self._device is not None:
            return self._device
        for v in self._d.values():
            return v.device
        return torch.device("cpu")

    def _map(self, fn, batch_size):
        out = TensorDict({}, batch_size=batch_size, device=None)
        for k, v in self._d.items():
            out._d[k] = fn(v)
        return out

    def clone(self, recurse=True):
        return self._map(lambda v: v.clone() if recurse else v, self._bs)

    def to(self, device, **kw):
        out = self._map(lambda v: v.to(device), self._bs)
        out._device = torch.device(device)
        return out

    def cpu(self):
        return self.to("cpu")

    def cuda(self):
        return self.to("cuda")

    def contiguous(self):
        return self._map(lambda v: v.contiguous(), self._bs)

    def detach(self):
        return self._map(lambda v: v.detach(), self._bs)

    # The next three act on the leading batch dims of every leaf and keep the
    # trailing dims: this is what makes rl4co.utils.ops.batchify/unbatchify work.
    def expand(self, *shape):
        if len(shape) == 1 and not isinstance(shape[0], int):
            shape = tuple(shape[0])
        nb = len(self._bs)
        return self._map(
            lambda v: v.expand(*shape, *v.shape[nb:]), torch.Size(shape)
        )

    def view(self, *shape):
        if len(shape) == 1 and not isinstance(shape[0], int):
            shape = tuple(shape[0])
        nb = len(self._bs)
        new_bs = torch.empty(self._bs, device="meta").view(*shape).shape
        return self._map(lambda v: v.view(*new_bs, *v.shape[nb:]), new_bs)

    def reshape(self, *shape):
        if len(shape) == 1 and not isinstance(shape[0], int):
            shape = tuple(shape[0])
        nb = len(self._bs)
        new_bs = torch.empty(self._bs, device="meta").reshape(*shape).shape
        return self._map(lambda v: v.reshape(*new_bs, *v.shape[nb:]), new_bs)

    def permute(self, *dims):
        if len(dims) == 1 and not isinstance(dims[0], int):
            dims = tuple(dims[0])
        nb = len(self._bs)
        new_bs = torch.Size(self._bs[d] for d in dims)
        return self._map(
            lambda v: v.permute(*dims, *range(nb, v.dim())), new_bs
        )

    def gather(self, dim, index):
        def g(v):
            idx = index.view(index.shape + (1,) * (v.dim() - index.dim()))
            idx = idx.expand(*index.shape, *v.shape[index.dim():])
            return v.gather(dim, idx)
        return self._map(g, index.shape)

    def squeeze(self, dim):
        nb = list(self._bs)
        if nb[dim] == 1:
            nb.pop(dim)
            return self._map(lambda v: v.squeeze(dim), torch.Size(nb))
        return self

    def __repr__(self):
        f = ", ".join(f"{k}: {tuple(v.shape)}" for k, v in self._d.items())
        return f"TensorDict(stub)({f}; batch_size={tuple(self._bs)})"
