"""TEST INFRASTRUCTURE ONLY - deterministic synthetic policy weights.

The encoder fixtures (tests/golden/enc_*.npz) hold the reference encoder's output `h` for weights that are
NOT stored: both sides regenerate them from a seed with numpy's legacy MT19937 `RandomState` (bit-stable across
machines and numpy versions), walking the state_dict in key order.  The generating script loads them into the
UNMODIFIED reference policy (`load_state_dict(strict=True)`), the tests into the mirror policy the same way, so
a key or shape mismatch between the two is itself a test failure.  Values follow the reference's own init
scales (nn.Linear: U(-1/sqrt(fan_in), 1/sqrt(fan_in))) with non-trivial norm affines and - for BatchNorm -
non-trivial running statistics, so eval-mode batch normalisation (nn/ops.py:46-47) is actually exercised.
"""
import math

import numpy as np
import torch


def synth_state_dict(template: dict, seed: int) -> dict:
    """template: ordered {name: tensor} (a state_dict); returns {name: tensor} of the same shapes / dtypes."""
    rs = np.random.RandomState(seed)
    out = {}
    for name, t in template.items():
        shape = tuple(t.shape)
        if name.endswith("num_batches_tracked"):
            out[name] = torch.tensor(7, dtype=t.dtype)
            continue
        if name.endswith("running_mean"):
            v = rs.normal(0.0, 0.2, size=shape)
        elif name.endswith("running_var"):
            v = rs.uniform(0.5, 1.5, size=shape)
        elif "normalizer.weight" in name:
            v = rs.uniform(0.6, 1.4, size=shape)
        elif "normalizer.bias" in name:
            v = rs.uniform(-0.3, 0.3, size=shape)
        elif name.endswith("W_placeholder"):
            v = rs.uniform(-1.0, 1.0, size=shape)
        else:
            fan_in = shape[-1] if len(shape) > 1 else shape[0]
            bound = 1.0 / math.sqrt(max(fan_in, 1))
            v = rs.uniform(-bound, bound, size=shape)
        out[name] = torch.from_numpy(np.asarray(v, dtype=np.float32)).to(t.dtype)
    return out


def signature(template: dict) -> str:
    """names and shapes in order - stored in the fixture, compared by the tests."""
    return ";".join(f"{k}:{'x'.join(map(str, v.shape))}" for k, v in template.items())
