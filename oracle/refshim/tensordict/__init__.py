"""TEST INFRASTRUCTURE ONLY — minimal stand-in for the `tensordict` package.

`tensordict` is not installed in this image and cannot be installed (no network).
The reference (/root/reference, RL4CO v0.6 fork) only uses it on the rollout hot
path as a dict-of-tensors with a leading batch shape.  This stub implements exactly
the container operations the hot path exercises so that the
reference's own arithmetic files run UNMODIFIED when generating golden vectors.
It contributes no arithmetic.
"""
from .tensordict import TensorDict  # noqa: F401

__version__ = "0.7.0"
