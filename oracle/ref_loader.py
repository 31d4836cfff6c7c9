"""TEST INFRASTRUCTURE ONLY — imports the UNMODIFIED reference from /root/reference.

Used in this container (never on the GPU box, where /root/reference does not exist)
to (a) validate `oracle/rollout_oracle.py` against the reference's own code and
(b) generate the golden fixtures under tests/golden/ (see oracle/make_golden.py).

Recipe follows SURVEY.md Appendix A: put the refshim stubs (tensordict / torchrl /
matplotlib — container bookkeeping only, no arithmetic) and /root/reference on
sys.path, and pre-register empty namespace modules for the packages whose
__init__.py import Lightning, so that the real sub-modules import normally.

Every arithmetic line executed through this loader is the reference's own file.
"""
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("RL4CO_REFERENCE_ROOT", "/root/reference")
_SHIM = os.path.join(os.path.dirname(os.path.abspath(__file__)), "refshim")

_loaded = None


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "rl4co"))


def load_reference():
    """Returns a namespace with the reference's TSPEnv, CVRPEnv, AttentionModelPolicy,
    decoding module and ops module."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not reference_available():
        raise RuntimeError(f"reference not found at {REFERENCE_ROOT}")
    for p in (REFERENCE_ROOT, _SHIM):
        if p not in sys.path:
            sys.path.insert(0, p)
    for name in (
        "rl4co.models",
        "rl4co.models.common",
        "rl4co.models.zoo",
        "rl4co.models.zoo.am",
        "rl4co.models.zoo.pomo",
        "rl4co.models.rl",
    ):
        if name not in sys.modules:
            m = types.ModuleType(name)
            m.__path__ = [os.path.join(REFERENCE_ROOT, *name.split("."))]
            sys.modules[name] = m
    import logging

    logging.getLogger("rl4co").setLevel(logging.ERROR)

    from rl4co.envs.routing.tsp.env import TSPEnv
    from rl4co.envs.routing.cvrp.env import CVRPEnv
    from rl4co.models.zoo.am.policy import AttentionModelPolicy
    from rl4co.models.rl.reinforce.baselines import SharedBaseline
    import rl4co.utils.decoding as decoding
    import rl4co.utils.ops as ops
    import rl4co.models.nn.attention as attention
    from tensordict import TensorDict

    ns = types.SimpleNamespace(
        TSPEnv=TSPEnv,
        CVRPEnv=CVRPEnv,
        AttentionModelPolicy=AttentionModelPolicy,
        SharedBaseline=SharedBaseline,
        decoding=decoding,
        ops=ops,
        attention=attention,
        TensorDict=TensorDict,
    )
    _loaded = ns
    return ns
