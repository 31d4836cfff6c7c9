"""TEST INFRASTRUCTURE ONLY — drives the UNMODIFIED reference (via ref_loader) for
oracle validation and golden-fixture generation.  Container-only (/root/reference)."""
import contextlib

import torch

from .ref_loader import load_reference
from . import rollout_oracle as O


def make_env(env_name: str, num_loc: int, check_solution: bool = True):
    R = load_reference()
    Env = R.TSPEnv if env_name == "tsp" else R.CVRPEnv
    return Env(generator_params=dict(num_loc=num_loc), check_solution=check_solution)


def make_policy(env_name: str, pomo: bool = False, seed: int = 4321):
    """c1–c3: AttentionModelPolicy defaults (am/policy.py:50-122);
    c4–c5: POMO defaults (pomo/model.py:58-62)."""
    R = load_reference()
    torch.manual_seed(seed)
    kw = dict(num_encoder_layers=6, normalization="instance", use_graph_context=False) if pomo else {}
    return R.AttentionModelPolicy(env_name=env_name, **kw).eval()


def decoder_weights(policy) -> dict:
    return {k[len("decoder."):]: v.detach().clone() for k, v in policy.state_dict().items()
            if k.startswith("decoder.")}


def inst_to_td(env_name: str, inst: dict):
    R = load_reference()
    B = inst["locs"].size(0)
    return R.TensorDict({k: v.clone() for k, v in inst.items()}, batch_size=[B])


@contextlib.contextmanager
def noise_sampling(seed: int, B: int, multistart: bool, instance_offset: int = 0,
                   noise: torch.Tensor = None):
    """Replace DecodingStrategy.sampling (decoding.py:403-417) by argmax(logp + g) with the
    counter-based Gumbel noise of rollout_oracle.noise_gumbel.  Everything else untouched."""
    R = load_reference()
    DS = R.decoding.DecodingStrategy
    orig = DS.__dict__["sampling"]
    state = {"col": 1 if multistart else 0}

    def sampling(logprobs, mask=None):
        rows, n = logprobs.shape
        r = torch.arange(rows, device=logprobs.device)
        if noise is not None:
            g = noise[:, state["col"]]
        else:
            g = O.noise_gumbel(seed, r % B + instance_offset, r // B, state["col"], n)
        state["col"] += 1
        sel = (logprobs + g).argmax(-1)
        if mask is not None:
            assert not (~mask).gather(1, sel.unsqueeze(-1)).any(), "infeasible action selected"
        return sel

    DS.sampling = staticmethod(sampling)
    try:
        yield
    finally:
        DS.sampling = orig


@contextlib.contextmanager
def trace_decoder(policy, trace: dict):
    """Record (logits, mask) the reference decoder emits each step."""
    orig = policy.decoder.forward

    def fwd(td, cached, num_starts=0):
        logits, mask = orig(td, cached, num_starts)
        trace.setdefault("logits", []).append(logits.detach().clone())
        trace.setdefault("mask", []).append(mask.detach().clone())
        return logits, mask

    policy.decoder.forward = fwd
    try:
        yield
    finally:
        policy.decoder.forward = orig


def run_reference(env_name, inst, policy, decode_type, num_starts=None, actions=None,
                  noise_seed=0, instance_offset=0, noise=None, trace=None, check_solution=True,
                  return_entropy=False):
    """policy(td, env, decode_type=...) of the reference, returns its out-dict + h."""
    env = make_env(env_name, inst["locs"].size(1), check_solution)
    td = env.reset(inst_to_td(env_name, inst))
    B = inst["locs"].size(0)
    kw = {}
    if num_starts is not None:
        kw["num_starts"] = num_starts
    ctxs = [noise_sampling(noise_seed, B, "multistart" in decode_type, instance_offset, noise)]
    if trace is not None:
        ctxs.append(trace_decoder(policy, trace))
    with torch.inference_mode(), contextlib.ExitStack() as st:
        for c in ctxs:
            st.enter_context(c)
        if actions is not None:
            out = policy(td.clone(), env, actions=actions, decode_type=decode_type,
                         return_hidden=True, return_entropy=return_entropy, **kw)
        else:
            out = policy(td.clone(), env, decode_type=decode_type, return_hidden=True,
                         return_sum_log_likelihood=True, return_entropy=return_entropy, **kw)
    return out
