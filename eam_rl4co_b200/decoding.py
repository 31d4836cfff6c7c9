"""Decode strategies, step-granular (mirror of rl4co/utils/decoding.py:193-465).

`DecodingStrategy.step(logits, mask, td, action=None)` keeps the reference's signature;
tanh-clip -> mask -> temperature -> log_softmax -> greedy / noise-driven sampling / evaluate
-> gather runs in ONE kernel (`rl4co_select`).  Sampling is argmax(logp + Gumbel) with either a
supplied noise tensor or the counter generator keyed (seed, instance, start, step, node)
(SURVEY.md §7.3-4); torch.multinomial's RNG stream is implementation-defined and is not
reproduced.  top-k / top-p and beam search are outside the accelerated path and raise.
"""
import ctypes as C
from typing import Optional

import torch

from . import _lib
from .tensordict_lite import batchify


def get_log_likelihood(logprobs, actions=None, mask=None, return_sum: bool = True):
    """decoding.py:38-64."""
    if actions is not None and logprobs.dim() == 3:
        logprobs = logprobs.gather(-1, actions.unsqueeze(-1)).squeeze(-1)
    if mask is not None:
        logprobs[~mask] = 0
    assert (logprobs > -1000).all(), "Logprobs should not be -inf, check sampling procedure!"
    return logprobs.sum(1) if return_sum else logprobs


def calculate_entropy(logprobs):
    """ops.py:108-116."""
    logprobs = torch.nan_to_num(logprobs, nan=0.0)
    entropy = -(logprobs.exp() * logprobs).sum(dim=-1).sum(dim=1)
    assert entropy.isfinite().all(), "Entropy is not finite"
    return entropy


def fresh_noise_seed() -> int:
    """Seed of the counter noise of ONE sampling rollout when the caller gives none: drawn from torch's global
    CPU generator, so every call samples fresh noise (as `torch.multinomial` does in the reference,
    decoding.py:403-417) while `torch.manual_seed` still makes a run reproducible.  Ranks that seed alike draw
    alike; the noise itself is keyed by the global instance index, so sharded and unsharded runs agree."""
    return int(torch.randint(0, 2 ** 31 - 1, (1,)).item())


class DecodingStrategy:
    name = "base"
    mode = None

    def __init__(self, temperature: float = 1.0, top_p: float = 0.0, top_k: int = 0,
                 mask_logits: bool = True, tanh_clipping: float = 0, num_starts: Optional[int] = None,
                 multistart: bool = False, select_start_nodes_fn=None, select_best: bool = False,
                 store_all_logp: bool = False, noise_seed: Optional[int] = None, instance_offset: int = 0,
                 noise: Optional[torch.Tensor] = None, num_samples=None, multisample: bool = False,
                 **unused):
        if top_p or top_k:
            raise NotImplementedError("top_k / top_p filtering is outside the accelerated path")
        if not mask_logits:
            raise NotImplementedError("mask_logits=False is outside the accelerated path")
        if multisample or (num_samples and num_samples > 1):
            raise NotImplementedError("multisample decoding is outside the accelerated path")
        self.temperature, self.tanh_clipping = temperature, tanh_clipping
        if num_starts is not None:
            multistart = num_starts > 1
        self.multistart = multistart
        self.num_starts = num_starts if multistart else None
        self.select_start_nodes_fn = select_start_nodes_fn
        self.select_best, self.store_all_logp = select_best, store_all_logp
        if noise_seed is None:       # only sampling consumes the generator
            noise_seed = fresh_noise_seed() if (self.mode == _lib.MODE_SAMPLING and noise is None) else 0
        self.noise_seed = int(noise_seed) & 0xFFFFFFFF
        self.instance_offset, self.noise = instance_offset, noise
        self.actions, self.logprobs = [], []
        self._B = None

    # decoding.py:284-332
    def pre_decoder_hook(self, td, env, action=None):
        self._B = td.batch_size[0]
        if self.multistart:
            if self.num_starts is None:
                self.num_starts = env.get_num_starts(td)
        else:
            self.num_starts = 0
        if self.num_starts >= 1 and self.multistart:
            if action is None:
                if self.select_start_nodes_fn is not None:
                    action = self.select_start_nodes_fn(td, env, self.num_starts)
                else:
                    action = env.select_start_nodes(td, num_starts=self.num_starts)
            td = batchify(td, self.num_starts)
            td.set("action", action)
            td = env.step(td)["next"]
            if self.store_all_logp:
                logprobs = torch.zeros_like(td["action_mask"])
            else:
                logprobs = torch.zeros_like(action, device=td.device)   # int64 zeros, as :324
            self.logprobs.append(logprobs)
            self.actions.append(action)
        return td, env, self.num_starts

    # decoding.py:334-344
    def post_decoder_hook(self, td, env):
        assert len(self.logprobs) > 0, (
            "No logprobs were collected because all environments were done. Check your initial state")
        logprobs = torch.stack(self.logprobs, 1)
        actions = torch.stack(self.actions, 1)
        if self.num_starts > 0 and self.select_best:
            logprobs, actions, td, env = self._select_best(logprobs, actions, td, env)
        return logprobs, actions, td, env

    # decoding.py:346-389
    def step(self, logits, mask, td=None, action=None, **kw):
        rows, n = logits.shape
        dev = logits.device
        col = len(self.actions)
        logp_all = torch.empty_like(logits) if self.store_all_logp else None
        act = torch.empty(rows, dtype=torch.int64, device=dev)
        lp = torch.empty(rows, dtype=torch.float32, device=dev)
        err = _lib.new_err(dev)
        g = self.noise[:, col].contiguous() if (self.noise is not None and self.mode == _lib.MODE_SAMPLING) else None
        rc = _lib.lib().rl4co_select(
            _lib.ptr(logits.contiguous()), _lib.ptr(mask.contiguous()), C.c_float(self.tanh_clipping),
            C.c_float(self.temperature), C.c_int32(self.mode), _lib.ptr(g), C.c_uint32(self.noise_seed),
            C.c_int64(self.instance_offset), C.c_int64(self._B if self._B else rows), C.c_int32(col),
            _lib.ptr(action.contiguous() if action is not None else None), _lib.ptr(logp_all), _lib.ptr(act),
            _lib.ptr(lp), _lib.ptr(err), C.c_int64(rows), C.c_int32(n), _lib.stream_ptr())
        _lib.check(rc, "rl4co_select")
        _lib.raise_on_err(err)          # decoding.py:397-399 (the reference syncs here too)
        assert td is not None, "td must be provided"
        td.set("action", act)
        self.actions.append(act)
        self.logprobs.append(logp_all if self.store_all_logp else lp)
        return td

    # decoding.py:419-427
    def _select_best(self, logprobs, actions, td, env):
        from .tensordict_lite import unbatchify
        rewards = env.get_reward(td, actions)
        _, max_idxs = unbatchify(rewards, self.num_starts).max(dim=-1)

        def pick(x):
            x = unbatchify(x, self.num_starts)
            idx = max_idxs.view(max_idxs.shape + (1,) * (x.dim() - 1)).expand(-1, 1, *x.shape[2:])
            return x.gather(1, idx).squeeze(1)

        td_best = type(td)({k: pick(v) for k, v in td.items()}, batch_size=[max_idxs.shape[0]])
        return pick(logprobs), pick(actions), td_best, env


class Greedy(DecodingStrategy):
    name, mode = "greedy", _lib.MODE_GREEDY


class Sampling(DecodingStrategy):
    name, mode = "sampling", _lib.MODE_SAMPLING


class Evaluate(DecodingStrategy):
    name, mode = "evaluate", _lib.MODE_EVALUATE


def get_decoding_strategy(decoding_strategy, **config):
    """decoding.py:17-35."""
    registry = {"greedy": Greedy, "sampling": Sampling, "multistart_greedy": Greedy,
                "multistart_sampling": Sampling, "evaluate": Evaluate}
    if decoding_strategy == "beam_search":
        raise NotImplementedError("beam_search is outside the accelerated path")
    if decoding_strategy not in registry:
        raise ValueError(f"Unknown decode type '{decoding_strategy}'. Available: {list(registry)}")
    if "multistart" in decoding_strategy:
        config["multistart"] = True
    return registry[decoding_strategy](**config)
