"""AttentionModelPolicy: drop-in for `policy(td, env, decode_type=...)`.

Mirror of rl4co/models/zoo/am/policy.py:50-122 + ConstructivePolicy.forward
(rl4co/models/common/constructive/base.py:157-275): same signature, kwargs and out-dict.
By default the whole decode loop (pre-hook, decoder step, process_logits, selection, env
step, reward, log-likelihood) is ONE persistent kernel launch (`rl4co_rollout`);
`fused=False` runs the reference's step-by-step loop on the step-granular kernels
(`decoder(td, cache, S)` -> `strategy.step` -> `env.step`), which is the same API surface the
reference exposes per step.
"""
import ctypes as C
import warnings
from typing import Optional

import torch
import torch.nn as nn

from . import _lib
from .decoder import AttentionModelDecoder
from .decoding import calculate_entropy, fresh_noise_seed, get_decoding_strategy, get_log_likelihood
from .encoder import AttentionModelEncoder
from .envs import RL4COEnvBase, get_env
from .tensordict_lite import unbatchify

_MODES = {"greedy": _lib.MODE_GREEDY, "sampling": _lib.MODE_SAMPLING, "evaluate": _lib.MODE_EVALUATE}


class AttentionModelPolicy(nn.Module):
    _warned = False

    def __init__(self, encoder: nn.Module = None, decoder: nn.Module = None, embed_dim: int = 128,
                 num_encoder_layers: int = 3, num_heads: int = 8, normalization: str = "batch",
                 feedforward_hidden: int = 512, env_name: str = "tsp", use_graph_context: bool = True,
                 mask_inner: bool = True, check_nan: bool = True, temperature: float = 1.0,
                 tanh_clipping: float = 10.0, mask_logits: bool = True,
                 train_decode_type: str = "sampling", val_decode_type: str = "greedy",
                 test_decode_type: str = "greedy", **unused):
        super().__init__()
        self.env_name = env_name
        self.encoder = encoder if encoder is not None else AttentionModelEncoder(
            embed_dim, env_name, num_heads, num_encoder_layers, normalization, feedforward_hidden)
        self.decoder = decoder if decoder is not None else AttentionModelDecoder(
            embed_dim, num_heads, env_name, mask_inner=mask_inner, use_graph_context=use_graph_context,
            check_nan=check_nan)
        self.temperature, self.tanh_clipping, self.mask_logits = temperature, tanh_clipping, mask_logits
        self.train_decode_type = train_decode_type
        self.val_decode_type = val_decode_type
        self.test_decode_type = test_decode_type

    # ------------------------------------------------------------------------------------
    def forward(self, td, env=None, phase: str = "train", calc_reward: bool = True,
                return_actions: bool = True, return_entropy: bool = False, return_hidden: bool = False,
                return_init_embeds: bool = False, return_sum_log_likelihood: bool = True, actions=None,
                max_steps=1_000_000, fused: bool = True, check: bool = True, hidden=None,
                **decoding_kwargs) -> dict:
        """`hidden=` (extension): a precomputed encoder output `h [B, N', E]`; the encoder is skipped (the training
        step encodes once, rolls out on `h.detach()` and differentiates the re-evaluation through the same `h`)."""
        use_fused = fused and decoding_kwargs.get("select_start_nodes_fn") is None
        if use_fused:
            # The fused rollout is a no-grad kernel: its log_likelihood is detached.  Differentiable log-likelihoods
            # of the sampled actions come from train.evaluate_log_likelihood / train.pomo_train_step.
            if torch.is_grad_enabled() and phase == "train" and actions is None and not AttentionModelPolicy._warned:
                AttentionModelPolicy._warned = True
                warnings.warn("policy(td, env, phase='train') on the fused kernel returns a DETACHED log_likelihood; "
                              "use eam_rl4co_b200.train.pomo_train_step / evaluate_log_likelihood for gradients")
            if hidden is not None:
                hidden, init_embeds = hidden.detach(), None
            else:
                with torch.no_grad(), _lib.nvtx_range("rl4co_b200.encoder"):
                    hidden, init_embeds = self.encoder(td)                       # base.py:195
        elif hidden is not None:
            init_embeds = None
        else:
            hidden, init_embeds = self.encoder(td)                               # base.py:195
        if isinstance(env, str) or env is None:
            env = get_env(self.env_name if env is None else env)
        decode_type = decoding_kwargs.pop("decode_type", None)                   # base.py:204-208
        if actions is not None:
            decode_type = "evaluate"
        elif decode_type is None:
            decode_type = getattr(self, f"{phase}_decode_type")
        cfg = dict(
            temperature=decoding_kwargs.pop("temperature", self.temperature),
            tanh_clipping=decoding_kwargs.pop("tanh_clipping", self.tanh_clipping),
            mask_logits=decoding_kwargs.pop("mask_logits", self.mask_logits),
            store_all_logp=decoding_kwargs.pop("store_all_logp", return_entropy),
            **decoding_kwargs)
        if use_fused:
            out = self._forward_fused(td, env, hidden, decode_type, actions, max_steps, check, cfg)
        else:
            out = self._forward_steps(td, env, hidden, decode_type, actions, max_steps, cfg)
        logprobs, acts, td, reward, cache = out
        if not calc_reward:
            reward = td.get("reward", None)
        outdict = {"reward": reward,
                   "log_likelihood": get_log_likelihood(logprobs, acts, None, return_sum_log_likelihood)
                   if check or logprobs.dim() == 3 else
                   (logprobs.sum(1) if return_sum_log_likelihood else logprobs)}
        if return_actions:
            outdict["actions"] = acts
        if return_entropy:
            outdict["entropy"] = calculate_entropy(logprobs)
        if return_hidden:
            outdict["hidden"] = cache
        if return_init_embeds:
            outdict["init_embeds"] = init_embeds
        return outdict

    # ---- reference loop on step kernels (base.py:221-257) ---------------------------------
    def _forward_steps(self, td, env, hidden, decode_type, actions, max_steps, cfg):
        strategy = get_decoding_strategy(decode_type, **cfg)
        start_action = None
        if actions is not None and strategy.multistart:
            start_action, actions = actions[..., 0], actions[..., 1:]
        td, env, num_starts = strategy.pre_decoder_hook(td, env, action=start_action)
        td, env, cache = self.decoder.pre_decoder_hook(td, env, hidden, num_starts)
        step = 0
        while not td["done"].all():
            logits, mask = self.decoder(td, cache, num_starts)
            td = strategy.step(logits, mask, td,
                               action=actions[..., step] if actions is not None else None)
            td = env.step(td)["next"]
            step += 1
            if step > max_steps:
                break
        logprobs, acts, td, env = strategy.post_decoder_hook(td, env)
        reward = env.get_reward(td, acts)
        return logprobs, acts, td, reward, cache

    # ---- one persistent kernel ------------------------------------------------------------
    @torch.no_grad()
    def _forward_fused(self, td, env, hidden, decode_type, actions, max_steps, check, cfg):
        if cfg.get("top_k") or cfg.get("top_p"):
            raise NotImplementedError("top_k / top_p filtering is outside the accelerated path")
        if not cfg.get("mask_logits", True):
            raise NotImplementedError("mask_logits=False is outside the accelerated path")
        if decode_type == "beam_search" or cfg.get("num_samples") or cfg.get("multisample"):
            raise NotImplementedError(f"{decode_type} / multisample is outside the accelerated path")
        base = decode_type.replace("multistart_", "")
        if base not in _MODES:
            raise ValueError(f"Unknown decode type '{decode_type}'")
        num_starts = cfg.get("num_starts")
        multistart = ("multistart" in decode_type) or cfg.get("multistart", False)
        if num_starts is not None:
            multistart = num_starts > 1
        env_id = _lib.ENV_TSP if env.name == "tsp" else _lib.ENV_CVRP
        B = td.batch_size[0]
        n = td["action_mask"].shape[-1]
        S = 0
        if multistart:
            S = env.get_num_starts(td) if num_starts is None else num_starts
        rows = B * max(S, 1)
        dev = hidden.device
        with _lib.nvtx_range("rl4co_b200.cache_precompute"):
            td, env, cache = self.decoder.pre_decoder_hook(td, env, hidden, S)
        if actions is not None:
            actions = actions.contiguous()
            T = actions.shape[1]
        else:
            T = n if env_id == _lib.ENV_TSP else 2 * n       # CVRP: T <= 2N
        T = int(min(T, max_steps + (1 if S > 1 else 0)))
        store_all = bool(cfg.get("store_all_logp"))
        acts = torch.zeros(rows, T, dtype=torch.int64, device=dev)
        logp = torch.zeros(rows, T, dtype=torch.float32, device=dev)
        logp_all = torch.zeros(rows, T, n, dtype=torch.float32, device=dev) if store_all else None
        reward = torch.empty(rows, dtype=torch.float32, device=dev)
        ll = torch.empty(rows, dtype=torch.float32, device=dev)
        lengths = torch.empty(rows, dtype=torch.int32, device=dev)
        meta = torch.zeros(3, dtype=torch.int32, device=dev)      # [max_len, err0, err1]
        if getattr(cache, "tc_err", None) is not None:
            meta[1:].bitwise_or_(cache.tc_err)                    # fp16-range flag of rl4co_tc_pack, no sync
        noise = cfg.get("noise")
        if noise is not None:
            if tuple(noise.shape) != (rows, T, n):
                raise ValueError(f"noise must be [rows={rows}, T={T}, n={n}], got {tuple(noise.shape)}")
            noise = noise.contiguous()
        locs = td["locs"].contiguous()
        demand = td["demand"].contiguous() if env_id == _lib.ENV_CVRP else None
        cap = td["vehicle_capacity"].contiguous() if env_id == _lib.ENV_CVRP else None
        cs = cache.cstruct()
        seed = cfg.get("noise_seed")
        if seed is None:             # fresh counter noise per sampling call (decoding.fresh_noise_seed)
            seed = fresh_noise_seed() if (base == "sampling" and noise is None) else 0
        seed = int(seed) & 0xFFFFFFFF
        self.last_noise_seed = seed
        # forced first actions follow env.select_start_nodes: `% env.generator.num_loc` (ops.py:147-161)
        start_mod = int(getattr(getattr(env, "generator", None), "num_loc", 0) or 0)
        args = _lib.RolloutArgs(
            env=env_id, mode=_MODES[base], num_starts=S, max_steps=T,
            tanh_clipping=float(cfg["tanh_clipping"]), temperature=float(cfg["temperature"]),
            seed=seed, inst_offset=int(cfg.get("instance_offset", 0)),
            locs=locs.data_ptr(), demand=demand.data_ptr() if demand is not None else None,
            vehicle_capacity=cap.data_ptr() if cap is not None else None,
            gumbel=noise.data_ptr() if noise is not None else None,
            actions_in=actions.data_ptr() if actions is not None else None,
            actions_in_stride=actions.stride(0) if actions is not None else 0,
            actions_in_T=actions.shape[1] if actions is not None else 0,
            actions=acts.data_ptr(), logp=logp.data_ptr(),
            logp_all=logp_all.data_ptr() if store_all else None, reward=reward.data_ptr(),
            log_likelihood=ll.data_ptr(), lengths=lengths.data_ptr(), max_len=meta.data_ptr(),
            err=meta.data_ptr() + 4, impl=int(cfg.get("impl", _lib.IMPL_AUTO)), start_mod=start_mod,
            phase_cycles=cfg["_phase_cycles"].data_ptr() if cfg.get("_phase_cycles") is not None else None)
        events = cfg.get("_events")           # bench.py: CUDA events around the kernel alone
        if events is not None:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        with _lib.nvtx_range("rl4co_b200.rollout"):
            rc = _lib.lib().rl4co_rollout(C.byref(cs), C.byref(args), _lib.stream_ptr())
        _lib.check(rc, "rl4co_rollout")
        if events is not None:
            e1.record()
            events.append((e0, e1))
        self.last_lengths = lengths
        self.last_status = meta              # device int32 [max_len, error bits, first offending row + 1]
        if env_id == _lib.ENV_CVRP or check:
            m = meta.tolist()                 # the ONE host sync of the rollout
            if check:
                _lib.raise_on_err(meta[1:]) if m[1] else None
            if env_id == _lib.ENV_CVRP and actions is None:
                T_used = max(m[0], 1)         # batch-max length, as `while not done.all()`
                acts, logp = acts[:, :T_used], logp[:, :T_used]
                if store_all:
                    logp_all = logp_all[:, :T_used]
        if check and env.check_solution:
            env.check_solution_validity(td, acts)
        out_td = td
        logprobs = logp_all if store_all else logp
        if S > 0 and cfg.get("select_best"):
            best = unbatchify(reward, S).max(dim=-1)[1]                     # decoding.py:419-427
            pick = lambda x: unbatchify(x, S).gather(
                1, best.view(-1, *([1] * x.dim())).expand(-1, 1, *x.shape[1:])).squeeze(1)
            logprobs, acts, reward = pick(logprobs), pick(acts), pick(reward)
        return logprobs, acts, out_td, reward, cache
