"""POMO / REINFORCE training step on top of the rollout kernels.

Reference flow (pomo/model.py:89-114, reinforce.py:73-115, baselines.py:57-61): one differentiable
`policy(td, env, phase="train")` call that samples a multistart rollout while recording log-probabilities,
then `loss = -((reward - reward.mean(over starts)) * log_likelihood).mean()`.

Here the sampling rollout runs on the fused kernel (no autograd), and the log-likelihood of the sampled
actions is re-evaluated by `evaluate_log_likelihood`: the same arithmetic as the reference's evaluate mode
(`decoding.py:452-465`, `constructive/base.py:204-206,221-229`), but teacher-forced over ALL decoding steps at
once - with the actions known, every step's context (first / current node, remaining capacity) and feasibility
mask follows from the action prefix, so the T sequential decoder steps become one batch of independent rows: on CUDA
the fused kernels `rl4co_eval_ll_fwd` / `rl4co_eval_ll_bwd` (csrc/evaluate_kernels.cu, the backward's reductions on
tcgen05) behind a `torch.autograd.Function`, elsewhere one batched masked attention in plain PyTorch.  Same actions =>
same log-likelihood => same gradient as the reference's sample-while-recording pass.

Multi-GPU: instances are sharded, the shared baseline is per instance and therefore local,
and the step ends with ONE flat-buffer all-reduce of gradients + logged sums (`sharding.allreduce_mean_`).
"""
import ctypes as C
import math
from typing import Dict, Optional

import torch
import torch.nn.functional as F

from . import _lib, sharding
from .tensordict_lite import batchify


def _to_bs(x: torch.Tensor, S: int) -> torch.Tensor:
    """start-major rows [S*B, ...] -> [B, S, ...] (ops.py:35-56)."""
    return x.view(S, x.shape[0] // S, *x.shape[1:]).transpose(0, 1)


@torch.no_grad()
def teacher_forced_states(env, td, actions: torch.Tensor, num_starts: int) -> Dict[str, torch.Tensor]:
    """Replay `actions` [S*B, T] through the env kernels (tsp/env.py:62-88, cvrp/env.py:68-144) and collect what the
    decoder sees before every step t >= 1: current node, remaining capacity (CVRP) and the feasibility mask.
    Column 0 is the multistart start action (decoding.py:305-327), taken before the first decoder step."""
    S = int(num_starts)
    if S <= 1:
        raise NotImplementedError("evaluate_log_likelihood covers the multistart form (num_starts > 1)")
    T = actions.shape[1]
    cvrp = env.name == "cvrp"
    if actions.is_cuda and td["action_mask"].shape[-1] <= 128:
        return _replay_states_kernel(td, actions.contiguous(), S, cvrp)
    tdb = batchify(td, S)
    tdb.set("action", actions[:, 0].contiguous())
    tdb = env.step(tdb)["next"]
    cur, rem, mask = [], [], []
    for t in range(1, T):
        cur.append(tdb["current_node"].reshape(-1))
        mask.append(tdb["action_mask"])
        if cvrp:
            rem.append((tdb["vehicle_capacity"] - tdb["used_capacity"]).reshape(-1))
        tdb.set("action", actions[:, t].contiguous())
        tdb = env.step(tdb)["next"]
    out = {"first": _to_bs(actions[:, 0], S), "cur": _to_bs(torch.stack(cur, 1), S),
           "mask": _to_bs(torch.stack(mask, 1), S), "taken": _to_bs(actions[:, 1:], S), "td": tdb}
    if cvrp:
        out["rem"] = _to_bs(torch.stack(rem, 1), S)
    return out


def _replay_states_kernel(td, actions: torch.Tensor, S: int, cvrp: bool) -> Dict[str, torch.Tensor]:
    """The same states from ONE launch of rl4co_replay_states (T-1 env steps per row inside the kernel), written
    directly in [B, S, T-1, ...] order."""
    rows, T = actions.shape
    B = rows // S
    n = td["action_mask"].shape[-1]
    dev = actions.device
    cur = torch.empty(B, S, T - 1, dtype=torch.int64, device=dev)
    mask = torch.empty(B, S, T - 1, n, dtype=torch.bool, device=dev)
    rem = torch.empty(B, S, T - 1, dtype=torch.float32, device=dev) if cvrp else None
    demand = td["demand"].reshape(B, n - 1).contiguous().float() if cvrp else None
    cap = td["vehicle_capacity"].reshape(B).contiguous().float() if cvrp else None
    rc = _lib.lib().rl4co_replay_states(
        C.c_int32(_lib.ENV_CVRP if cvrp else _lib.ENV_TSP), _lib.ptr(actions), C.c_int64(rows), C.c_int32(T),
        C.c_int64(B), C.c_int32(n), _lib.ptr(demand), _lib.ptr(cap), _lib.ptr(cur), _lib.ptr(mask), _lib.ptr(rem),
        _lib.stream_ptr())
    _lib.check(rc, "rl4co_replay_states")
    out = {"first": _to_bs(actions[:, 0], S), "cur": cur, "mask": mask, "taken": _to_bs(actions[:, 1:], S), "td": None}
    if cvrp:
        out["rem"] = rem
    return out


class _FusedEvaluate(torch.autograd.Function):
    """log p(taken | row) [B,R] from the fused kernels rl4co_eval_ll_fwd / rl4co_eval_ll_bwd (csrc/evaluate_kernels.cu):
    masked 8-head glimpse + pointer + log_softmax + gather per row, tables of an instance in shared memory; the backward
    recomputes the probabilities from the saved log-sum-exps (row pass) and reduces over rows on the tensor cores
    (dgk, dgv, dlkf with their accumulators in TMEM).  Differentiable in q, gk, gv, lkf."""

    @staticmethod
    def forward(ctx, q, gk, gv, lkf, mask, taken, tanh_clipping, temperature):
        B, R, E = q.shape
        n = gk.shape[1]
        q, gk, gv, lkf = (t.contiguous().float() for t in (q, gk, gv, lkf))
        mask = mask.reshape(B, R, n).contiguous()
        taken = taken.reshape(B, R).contiguous()
        dev = q.device
        logp = torch.empty(B, R, device=dev)
        heads = torch.empty(B, R, E, device=dev)
        lse_h = torch.empty(B, R, 8, device=dev)
        lse_z = torch.empty(B, R, device=dev)
        maskw = torch.empty(B, R, 4, dtype=torch.int32, device=dev)
        rc = _lib.lib().rl4co_eval_ll_fwd(_lib.ptr(q), _lib.ptr(gk), _lib.ptr(gv), _lib.ptr(lkf), _lib.ptr(mask),
                                          _lib.ptr(taken), _lib.ptr(logp), _lib.ptr(heads), _lib.ptr(lse_h), _lib.ptr(lse_z),
                                          _lib.ptr(maskw), C.c_int64(B), C.c_int64(R), C.c_int32(n), C.c_float(tanh_clipping),
                                          C.c_float(temperature), _lib.stream_ptr())
        _lib.check(rc, "rl4co_eval_ll_fwd")
        ctx.save_for_backward(q, gk, gv, lkf, maskw, taken, heads, lse_h, lse_z)
        ctx.scalars = (float(tanh_clipping), float(temperature))
        return logp

    @staticmethod
    def backward(ctx, dlogp):
        q, gk, gv, lkf, maskw, taken, heads, lse_h, lse_z = ctx.saved_tensors
        B, R, E = q.shape
        n = gk.shape[1]
        dlogp = dlogp.contiguous().float()
        dheads, du_t, dd = torch.empty_like(heads), torch.empty_like(heads), torch.empty_like(lse_h)   # scratch
        scale_ws = torch.empty(2, device=q.device)                                                     # scratch
        dq = torch.empty_like(q)
        dgk, dgv, dlkf = torch.zeros_like(gk), torch.zeros_like(gv), torch.zeros_like(lkf)
        rc = _lib.lib().rl4co_eval_ll_bwd(_lib.ptr(q), _lib.ptr(gk), _lib.ptr(gv), _lib.ptr(lkf), _lib.ptr(maskw),
                                          _lib.ptr(taken), _lib.ptr(heads), _lib.ptr(lse_h), _lib.ptr(lse_z), _lib.ptr(dlogp),
                                          _lib.ptr(dheads), _lib.ptr(du_t), _lib.ptr(dd), _lib.ptr(scale_ws), _lib.ptr(dq),
                                          _lib.ptr(dgk), _lib.ptr(dgv), _lib.ptr(dlkf),
                                          C.c_int64(B), C.c_int64(R), C.c_int32(n), C.c_float(ctx.scalars[0]),
                                          C.c_float(ctx.scalars[1]), _lib.stream_ptr())
        _lib.check(rc, "rl4co_eval_ll_bwd")
        return dq, dgk, dgv, dlkf, None, None, None, None


class _FusedQuery(torch.autograd.Function):
    """Glimpse query q [B, S*Tm, 128] of all teacher-forced rows from rl4co_eval_query_fwd / rl4co_eval_query_bwd
    (csrc/evaluate_query.cu): pb[b, cur] (+ pa[b, first]) (+ rem * wcap) (+ gctx[b]) in one streaming pass; the backward
    sums dq over the rows of each node (counting sort by node inside a CTA, register sums, one red.global per 64 rows and
    channel) where PyTorch runs index_add_ with 128 colliding atomics per row.  Differentiable in pb, pa, wcap, gctx."""

    @staticmethod
    def forward(ctx, pb, pa, wcap, gctx, cur, first, rem):
        B, n, E = pb.shape
        S, Tm = cur.shape[1], cur.shape[2]
        f32 = lambda t: None if t is None else t.contiguous().float()
        pb, pa, wcap, gctx = f32(pb), f32(pa), f32(wcap), f32(gctx)
        cur = cur.contiguous()
        first = first.contiguous() if pa is not None else None
        rem = f32(rem) if wcap is not None else None
        q = torch.empty(B, S * Tm, E, device=pb.device)
        rc = _lib.lib().rl4co_eval_query_fwd(_lib.ptr(pb), _lib.ptr(pa), _lib.ptr(wcap), _lib.ptr(gctx), _lib.ptr(cur),
                                             _lib.ptr(first), _lib.ptr(rem), _lib.ptr(q), C.c_int64(B), C.c_int64(S),
                                             C.c_int64(Tm), C.c_int32(n), _lib.stream_ptr())
        _lib.check(rc, "rl4co_eval_query_fwd")
        ctx.save_for_backward(cur, first, rem)
        ctx.meta = (B, n, S, Tm, gctx is not None)
        return q

    @staticmethod
    def backward(ctx, dq):
        cur, first, rem = ctx.saved_tensors
        B, n, S, Tm, has_gctx = ctx.meta
        dq = dq.contiguous().float()
        dpb = torch.zeros(B, n, dq.shape[-1], device=dq.device)
        dpa = torch.zeros_like(dpb) if first is not None else None
        dwcap = torch.zeros(dq.shape[-1], device=dq.device) if rem is not None else None
        rc = _lib.lib().rl4co_eval_query_bwd(_lib.ptr(dq), _lib.ptr(cur), _lib.ptr(first), _lib.ptr(rem), _lib.ptr(dpb),
                                             _lib.ptr(dpa), _lib.ptr(dwcap), C.c_int64(B), C.c_int64(S), C.c_int64(Tm),
                                             C.c_int32(n), _lib.stream_ptr())
        _lib.check(rc, "rl4co_eval_query_bwd")
        return dpb, dpa, dwcap, (dpb.sum(1) if has_gctx else None), None, None, None


def fused_evaluate_supported(h: torch.Tensor, n: int) -> bool:
    """The fused kernels cover CUDA fp32 embeddings of width 128 (8 heads) and n <= rl4co_eval_ll_max_nodes()."""
    return h.is_cuda and h.shape[-1] == 128 and n <= int(_lib.lib().rl4co_eval_ll_max_nodes())


def batched_log_likelihood(decoder, h: torch.Tensor, states: Dict[str, torch.Tensor], tanh_clipping: float = 10.0,
                           temperature: float = 1.0, fused: Optional[bool] = None) -> torch.Tensor:
    """Differentiable log p(action_t | prefix) for all (instance, start, step) at once -> [B, S, T-1].

    am/decoder.py:133-198 (context -> glimpse query), attention.py:282-328 (masked MHA glimpse, project_out,
    logits / sqrt(E)), decoding.py:140-190 (tanh clip, mask, temperature, log_softmax).  `h` is the encoder output with its
    graph attached.  The per-instance table products are PyTorch (autograd); the query of every row (`_FusedQuery`) and
    the masked attention, pointer and log-softmax of the B*S*(T-1) rows (`_FusedEvaluate`) run on the fused forward /
    backward kernels on CUDA, and as plain PyTorch ops (`fused=False`, CPU tensors, n > 128) otherwise - the form the host
    tests pin against the oracle."""
    B, n, E = h.shape
    H = decoder.num_heads
    cur, mask, taken = states["cur"], states["mask"], states["taken"]
    S, Tm = cur.shape[1], cur.shape[2]
    R = S * Tm
    Wk, Wv, Wl = decoder.project_node_embeddings.weight.chunk(3, dim=0)           # decoder.py:206-235
    Wc = decoder.context_embedding.project_context.weight
    gk, gv, lk = F.linear(h, Wk), F.linear(h, Wv), F.linear(h, Wl)
    if fused is None:
        fused = fused_evaluate_supported(h, n) and H == 8
    tsp = decoder.env_name == "tsp"
    # project_context is linear: apply its halves to the n node embeddings, the rows gather the products
    pa = F.linear(h, Wc[:, :E]) if tsp else None                                   # context.py:118-140
    pb = F.linear(h, Wc[:, E:] if tsp else Wc[:, :E])                              # context.py:70-74,143-157
    gctx = (F.linear(h.mean(1), decoder.project_fixed_context.weight)              # decoder.py:223-224
            if decoder.use_graph_context else None)
    if fused:
        q = _FusedQuery.apply(pb, pa, None if tsp else Wc[:, E], gctx, cur, states["first"] if tsp else None,
                              None if tsp else states["rem"])
    else:
        # row gathers tab[b, i[b, ...]] as ONE index_select over the flattened table (a gather with an index expanded to
        # the 128 channels reads 8 bytes of index per element; its backward is the same scatter-add either way)
        boff = (torch.arange(B, device=h.device) * n).view(B, 1)
        idx = lambda tab, i: tab.reshape(B * n, E).index_select(0, (i.reshape(B, -1) + boff).reshape(-1)).view(B, -1, E)
        q = idx(pb, cur).view(B, S, Tm, E)
        q = q + (idx(pa, states["first"]).view(B, S, 1, E) if tsp else states["rem"].unsqueeze(-1) * Wc[:, E])
        if gctx is not None:
            q = q + gctx.view(B, 1, 1, E)
    if fused:
        # logits = (glimpse W_out^T) lk^T = glimpse (lk W_out)^T: the fold the rollout kernels use (decoder.py cache)
        lkf = torch.matmul(lk, decoder.pointer.project_out.weight)
        logp = _FusedEvaluate.apply(q.reshape(B, R, E), gk, gv, lkf, mask, taken, float(tanh_clipping), float(temperature))
        return logp.view(B, S, Tm)
    heads = lambda x, L: x.reshape(B, L, H, E // H).transpose(1, 2)
    am = mask.reshape(B, 1, R, n)
    glimpse = F.scaled_dot_product_attention(heads(q, R), heads(gk, n), heads(gv, n), attn_mask=am)
    glimpse = F.linear(glimpse.transpose(1, 2).reshape(B, R, E), decoder.pointer.project_out.weight)
    logits = torch.bmm(glimpse, lk.transpose(1, 2)) / math.sqrt(E)                 # attention.py:299-301
    if tanh_clipping > 0:
        logits = torch.tanh(logits) * tanh_clipping
    logits = logits.masked_fill(~mask.reshape(B, R, n), float("-inf")) / temperature
    logp = F.log_softmax(logits, dim=-1)
    return logp.gather(-1, taken.reshape(B, R, 1)).view(B, S, Tm)


def evaluate_log_likelihood(policy, td, env, actions: torch.Tensor, num_starts: int,
                            states: Optional[Dict[str, torch.Tensor]] = None,
                            hidden: Optional[torch.Tensor] = None) -> torch.Tensor:
    """log-likelihood [S*B] (start-major, like `policy(...)["log_likelihood"]`) of given multistart `actions`,
    with the autograd graph to every policy parameter (encoder included: the PyTorch modules run here, or the
    caller's `hidden` with its graph)."""
    if states is None:
        states = teacher_forced_states(env, td, actions, num_starts)
    with torch.enable_grad():
        h = hidden if hidden is not None else policy.encoder(td)[0]
        logp = batched_log_likelihood(policy.decoder, h, states, policy.tanh_clipping, policy.temperature)
    return logp.sum(-1).transpose(0, 1).reshape(-1)


def reinforce_loss(reward: torch.Tensor, log_likelihood: torch.Tensor, num_starts: int):
    """pomo/model.py:106-112 + baselines.py:57-61 (shared baseline = mean over the starts of an instance) +
    reinforce.py:97-106.  reward, log_likelihood: [S*B] start-major.  Returns (loss, advantage [B,S])."""
    r, ll = _to_bs(reward, num_starts), _to_bs(log_likelihood, num_starts)
    adv = r - r.mean(dim=1, keepdim=True)
    return -(adv * ll).mean(), adv


def graphed_encoder(policy, td, num_warmup_iters: int = 3):
    """The differentiable (PyTorch-module) encoder with its forward and its backward each captured as ONE CUDA graph
    (torch.cuda.make_graphed_callables): at the c5 size the ~600 launches of small fp32 GEMM / attention / normalisation
    kernels per step are CPU-bound, replayed from two graphs they are not.  Returns `encode(td) -> h [B, n, E]` (with
    its autograd graph) for instances of the shapes of `td`; pass it as `pomo_train_step(..., encoder=encode)`.  The
    output and the parameter gradients live in the graphs' static buffers: valid until the next call / backward.
    Buffers (BatchNorm running statistics) are restored after the capture's warm-up passes.  (PyTorch warns once that the
    parameters' AccumulateGrad nodes were created on the capture's side stream: the engine then synchronises the two
    streams when it accumulates the encoder gradients, which is what the step needs anyway.)"""
    keys = [k for k in ("locs", "demand") if k in td.keys()]

    class _Encode(torch.nn.Module):
        def __init__(self, enc):
            super().__init__()
            self.enc = enc

        def forward(self, *xs):
            return self.enc(dict(zip(keys, xs)))[0]

    mod = _Encode(policy.encoder)
    mod.train(policy.encoder.training)
    saved = {k: v.clone() for k, v in mod.named_buffers()}
    graphed = torch.cuda.make_graphed_callables(mod, tuple(td[k].detach().clone() for k in keys),
                                                num_warmup_iters=num_warmup_iters)
    with torch.no_grad():
        for k, v in mod.named_buffers():
            v.copy_(saved[k])
    return lambda td_: graphed(*(td_[k] for k in keys))


def pomo_train_step(policy, env, td, optimizer=None, num_starts: Optional[int] = None, noise_seed: Optional[int] = None,
                    instance_offset: int = 0, max_grad_norm: Optional[float] = 1.0, group=None,
                    encoder=None) -> Dict[str, float]:
    """One POMO training step on this rank's shard `td`: sampled multistart rollout (fused kernel), differentiable
    re-evaluation, shared-baseline REINFORCE loss, backward, one all-reduce, clip, optimiser step.
    `encoder`: an `encode(td) -> h` replacing `policy.encoder(td)[0]` (see `graphed_encoder`).
    `noise_seed=None` draws a fresh seed per step from torch's global generator (decoding.fresh_noise_seed); ranks
    seeded alike draw alike and the noise is keyed by `instance_offset + b`, so shards never share noise."""
    kw = {} if num_starts is None else {"num_starts": num_starts}
    # ONE encoder pass per step, as in the reference's single differentiable forward (train-mode BatchNorm statistics
    # are updated once, and the rollout samples from exactly the embeddings the loss differentiates)
    with torch.enable_grad():
        h = encoder(td) if encoder is not None else policy.encoder(td)[0]
    with torch.no_grad():
        out = policy(td, env, decode_type="multistart_sampling", noise_seed=noise_seed,
                     instance_offset=instance_offset, hidden=h, **kw)
    S = out["actions"].shape[0] // td.batch_size[0]
    ll = evaluate_log_likelihood(policy, td, env, out["actions"], S, hidden=h)
    loss, _ = reinforce_loss(out["reward"], ll, S)
    params = [p for p in policy.parameters() if p.requires_grad]
    for p in params:
        p.grad = None
    loss.backward()
    grads = []
    for p in params:
        if p.grad is None:
            p.grad = torch.zeros_like(p)
        grads.append(p.grad)
    r = _to_bs(out["reward"], S)
    extras = torch.stack([loss.detach(), r.mean(), r.max(1)[0].mean()])      # logged means (rl/common/base.py:233-240)
    extras = sharding.allreduce_mean_(grads, extras, group=group)
    if max_grad_norm is not None:
        torch.nn.utils.clip_grad_norm_(params, max_grad_norm)
    if optimizer is not None:
        optimizer.step()
    ll_gap = (ll.detach() - out["log_likelihood"]).abs().max()
    return {"loss": extras[0].item(), "reward": extras[1].item(), "max_reward": extras[2].item(),
            "ll_gap": ll_gap.item()}
