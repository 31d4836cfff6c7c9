"""TEST INFRASTRUCTURE ONLY — CPU restatement of the reference's construction-rollout path.

This file is the ORACLE: a plain-torch (fp32, op-for-op) restatement of the reference's
TSP/CVRP env step + AttentionModel/POMO decoder step + decode strategies, without
tensordict/torchrl/lightning so that it can travel to the GPU box (where
/root/reference does not exist).  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import it; the product path
(eam_rl4co_b200/) never does.

Parity pinning: the reference's test-suite holds NO golden vectors for this path
(SURVEY.md §4, §8c).  The oracle is therefore pinned against OUTPUTS OF THE REFERENCE
ITSELF, run unmodified in the build container through oracle/ref_loader.py:
  * tests/test_oracle_vs_reference.py compares every function here with the reference
    live (skipped where /root/reference is absent);
  * oracle/make_golden.py dumps reference outputs to tests/golden/*.npz, and
    tests/test_oracle_golden.py checks the oracle against those fixtures everywhere.

All file:line citations are relative to /root/reference.
State is a plain dict of tensors with the reference's keys, dtypes and shapes.
"""
import math

import torch
import torch.nn.functional as F

# --------------------------------------------------------------------------------------
# Counter-based noise (SURVEY.md §7.3-4, §8c "Noise-driven sampling in the oracle").
# torch.multinomial (rl4co/utils/decoding.py:407) consumes RNG in an implementation-
# defined way, so bit-exact sampling is defined with SUPPLIED noise:
#   action = argmax_n(logp[n] + g[n]),  g = -log(-log(u)),  u = U(seed, b, s, t, n)
# The integer hash below is implemented identically in csrc/noise.cuh.
# --------------------------------------------------------------------------------------
_M32 = 0xFFFFFFFF


def _mix32(x: torch.Tensor) -> torch.Tensor:
    """lowbias32 integer finaliser on int64 tensors holding uint32 values."""
    x = x & _M32
    x = x ^ (x >> 16)
    x = (x * 0x7FEB352D) & _M32
    x = x ^ (x >> 15)
    x = (x * 0x846CA68B) & _M32
    x = x ^ (x >> 16)
    return x


def noise_uniform(seed: int, b: torch.Tensor, s: torch.Tensor, t: int, n_nodes: int):
    """u in (0,1), fp32-exact: ((hash >> 9) + 0.5) * 2^-23.  b, s: int64 [rows]."""
    h = _mix32((seed & _M32) ^ ((b * 0x9E3779B1) & _M32))
    h = _mix32(h ^ ((s * 0x85EBCA77) & _M32))
    h = _mix32(h ^ ((t * 0xC2B2AE3D) & _M32))
    n = torch.arange(n_nodes, dtype=torch.int64, device=b.device)
    h = _mix32(h[:, None] ^ ((n[None, :] * 0x27D4EB2F) & _M32))
    return ((h >> 9).to(torch.float32) + 0.5) * (2.0 ** -23)


def noise_gumbel(seed: int, b, s, t: int, n_nodes: int) -> torch.Tensor:
    u = noise_uniform(seed, b, s, t, n_nodes)
    return -torch.log(-torch.log(u))


# --------------------------------------------------------------------------------------
# rl4co/utils/ops.py
# --------------------------------------------------------------------------------------
def batchify(x: torch.Tensor, repeats: int) -> torch.Tensor:
    """ops.py:13-32 — start-major copy: row = s*B + b."""
    if repeats <= 0:
        return x
    s = x.shape
    return x.expand(repeats, *s).contiguous().view(s[0] * repeats, *s[1:])


def unbatchify(x: torch.Tensor, repeats: int) -> torch.Tensor:
    """ops.py:35-56 — [S*B, ...] -> [B, S, ...] (non-contiguous view)."""
    if repeats <= 0:
        return x
    s = x.shape
    return x.view(repeats, s[0] // repeats, *s[1:]).permute(1, 0, *range(2, len(s) + 1))


def gather_by_index(src, idx, dim=1, squeeze=True):
    """ops.py:59-71."""
    expanded_shape = list(src.shape)
    expanded_shape[dim] = -1
    idx = idx.view(idx.shape + (1,) * (src.dim() - idx.dim())).expand(expanded_shape)
    squeeze = idx.size(dim) == 1 and squeeze
    return src.gather(dim, idx).squeeze(dim) if squeeze else src.gather(dim, idx)


def get_tour_length(ordered_locs):
    """ops.py:82-95."""
    nxt = torch.roll(ordered_locs, -1, dims=-2)
    return (nxt - ordered_locs).norm(p=2, dim=-1).sum(-1)


def select_start_nodes(env_name: str, B: int, num_starts: int, num_loc: int, device="cpu"):
    """ops.py:133-169 (TSP branch :147-151, depot branch :155-161)."""
    sel = torch.arange(num_starts, device=device).repeat_interleave(B) % num_loc
    return sel if env_name == "tsp" else sel + 1


def get_num_starts(env_name: str, action_mask: torch.Tensor) -> int:
    """ops.py:120-130."""
    n = action_mask.shape[-1]
    return n - 1 if env_name == "cvrp" else n


# --------------------------------------------------------------------------------------
# rl4co/envs/routing/tsp/env.py
# --------------------------------------------------------------------------------------
def tsp_reset(locs: torch.Tensor) -> dict:
    """tsp/env.py:90-115 (+ done from envs/common/base.py:144-148)."""
    B, N, _ = locs.shape
    dev = locs.device
    cur = torch.zeros(B, dtype=torch.int64, device=dev)
    return {
        "locs": locs,
        "first_node": cur,
        "current_node": cur,
        "i": torch.zeros(B, 1, dtype=torch.int64, device=dev),
        "action_mask": torch.ones(B, N, dtype=torch.bool, device=dev),
        "reward": torch.zeros(B, 1, dtype=torch.float32, device=dev),
        "done": torch.zeros(B, 1, dtype=torch.bool, device=dev),
    }


def tsp_step(td: dict) -> dict:
    """tsp/env.py:62-88."""
    current_node = td["action"]
    first_node = current_node if td["i"].all() == 0 else td["first_node"]
    available = td["action_mask"].scatter(
        -1, current_node.unsqueeze(-1).expand_as(td["action_mask"]), 0
    )
    done = torch.sum(available, dim=-1) == 0
    td.update(
        first_node=first_node,
        current_node=current_node,
        i=td["i"] + 1,
        action_mask=available,
        reward=torch.zeros_like(done),
        done=done,
    )
    return td


def tsp_reward(locs: torch.Tensor, actions: torch.Tensor) -> torch.Tensor:
    """tsp/env.py:152-158."""
    return -get_tour_length(gather_by_index(locs, actions))


def tsp_check_solution(actions: torch.Tensor) -> None:
    """tsp/env.py:160-168."""
    n = actions.size(1)
    ok = (torch.arange(n, device=actions.device).view(1, -1).expand_as(actions)
          == actions.sort(1)[0]).all()
    assert ok, "Invalid tour"


# --------------------------------------------------------------------------------------
# rl4co/envs/routing/cvrp/env.py
# --------------------------------------------------------------------------------------
def cvrp_action_mask(td: dict) -> torch.Tensor:
    """cvrp/env.py:132-144."""
    exceeds_cap = td["demand"] + td["used_capacity"] > td["vehicle_capacity"] + 1e-5
    mask_loc = td["visited"][..., 1:].to(exceeds_cap.dtype) | exceeds_cap
    mask_depot = (td["current_node"] == 0) & ((mask_loc == 0).int().sum(-1) > 0)[:, None]
    return ~torch.cat((mask_depot, mask_loc), -1)


def cvrp_reset(locs: torch.Tensor, depot: torch.Tensor, demand: torch.Tensor,
               vehicle_capacity: float = 1.0) -> dict:
    """cvrp/env.py:102-130. `demand` is already divided by capacity (generator.py:143)."""
    B, N, _ = locs.shape
    dev = locs.device
    td = {
        "locs": torch.cat((depot[:, None, :], locs), -2),
        "demand": demand,
        "current_node": torch.zeros(B, 1, dtype=torch.long, device=dev),
        "used_capacity": torch.zeros(B, 1, device=dev),
        "vehicle_capacity": torch.full((B, 1), vehicle_capacity, device=dev),
        "visited": torch.zeros(B, N + 1, dtype=torch.uint8, device=dev),
        "done": torch.zeros(B, 1, dtype=torch.bool, device=dev),
    }
    td["action_mask"] = cvrp_action_mask(td)
    return td


def cvrp_step(td: dict) -> dict:
    """cvrp/env.py:68-100."""
    current_node = td["action"][:, None]
    n_loc = td["demand"].size(-1)
    selected_demand = gather_by_index(
        td["demand"], torch.clamp(current_node - 1, 0, n_loc - 1), squeeze=False
    )
    used_capacity = (td["used_capacity"] + selected_demand) * (current_node != 0).float()
    visited = td["visited"].scatter(-1, current_node, 1)
    done = visited.sum(-1) == visited.size(-1)
    td.update(
        current_node=current_node,
        used_capacity=used_capacity,
        visited=visited,
        reward=torch.zeros_like(done),
        done=done,
    )
    td["action_mask"] = cvrp_action_mask(td)
    return td


def cvrp_reward(locs: torch.Tensor, actions: torch.Tensor) -> torch.Tensor:
    """cvrp/env.py:146-155. `locs` includes the depot at index 0."""
    ordered = torch.cat([locs[..., 0:1, :], gather_by_index(locs, actions)], dim=1)
    return -get_tour_length(ordered)


def cvrp_check_solution(td: dict, actions: torch.Tensor) -> None:
    """cvrp/env.py:157-185."""
    batch_size, graph_size = td["demand"].size()
    sorted_pi = actions.sort(1)[0]
    ok = (torch.arange(1, graph_size + 1, device=actions.device).view(1, -1)
          .expand(batch_size, graph_size) == sorted_pi[:, -graph_size:]).all() and (
        sorted_pi[:, :-graph_size] == 0).all()
    assert ok, "Invalid tour"
    demand_with_depot = torch.cat((-td["vehicle_capacity"], td["demand"]), 1)
    d = demand_with_depot.gather(1, actions)
    used_cap = torch.zeros_like(td["demand"][:, 0])
    for i in range(actions.size(1)):
        used_cap += d[:, i]
        used_cap[used_cap < 0] = 0
        assert (used_cap <= td["vehicle_capacity"][:, 0] + 1e-5).all(), "Used more than capacity"


# --------------------------------------------------------------------------------------
# Decoder: rl4co/models/zoo/am/decoder.py, nn/env_embeddings/context.py, nn/attention.py
# Weights are passed as a dict with the reference's state_dict names (decoder.* prefix
# stripped):
#   "project_node_embeddings.weight" [3E,E], "project_fixed_context.weight" [E,E],
#   "context_embedding.project_context.weight" [E,2E] (tsp) / [E,E+1] (cvrp),
#   "context_embedding.W_placeholder" [2E] (tsp only), "pointer.project_out.weight" [E,E]
# --------------------------------------------------------------------------------------
def precompute_cache(h: torch.Tensor, w: dict, use_graph_context: bool = True) -> dict:
    """am/decoder.py:206-235.  K,V,L are chunk() VIEWS of one [B,N',3E] buffer."""
    k, v, l = F.linear(h, w["project_node_embeddings.weight"]).chunk(3, dim=-1)
    if use_graph_context:
        gc = F.linear(h.mean(1), w["project_fixed_context.weight"])
    else:
        gc = 0
    return {"node_embeddings": h, "graph_context": gc, "glimpse_key": k,
            "glimpse_val": v, "logit_key": l}


def tsp_context(h: torch.Tensor, td: dict, w: dict, batch_dims: int) -> torch.Tensor:
    """context.py:118-140 (TSPContext.forward)."""
    B = h.size(0)
    first, cur = td["first_node"], td["current_node"]
    node_dim = (-1,) if first.dim() == 1 else (first.size(-1), -1)
    if td["i"][(0,) * td["i"].dim()].item() < 1:
        wp = w["context_embedding.W_placeholder"]
        if batch_dims < 2:
            ctx = wp[None, :].expand(B, wp.size(-1))
        else:
            ctx = wp[None, None, :].expand(B, first.size(1), wp.size(-1))
    else:
        ctx = gather_by_index(h, torch.stack([first, cur], -1).view(B, -1)).view(B, *node_dim)
    return F.linear(ctx, w["context_embedding.project_context.weight"])


def cvrp_context(h: torch.Tensor, td: dict, w: dict) -> torch.Tensor:
    """context.py:70-74 (EnvContext.forward) + :155-157 (VRPContext._state_embedding)."""
    cur_emb = gather_by_index(h, td["current_node"])
    state = td["vehicle_capacity"] - td["used_capacity"]
    return F.linear(torch.cat([cur_emb, state], -1),
                    w["context_embedding.project_context.weight"])


def _make_heads(v: torch.Tensor, num_heads: int) -> torch.Tensor:
    """attention.py:324 — '... g (h s) -> ... h g s'."""
    *lead, g, e = v.shape
    return v.view(*lead, g, num_heads, e // num_heads).transpose(-3, -2)


def sdpa_simple(q, k, v, attn_mask=None):
    """attention.py:19-52."""
    scores = torch.matmul(q, k.transpose(-2, -1)) / (k.size(-1) ** 0.5)
    if attn_mask is not None:
        scores.masked_fill_(~attn_mask, float("-inf"))
    return torch.matmul(F.softmax(scores, dim=-1), v)


def pointer_attention(query, key, value, logit_key, attn_mask, w: dict, num_heads: int = 8,
                      mask_inner: bool = True, sdpa: str = "default", check_nan: bool = True):
    """attention.py:282-328 (PointerAttention.forward/_inner_mha/_project_out)."""
    q, k, v = (_make_heads(x, num_heads) for x in (query, key, value))
    if mask_inner:
        am = attn_mask.unsqueeze(1) if attn_mask.ndim == 3 else attn_mask.unsqueeze(1).unsqueeze(2)
    else:
        am = None
    if sdpa == "default":
        heads = F.scaled_dot_product_attention(q, k, v, attn_mask=am)
    else:
        heads = sdpa_simple(q, k, v, attn_mask=am)
    heads = heads.transpose(-3, -2).reshape(*heads.shape[:-3], heads.shape[-2], -1)
    glimpse = F.linear(heads, w["pointer.project_out.weight"])
    logits = torch.bmm(glimpse, logit_key.squeeze(-2).transpose(-2, -1)).squeeze(-2) / math.sqrt(
        glimpse.size(-1))
    if check_nan:
        assert not torch.isnan(logits).any(), "Logits contain NaNs"
    return logits


_STATE_KEYS_TSP = ("first_node", "current_node", "i", "action_mask")
_STATE_KEYS_CVRP = ("current_node", "used_capacity", "vehicle_capacity", "action_mask")


def decoder_step(env_name: str, td: dict, cache: dict, w: dict, num_starts: int = 0,
                 num_heads: int = 8, sdpa: str = "default"):
    """am/decoder.py:161-198 (static embeddings: TSP/CVRP).  Returns (logits, mask) [rows,N']."""
    keys = _STATE_KEYS_TSP if env_name == "tsp" else _STATE_KEYS_CVRP
    if num_starts > 1:
        view = {k: unbatchify(td[k], num_starts) for k in keys}     # decoder.py:183-184
        batch_dims = 2
    else:
        view = {k: td[k] for k in keys}
        batch_dims = 1
    h, gc = cache["node_embeddings"], cache["graph_context"]
    if batch_dims == 2 and isinstance(gc, torch.Tensor):
        gc = gc.unsqueeze(1)                                         # decoder.py:137-138
    if env_name == "tsp":
        ctx = tsp_context(h, view, w, batch_dims)
    else:
        ctx = cvrp_context(h, view, w)
    glimpse_q = ctx + gc
    glimpse_q = glimpse_q.unsqueeze(1) if glimpse_q.ndim == 2 else glimpse_q
    mask = view["action_mask"]
    logits = pointer_attention(glimpse_q, cache["glimpse_key"], cache["glimpse_val"],
                               cache["logit_key"], mask, w, num_heads=num_heads, sdpa=sdpa)
    if num_starts > 1:
        # rearrange "b s l -> (s b) l"  (decoder.py:195-197)
        logits = logits.permute(1, 0, 2).reshape(-1, logits.size(-1))
        mask = mask.permute(1, 0, 2).reshape(-1, mask.size(-1))
    return logits, mask


# --------------------------------------------------------------------------------------
# rl4co/utils/decoding.py
# --------------------------------------------------------------------------------------
def process_logits(logits, mask, temperature=1.0, tanh_clipping=10.0, mask_logits=True):
    """decoding.py:140-190 (top-k / top-p = 0)."""
    if tanh_clipping > 0:
        logits = torch.tanh(logits) * tanh_clipping
    if mask_logits:
        logits[~mask] = float("-inf")
    logits = logits / temperature
    return F.log_softmax(logits, dim=-1)


def select_greedy(logprobs, mask):
    """decoding.py:391-401 — argmax returns the FIRST max index on ties."""
    selected = logprobs.argmax(dim=-1)
    assert not (~mask).gather(1, selected.unsqueeze(-1)).any(), "infeasible action selected"
    return selected


def select_noise(logprobs, gumbel):
    """Supplied-noise replacement for decoding.py:403-417 (see noise_* above)."""
    return (logprobs + gumbel).argmax(dim=-1)


def rollout(env_name: str, inst: dict, h: torch.Tensor, w: dict, decode_type: str = "greedy",
            num_starts: int = None, actions: torch.Tensor = None, use_graph_context: bool = True,
            temperature: float = 1.0, tanh_clipping: float = 10.0, noise_seed: int = 0,
            instance_offset: int = 0, noise: torch.Tensor = None, num_heads: int = 8,
            sdpa: str = "default", store_all_logp: bool = False, check_solution: bool = True,
            max_steps: int = 1_000_000, trace: dict = None) -> dict:
    """ConstructivePolicy.forward after the encoder (constructive/base.py:197-275).

    inst: {"locs"} (tsp) or {"locs","depot","demand"} (cvrp) — generator output.
    h: encoder output [B,N',E].  decode_type in greedy|sampling|multistart_greedy|
    multistart_sampling|evaluate (with actions=[rows,T]; multistart evaluate when num_starts>1,
    following the fork's start_action split, base.py:221-229).
    Sampling uses supplied noise: `noise` [rows,T,N'] Gumbel tensor if given, else the
    counter generator keyed (noise_seed, instance_offset+b, s, t, n), t = action column.
    trace: optional dict that receives per-step lists (logits, mask, logprobs) for fixtures.
    """
    if actions is not None:
        decode_type = "evaluate"                 # base.py:205-206
    multistart = "multistart" in decode_type or (num_starts is not None and num_starts > 1)
    mode = decode_type.replace("multistart_", "")
    td = tsp_reset(inst["locs"]) if env_name == "tsp" else cvrp_reset(
        inst["locs"], inst["depot"], inst["demand"])
    B = td["locs"].size(0)
    n_nodes = td["action_mask"].size(-1)
    step_fn = tsp_step if env_name == "tsp" else cvrp_step
    out_actions, out_logp = [], []

    # decoding.py:284-332 pre_decoder_hook
    S = 0
    if multistart:
        S = get_num_starts(env_name, td["action_mask"]) if num_starts is None else num_starts
        if S is not None and S <= 1:      # decoding.py:254-255: num_starts<=1 => not multistart
            multistart, S = False, 0
    if multistart:
        if actions is not None:
            start_action, actions = actions[..., 0], actions[..., 1:]
        else:
            num_loc = n_nodes if env_name == "tsp" else n_nodes - 1
            start_action = select_start_nodes(env_name, B, S, num_loc, td["locs"].device)
        td = {k: batchify(v, S) for k, v in td.items()}
        td["action"] = start_action
        td = step_fn(td)
        out_logp.append(torch.zeros_like(td["action_mask"]) if store_all_logp
                        else torch.zeros_like(start_action))          # int64 zeros quirk (:324)
        out_actions.append(start_action)
    rows = td["action_mask"].size(0)
    row_b = torch.arange(rows, device=h.device) % B
    row_s = torch.arange(rows, device=h.device) // B

    cache = precompute_cache(h, w, use_graph_context)

    step = 0
    while not td["done"].all():
        logits, mask = decoder_step(env_name, td, cache, w, S, num_heads, sdpa)
        if trace is not None:
            trace.setdefault("logits", []).append(logits.clone())
            trace.setdefault("mask", []).append(mask.clone())
        logprobs = process_logits(logits, mask, temperature, tanh_clipping)
        col = len(out_actions)
        if mode == "greedy":
            selected = select_greedy(logprobs, mask)
        elif mode == "sampling":
            g = noise[:, col] if noise is not None else noise_gumbel(
                noise_seed, row_b + instance_offset, row_s, col, n_nodes)
            selected = select_noise(logprobs, g)
            assert not (~mask).gather(1, selected.unsqueeze(-1)).any(), "infeasible action selected"
        elif mode == "evaluate":
            selected = actions[..., step]
        else:
            raise ValueError(decode_type)
        if trace is not None:
            trace.setdefault("logprobs", []).append(logprobs.clone())
        out_logp.append(logprobs if store_all_logp else gather_by_index(logprobs, selected, dim=1))
        out_actions.append(selected)
        td["action"] = selected
        td = step_fn(td)
        step += 1
        if step > max_steps:
            break

    logprobs = torch.stack(out_logp, 1)         # decoding.py:340-341
    acts = torch.stack(out_actions, 1)
    if check_solution:
        tsp_check_solution(acts) if env_name == "tsp" else cvrp_check_solution(td, acts)
    reward = tsp_reward(td["locs"], acts) if env_name == "tsp" else cvrp_reward(td["locs"], acts)
    lp = logprobs
    if lp.dim() == 3:                            # decoding.py:49-50
        lp = lp.gather(-1, acts.unsqueeze(-1)).squeeze(-1)
    assert (lp > -1000).all(), "Logprobs should not be -inf, check sampling procedure!"
    out = {"reward": reward, "log_likelihood": lp.sum(1), "actions": acts, "logprobs": lp,
           "num_starts": S, "td": td}
    if store_all_logp:
        l = torch.nan_to_num(logprobs, nan=0.0)   # ops.py:108-116 calculate_entropy
        out["entropy"] = (-(l.exp() * l).sum(-1)).sum(1)
    return out


# --------------------------------------------------------------------------------------
# Callers ("next" rows of SURVEY.md §8f)
# --------------------------------------------------------------------------------------
def reinforce_shared_baseline_loss(reward, log_likelihood, num_starts: int):
    """pomo/model.py:106-112 + reinforce.py:97-106 + baselines.py:57-61.
    reward, ll: [S*B] start-major.  Returns (loss, advantage [B,S])."""
    r = unbatchify(reward, num_starts)
    ll = unbatchify(log_likelihood, num_starts)
    bl = r.mean(dim=1, keepdim=True)
    adv = r - bl
    return -(adv * ll).mean(), adv


# --------------------------------------------------------------------------------------
# Generators (input spec): tsp/generator.py:53-62, cvrp/generator.py:120-147
# --------------------------------------------------------------------------------------
CAPACITIES = {10: 20.0, 15: 25.0, 20: 30.0, 30: 33.0, 40: 37.0, 50: 40.0, 60: 43.0, 75: 45.0,
              100: 50.0, 125: 55.0, 150: 60.0, 200: 70.0, 500: 100.0, 1000: 150.0}


def generate_tsp(B: int, N: int) -> dict:
    return {"locs": torch.distributions.Uniform(0.0, 1.0).sample((B, N, 2))}


def generate_cvrp(B: int, N: int) -> dict:
    U = torch.distributions.Uniform
    locs = U(0.0, 1.0).sample((B, N + 1, 2))
    depot, locs = locs[..., 0, :], locs[..., 1:, :]
    demand = U(0.0, 9.0).sample((B, N))          # min_demand-1=0, max_demand=9 (generator.py:96-99)
    demand = (demand.int() + 1).float()
    cap = CAPACITIES[min(CAPACITIES, key=lambda k: abs(k - N))]
    return {"locs": locs, "depot": depot, "demand": demand / cap}
