"""GPU: the tcgen05 / TMEM path (csrc/rollout_tc.cu) against the fp32 CUDA-core kernels, a torch fp32
restatement of each stage (scores, heads, logits) and the oracle.  Stage-level checks come first so a
descriptor / layout mistake is localised."""
import math

import pytest
import torch

import eam_rl4co_b200 as P
from eam_rl4co_b200 import _lib
from oracle import rollout_oracle as O
from helpers import LL_RTOL, compare_rollout, decoder_weights, inst_to_env_td, rel_close

pytestmark = pytest.mark.gpu
DEV = "cuda"
LOGIT_ATOL = 2e-5


def _setup(env_name, N, B, seed=0):
    torch.manual_seed(1234 + seed)
    inst = O.generate_tsp(B, N) if env_name == "tsp" else O.generate_cvrp(B, N)
    torch.manual_seed(4321)
    pol = P.POMOPolicy(env_name).eval()
    w = decoder_weights(pol)
    pol = pol.to(DEV)
    env, td = inst_to_env_td(P, env_name, inst, DEV)
    with torch.no_grad():
        h, _ = pol.encoder(td)
    return inst, pol, w, env, td, h


def _mid_rollout_state(pol, env, td, h, S, steps):
    """Reference-shaped loop on the fp32 step kernels for `steps` decode steps -> (td, cache)."""
    strategy = P.get_decoding_strategy("multistart_greedy", num_starts=S)
    td, env, num_starts = strategy.pre_decoder_hook(td, env)
    td, env, cache = pol.decoder.pre_decoder_hook(td, env, h, num_starts)
    for _ in range(steps):
        logits, mask = pol.decoder(td, cache, num_starts, impl=_lib.IMPL_FP32)
        td = strategy.step(logits, mask, td)
        td = env.step(td)["next"]
    return td, cache, num_starts


def _torch_stages(env_name, cache, td, B, S):
    """fp32 torch restatement of the folded algebra the kernels use (GPU, highest precision)."""
    n, E = cache.glimpse_key.shape[1:]
    unb = lambda x: P.unbatchify(x, S)
    cur = unb(td["current_node"].reshape(-1))
    idx = lambda t, i: t.gather(1, i[..., None].expand(B, S, E))
    if env_name == "tsp":
        q = idx(cache.pa, unb(td["first_node"].reshape(-1))) + idx(cache.pb, cur)
    else:
        rem = unb((td["vehicle_capacity"] - td["used_capacity"]).reshape(-1))
        q = idx(cache.pb, cur) + rem[..., None] * cache.wcap
    mask = unb(td["action_mask"])                                        # [B,S,n]
    qh = q.view(B, S, 8, 16).transpose(1, 2).double()
    kh = cache.glimpse_key.view(B, n, 8, 16).transpose(1, 2).double()
    vh = cache.glimpse_val.view(B, n, 8, 16).transpose(1, 2).double()
    scores = qh @ kh.transpose(-1, -2) * 0.25                           # [B,8,S,n]
    p = scores.masked_fill(~mask[:, None], float("-inf")).softmax(-1)
    heads = (p @ vh).transpose(1, 2).reshape(B, S, E)
    logits = heads @ cache.logit_key_folded.double().transpose(1, 2) / math.sqrt(E)
    to_rows = lambda x: x.permute(1, 0, *range(2, x.dim())).reshape(S * B, *x.shape[2:]).float()
    return to_rows(scores.transpose(1, 2)), to_rows(heads), to_rows(logits)   # rows = s*B + b


@pytest.mark.parametrize("env_name,N,B,steps", [("tsp", 20, 3, 2), ("tsp", 50, 2, 7), ("tsp", 100, 3, 11),
                                                ("cvrp", 20, 3, 3), ("cvrp", 100, 2, 9)])
def test_tc_decode_step_stages(env_name, N, B, steps):
    torch.backends.cuda.matmul.allow_tf32 = False
    inst, pol, w, env, td, h = _setup(env_name, N, B)
    S = N
    td, cache, num_starts = _mid_rollout_state(pol, env, td.clone(), h, S, steps)
    assert cache.tc_img is not None
    dbg = {}
    lg_tc, mask = pol.decoder(td, cache, num_starts, impl=_lib.IMPL_TC, debug=dbg)
    lg_32, _ = pol.decoder(td, cache, num_starts, impl=_lib.IMPL_FP32)
    sc_ref, hd_ref, lg_ref = _torch_stages(env_name, cache, td, B, S)
    e_sc = (dbg["scores"] - sc_ref).abs().max().item()
    e_hd = (dbg["heads"] - hd_ref).abs().max().item()
    e_lg = (lg_tc - lg_ref).abs().max().item()
    e_32 = (lg_tc - lg_32).abs().max().item()
    print(f"{env_name}{N}: |dscore|={e_sc:.2e} |dheads|={e_hd:.2e} |dlogit| vs torch={e_lg:.2e} vs fp32 kernel={e_32:.2e}")
    assert e_sc < 2e-5, f"scores stage off by {e_sc}"
    assert e_hd < 2e-5, f"heads stage off by {e_hd}"
    assert e_lg < LOGIT_ATOL and e_32 < LOGIT_ATOL, (e_lg, e_32)


@pytest.mark.parametrize("env_name,N,B", [("tsp", 100, 2), ("cvrp", 50, 2)])
def test_tc_peaky_attention_takes_the_rescale_path(env_name, N, B):
    """Glimpse keys scaled x12: attention scores spread over tens of units, so the one-pass softmax must raise
    its running maximum (and rescale the chunks already written) many times.  Same answers as the fp32 kernel."""
    inst, pol, w, env, td, h = _setup(env_name, N, B, seed=7)
    with torch.no_grad():
        pol.decoder.project_node_embeddings.weight[:128].mul_(12.0)
    td, cache, num_starts = _mid_rollout_state(pol, env, td.clone(), h, N, 5)
    dbg = {}
    lg_tc, _ = pol.decoder(td, cache, num_starts, impl=_lib.IMPL_TC, debug=dbg)
    lg_32, _ = pol.decoder(td, cache, num_starts, impl=_lib.IMPL_FP32)
    sc_ref, hd_ref, lg_ref = _torch_stages(env_name, cache, td, B, N)
    spread = (sc_ref.max() - sc_ref.min()).item()
    e_hd = (dbg["heads"] - hd_ref).abs().max().item()
    e_lg = (lg_tc - lg_ref).abs().max().item()
    print(f"{env_name}{N} peaky: score spread={spread:.1f} |dheads|={e_hd:.2e} |dlogit|={e_lg:.2e} "
          f"vs fp32 kernel={(lg_tc - lg_32).abs().max().item():.2e}")
    assert spread > 30
    assert e_hd < 5e-5 and e_lg < 5e-5


@pytest.mark.parametrize("env_name,N,B", [("tsp", 20, 5), ("tsp", 100, 3), ("cvrp", 20, 5), ("cvrp", 100, 3)])
@pytest.mark.parametrize("decode_type", ["multistart_greedy", "multistart_sampling"])
def test_tc_rollout_matches_fp32_rollout_and_oracle(env_name, N, B, decode_type):
    inst, pol, w, env, td, h = _setup(env_name, N, B, seed=1)
    a = pol(td.clone(), env, decode_type=decode_type, noise_seed=77, impl=_lib.IMPL_TC)
    b = pol(td.clone(), env, decode_type=decode_type, noise_seed=77, impl=_lib.IMPL_FP32)
    T = max(a["actions"].shape[1], b["actions"].shape[1])
    pad = lambda x: torch.nn.functional.pad(x, (0, T - x.shape[1]))
    same = (pad(a["actions"]) == pad(b["actions"])).all(1)
    assert same.float().mean().item() >= 0.97
    assert rel_close(a["reward"][same], b["reward"][same], LL_RTOL)
    assert rel_close(a["log_likelihood"][same], b["log_likelihood"][same], LL_RTOL)
    compare_rollout(env_name, inst, h.cpu(), w, False, a, decode_type, noise_seed=77)


# sizes on the edges of the tensor path: S = 8 (smallest multistart the path takes), node counts on both sides of the
# 16-node MMA granule and of the 32-node mask words, tails of <= 12 real nodes in the last chunk, and the largest
# tile (112 nodes; CVRP: 111 customers + depot, S = 111)
@pytest.mark.parametrize("env_name,N", [("tsp", 8), ("tsp", 13), ("tsp", 16), ("tsp", 17), ("tsp", 33), ("tsp", 96),
                                        ("tsp", 97), ("tsp", 112), ("cvrp", 8), ("cvrp", 15), ("cvrp", 16),
                                        ("cvrp", 111)])
def test_tc_rollout_boundary_sizes(env_name, N):
    B = 2
    inst, pol, w, env, td, h = _setup(env_name, N, B, seed=3)
    a = pol(td.clone(), env, decode_type="multistart_greedy", impl=_lib.IMPL_TC)
    b = pol(td.clone(), env, decode_type="multistart_greedy", impl=_lib.IMPL_FP32)
    assert a["actions"].shape[0] == B * N
    T = max(a["actions"].shape[1], b["actions"].shape[1])
    pad = lambda x: torch.nn.functional.pad(x, (0, T - x.shape[1]))
    same = (pad(a["actions"]) == pad(b["actions"])).all(1)
    assert same.float().mean().item() >= 0.95
    assert rel_close(a["reward"][same], b["reward"][same], LL_RTOL)
    assert rel_close(a["log_likelihood"][same], b["log_likelihood"][same], LL_RTOL)
    compare_rollout(env_name, inst, h.cpu(), w, False, a, "multistart_greedy")
    s = pol(td.clone(), env, decode_type="multistart_sampling", noise_seed=5, impl=_lib.IMPL_TC)
    compare_rollout(env_name, inst, h.cpu(), w, False, s, "multistart_sampling", noise_seed=5)


@pytest.mark.parametrize("env_name", ["tsp", "cvrp"])
def test_tc_evaluate_and_store_all_logp(env_name):
    N, B = 20, 4
    inst, pol, w, env, td, h = _setup(env_name, N, B, seed=2)
    out = pol(td.clone(), env, decode_type="multistart_sampling", noise_seed=3, impl=_lib.IMPL_TC,
              return_entropy=True)
    ev = pol(td.clone(), env, actions=out["actions"], num_starts=N, impl=_lib.IMPL_TC)
    assert torch.equal(ev["actions"], out["actions"])
    assert rel_close(ev["log_likelihood"], out["log_likelihood"], 1e-6)
    with torch.inference_mode():
        ref = O.rollout(env_name, inst, h.cpu(), w, actions=out["actions"].cpu(), num_starts=N,
                        use_graph_context=False, store_all_logp=True)
    assert rel_close(ev["log_likelihood"], ref["log_likelihood"], LL_RTOL)
    assert torch.allclose(out["entropy"].cpu(), ref["entropy"], rtol=1e-4, atol=1e-4)


def test_tc_pack_flags_fp16_range():
    inst, pol, w, env, td, h = _setup("tsp", 20, 2, seed=3)
    with torch.no_grad():
        pol.decoder.project_node_embeddings.weight.mul_(1e6)
    with pytest.raises(AssertionError, match="fp16 range"):
        pol(td.clone(), env, decode_type="multistart_greedy", impl=_lib.IMPL_TC)


def test_tc_full_size_properties():
    """BASELINE shape (TSP100, S=100) on a batch that needs several instances per CTA (B > #SMs)."""
    N, B = 100, 160
    inst, pol, w, env, td, h = _setup("tsp", N, B, seed=4)
    out = pol(td.clone(), env, decode_type="multistart_greedy", impl=_lib.IMPL_TC)
    a = out["actions"]
    assert a.shape == (B * N, N)
    assert bool((a.sort(1)[0] == torch.arange(N, device=DEV)[None]).all())
    assert rel_close(out["reward"], env.get_reward(td, a), 1e-6)
    ref = pol(td.clone(), env, decode_type="multistart_greedy", impl=_lib.IMPL_FP32)
    same = (a == ref["actions"]).all(1).float().mean().item()
    assert same >= 0.99, same
    ev = pol(td.clone(), env, actions=a, num_starts=N, impl=_lib.IMPL_FP32)
    assert rel_close(ev["log_likelihood"], out["log_likelihood"], LL_RTOL)


@pytest.mark.parametrize("decode_type", ["multistart_greedy", "multistart_sampling"])
def test_tc_full_size_properties_cvrp(decode_type):
    """c5 shape (CVRP100, S=100) with several instances per CTA: size-independent properties of the tours (every
    customer once, capacity never exceeded, reward = recomputed length) and the log-likelihood of the same actions
    re-evaluated teacher-forced by the fp32 kernels.  Sampling runs the exponential-race selection and, with rows
    returning to the depot at random, the feasible-set terms of the softmax reference."""
    N, B = 100, 160
    inst, pol, w, env, td, h = _setup("cvrp", N, B, seed=5)
    out = pol(td.clone(), env, decode_type=decode_type, noise_seed=11, impl=_lib.IMPL_TC)
    a = out["actions"]
    assert a.shape[0] == B * N
    # cvrp/env.py:157-185 (every customer exactly once, demand served per route <= capacity) and :146-155
    assert rel_close(out["reward"], env.get_reward(td, a, check_solution=True), 1e-6)
    ev = pol(td.clone(), env, actions=a, num_starts=N, impl=_lib.IMPL_FP32)
    assert rel_close(ev["log_likelihood"], out["log_likelihood"], LL_RTOL)
    assert rel_close(ev["reward"], out["reward"], 1e-6)


@pytest.mark.parametrize("env_name", ["tsp", "cvrp"])
@pytest.mark.parametrize("temperature", [0.1, 2.0])
def test_tc_rollout_temperature(env_name, temperature):
    """the constant log-sum-exp reference clip/temperature underflows at low temperature (every term of
    the sum flushes to zero); beyond clip/temperature = 40 the row maximum is exchanged instead.  Both decode modes
    against the oracle at temperature != 1."""
    import eam_rl4co_b200 as P
    from helpers import compare_rollout, decoder_weights, inst_to_env_td
    N, B = 20, 8
    torch.manual_seed(77)
    inst = O.generate_tsp(B, N) if env_name == "tsp" else O.generate_cvrp(B, N)
    torch.manual_seed(4321)
    pol = P.POMOPolicy(env_name, temperature=temperature).eval()
    w = decoder_weights(pol)
    pol = pol.to("cuda")
    env, td = inst_to_env_td(P, env_name, inst, "cuda")
    with torch.no_grad():
        h, _ = pol.encoder(td)
    for dt in ("multistart_greedy", "multistart_sampling"):
        out = pol(td.clone(), env, decode_type=dt, noise_seed=11, impl=P._lib.IMPL_TC)
        assert torch.isfinite(out["log_likelihood"]).all()
        with torch.inference_mode():
            ref = O.rollout(env_name, inst, h.cpu(), w, dt, use_graph_context=False, noise_seed=11,
                            temperature=temperature)
            ev = O.rollout(env_name, inst, h.cpu(), w, actions=out["actions"].cpu(), num_starts=N,
                           use_graph_context=False, temperature=temperature)
        assert rel_close(out["log_likelihood"], ev["log_likelihood"], 2e-5)
        T = min(ref["actions"].shape[1], out["actions"].shape[1])
        same = (ref["actions"][:, :T] == out["actions"].cpu()[:, :T]).all(1).float().mean().item()
        assert same >= 0.9, (dt, temperature, same)


@pytest.mark.parametrize("env_name", ["tsp", "cvrp"])
@pytest.mark.parametrize("N,B", [(20, 9), (50, 5), (100, 3)])
@pytest.mark.parametrize("decode_type", ["multistart_greedy", "multistart_sampling"])
def test_streaming_kernel_matches_oracle_and_resident_kernel(env_name, N, B, decode_type):
    """rollout_tc2_kernel (operand image v2 streamed from L2, two CTAs per SM; the default when B exceeds the SM count) forced
    at a small batch: against the oracle (tie-aware) and against the resident kernel; then evaluate mode on its actions."""
    import eam_rl4co_b200 as P
    from helpers import compare_rollout, decoder_weights, inst_to_env_td
    torch.manual_seed(31 + N)
    inst = O.generate_tsp(B, N) if env_name == "tsp" else O.generate_cvrp(B, N)
    torch.manual_seed(4321)
    pol = P.POMOPolicy(env_name).eval()
    w = decoder_weights(pol)
    pol = pol.to("cuda")
    env, td = inst_to_env_td(P, env_name, inst, "cuda")
    with torch.no_grad():
        h, _ = pol.encoder(td)
    pol.decoder.tc_layout = _lib.TC_LAYOUT_STREAM
    out = pol(td.clone(), env, decode_type=decode_type, noise_seed=5, impl=_lib.IMPL_TC)
    st = compare_rollout(env_name, inst, h.cpu(), w, False, out, decode_type, noise_seed=5, min_exact=0.97)
    ev = pol(td.clone(), env, actions=out["actions"], num_starts=N, impl=_lib.IMPL_TC)
    assert rel_close(ev["log_likelihood"], out["log_likelihood"], 2e-6)
    pol.decoder.tc_layout = _lib.TC_LAYOUT_RESIDENT
    ref = pol(td.clone(), env, decode_type=decode_type, noise_seed=5, impl=_lib.IMPL_TC)
    T = min(ref["actions"].shape[1], out["actions"].shape[1])
    same = (ref["actions"][:, :T] == out["actions"][:, :T]).all(1)
    assert same.float().mean().item() >= 0.97
    assert rel_close(out["log_likelihood"][same], ref["log_likelihood"][same], 2e-6)
    assert rel_close(out["reward"][same], ref["reward"][same], 1e-6)
    print(f"{env_name}{N} {decode_type}: exact vs oracle {st['exact']:.4f}, identical to resident kernel {same.float().mean().item():.4f}")
