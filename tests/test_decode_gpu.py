"""GPU: decoder step (pointer attention logits) and selection kernels vs the oracle."""
import pytest
import torch

import eam_rl4co_b200 as P
from oracle import rollout_oracle as O
from helpers import decoder_weights, inst_to_env_td, golden_files, load_golden

pytestmark = pytest.mark.gpu
DEV = "cuda"
LOGIT_ATOL = 2e-5   # fp32, different summation order


def _policy(env_name, pomo):
    torch.manual_seed(4321)
    pol = (P.POMOPolicy(env_name) if pomo else P.AttentionModelPolicy(env_name=env_name)).eval()
    return pol


@pytest.mark.parametrize("env_name", ["tsp", "cvrp"])
@pytest.mark.parametrize("pomo,N,B", [(False, 20, 16), (True, 50, 4), (True, 100, 2)])
def test_decode_step_logits_match_oracle_along_a_rollout(env_name, pomo, N, B):
    torch.manual_seed(N + B)
    inst = O.generate_tsp(B, N) if env_name == "tsp" else O.generate_cvrp(B, N)
    pol = _policy(env_name, pomo)
    w = decoder_weights(pol)
    pol = pol.to(DEV)
    env, td = inst_to_env_td(P, env_name, inst, DEV)
    with torch.no_grad():
        h, _ = pol.encoder(td)
    h_cpu = h.cpu()
    S = N if pomo else 0
    ref = O.tsp_reset(inst["locs"]) if env_name == "tsp" else O.cvrp_reset(
        inst["locs"], inst["depot"], inst["demand"])
    step_fn = O.tsp_step if env_name == "tsp" else O.cvrp_step
    if S:
        start = O.select_start_nodes(env_name, B, S, N)
        ref = {k: O.batchify(v, S) for k, v in ref.items()}
        ref["action"] = start
        ref = step_fn(ref)
        td = P.batchify(td, S)
        td.set("action", start.to(DEV))
        td = env.step(td)["next"]
    cache_ref = O.precompute_cache(h_cpu, w, not pomo)
    _, _, cache = pol.decoder.pre_decoder_hook(td, env, h, S)
    steps = 0
    while not ref["done"].all() and steps < 12:
        lg_ref, mask_ref = O.decoder_step(env_name, ref, cache_ref, w, S)
        lg, mask = pol.decoder(td, cache, S)
        assert torch.equal(mask.cpu(), mask_ref)
        assert torch.allclose(lg.cpu(), lg_ref, atol=LOGIT_ATOL, rtol=1e-5), (lg.cpu() - lg_ref).abs().max()
        a = torch.multinomial(mask_ref.float(), 1).squeeze(1)
        ref["action"] = a
        ref = step_fn(ref)
        td.set("action", a.to(DEV))
        td = env.step(td)["next"]
        steps += 1


@pytest.mark.parametrize("n", [1, 5, 20, 33, 101, 128])
def test_select_matches_process_logits_and_strategies(n):
    torch.manual_seed(n)
    rows, B = 96, 32
    logits = torch.randn(rows, n) * 3
    mask = torch.rand(rows, n) > 0.4
    mask[:, torch.randint(0, n, (1,))] = True
    mask[torch.arange(rows), torch.randint(0, n, (rows,))] = True
    lp_ref = O.process_logits(logits.clone(), mask, 1.3, 10.0)
    for mode, name in ((0, "greedy"), (1, "sampling"), (2, "evaluate")):
        strat = P.get_decoding_strategy(name, temperature=1.3, tanh_clipping=10.0, store_all_logp=False,
                                        noise_seed=77, instance_offset=5)
        strat._B = B
        given = torch.multinomial(mask.float(), 1).squeeze(1)
        td = P.TensorDict({}, batch_size=[rows])
        strat.actions = [None] * 4          # pretend we are at column 4
        strat.logprobs = [None] * 4
        td = strat.step(logits.to(DEV), mask.to(DEV), td, action=given.to(DEV) if mode == 2 else None)
        act = td["action"].cpu()
        lp_sel = strat.logprobs[-1].cpu()
        if mode == 0:
            exp = lp_ref.argmax(-1)
        elif mode == 1:
            r = torch.arange(rows)
            g = O.noise_gumbel(77, r % B + 5, r // B, 4, n)
            exp = (lp_ref + g).argmax(-1)
        else:
            exp = given
        mism = (act != exp).nonzero().flatten()
        for i in mism.tolist():   # only exact ties / 1-ulp noise differences may differ
            assert abs(lp_ref[i, act[i]] - lp_ref[i, exp[i]]) < 1e-4 or mode == 1
        assert len(mism) <= (2 if mode == 1 else 0)
        assert torch.allclose(lp_sel, lp_ref.gather(1, act[:, None]).squeeze(1), atol=2e-6, rtol=1e-6)
    strat = P.get_decoding_strategy("greedy", temperature=1.3, tanh_clipping=10.0, store_all_logp=True)
    td = strat.step(logits.to(DEV), mask.to(DEV), P.TensorDict({}, batch_size=[rows]))
    full = strat.logprobs[-1].cpu()
    assert torch.equal(torch.isinf(full), torch.isinf(lp_ref))
    fin = ~torch.isinf(lp_ref)
    assert torch.allclose(full[fin], lp_ref[fin], atol=2e-6, rtol=1e-6)


def test_counter_noise_matches_oracle_generator():
    import ctypes as C
    from eam_rl4co_b200 import _lib
    rows, n, B = 40, 101, 8
    out = torch.empty(rows, n, device=DEV)
    rc = _lib.lib().rl4co_noise_gumbel(C.c_uint32(2026), C.c_int64(3), C.c_int64(B), C.c_int32(9),
                                       _lib.ptr(out), C.c_int64(rows), C.c_int32(n), _lib.stream_ptr())
    assert rc == 0
    r = torch.arange(rows)
    g = O.noise_gumbel(2026, r % B + 3, r // B, 9, n)
    assert torch.allclose(out.cpu(), g, atol=1e-5, rtol=1e-5)     # same uniforms; logf differs by ulps


def test_infeasible_and_nan_raise_reference_messages():
    rows, n = 4, 10
    logits = torch.randn(rows, n, device=DEV)
    mask = torch.ones(rows, n, dtype=torch.bool, device=DEV)
    mask[1] = False
    strat = P.get_decoding_strategy("greedy", tanh_clipping=10.0)
    with pytest.raises(AssertionError, match="infeasible action selected"):
        strat.step(logits, mask, P.TensorDict({}, batch_size=[rows]))


@pytest.mark.parametrize("path", golden_files(), ids=lambda p: p.split("/")[-1])
def test_decode_step_vs_reference_fixture_logits(path):
    """First decoder steps of the UNMODIFIED reference (fixtures) vs the CUDA decoder."""
    g = load_golden(path)
    z = g["z"]
    env_name = g["env"]
    pol = (P.AttentionModelPolicy(env_name=env_name) if g["use_graph_context"] else P.POMOPolicy(env_name)).eval()
    pol.decoder.load_state_dict(g["w"])
    pol = pol.to(DEV)
    inst = g["inst"]
    N = inst["locs"].shape[1]
    for dt in ("greedy", "multistart_greedy"):
        if f"{dt}.logits.0" not in z.files:
            continue
        env, td = inst_to_env_td(P, env_name, inst, DEV)
        S = N if dt.startswith("multistart") else 0
        actions = torch.from_numpy(z[f"{dt}.actions"]).long().to(DEV)
        col = 0
        if S:
            td = P.batchify(td, S)
            td.set("action", actions[:, 0])
            td = env.step(td)["next"]
            col = 1
        _, _, cache = pol.decoder.pre_decoder_hook(td, env, g["h"].to(DEV), S)
        for t in range(3):
            lg, mask = pol.decoder(td, cache, S)
            assert torch.equal(mask.cpu(), torch.from_numpy(z[f"{dt}.mask.{t}"]))
            ref = torch.from_numpy(z[f"{dt}.logits.{t}"])
            assert torch.allclose(lg.cpu(), ref, atol=LOGIT_ATOL, rtol=1e-5), (lg.cpu() - ref).abs().max()
            td.set("action", actions[:, col + t])
            td = env.step(td)["next"]
