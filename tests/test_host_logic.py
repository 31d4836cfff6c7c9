"""CPU: host-side logic that needs no GPU."""
import pytest
import torch

import eam_rl4co_b200 as P
from oracle import rollout_oracle as O


def test_batchify_unbatchify_roundtrip():
    # reference tests/test_utils.py:12-30
    for a, b in [(3, 2), (5, 1), (4, 7)]:
        x = torch.rand(a, b, 6)
        assert torch.equal(P.unbatchify(P.batchify(x, 5), 5)[:, 0], x)
        assert torch.equal(P.batchify(x, 5), O.batchify(x, 5))
    td = P.TensorDict({"x": torch.arange(6).view(3, 2)}, batch_size=[3])
    tb = P.batchify(td, 4)
    assert tb.batch_size[0] == 12 and torch.equal(tb["x"][3:6], td["x"])
    assert torch.equal(P.unbatchify(tb, 4)["x"][:, 2], td["x"])


def test_generators_follow_reference_spec():
    torch.manual_seed(1234)
    a = P.TSPGenerator(num_loc=20)(4)
    torch.manual_seed(1234)
    b = O.generate_tsp(4, 20)
    assert torch.equal(a["locs"], b["locs"])
    torch.manual_seed(1234)
    a = P.CVRPGenerator(num_loc=100)(4)
    torch.manual_seed(1234)
    b = O.generate_cvrp(4, 100)
    for k in ("locs", "depot", "demand"):
        assert torch.equal(a[k], b[k])
    d = a["demand"] * 50.0
    assert bool(((d.round() - d).abs() < 1e-4).all()) and d.min() >= 1 - 1e-4 and d.max() <= 9 + 1e-4


def test_parameter_names_match_reference_state_dict():
    pol = P.AttentionModelPolicy(env_name="tsp")
    keys = set(pol.state_dict().keys())
    for k in ["decoder.context_embedding.project_context.weight", "decoder.context_embedding.W_placeholder",
              "decoder.pointer.project_out.weight", "decoder.project_node_embeddings.weight",
              "decoder.project_fixed_context.weight", "encoder.init_embedding.init_embed.weight",
              "encoder.net.layers.0.0.module.Wqkv.weight", "encoder.net.layers.2.3.normalizer.weight"]:
        assert k in keys, k
    pol = P.POMOPolicy("cvrp")
    assert pol.decoder.context_embedding.project_context.weight.shape == (128, 129)
    assert len(pol.encoder.net.layers) == 6 and not pol.decoder.use_graph_context


def test_unsupported_options_raise_loudly():
    with pytest.raises(NotImplementedError):
        P.get_decoding_strategy("beam_search")
    with pytest.raises(NotImplementedError):
        P.get_decoding_strategy("sampling", top_k=5)
    with pytest.raises(NotImplementedError):
        P.get_decoding_strategy("greedy", mask_logits=False)
    with pytest.raises(NotImplementedError):
        P.AttentionModelDecoder(embed_dim=64)
    with pytest.raises(NotImplementedError):
        P.AttentionModelDecoder(env_name="op")
    s = P.get_decoding_strategy("multistart_greedy", tanh_clipping=10.0)
    assert s.multistart and s.mode == 0
    s = P.get_decoding_strategy("evaluate", num_starts=5)
    assert s.multistart and s.mode == 2


def test_folded_tables_equal_reference_cache_algebra():
    """pre_decoder_hook's folded tables are algebraically the reference's cache (fp32 CPU check)."""
    torch.manual_seed(0)
    for env_name in ("tsp", "cvrp"):
        pol = P.AttentionModelPolicy(env_name=env_name).eval()
        dec = pol.decoder
        h = torch.randn(3, 11, 128)
        c = dec._precompute_cache(h)
        w = {k: v for k, v in dec.state_dict().items()}
        ref = O.precompute_cache(h, w, True)
        assert torch.allclose(c.glimpse_key, ref["glimpse_key"], atol=1e-5)
        assert torch.allclose(c.glimpse_val, ref["glimpse_val"], atol=1e-5)
        heads = torch.randn(3, 128)
        glimpse = torch.nn.functional.linear(heads, w["pointer.project_out.weight"])
        lg_ref = torch.einsum("be,bne->bn", glimpse, ref["logit_key"])
        lg = torch.einsum("be,bne->bn", heads, c.logit_key_folded)
        assert torch.allclose(lg, lg_ref, atol=1e-4, rtol=1e-4)
        assert torch.allclose(c.logit_key, ref["logit_key"], atol=1e-5)


def test_cache_stamp_follows_optimizer_steps_and_split_k_linear_on_cpu():
    """Host logic of two training-path pieces: (1) every torch.optim step moves the stamp the packed-weight caches are keyed
    by (fused optimisers do not move a parameter's `_version`); (2) the split-K weight gradient of the encoder's Linear
    layers equals nn.Linear's autograd (pure torch ops, checked here on CPU)."""
    import torch
    from eam_rl4co_b200 import encoder as E
    w = torch.nn.Parameter(torch.randn(8, 8))
    before = E.WeightPacker.stamp(w)
    opt = torch.optim.SGD([w], lr=0.1)
    w.grad = torch.ones_like(w)
    opt.step()
    assert E.WeightPacker.stamp(w) != before and E.WeightPacker.stamp(w)[3] == before[3] + 1
    assert E._split_chunks(64 * 101) == 32 and E._split_chunks(101) == 1 and E._split_chunks(512) == 4
    torch.manual_seed(0)
    lin = torch.nn.Linear(16, 24)
    x = torch.randn(8, 64, 16, requires_grad=True)
    cot = torch.randn(8, 64, 24)
    (E._SplitKLinearFn.apply(x, lin.weight, lin.bias) * cot).sum().backward()
    got = (x.grad.clone(), lin.weight.grad.clone(), lin.bias.grad.clone())
    x.grad = lin.weight.grad = lin.bias.grad = None
    (lin(x) * cot).sum().backward()
    for a, b in zip(got, (x.grad, lin.weight.grad, lin.bias.grad)):
        assert torch.allclose(a, b, rtol=1e-5, atol=1e-5)
