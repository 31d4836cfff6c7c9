"""GPU: the tcgen05 encoder kernels (csrc/encoder_tc.cu) against the PyTorch modules they replace at inference
(eam_rl4co_b200/encoder.py; reference nn/graph/attnnet.py:16-103, nn/attention.py:66-136, nn/ops.py:32-56)."""
import ctypes as C

import pytest
import torch
import torch.nn.functional as F

import eam_rl4co_b200 as P
from eam_rl4co_b200 import _lib
from oracle import rollout_oracle as O
from helpers import inst_to_env_td

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _net(seed=0):
    torch.manual_seed(seed)
    pol = P.POMOPolicy("tsp").eval().to(DEV)
    return pol.encoder.net


def _lin64(lin, x):
    return F.linear(x.double(), lin.weight.double(), lin.bias.double())


def _norm64(norm, x):           # nn/ops.py:48-49 in float64
    nrm = norm.normalizer
    xc = x - x.mean(1, keepdim=True)
    var = (xc * xc).mean(1, keepdim=True)
    return xc * torch.rsqrt(var + nrm.eps) * nrm.weight.double() + nrm.bias.double()


def _rel(a, b):
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-12)).item()


@pytest.mark.parametrize("n,B", [(100, 3), (20, 5), (128, 2), (51, 150)])
@pytest.mark.parametrize("which", ["qkv", "ffn1", "out_norm", "ffn2_norm"])
def test_enc_linear_matches_torch(which, n, B):
    torch.backends.cuda.matmul.allow_tf32 = False
    net = _net()
    layer = net.layers[1]
    mha, norm1, mlp, norm2 = layer[0].module, layer[1], layer[2].module, layer[3]
    with torch.no_grad():
        for nrm in (norm1, norm2):      # non-trivial affine parameters
            nrm.normalizer.weight.uniform_(0.5, 1.5)
            nrm.normalizer.bias.uniform_(-0.5, 0.5)
    g = torch.Generator(device=DEV).manual_seed(n * 1000 + B)
    x = torch.randn(B, n, 128, device=DEV, generator=g)
    with torch.no_grad():
        if which == "qkv":
            got = net._linear(x, mha.Wqkv, _lib.ENC_EPI_PLAIN, 384)
            ref = _lin64(mha.Wqkv, x).float()
        elif which == "ffn1":
            got = net._linear(x, mlp.lins[0], _lib.ENC_EPI_RELU, 512)
            ref = F.relu(_lin64(mlp.lins[0], x)).float()
        elif which == "out_norm":
            res = torch.randn(B, n, 128, device=DEV, generator=g)
            got = net._linear(x, mha.out_proj, _lib.ENC_EPI_RES_NORM, 128, residual=res, norm=norm1)
            ref = _norm64(norm1, res.double() + _lin64(mha.out_proj, x)).float()
        else:
            f = torch.randn(B, n, 512, device=DEV, generator=g).relu()
            res = torch.randn(B, n, 128, device=DEV, generator=g)
            got = net._linear(f, mlp.lins[1], _lib.ENC_EPI_RES_NORM, 128, residual=res, norm=norm2)
            ref = _norm64(norm2, res.double() + _lin64(mlp.lins[1], f)).float()
    assert got.shape == ref.shape
    err = _rel(got, ref)
    print(f"{which} n={n} B={B}: rel err {err:.2e}")
    assert err < 2e-6, err           # fp32-level agreement (tolerance written here: 2e-6 of the output range)


@pytest.mark.parametrize("n,B", [(100, 3), (21, 4), (128, 2), (51, 300), (112, 2), (101, 150), (113, 2), (5, 3)])
def test_enc_attention_matches_sdpa(n, B):
    g = torch.Generator(device=DEV).manual_seed(n + B)
    qkv = torch.randn(B, n, 384, device=DEV, generator=g) * 1.5
    out = torch.empty(B, n, 128, device=DEV)
    L = _lib.lib()
    _lib.check(L.rl4co_enc_attention(_lib.ptr(qkv), _lib.ptr(out), C.c_int64(B), n, _lib.stream_ptr()), "attn")
    q, k, v = qkv.double().view(B, n, 3, 8, 16).permute(2, 0, 3, 1, 4).unbind(0)
    ref = F.scaled_dot_product_attention(q, k, v).transpose(1, 2).reshape(B, n, 128).float()
    err = _rel(out, ref)
    print(f"attention n={n} B={B}: rel err {err:.2e}")
    assert err < 2e-6, err


@pytest.mark.parametrize("env_name,N,B", [("tsp", 100, 4), ("tsp", 20, 7), ("cvrp", 50, 3), ("cvrp", 100, 160)])
def test_encoder_kernels_match_torch_encoder(env_name, N, B):
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(1234)
    inst = O.generate_tsp(B, N) if env_name == "tsp" else O.generate_cvrp(B, N)
    torch.manual_seed(4321)
    pol = P.POMOPolicy(env_name).eval().to(DEV)
    env, td = inst_to_env_td(P, env_name, inst, DEV)
    net = pol.encoder.net
    with torch.no_grad():
        assert net.kernels_apply(pol.encoder.init_embedding(td))
        h_k, _ = pol.encoder(td)
        net.use_kernels = False
        h_t, _ = pol.encoder(td)
        net.use_kernels = True
    with torch.enable_grad():           # training keeps the differentiable modules
        assert not net.kernels_apply(pol.encoder.init_embedding(td))
    err = (h_k - h_t).abs().max().item()
    print(f"{env_name}{N} B={B}: max |h_kernels - h_torch| = {err:.2e} (|h| max {h_t.abs().max().item():.2f})")
    assert err < 5e-5, err           # six layers of fp32-level differences through the instance norms


def test_rollout_with_kernel_encoder_matches_torch_encoder():
    torch.backends.cuda.matmul.allow_tf32 = False
    N, B = 100, 8
    torch.manual_seed(7)
    inst = O.generate_tsp(B, N)
    torch.manual_seed(4321)
    pol = P.POMOPolicy("tsp").eval().to(DEV)
    env, td = inst_to_env_td(P, "tsp", inst, DEV)
    a = pol(td.clone(), env, decode_type="multistart_greedy")
    pol.encoder.net.use_kernels = False
    b = pol(td.clone(), env, decode_type="multistart_greedy")
    same = (a["actions"] == b["actions"]).all(1).float().mean().item()
    print(f"identical tours with kernel vs torch encoder: {same:.4f}")
    assert same >= 0.97
    best_a = a["reward"].view(N, B).max(0)[0]
    best_b = b["reward"].view(N, B).max(0)[0]
    assert torch.allclose(best_a, best_b, rtol=2e-3)


# ---- pinned to the REFERENCE encoder (tests/golden/enc_*.npz, oracle/make_golden.py encoder_fixtures) -----------
from helpers import encoder_golden_files, load_encoder_golden  # noqa: E402


@pytest.mark.parametrize("path", encoder_golden_files(), ids=lambda p: p.split("/")[-1])
def test_encoder_kernels_match_reference_fixture(path):
    """enc_* kernels (instance norm for POMO, the eval-BatchNorm affine epilogue for the AttentionModel) against
    the h the unmodified reference produced; then the full policy on top against the reference's greedy tours."""
    torch.backends.cuda.matmul.allow_tf32 = False
    g = load_encoder_golden(path, P, DEV)
    pol, env_name = g["policy"], g["env"]
    env, td = inst_to_env_td(P, env_name, g["inst"], DEV)
    with torch.no_grad():
        assert pol.encoder.net.kernels_apply(pol.encoder.init_embedding(td)), "kernel encoder not taken"
        h, _ = pol.encoder(td)
    err = (h.cpu() - g["h"]).abs().max().item()
    print(f"{g['name']}: max |h_kernels - h_reference| = {err:.2e} (|h| max {g['h'].abs().max().item():.2f})")
    assert err < 5e-5, err
    dt = "multistart_greedy" if g["pomo"] else "greedy"
    out = pol(td.clone(), env, decode_type=dt)
    ref_a = torch.from_numpy(g["z"][f"{dt}.actions"]).long()
    T = max(ref_a.shape[1], out["actions"].shape[1])
    pad = lambda a: F.pad(a, (0, T - a.shape[1]))
    same = (pad(out["actions"].cpu()) == pad(ref_a)).all(1)
    print(f"   identical greedy tours end to end: {same.float().mean().item():.4f}")
    assert same.float().mean().item() >= 0.9
    r_ref = torch.from_numpy(g["z"][f"{dt}.reward"])
    ll_ref = torch.from_numpy(g["z"][f"{dt}.log_likelihood"])
    assert torch.allclose(out["reward"].cpu()[same], r_ref[same], rtol=1e-5, atol=1e-5)
    assert torch.allclose(out["log_likelihood"].cpu()[same], ll_ref[same], rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("n,B", [(100, 3), (20, 5), (128, 2), (51, 300)])
@pytest.mark.parametrize("norm", ["instance", "batch"])
def test_fused_ffn_matches_fp64(n, B, norm):
    """rl4co_enc_ffn: Norm(x + W2 relu(W1 x + b1) + b2) with the hidden activations kept in TMEM, against float64."""
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(5)
    pol = (P.POMOPolicy("tsp") if norm == "instance" else P.AttentionModelPolicy(env_name="tsp")).eval().to(DEV)
    net = pol.encoder.net
    layer = net.layers[1]
    mlp, norm2 = layer[2].module, layer[3]
    with torch.no_grad():
        norm2.normalizer.weight.uniform_(0.5, 1.5)
        norm2.normalizer.bias.uniform_(-0.5, 0.5)
        if norm == "batch":
            norm2.normalizer.running_mean.normal_(0, 0.3)
            norm2.normalizer.running_var.uniform_(0.5, 1.5)
    g = torch.Generator(device=DEV).manual_seed(n * 7 + B)
    x = torch.randn(B, n, 128, device=DEV, generator=g)
    with torch.no_grad():
        got = net._ffn(x, mlp, norm2)
        f = F.relu(_lin64(mlp.lins[0], x))
        pre = x.double() + F.linear(f, mlp.lins[1].weight.double(), mlp.lins[1].bias.double())
        if norm == "instance":
            ref = _norm64(norm2, pre)
        else:
            bn = norm2.normalizer
            ref = (pre - bn.running_mean.double()) * torch.rsqrt(bn.running_var.double() + bn.eps) * bn.weight.double() \
                + bn.bias.double()
    err = _rel(got, ref.float())
    print(f"fused ffn {norm} n={n} B={B}: rel err {err:.2e}")
    assert err < 2e-6, err


@pytest.mark.parametrize("n,B", [(100, 3), (20, 5), (112, 2), (51, 300), (101, 4)])
@pytest.mark.parametrize("norm", ["instance", "batch"])
def test_fused_attention_block_matches_fp64(n, B, norm):
    """rl4co_enc_attn_block: Norm(x + MHA(x) Wo^T + bo) with q, k, v^T and the heads kept on chip, against float64."""
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(6)
    pol = (P.POMOPolicy("tsp") if norm == "instance" else P.AttentionModelPolicy(env_name="tsp")).eval().to(DEV)
    net = pol.encoder.net
    layer = net.layers[0]
    mha, norm1 = layer[0].module, layer[1]
    with torch.no_grad():
        norm1.normalizer.weight.uniform_(0.5, 1.5)
        norm1.normalizer.bias.uniform_(-0.5, 0.5)
        mha.Wqkv.bias.uniform_(-0.3, 0.3)
        mha.out_proj.bias.uniform_(-0.3, 0.3)
        if norm == "batch":
            norm1.normalizer.running_mean.normal_(0, 0.3)
            norm1.normalizer.running_var.uniform_(0.5, 1.5)
    g = torch.Generator(device=DEV).manual_seed(n * 3 + B)
    x = torch.randn(B, n, 128, device=DEV, generator=g)
    with torch.no_grad():
        got = net._attn_block(x, mha, norm1)
        qkv = _lin64(mha.Wqkv, x)
        q, k, v = qkv.view(B, n, 3, 8, 16).permute(2, 0, 3, 1, 4).unbind(0)
        a = F.scaled_dot_product_attention(q, k, v).transpose(1, 2).reshape(B, n, 128)
        pre = x.double() + F.linear(a, mha.out_proj.weight.double(), mha.out_proj.bias.double())
        if norm == "instance":
            ref = _norm64(norm1, pre)
        else:
            bn = norm1.normalizer
            ref = (pre - bn.running_mean.double()) * torch.rsqrt(bn.running_var.double() + bn.eps) * bn.weight.double() \
                + bn.bias.double()
    err = _rel(got, ref.float())
    print(f"fused attention block {norm} n={n} B={B}: rel err {err:.2e}")
    assert err < 2e-6, err
