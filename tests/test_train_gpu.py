"""GPU: the POMO / REINFORCE training step (eam_rl4co_b200/train.py): states replayed through
the env kernels, log-likelihood of the kernel-sampled actions re-evaluated with autograd, loss and gradients against
autograd through the oracle's sequential evaluate (reference: pomo/model.py:89-114, reinforce.py:73-115)."""
import pytest
import torch

import eam_rl4co_b200 as P
from eam_rl4co_b200 import train
from oracle import rollout_oracle as O
from helpers import LL_RTOL, inst_to_env_td, rel_close
from test_train_host import _oracle_states

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _setup(env_name, B, N, seed=0):
    torch.manual_seed(100 + seed)
    inst = O.generate_tsp(B, N) if env_name == "tsp" else O.generate_cvrp(B, N)
    torch.manual_seed(200 + seed)
    pol = P.POMOPolicy(env_name).to(DEV)
    env, td = inst_to_env_td(P, env_name, inst, DEV)
    return inst, pol, env, td


@pytest.mark.parametrize("env_name,B,N", [("tsp", 4, 20), ("cvrp", 4, 20), ("tsp", 2, 100), ("cvrp", 2, 50)])
def test_teacher_forced_states_are_bit_exact(env_name, B, N):
    inst, pol, env, td = _setup(env_name, B, N)
    with torch.no_grad():
        out = pol(td.clone(), env, decode_type="multistart_sampling", noise_seed=3)
    S = out["actions"].shape[0] // B
    st = train.teacher_forced_states(env, td.clone(), out["actions"], S)
    ref = _oracle_states(env_name, inst, out["actions"].cpu(), S)
    for k in ("first", "cur", "mask", "taken"):
        assert torch.equal(st[k].cpu(), ref[k]), k
    if env_name == "cvrp":
        assert torch.equal(st["rem"].cpu(), ref["rem"])


@pytest.mark.parametrize("env_name,B,N", [("tsp", 4, 20), ("cvrp", 3, 20), ("tsp", 2, 100), ("cvrp", 2, 100)])
def test_evaluate_log_likelihood_matches_the_sampling_kernel(env_name, B, N):
    torch.backends.cuda.matmul.allow_tf32 = False
    inst, pol, env, td = _setup(env_name, B, N, seed=1)
    with torch.no_grad():
        out = pol(td.clone(), env, decode_type="multistart_sampling", noise_seed=4)
    S = out["actions"].shape[0] // B
    ll = train.evaluate_log_likelihood(pol, td.clone(), env, out["actions"], S)
    assert ll.requires_grad and ll.shape == out["log_likelihood"].shape
    assert rel_close(ll.detach(), out["log_likelihood"], LL_RTOL)


@pytest.mark.parametrize("env_name", ["tsp", "cvrp"])
def test_train_step_gradients_match_oracle_autograd(env_name):
    torch.backends.cuda.matmul.allow_tf32 = False
    B, N = 3, 12
    inst, pol, env, td = _setup(env_name, B, N, seed=2)
    stats = train.pomo_train_step(pol, env, td.clone(), optimizer=None, noise_seed=6, max_grad_norm=None)
    assert stats["ll_gap"] < 1e-4 and stats["loss"] == stats["loss"]
    # the same step on the CPU: actions from the kernel, autograd through the oracle (encoder = the torch modules)
    with torch.no_grad():
        out = pol(td.clone(), env, decode_type="multistart_sampling", noise_seed=6)
    S = out["actions"].shape[0] // B
    cpu = P.POMOPolicy(env_name)
    cpu.load_state_dict(pol.state_dict())
    enc_in = dict(inst)
    if env_name == "cvrp":
        enc_in = {"locs": torch.cat((inst["depot"][:, None, :], inst["locs"]), 1), "demand": inst["demand"]}
    h, _ = cpu.encoder(enc_in)
    w = dict(cpu.decoder.named_parameters())
    ev = O.rollout(env_name, inst, h, w, actions=out["actions"].cpu(), num_starts=S, use_graph_context=False)
    loss, _ = O.reinforce_shared_baseline_loss(out["reward"].cpu(), ev["log_likelihood"], S)
    loss.backward()
    assert abs(loss.item() - stats["loss"]) <= 1e-4 * max(1.0, abs(loss.item()))
    checked = 0
    gmax = max(p.grad.abs().max().item() for p in cpu.parameters() if p.grad is not None)
    for (k, p_gpu), (_, p_cpu) in zip(pol.named_parameters(), cpu.named_parameters()):
        if p_cpu.grad is None:
            continue
        g, r = p_gpu.grad.cpu(), p_cpu.grad
        # fp32 GPU vs fp32 CPU reductions; parameters whose exact gradient is zero (biases in front of an
        # instance norm) hold rounding noise only, hence the floor relative to the largest gradient
        scale = max(r.abs().max().item(), 1e-3 * gmax)
        assert (g - r).abs().max().item() / scale < 2e-3, k
        checked += 1
    assert checked >= 20


def test_train_step_with_optimizer_improves_the_sampled_objective():
    inst, pol, env, td = _setup("tsp", 16, 20, seed=3)
    opt = torch.optim.Adam(pol.parameters(), lr=1e-3)
    before = [p.detach().clone() for p in pol.parameters()]
    hist = [train.pomo_train_step(pol, env, td.clone(), optimizer=opt, noise_seed=i) for i in range(12)]
    assert all(h["loss"] == h["loss"] and h["ll_gap"] < 1e-3 for h in hist)
    assert any(not torch.equal(a, b) for a, b in zip(before, pol.parameters()))
    # the encoder kernels pick up the updated weights (packed tiles are cached per parameter version)
    with torch.no_grad():
        net = pol.encoder.net
        hk, _ = pol.encoder(td)
        net.use_kernels = False
        ht, _ = pol.encoder(td)
        net.use_kernels = True
    assert (hk - ht).abs().max().item() < 1e-4
    assert sum(h["reward"] for h in hist[-4:]) / 4 > sum(h["reward"] for h in hist[:4]) / 4 - 0.05


@pytest.mark.parametrize("env_name,B,N,clip,temp", [("tsp", 3, 20, 10.0, 1.0), ("cvrp", 3, 20, 10.0, 1.0),
                                                    ("tsp", 2, 100, 10.0, 1.0), ("cvrp", 2, 100, 10.0, 1.0),
                                                    ("tsp", 2, 33, 0.0, 1.0), ("cvrp", 2, 20, 10.0, 0.7)])
def test_fused_evaluate_kernels_match_eager_autograd(env_name, B, N, clip, temp):
    """rl4co_eval_ll_fwd / rl4co_eval_ll_bwd against the plain-PyTorch form of the same rows (decoding.py:452-465,
    attention.py:282-328, decoding.py:140-190): log-probabilities and the gradients w.r.t. the embeddings and every
    decoder weight, under a random cotangent."""
    torch.backends.cuda.matmul.allow_tf32 = False
    inst, pol, env, td = _setup(env_name, B, N, seed=5)
    with torch.no_grad():
        out = pol(td.clone(), env, decode_type="multistart_sampling", noise_seed=9)
        h0, _ = pol.encoder(td)
    S = out["actions"].shape[0] // B
    states = train.teacher_forced_states(env, td.clone(), out["actions"], S)
    torch.manual_seed(7)
    cot0 = torch.randn(B, S, states["cur"].shape[2], device=DEV)
    for cot_scale in (1.0, 1.5e-4):                               # O(1), and the advantage / (B S) of a c5 training step
        cot = cot0 * cot_scale
        res = {}
        for fused in (True, False):
            h = h0.clone().requires_grad_(True)
            for p in pol.decoder.parameters():
                p.grad = None
            logp = train.batched_log_likelihood(pol.decoder, h, states, clip, temp, fused=fused)
            (logp * cot).sum().backward()
            grads = {k: p.grad.clone() for k, p in pol.decoder.named_parameters() if p.grad is not None}
            res[fused] = (logp.detach(), h.grad.clone(), grads)
        lf, hf, gf = res[True]
        le, he, ge = res[False]
        assert torch.isfinite(lf).all()
        assert (lf - le).abs().max().item() < 2e-5 * max(1.0, le.abs().max().item())
        tol = lambda a, b: (a - b).abs().max().item() / max(b.abs().max().item(), 1e-6 * cot_scale)
        assert tol(hf, he) < 2e-4
        assert set(gf) == set(ge) and len(gf) >= 3
        for k in ge:
            assert tol(gf[k], ge[k]) < 2e-4, (k, cot_scale)


def test_fused_evaluate_rejects_unsupported_sizes():
    from eam_rl4co_b200 import _lib
    import ctypes as C
    L = _lib.lib()
    assert L.rl4co_eval_ll_max_nodes() == 128
    z = torch.zeros(8, device=DEV)
    p = C.c_void_p(z.data_ptr())
    rc = L.rl4co_eval_ll_fwd(p, p, p, p, p, p, p, p, p, p, p, C.c_int64(1), C.c_int64(1), C.c_int32(129), C.c_float(10.0),
                             C.c_float(1.0), None)
    assert rc != 0 and b"n_nodes" in L.rl4co_last_error()
    h = torch.zeros(1, 130, 128, device=DEV)
    assert not train.fused_evaluate_supported(h, 130) and train.fused_evaluate_supported(h, 128)


def _eager_rows(q, gk, gv, lkf, mask, taken, clip, temp):
    """fp64 PyTorch statement of one evaluate row batch (attention.py:308-328, decoding.py:140-190,452-465)."""
    B, R, E = q.shape
    n = gk.shape[1]
    H = 8
    hd = lambda x, L: x.reshape(B, L, H, E // H).transpose(1, 2)
    s = torch.matmul(hd(q, R), hd(gk, n).transpose(-1, -2)) / 4.0
    s = s.masked_fill(~mask[:, None], float("-inf"))
    o = torch.matmul(torch.softmax(s, -1), hd(gv, n)).transpose(1, 2).reshape(B, R, E)
    u = torch.matmul(o, lkf.transpose(1, 2)) / (E ** 0.5)
    z = (torch.tanh(u) * clip if clip > 0 else u).masked_fill(~mask, float("-inf")) / temp
    return torch.log_softmax(z, -1).gather(-1, taken[..., None])[..., 0]


@pytest.mark.parametrize("B,R,n,clip,temp", [(3, 37, 101, 10.0, 1.0), (2, 5, 2, 10.0, 1.0), (1, 1, 112, 0.0, 0.5), (2, 131, 128, 10.0, 1.0),
                                              (5, 130, 64, 10.0, 2.0), (70, 9, 21, 10.0, 1.0)])
def test_fused_evaluate_rows_against_fp64(B, R, n, clip, temp):
    _fused_rows_against_fp64(B, R, n, clip, temp, 1.0)


@pytest.mark.parametrize("cot_scale", [1e-4, 1e-8, 1e-20, 3e4])
def test_fused_evaluate_gradients_at_any_loss_scale(cot_scale):
    """dlogp of a training step is advantage / (B S) ~ 1e-4 and below: the backward's fp16 hi / lo operands must not lose the
    gradient to fp16's exponent range (rl4co_eval_ll_bwd rescales by a power of two)."""
    _fused_rows_against_fp64(3, 37, 101, 10.0, 1.0, cot_scale)


def _fused_rows_against_fp64(B, R, n, clip, temp, cot_scale):
    """Ragged row counts (not a multiple of the four-row groups or of the CTA split), the smallest and largest node
    counts, rows with a single feasible node, more instances than SMs / 2."""
    g = torch.Generator().manual_seed(B * 1000 + R * 10 + n)
    rnd = lambda *s: torch.randn(*s, generator=g)
    q, gk, gv, lkf = rnd(B, R, 128), rnd(B, n, 128), rnd(B, n, 128), rnd(B, n, 128) * 0.3
    mask = torch.rand(B, R, n, generator=g) < 0.4
    taken = torch.randint(0, n, (B, R), generator=g)
    mask[:, ::3] = False                                          # every third row: only the taken node is feasible
    mask.scatter_(2, taken[..., None], True)
    cot = rnd(B, R) * cot_scale
    args64 = [t.double().to(DEV).requires_grad_(True) for t in (q, gk, gv, lkf)]
    ref = _eager_rows(*args64, mask.to(DEV), taken.to(DEV), clip, temp)
    (ref * cot.double().to(DEV)).sum().backward()
    args32 = [t.to(DEV).requires_grad_(True) for t in (q, gk, gv, lkf)]
    got = train._FusedEvaluate.apply(*args32, mask.to(DEV), taken.to(DEV), clip, temp)
    (got * cot.to(DEV)).sum().backward()
    assert (got.double() - ref).abs().max().item() < 2e-5 * max(1.0, ref.abs().max().item())
    assert (got[:, ::3].abs() < 1e-6).all()                       # a single feasible node has probability one
    for a32, a64, name in zip(args32, args64, ("q", "gk", "gv", "lkf")):
        scale = max(a64.grad.abs().max().item(), 1e-6 * cot_scale)
        assert (a32.grad.double() - a64.grad).abs().max().item() / scale < 1e-4, name


@pytest.mark.parametrize("B,S,Tm,n,tsp,skew", [(3, 7, 11, 21, True, False), (2, 5, 9, 101, False, False),
                                                (70, 3, 5, 2, False, False), (1, 1, 1, 128, True, False),
                                                (2, 100, 111, 101, False, True), (1, 2, 140000, 50, True, True)])
def test_fused_query_against_fp64(B, S, Tm, n, tsp, skew):
    """rl4co_eval_query_fwd / rl4co_eval_query_bwd against the PyTorch statement of context.py:118-157 + decoder.py:223-224
    in fp64: ragged sizes, more instances than SMs, one node collecting most rows (a CVRP depot), and a row range longer
    than one sorting pass of the backward kernel."""
    g = torch.Generator().manual_seed(B * 7 + S * 5 + Tm * 3 + n)
    rnd = lambda *s: torch.randn(*s, generator=g)
    pb, pa, wcap, gctx = rnd(B, n, 128), rnd(B, n, 128), rnd(128), rnd(B, 128)
    cur = torch.randint(0, n, (B, S, Tm), generator=g)
    if skew:
        cur[torch.rand(B, S, Tm, generator=g) < 0.7] = 0
    first = torch.randint(0, n, (B, S), generator=g)
    rem = torch.rand(B, S, Tm, generator=g)
    cot = rnd(B, S * Tm, 128)
    cur, first, rem, cot = cur.to(DEV), first.to(DEV), rem.to(DEV), cot.to(DEV)

    def eager(pb, pa, wcap, gctx):
        bi = torch.arange(B, device=DEV).view(B, 1, 1)
        q = pb[bi, cur] + gctx.view(B, 1, 1, 128)
        q = q + (pa[bi[:, :, 0], first].view(B, S, 1, 128) if tsp else rem.unsqueeze(-1).to(pb.dtype) * wcap)
        return q.reshape(B, S * Tm, 128)

    a64 = [t.double().to(DEV).requires_grad_(True) for t in (pb, pa, wcap, gctx)]
    ref = eager(*a64)
    (ref * cot.double()).sum().backward()
    a32 = [t.to(DEV).requires_grad_(True) for t in (pb, pa, wcap, gctx)]
    got = train._FusedQuery.apply(a32[0], a32[1] if tsp else None, None if tsp else a32[2], a32[3], cur,
                                  first if tsp else None, None if tsp else rem)
    (got * cot).sum().backward()
    assert (got.double() - ref).abs().max().item() < 1e-5
    used = [(0, "pb"), (1, "pa"), (3, "gctx")] if tsp else [(0, "pb"), (2, "wcap"), (3, "gctx")]
    for i, name in used:
        scale = max(a64[i].grad.abs().max().item(), 1e-6)
        assert (a32[i].grad.double() - a64[i].grad).abs().max().item() / scale < 2e-5, name
    assert a32[2 if tsp else 1].grad is None


@pytest.mark.parametrize("env_name", ["tsp", "cvrp"])
def test_graphed_encoder_step_equals_the_eager_step(env_name):
    """train.graphed_encoder: encoder forward / backward replayed from CUDA graphs give the loss and the gradients of the
    module-by-module step (same kernels, same order), on fresh instances of the captured shape, and leave the
    normalisation buffers as they were."""
    import copy
    inst, pol, env, td = _setup(env_name, 4, 20, seed=11)
    pol.train()
    pol2 = copy.deepcopy(pol)
    bufs = {k: v.clone() for k, v in pol2.encoder.named_buffers()}
    encode = train.graphed_encoder(pol2, td)
    for k, v in pol2.encoder.named_buffers():
        assert torch.equal(v, bufs[k]), k
    for seed in (3, 4):
        inst_s, _, _, td_s = _setup(env_name, 4, 20, seed=20 + seed)
        a = train.pomo_train_step(pol, env, td_s.clone(), optimizer=None, noise_seed=seed, max_grad_norm=None)
        b = train.pomo_train_step(pol2, env, td_s.clone(), optimizer=None, noise_seed=seed, max_grad_norm=None,
                                  encoder=encode)
        assert abs(a["loss"] - b["loss"]) < 1e-6 * max(1.0, abs(a["loss"])) and a["reward"] == b["reward"]
        gmax = max(p.grad.abs().max().item() for p in pol.parameters())
        for (k, p), (_, p2) in zip(pol.named_parameters(), pol2.named_parameters()):
            # (biases in front of a normalisation layer have gradients that are rounding noise: floor the scale)
            scale = max(p.grad.abs().max().item(), 1e-2 * gmax)
            assert (p.grad - p2.grad).abs().max().item() / scale < 1e-4, k


@pytest.mark.parametrize("B,S,Tm,n", [(3, 1, 37, 101), (2, 131, 1, 128), (2, 5, 1, 2), (5, 13, 10, 64)])
def test_evaluate_entry_points_stay_inside_their_buffers(B, S, Tm, n):
    """Every output / scratch buffer of rl4co_eval_query_* and rl4co_eval_ll_* sits between two guard bands filled with a
    sentinel; after forward and backward the bands are untouched (ragged tiles, n below / at the tile width)."""
    import ctypes as C
    from eam_rl4co_b200 import _lib
    L = _lib.lib()
    R, G, SENT = S * Tm, 4096, -7.25
    g = torch.Generator().manual_seed(B + R + n)
    bufs = []

    def guarded(*shape, dtype=torch.float32):
        numel = 1
        for d in shape:
            numel *= d
        flat = torch.full((numel + 2 * G,), SENT if dtype.is_floating_point else -7, dtype=dtype, device=DEV)
        bufs.append((flat, numel))
        return flat[G:G + numel].view(*shape)

    dev = lambda t: t.to(DEV).contiguous()
    pb, gctx, wcap = dev(torch.randn(B, n, 128, generator=g)), dev(torch.randn(B, 128, generator=g)), dev(torch.randn(128, generator=g))
    gk, gv, lkf = (dev(torch.randn(B, n, 128, generator=g)) for _ in range(3))
    cur, rem = dev(torch.randint(0, n, (B, S, Tm), generator=g)), dev(torch.rand(B, S, Tm, generator=g))
    mask = torch.rand(B, R, n, generator=g) < 0.5
    taken = torch.randint(0, n, (B, R), generator=g)
    mask.scatter_(2, taken[..., None], True)
    mask, taken = dev(mask), dev(taken)
    dlogp = dev(torch.randn(B, R, generator=g))
    q, heads, dheads, du_t, dq = (guarded(B, R, 128) for _ in range(5))
    logp, lse_z, lse_h, dd = guarded(B, R), guarded(B, R), guarded(B, R, 8), guarded(B, R, 8)
    maskw, ws = guarded(B, R, 4, dtype=torch.int32), guarded(2)
    dgk, dgv, dlkf, dpb, dwcap = guarded(B, n, 128), guarded(B, n, 128), guarded(B, n, 128), guarded(B, n, 128), guarded(128)
    for t in (dgk, dgv, dlkf, dpb, dwcap):
        t.zero_()
    i64, i32, f32, p = C.c_int64, C.c_int32, C.c_float, _lib.ptr
    st = _lib.stream_ptr()
    _lib.check(L.rl4co_eval_query_fwd(p(pb), None, p(wcap), p(gctx), p(cur), None, p(rem), p(q), i64(B), i64(S), i64(Tm), i32(n), st), "q fwd")
    _lib.check(L.rl4co_eval_ll_fwd(p(q), p(gk), p(gv), p(lkf), p(mask), p(taken), p(logp), p(heads), p(lse_h), p(lse_z), p(maskw),
                                   i64(B), i64(R), i32(n), f32(10.0), f32(1.0), st), "fwd")
    _lib.check(L.rl4co_eval_ll_bwd(p(q), p(gk), p(gv), p(lkf), p(maskw), p(taken), p(heads), p(lse_h), p(lse_z), p(dlogp), p(dheads),
                                   p(du_t), p(dd), p(ws), p(dq), p(dgk), p(dgv), p(dlkf), i64(B), i64(R), i32(n), f32(10.0), f32(1.0), st),
               "bwd")
    _lib.check(L.rl4co_eval_query_bwd(p(dq), p(cur), None, p(rem), p(dpb), None, p(dwcap), i64(B), i64(S), i64(Tm), i32(n), st), "q bwd")
    torch.cuda.synchronize()
    for flat, numel in bufs:
        sent = SENT if flat.dtype.is_floating_point else -7
        assert bool((flat[:G] == sent).all()) and bool((flat[G + numel:] == sent).all())
        assert bool(torch.isfinite(flat[G:G + numel].float()).all())
    assert bool((q != SENT).all()) and bool((logp <= 1e-6).all())


@pytest.mark.parametrize("B,N", [(3, 21), (70, 101), (2, 128), (4, 2), (5, 9)])
def test_instance_norm_kernels_against_fp64(B, N):
    """rl4co_instnorm_fwd / rl4co_instnorm_bwd against nn/ops.py:48-49 (nn.InstanceNorm1d over the node axis, affine)
    in fp64: output and the gradients w.r.t. input, weight and bias."""
    from eam_rl4co_b200 import encoder as E
    g = torch.Generator().manual_seed(B * 31 + N)
    x = torch.randn(B, N, 128, generator=g) * 3 + 0.5
    w, b = torch.randn(128, generator=g), torch.randn(128, generator=g)
    cot = torch.randn(B, N, 128, generator=g)
    ref_mod = torch.nn.InstanceNorm1d(128, affine=True).double().to(DEV)
    ref_mod.weight.data.copy_(w)
    ref_mod.bias.data.copy_(b)
    x64 = x.double().to(DEV).requires_grad_(True)
    ref = ref_mod(x64.permute(0, 2, 1)).permute(0, 2, 1)
    (ref * cot.double().to(DEV)).sum().backward()
    x32, w32, b32 = (t.to(DEV).requires_grad_(True) for t in (x, w, b))
    assert E.instance_norm_kernel_applies(x32)
    got = E._InstanceNormFn.apply(x32, w32, b32, ref_mod.eps)
    (got * cot.to(DEV)).sum().backward()
    tol = lambda a, r: (a.double() - r).abs().max().item() / max(r.abs().max().item(), 1e-6)
    assert tol(got, ref) < 1e-5
    assert tol(x32.grad, x64.grad) < 1e-4
    assert tol(w32.grad, ref_mod.weight.grad) < 1e-5 and tol(b32.grad, ref_mod.bias.grad) < 1e-5
    # the module routes its differentiable CUDA path through the kernel and equals its own PyTorch statement
    mod = E.Normalization(128, "instance").to(DEV)
    xa = x.to(DEV)
    with torch.no_grad():
        plain = mod(xa)
    assert tol(mod(xa.clone().requires_grad_(True)), plain.double()) < 1e-5


def test_split_k_linear_gradients_equal_nn_linear():
    """encoder._SplitKLinearFn (weight gradient as a split-K bmm) against nn.Linear's own autograd on the encoder's
    feed-forward block at the c5 token count."""
    from eam_rl4co_b200 import encoder as E
    torch.manual_seed(3)
    mlp = E.MLP(128, 512).to(DEV)
    x = torch.randn(64, 101, 128, device=DEV)
    cot = torch.randn(64, 101, 128, device=DEV)
    res = {}
    for flag in (True, False):
        E.SPLITK_LINEAR = flag
        try:
            xi = x.clone().requires_grad_(True)
            for p in mlp.parameters():
                p.grad = None
            (mlp(xi) * cot).sum().backward()
            res[flag] = [xi.grad.clone()] + [p.grad.clone() for p in mlp.parameters()]
        finally:
            E.SPLITK_LINEAR = True
    assert E._split_chunks(64 * 101) == 32
    for a, b in zip(res[True], res[False]):
        assert (a - b).abs().max().item() < 2e-5 * max(b.abs().max().item(), 1e-6)


def test_fused_optimizer_does_not_leave_stale_packed_weights():
    """torch.optim.Adam(fused=True) updates the parameters without moving their `_version`: the packed fp16 copies of the
    decoder / encoder weights must still follow (the caches are stamped with a global optimiser-step epoch).  Measured
    before the fix: the rollout of the second step ran on the first step's weights, ll_gap = 21."""
    inst, pol, env, td = _setup("cvrp", 4, 20, seed=13)
    pol.train()
    opt = torch.optim.Adam(pol.parameters(), lr=1e-3, fused=True)
    for it in range(4):
        stats = train.pomo_train_step(pol, env, td.clone(), optimizer=opt, noise_seed=it)
        assert stats["ll_gap"] < 1e-3, (it, stats)
    pol.eval()
    net = pol.encoder.net
    with torch.no_grad():
        hk, _ = pol.encoder(td)
        net.use_kernels = False
        ht, _ = pol.encoder(td)
        net.use_kernels = True
    assert (hk - ht).abs().max().item() < 1e-4
