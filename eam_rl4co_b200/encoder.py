"""Graph-attention encoder of the AttentionModel.

It runs once per rollout and produces the node embeddings `h [B, N', E]` the rollout kernels consume
(§8f rank 3 is its kernel path).  Module / parameter names follow the reference
(rl4co/models/zoo/am/encoder.py:70-91, nn/graph/attnnet.py:16-103, nn/attention.py:66-136,
nn/ops.py:32-56, nn/mlp.py, nn/env_embeddings/init.py:55-68,115-138) so that reference
state_dicts load unchanged.

Two executions of the same arithmetic:
  * the plain-PyTorch modules below (differentiable: training, and any configuration the kernels do not cover);
  * `GraphAttentionNetwork._forward_kernels`: under `torch.no_grad()` on CUDA with instance normalisation (POMO)
    or eval-mode batch normalisation (the AttentionModel default, running statistics = a per-channel affine),
    embed_dim 128, 8 heads and N' <= 128 the layers run on the tcgen05 kernels of csrc/encoder_tc.cu (split-fp16
    GEMMs with the residual + norm fused into their epilogues, tcgen05 attention) - 5 launches per layer.
"""
import ctypes as C

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib


class TSPInitEmbedding(nn.Module):
    def __init__(self, embed_dim):
        super().__init__()
        self.init_embed = nn.Linear(2, embed_dim, True)

    def forward(self, td):
        return self.init_embed(td["locs"])


class VRPInitEmbedding(nn.Module):
    def __init__(self, embed_dim):
        super().__init__()
        self.init_embed = nn.Linear(3, embed_dim, True)
        self.init_embed_depot = nn.Linear(2, embed_dim, True)

    def forward(self, td):
        depot, cities = td["locs"][:, :1, :], td["locs"][:, 1:, :]
        return torch.cat((self.init_embed_depot(depot),
                          self.init_embed(torch.cat((cities, td["demand"][..., None]), -1))), -2)


SPLITK_LINEAR = True      # training path on CUDA: weight gradients of the encoder's Linear layers as split-K batched products


def _split_chunks(tokens: int) -> int:
    for c in (32, 16, 8, 4, 2):
        if tokens % c == 0 and tokens // c >= 128:
            return c
    return 1


class _SplitKLinearFn(torch.autograd.Function):
    """y = x W^T + b (nn.Linear) whose weight gradient dW = dy^T x is a split-K batched product.  At the training sizes
    (6 464 tokens, 128 .. 512 features, fp32) cuBLAS runs the single GEMM on 12 CTAs of one tile each (38 us, 24 of them
    per step); 32 token chunks as one bmm plus a sum fill the GPU.  Same fp32 arithmetic, another summation order."""

    @staticmethod
    def forward(ctx, x, w, b):
        ctx.save_for_backward(x, w)
        return F.linear(x, w, b)

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        x2, dy2 = x.reshape(-1, x.shape[-1]), dy.reshape(-1, dy.shape[-1])
        dx = (dy2 @ w).view_as(x) if ctx.needs_input_grad[0] else None
        tokens = x2.shape[0]
        c = _split_chunks(tokens)
        if c > 1:
            dw = torch.bmm(dy2.view(c, tokens // c, -1).transpose(1, 2), x2.view(c, tokens // c, -1)).sum(0)
        else:
            dw = dy2.t() @ x2
        return dx, dw, dy2.sum(0)


def _linear(mod: nn.Linear, x: torch.Tensor) -> torch.Tensor:
    if SPLITK_LINEAR and x.is_cuda and torch.is_grad_enabled() and mod.weight.requires_grad and mod.bias is not None:
        return _SplitKLinearFn.apply(x, mod.weight, mod.bias)
    return mod(x)


class MultiHeadAttention(nn.Module):
    def __init__(self, embed_dim, num_heads):
        super().__init__()
        self.num_heads = num_heads
        self.Wqkv = nn.Linear(embed_dim, 3 * embed_dim, bias=True)
        self.out_proj = nn.Linear(embed_dim, embed_dim, bias=True)

    def forward(self, x):
        B, S, E = x.shape
        qkv = _linear(self.Wqkv, x).view(B, S, 3, self.num_heads, E // self.num_heads)
        q, k, v = qkv.permute(2, 0, 3, 1, 4).unbind(0)
        out = F.scaled_dot_product_attention(q, k, v)
        return _linear(self.out_proj, out.transpose(1, 2).reshape(B, S, E))


class SkipConnection(nn.Module):
    def __init__(self, module):
        super().__init__()
        self.module = module

    def forward(self, x):
        return x + self.module(x)


class _InstanceNormFn(torch.autograd.Function):
    """nn.InstanceNorm1d over the node axis of [B, N, 128] (nn/ops.py:48-49) as one kernel each way
    (rl4co_instnorm_fwd / rl4co_instnorm_bwd, csrc/norm_kernels.cu): the training path's module-by-module statement is
    ~9 launches forward and ~20 backward per call, 12 calls per POMO encode."""

    @staticmethod
    def forward(ctx, x, weight, bias, eps):
        import ctypes as C
        B, N, E = x.shape
        x, weight, bias = x.contiguous(), weight.contiguous(), bias.contiguous()
        y = torch.empty_like(x)
        mean = torch.empty(B, E, device=x.device)
        rstd = torch.empty(B, E, device=x.device)
        _lib.check(_lib.lib().rl4co_instnorm_fwd(_lib.ptr(x), _lib.ptr(weight), _lib.ptr(bias), _lib.ptr(y), _lib.ptr(mean),
                                                 _lib.ptr(rstd), C.c_int64(B), C.c_int32(N), C.c_float(eps),
                                                 _lib.stream_ptr()), "rl4co_instnorm_fwd")
        ctx.save_for_backward(x, weight, mean, rstd)
        return y

    @staticmethod
    def backward(ctx, dy):
        import ctypes as C
        x, weight, mean, rstd = ctx.saved_tensors
        B, N, E = x.shape
        dy = dy.contiguous()
        dx = torch.empty_like(x)
        dw, db = torch.zeros_like(weight), torch.zeros_like(weight)
        _lib.check(_lib.lib().rl4co_instnorm_bwd(_lib.ptr(x), _lib.ptr(weight), _lib.ptr(mean), _lib.ptr(rstd), _lib.ptr(dy),
                                                 _lib.ptr(dx), _lib.ptr(dw), _lib.ptr(db), C.c_int64(B), C.c_int32(N),
                                                 _lib.stream_ptr()), "rl4co_instnorm_bwd")
        return dx, dw, db, None


def instance_norm_kernel_applies(x: torch.Tensor) -> bool:
    """The fused normalisation serves the differentiable (training) path on CUDA fp32 [B, N <= 128, 128]; under
    no_grad the plain PyTorch statement stays the reference the kernel tests compare against."""
    return (x.is_cuda and x.dtype == torch.float32 and x.dim() == 3 and x.shape[-1] == 128 and x.shape[1] <= 128
            and torch.is_grad_enabled())


class Normalization(nn.Module):
    def __init__(self, embed_dim, normalization="batch"):
        super().__init__()
        cls = {"batch": nn.BatchNorm1d, "instance": nn.InstanceNorm1d}[normalization]
        self.normalizer = cls(embed_dim, affine=True)

    def forward(self, x):
        if isinstance(self.normalizer, nn.BatchNorm1d):
            return self.normalizer(x.reshape(-1, x.size(-1))).view(*x.size())
        # nn/ops.py:48-49: InstanceNorm1d over the node axis of [B, N, E].  Same statistics (biased variance,
        # eps inside the sqrt, affine), computed in place on the [B, N, E] layout: the permute -> cuDNN
        # batch-norm route of nn.InstanceNorm1d costs ~1 ms per call at B=1024 on a B200, 12 calls per encode.
        nrm = self.normalizer
        if instance_norm_kernel_applies(x):
            return _InstanceNormFn.apply(x, nrm.weight, nrm.bias, float(nrm.eps))
        mean = x.mean(dim=1, keepdim=True)
        xc = x - mean
        var = (xc * xc).mean(dim=1, keepdim=True)
        return xc * torch.rsqrt(var + nrm.eps) * nrm.weight + nrm.bias


class MLP(nn.Module):
    def __init__(self, embed_dim, hidden):
        super().__init__()
        self.lins = nn.ModuleList([nn.Linear(embed_dim, hidden), nn.Linear(hidden, embed_dim)])

    def forward(self, x):
        return _linear(self.lins[1], F.relu(_linear(self.lins[0], x)))


# Derived copies of the weights (packed fp16 MMA tiles, folded BatchNorm affines) are cached.  A parameter's `_version`
# counter does NOT see every in-place update: PyTorch's fused optimisers (`torch.optim.Adam(..., fused=True)`) write the
# parameters from a multi-tensor kernel without touching it - measured: after one fused Adam step the rollout ran on the
# previous step's packed decoder weights (log-likelihood 21 off its re-evaluation).  Every `torch.optim` step therefore
# moves a global epoch that is part of every cache stamp.
_OPTIMIZER_EPOCH = [0]


def _bump_optimizer_epoch(*_args, **_kwargs):
    _OPTIMIZER_EPOCH[0] += 1


try:
    from torch.optim.optimizer import register_optimizer_step_post_hook as _register_step_hook
    _register_step_hook(_bump_optimizer_epoch)
except ImportError:                                   # pragma: no cover - torch < 2.0
    pass


class WeightPacker:
    """fp32 weight [n_out, k] -> fp16 hi/lo MMA tiles of csrc/encoder_tc.cu, cached per (key, version, storage,
    optimiser epoch)."""

    def __init__(self):
        self.cache = {}

    def clear(self):
        self.cache.clear()

    @staticmethod
    def stamp(w):
        # `_version` moves with in-place updates autograd can see, data_ptr with re-allocation (`.to`,
        # `load_state_dict(assign=True)`), the epoch with every torch.optim step; writes through `.data` outside an
        # optimiser that keep all three need invalidate_packed()
        return (w._version, w.data_ptr(), w.device, _OPTIMIZER_EPOCH[0])

    def pack(self, w, key=None, stamp=None):
        key = id(w) if key is None else key
        ver = self.stamp(w) if stamp is None else stamp
        hit = self.cache.get(key)
        if hit is not None and hit[0] == ver:
            return hit[1]
        n_out, k = w.shape
        L = _lib.lib()
        buf = torch.empty(L.rl4co_enc_packed_bytes(n_out, k), dtype=torch.uint8, device=w.device)
        _lib.check(L.rl4co_enc_pack_weight(_lib.ptr(w.detach().contiguous()), n_out, k, _lib.ptr(buf), _lib.stream_ptr()),
                   "rl4co_enc_pack_weight")
        self.cache[key] = (ver, buf)
        return buf


def kernel_linear(x, packed, bias, out_ch, epi=0, residual=None, gamma=None, beta=None, eps=0.0):
    """Y [B, n, out_ch] = epilogue(X [B, n, k] W^T + bias) on the tcgen05 linear kernel (rl4co_enc_linear)."""
    B, n, k = x.shape
    y = torch.empty(B, n, out_ch, dtype=torch.float32, device=x.device)
    _lib.check(_lib.lib().rl4co_enc_linear(
        _lib.ptr(x), _lib.ptr(packed), _lib.ptr(bias), _lib.ptr(residual), _lib.ptr(gamma), _lib.ptr(beta),
        C.c_float(eps), _lib.ptr(y), C.c_int64(B), n, k, out_ch, epi, _lib.stream_ptr()), "rl4co_enc_linear")
    return y


class GraphAttentionNetwork(nn.Module):
    def __init__(self, num_heads, embed_dim, num_layers, normalization, feedforward_hidden):
        super().__init__()
        self.layers = nn.Sequential(*(
            nn.Sequential(
                SkipConnection(MultiHeadAttention(embed_dim, num_heads)),
                Normalization(embed_dim, normalization),
                SkipConnection(MLP(embed_dim, feedforward_hidden)),
                Normalization(embed_dim, normalization),
            ) for _ in range(num_layers)))

        self.num_heads, self.embed_dim, self.normalization = num_heads, embed_dim, normalization
        self.feedforward_hidden = feedforward_hidden
        self.use_kernels = True        # set False to force the PyTorch modules
        self.fuse_ffn = True           # FFN1 + FFN2 + norm in one kernel (False: two enc_linear launches)
        self.fuse_attention = True     # QKV + attention + out-proj + norm in one kernel (n <= 112)
        self._packer = WeightPacker()  # id(weight) -> (stamp, packed tiles)
        self._packed = self._packer.cache

    def kernels_apply(self, x):
        norm_ok = self.normalization == "instance" or (self.normalization == "batch" and not self.training)
        return (self.use_kernels and x.is_cuda and not torch.is_grad_enabled() and x.dtype == torch.float32
                and norm_ok and self.embed_dim == 128 and self.num_heads == 8
                and self.feedforward_hidden % 128 == 0 and self.feedforward_hidden <= 512 and x.shape[1] <= 128)

    def invalidate_packed(self):
        """Drop the packed fp16 weight tiles (they are rebuilt on the next kernel forward).  Called by
        `load_state_dict` / `_apply`; call it yourself after writing weights through `.data` in a way that keeps
        the storage and the version counter (EMA / SWA helpers)."""
        self._packed.clear()

    def _apply(self, fn, *a, **kw):
        self._packed.clear()
        return super()._apply(fn, *a, **kw)

    def _load_from_state_dict(self, *a, **kw):
        self._packed.clear()
        return super()._load_from_state_dict(*a, **kw)

    _stamp = staticmethod(WeightPacker.stamp)

    def _pack(self, w):
        """nn.Linear weight -> MMA tiles; cached per (parameter version, storage)."""
        return self._packer.pack(w)

    def _bn_affine(self, bn):
        """Eval-mode BatchNorm1d (nn/ops.py:46-47) as a per-channel affine; cached per version of its four tensors."""
        key = ("bn", id(bn))
        ver = tuple(self._stamp(t) for t in (bn.weight, bn.bias, bn.running_mean, bn.running_var))
        hit = self._packed.get(key)
        if hit is not None and hit[0] == ver:
            return hit[1]
        g = (bn.weight.detach() * torch.rsqrt(bn.running_var + bn.eps)).contiguous()
        b = (bn.bias.detach() - bn.running_mean * g).contiguous()
        self._packed[key] = (ver, (g, b))
        return g, b

    def _linear(self, x, lin, epi, out_ch, residual=None, norm=None):
        g = b = None
        eps = 0.0
        if norm is not None:
            nrm = norm.normalizer
            if isinstance(nrm, nn.BatchNorm1d):
                epi = _lib.ENC_EPI_RES_AFFINE
                g, b = self._bn_affine(nrm)
            else:
                g, b, eps = nrm.weight.detach(), nrm.bias.detach(), nrm.eps
        return kernel_linear(x, self._pack(lin.weight), lin.bias.detach(), out_ch, epi, residual, g, b, eps)

    def _forward_kernels(self, x):
        L = _lib.lib()
        x = x.contiguous()
        B, n, E = x.shape
        for layer in self.layers:
            mha, norm1, mlp, norm2 = layer[0].module, layer[1], layer[2].module, layer[3]
            if self.fuse_attention and n <= 112:
                x = self._attn_block(x, mha, norm1)
            else:
                qkv = self._linear(x, mha.Wqkv, _lib.ENC_EPI_PLAIN, 3 * E)
                heads = torch.empty(B, n, E, dtype=torch.float32, device=x.device)
                _lib.check(L.rl4co_enc_attention(_lib.ptr(qkv), _lib.ptr(heads), C.c_int64(B), n, _lib.stream_ptr()),
                           "rl4co_enc_attention")
                x = self._linear(heads, mha.out_proj, _lib.ENC_EPI_RES_NORM, E, residual=x, norm=norm1)
            if self.fuse_ffn:
                x = self._ffn(x, mlp, norm2)
            else:
                f = self._linear(x, mlp.lins[0], _lib.ENC_EPI_RELU, self.feedforward_hidden)
                x = self._linear(f, mlp.lins[1], _lib.ENC_EPI_RES_NORM, E, residual=x, norm=norm2)
        return x

    def _norm_args(self, norm):
        nrm = norm.normalizer
        if isinstance(nrm, nn.BatchNorm1d):
            g, b = self._bn_affine(nrm)
            return _lib.ENC_EPI_RES_AFFINE, g, b, 0.0
        return _lib.ENC_EPI_RES_NORM, nrm.weight.detach(), nrm.bias.detach(), nrm.eps

    def _attn_block(self, x, mha, norm):
        """Norm(x + MHA(x) Wo^T + bo) in ONE kernel (rl4co_enc_attn_block): q, k, v and the heads stay on chip."""
        B, n, E = x.shape
        y = torch.empty_like(x)
        epi, g, b, eps = self._norm_args(norm)
        _lib.check(_lib.lib().rl4co_enc_attn_block(
            _lib.ptr(x), _lib.ptr(self._pack(mha.Wqkv.weight)), _lib.ptr(mha.Wqkv.bias.detach()),
            _lib.ptr(self._pack(mha.out_proj.weight)), _lib.ptr(mha.out_proj.bias.detach()), _lib.ptr(g), _lib.ptr(b),
            C.c_float(eps), _lib.ptr(y), C.c_int64(B), n, epi, _lib.stream_ptr()), "rl4co_enc_attn_block")
        return y

    def _ffn(self, x, mlp, norm):
        """Norm(x + W2 relu(W1 x + b1) + b2) in ONE kernel (rl4co_enc_ffn): the hidden activations stay in TMEM."""
        B, n, E = x.shape
        y = torch.empty_like(x)
        epi, g, b, eps = self._norm_args(norm)
        l1, l2 = mlp.lins
        _lib.check(_lib.lib().rl4co_enc_ffn(
            _lib.ptr(x), _lib.ptr(self._pack(l1.weight)), _lib.ptr(l1.bias.detach()), _lib.ptr(self._pack(l2.weight)),
            _lib.ptr(l2.bias.detach()), _lib.ptr(g), _lib.ptr(b), C.c_float(eps), _lib.ptr(y), C.c_int64(B), n,
            self.feedforward_hidden, epi, _lib.stream_ptr()), "rl4co_enc_ffn")
        return y

    def forward(self, x):
        if self.kernels_apply(x):
            return self._forward_kernels(x)
        return self.layers(x)


class AttentionModelEncoder(nn.Module):
    def __init__(self, embed_dim=128, env_name="tsp", num_heads=8, num_layers=3,
                 normalization="batch", feedforward_hidden=512):
        super().__init__()
        self.init_embedding = (TSPInitEmbedding if env_name == "tsp" else VRPInitEmbedding)(embed_dim)
        self.net = GraphAttentionNetwork(num_heads, embed_dim, num_layers, normalization,
                                         feedforward_hidden)

    def forward(self, td):
        init_h = self.init_embedding(td)
        return self.net(init_h), init_h
