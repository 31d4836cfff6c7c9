"""AttentionModel decoder, host side (mirror of rl4co/models/zoo/am/decoder.py).

Parameter names match the reference so its checkpoints load unchanged:
  context_embedding.project_context.weight [E, 2E | E+1], context_embedding.W_placeholder [2E]
  (TSP), pointer.project_out.weight [E,E], project_node_embeddings.weight [3E,E],
  project_fixed_context.weight [E,E].
`pre_decoder_hook` builds the per-instance tables with ordinary GEMMs once per rollout
(am/decoder.py:206-235 stays in torch); `forward` is one decoder step on the
sm_100a kernel `rl4co_decode_step`.
"""
import ctypes as C
from dataclasses import dataclass
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib


class TSPContext(nn.Module):  # context.py:105-140
    def __init__(self, embed_dim):
        super().__init__()
        self.embed_dim = embed_dim
        self.project_context = nn.Linear(2 * embed_dim, embed_dim, bias=False)
        self.W_placeholder = nn.Parameter(torch.Tensor(2 * embed_dim).uniform_(-1, 1))


class VRPContext(nn.Module):  # context.py:143-157
    def __init__(self, embed_dim):
        super().__init__()
        self.embed_dim = embed_dim
        self.project_context = nn.Linear(embed_dim + 1, embed_dim, bias=False)


class PointerAttention(nn.Module):  # attention.py:224-328
    def __init__(self, embed_dim, num_heads, mask_inner=True, check_nan=True):
        super().__init__()
        self.num_heads = num_heads
        self.mask_inner = mask_inner
        self.check_nan = check_nan
        self.project_out = nn.Linear(embed_dim, embed_dim, bias=False)


@dataclass
class PrecomputedCache:
    """am/decoder.py:22-41 plus the folded tables the kernels read (include/rl4co_b200.h)."""
    node_embeddings: torch.Tensor
    graph_context: object
    glimpse_key: torch.Tensor
    glimpse_val: torch.Tensor
    logit_key_folded: torch.Tensor        # L @ W_out
    pa: Optional[torch.Tensor]
    pb: torch.Tensor
    q0: Optional[torch.Tensor]
    wcap: Optional[torch.Tensor]
    env_id: int
    _logit_key_w: Optional[torch.Tensor] = None
    tc_img: Optional[torch.Tensor] = None   # [B, rl4co_tc_image_bytes(n)] uint8: tcgen05 operand tiles
    tc_err: Optional[torch.Tensor] = None
    tc_layout: int = 0                      # _lib.TC_LAYOUT_*: resident tiles / head-sliced for the streaming kernel

    @property
    def logit_key(self):
        return F.linear(self.node_embeddings, self._logit_key_w)

    def cstruct(self) -> _lib.Cache:
        B, n, E = self.node_embeddings.shape
        p = lambda t: (t.data_ptr() if t is not None else None)
        return _lib.Cache(p(self.glimpse_key), p(self.glimpse_val), p(self.logit_key_folded), p(self.pa),
                          p(self.pb), p(self.q0), p(self.wcap), p(self.tc_img), B, n, E, 8, self.tc_layout)


class AttentionModelDecoder(nn.Module):
    def __init__(self, embed_dim: int = 128, num_heads: int = 8, env_name: str = "tsp",
                 mask_inner: bool = True, use_graph_context: bool = True, check_nan: bool = True,
                 linear_bias: bool = False, out_bias_pointer_attn: bool = False, **unused):
        super().__init__()
        if embed_dim != 128 or num_heads != 8:
            raise NotImplementedError("rl4co_b200 kernels are built for embed_dim=128, num_heads=8")
        if linear_bias or out_bias_pointer_attn or not mask_inner:
            raise NotImplementedError("decoder biases / mask_inner=False are not supported by the fused path")
        if env_name not in ("tsp", "cvrp"):
            raise NotImplementedError(f"env '{env_name}' is outside the accelerated path (tsp, cvrp)")
        self.env_name, self.embed_dim, self.num_heads = env_name, embed_dim, num_heads
        self.context_embedding = TSPContext(embed_dim) if env_name == "tsp" else VRPContext(embed_dim)
        self.pointer = PointerAttention(embed_dim, num_heads, mask_inner, check_nan)
        self.project_node_embeddings = nn.Linear(embed_dim, 3 * embed_dim, bias=False)
        self.project_fixed_context = nn.Linear(embed_dim, embed_dim, bias=False)
        self.use_graph_context = use_graph_context
        self.is_dynamic_embedding = False
        self.use_kernels = True        # cache precompute on the tcgen05 linear kernel (False: torch GEMM)
        self.tc_layout = None          # None: by batch size; _lib.TC_LAYOUT_RESIDENT / _STREAM force a rollout kernel
        self._packer, self._zero_bias = None, None

    def pre_decoder_hook(self, td, env, embeddings, num_starts: int = 0):
        return td, env, self._precompute_cache(embeddings, num_starts=num_starts)

    @torch.no_grad()
    def _precompute_cache(self, h: torch.Tensor, num_starts: int = 0) -> PrecomputedCache:
        E = self.embed_dim
        B, n, _ = h.shape
        Wk, Wv, Wl = self.project_node_embeddings.weight.chunk(3, dim=0)
        Wo = self.pointer.project_out.weight
        Wc = self.context_embedding.project_context.weight
        Wl_folded = Wo.t() @ Wl                     # L @ W_out == h @ (W_out^T W_l)^T
        if self.use_graph_context:
            gc = F.linear(h.mean(1), self.project_fixed_context.weight)   # decoder.py:223-224
        else:
            gc = 0
        tsp = self.env_name == "tsp"
        ws = [Wk, Wv, Wl_folded, Wc[:, :E]] + ([Wc[:, E:]] if tsp else [])
        if self.use_kernels and h.is_cuda and h.dtype == torch.float32 and n <= 128:
            # the five [n,128] x [128,128] products on the tcgen05 linear kernel of csrc/encoder_tc.cu (split-fp16,
            # fp32-accurate) instead of an fp32 SIMT GEMM; weight tiles are packed once per parameter version
            tabs = self._precompute_kernels(h.contiguous(), ws)
        else:
            W = torch.stack([w.contiguous() for w in ws])                   # [k, E, E]
            tabs = torch.matmul(h.reshape(1, B * n, E), W.transpose(1, 2)).view(len(ws), B, n, E)
        gk, gv, lkf = tabs[0], tabs[1], tabs[2]
        gcb = gc[:, None, :] if isinstance(gc, torch.Tensor) else 0
        if tsp:
            pa, pb = tabs[3], (tabs[4] + gcb).contiguous()
            q0 = F.linear(self.context_embedding.W_placeholder, Wc)[None, :].expand(B, E)
            q0 = (q0 + gc).contiguous()
            wcap = None
        else:
            pa, pb, q0 = None, (tabs[3] + gcb).contiguous(), None
            wcap = Wc[:, E].contiguous()
        cache = PrecomputedCache(h, gc, gk, gv, lkf, pa, pb, q0, wcap,
                                 _lib.ENV_TSP if tsp else _lib.ENV_CVRP, Wl)
        if h.is_cuda and 8 <= num_starts <= 128:
            self._pack_tensor_tiles(cache, self.tc_layout)
        return cache

    def _precompute_kernels(self, h, ws):
        from .encoder import WeightPacker, kernel_linear
        if self._packer is None:
            self._packer = WeightPacker()
        params = (self.project_node_embeddings.weight, self.pointer.project_out.weight,
                  self.context_embedding.project_context.weight)
        stamp = tuple(WeightPacker.stamp(p) for p in params)
        if self._zero_bias is None or self._zero_bias.device != h.device:
            self._zero_bias = torch.zeros(128, dtype=torch.float32, device=h.device)
        return [kernel_linear(h, self._packer.pack(w.contiguous(), key=("pre", i), stamp=stamp), self._zero_bias, 128)
                for i, w in enumerate(ws)]

    def _apply(self, fn, *a, **kw):
        if getattr(self, "_packer", None) is not None:
            self._packer.clear()
        return super()._apply(fn, *a, **kw)

    def _load_from_state_dict(self, *a, **kw):
        if getattr(self, "_packer", None) is not None:
            self._packer.clear()
        return super()._load_from_state_dict(*a, **kw)

    @staticmethod
    def _pack_tensor_tiles(cache: PrecomputedCache, layout: Optional[int] = None) -> None:
        """Operand image of the tcgen05 kernels (multistart rows): fp16 hi/lo tiles of K, L, V^T.  `layout` None: the
        library's choice for this batch size (streaming kernel, two instances per SM, when B exceeds the SM count)."""
        B, n, _ = cache.node_embeddings.shape
        cache.tc_layout = int(_lib.lib().rl4co_tc_preferred_layout(C.c_int64(B))) if layout is None else int(layout)
        nbytes = _lib.lib().rl4co_tc_image_bytes(C.c_int32(n))
        if nbytes <= 0:
            return                      # n_nodes beyond the tensor path: the fp32 kernels take the rows
        img = torch.empty(B, nbytes, dtype=torch.uint8, device=cache.node_embeddings.device)
        err = _lib.new_err(img.device)
        cs = cache.cstruct()
        _lib.check(_lib.lib().rl4co_tc_pack(C.byref(cs), _lib.ptr(img), _lib.ptr(err), _lib.stream_ptr()),
                   "rl4co_tc_pack")
        cache.tc_img = img
        cache.tc_err = err              # read together with the rollout's error word

    @torch.no_grad()
    def forward(self, td, cached: PrecomputedCache, num_starts: int = 0, impl: int = _lib.IMPL_AUTO,
                debug: Optional[dict] = None):
        """am/decoder.py:161-198 -> (logits [rows,N'], mask [rows,N']) in start-major row order.
        Multistart rows (8 <= S <= 128) past the first step run on the tcgen05 kernel."""
        mask = td["action_mask"].contiguous()
        rows, n = mask.shape
        dev = mask.device
        logits = torch.empty(rows, n, dtype=torch.float32, device=dev)
        err = _lib.new_err(dev)
        cs = cached.cstruct()
        cur = td["current_node"].contiguous()
        use_tc = (impl != _lib.IMPL_FP32 and cached.tc_img is not None and 8 <= rows // cs.B <= 128
                  and cached.tc_layout == _lib.TC_LAYOUT_RESIDENT)
        if impl == _lib.IMPL_TC and not use_tc:
            raise RuntimeError("tensor-core decode step needs multistart rows (8 <= S <= 128) and a packed cache")
        if use_tc:
            flat = lambda t: t.reshape(rows).contiguous()
            first = flat(td["first_node"]) if self.env_name == "tsp" else None
            used = flat(td["used_capacity"]) if self.env_name == "cvrp" else None
            cap = flat(td["vehicle_capacity"]) if self.env_name == "cvrp" else None
            ds = dh = None
            if debug is not None:
                ds = torch.zeros(rows, 8, n, dtype=torch.float32, device=dev)
                dh = torch.zeros(rows, 128, dtype=torch.float32, device=dev)
                debug["scores"], debug["heads"] = ds, dh
            rc = _lib.lib().rl4co_tc_decode_step(
                C.c_int32(cached.env_id), C.byref(cs), _lib.ptr(first), _lib.ptr(flat(cur)), _lib.ptr(used),
                _lib.ptr(cap), _lib.ptr(mask), _lib.ptr(logits), _lib.ptr(ds), _lib.ptr(dh), _lib.ptr(err),
                C.c_int64(rows), _lib.stream_ptr())
            _lib.check(rc, "rl4co_tc_decode_step")
            if self.pointer.check_nan:
                _lib.raise_on_err(err)
            return logits, mask
        if self.env_name == "tsp":
            first, i_step, used, cap = td["first_node"].contiguous(), td["i"].contiguous(), None, None
        else:
            first, i_step = None, None
            used, cap = td["used_capacity"].contiguous(), td["vehicle_capacity"].contiguous()
        rc = _lib.lib().rl4co_decode_step(
            C.c_int32(cached.env_id), C.byref(cs), _lib.ptr(first), _lib.ptr(cur), _lib.ptr(i_step),
            _lib.ptr(used), _lib.ptr(cap), _lib.ptr(mask), _lib.ptr(logits), _lib.ptr(err),
            C.c_int64(rows), _lib.stream_ptr())
        _lib.check(rc, "rl4co_decode_step")
        if self.pointer.check_nan:
            _lib.raise_on_err(err)      # attention.py:303-304 (the reference syncs here too)
        return logits, mask
