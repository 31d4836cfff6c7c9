"""ctypes binding of librl4co_b200.so (the C ABI declared in include/rl4co_b200.h).

There is NO fallback: if the shared library is missing or a call fails, this raises.
"""
import ctypes as C
import os
import subprocess

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# RL4CO_B200_LIB: experiment builds of the same sources (profiles/variants.sh); the product path is the in-tree library
LIB_PATH = os.environ.get("RL4CO_B200_LIB") or os.path.join(_HERE, "librl4co_b200.so")
CSRC = os.path.join(_HERE, "csrc")
SOURCES = ["capi.cu", "env_kernels.cu", "decode_kernels.cu", "rollout_kernel.cu", "rollout_coop.cu", "rollout_tc.cu", "encoder_tc.cu", "reduce_kernels.cu", "data_kernels.cu", "evaluate_kernels.cu", "evaluate_query.cu", "norm_kernels.cu"]

ENV_TSP, ENV_CVRP = 0, 1
MODE_GREEDY, MODE_SAMPLING, MODE_EVALUATE = 0, 1, 2
ERR_MESSAGES = [  # (bit, reference assertion text); first match is raised
    (64, "decoder cache value outside the fp16 range of the tensor-core path"),
    (1, "Logits contain NaNs"),
    (2, "infeasible action selected"),
    (4, "Logprobs should not be -inf, check sampling procedure!"),
    (8, "Invalid tour"),
    (16, "Used more than capacity"),
    (32, "Exceeded maximum number of steps during decoding"),
]
IMPL_AUTO, IMPL_FP32, IMPL_TC, IMPL_STREAM, IMPL_COOP = 0, 1, 2, 3, 4
TC_LAYOUT_RESIDENT, TC_LAYOUT_STREAM = 0, 1
ENC_EPI_PLAIN, ENC_EPI_RELU, ENC_EPI_RES_NORM, ENC_EPI_RES_AFFINE = 0, 1, 2, 3
ABI_VERSION = 3

EXPORTS = [
    "rl4co_b200_abi_version", "rl4co_last_error", "rl4co_tsp_step", "rl4co_cvrp_step_mask",
    "rl4co_cvrp_action_mask", "rl4co_tour_reward", "rl4co_select", "rl4co_decode_step",
    "rl4co_rollout", "rl4co_rollout_noalloc", "rl4co_noalloc_max_instances", "rl4co_noise_gumbel", "rl4co_tc_image_bytes", "rl4co_tc_preferred_layout", "rl4co_tc_pack", "rl4co_tc_decode_step",
    "rl4co_enc_packed_bytes", "rl4co_enc_pack_weight", "rl4co_enc_linear", "rl4co_enc_ffn", "rl4co_enc_attn_block", "rl4co_enc_attention", "rl4co_best_of_starts", "rl4co_generate_instances", "rl4co_replay_states",
    "rl4co_eval_ll_max_nodes", "rl4co_eval_ll_fwd", "rl4co_eval_ll_bwd", "rl4co_eval_query_fwd", "rl4co_eval_query_bwd", "rl4co_instnorm_max_nodes", "rl4co_instnorm_fwd", "rl4co_instnorm_bwd",
]


def build(verbose: bool = False) -> str:
    """Compile the CUDA sources for sm_100a into the in-tree shared library."""
    srcs = [os.path.join(CSRC, s) for s in SOURCES]
    extra = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    extra.append(os.path.join(_HERE, "..", "include", "rl4co_b200.h"))
    if os.path.exists(LIB_PATH):
        newest = max(os.path.getmtime(p) for p in srcs + extra)
        if os.path.getmtime(LIB_PATH) >= newest:
            return LIB_PATH
    cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
           "-Xcompiler", "-fPIC", "-shared", "-o", LIB_PATH] + srcs
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    subprocess.run(cmd, check=True)
    return LIB_PATH


class Cache(C.Structure):
    _fields_ = [("gk", C.c_void_p), ("gv", C.c_void_p), ("lk", C.c_void_p), ("pa", C.c_void_p),
                ("pb", C.c_void_p), ("q0", C.c_void_p), ("wcap", C.c_void_p), ("tc_img", C.c_void_p),
                ("B", C.c_int64),
                ("n_nodes", C.c_int32), ("embed_dim", C.c_int32), ("num_heads", C.c_int32), ("tc_layout", C.c_int32)]


class RolloutArgs(C.Structure):
    _fields_ = [("env", C.c_int32), ("mode", C.c_int32), ("num_starts", C.c_int32),
                ("max_steps", C.c_int32), ("tanh_clipping", C.c_float), ("temperature", C.c_float),
                ("seed", C.c_uint32), ("inst_offset", C.c_int64), ("locs", C.c_void_p),
                ("demand", C.c_void_p), ("vehicle_capacity", C.c_void_p), ("gumbel", C.c_void_p),
                ("actions_in", C.c_void_p), ("actions_in_stride", C.c_int64),
                ("actions_in_T", C.c_int32), ("actions", C.c_void_p), ("logp", C.c_void_p),
                ("logp_all", C.c_void_p), ("reward", C.c_void_p), ("log_likelihood", C.c_void_p),
                ("lengths", C.c_void_p), ("max_len", C.c_void_p), ("err", C.c_void_p), ("impl", C.c_int32), ("start_mod", C.c_int32),
                ("phase_cycles", C.c_void_p)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'`."
                " There is no CPU or PyTorch fallback for this path.")
        L = C.CDLL(LIB_PATH)
        L.rl4co_last_error.restype = C.c_char_p
        for name in EXPORTS:
            getattr(L, name)  # fail loudly on a missing symbol
        L.rl4co_tc_image_bytes.restype = C.c_int64
        L.rl4co_noalloc_max_instances.restype = C.c_int64
        L.rl4co_enc_packed_bytes.restype = C.c_int64
        if L.rl4co_b200_abi_version() != ABI_VERSION:
            raise RuntimeError("librl4co_b200.so ABI version mismatch")
        _lib = L
    return _lib


def ptr(t):
    """Device pointer of a tensor (None -> NULL). Tensors must be CUDA + contiguous."""
    if t is None:
        return C.c_void_p(0)
    if not t.is_cuda:
        raise RuntimeError("rl4co_b200 kernels need CUDA tensors (no CPU fallback)")
    if not t.is_contiguous():
        raise RuntimeError("rl4co_b200 kernels need contiguous tensors")
    return C.c_void_p(t.data_ptr())


def stream_ptr():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def check(rc: int, what: str):
    if rc != 0:
        msg = lib().rl4co_last_error().decode()
        raise RuntimeError(f"{what} failed (code {rc}): {msg}")


def new_err(device):
    return torch.zeros(2, dtype=torch.int32, device=device)


def raise_on_err(err):
    """One device->host read per rollout; raises the reference's AssertionError texts."""
    e = err.tolist()
    if e[0]:
        for bit, msg in ERR_MESSAGES:
            if e[0] & bit:
                raise AssertionError(f"{msg} (row {e[1] - 1})")


class nvtx_range:
    """NVTX range around a phase of the path (encoder / cache precompute / rollout): nsys and ncu captures of a
    `policy(td, env, ...)` call are self-describing.  A no-op when CUDA is not initialised."""

    def __init__(self, name: str):
        self.name, self.on = name, False

    def __enter__(self):
        if torch.cuda.is_available() and torch.cuda.is_initialized():
            torch.cuda.nvtx.range_push(self.name)
            self.on = True
        return self

    def __exit__(self, *exc):
        if self.on:
            torch.cuda.nvtx.range_pop()
        return False
