"""CPU: the C-ABI library builds, loads and exports every symbol include/rl4co_b200.h declares.
No compute calls here (no GPU); argument validation paths that return before any launch are ok."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "rl4co_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rl4co_[a-z0-9_]+)\s*\(", text)))


def test_library_builds_and_exports_header_symbols():
    from eam_rl4co_b200 import _lib
    _lib.build()
    L = _lib.lib()
    syms = _declared_symbols()
    assert len(syms) >= 10
    for s in syms:
        assert hasattr(L, s), f"{s} declared in the header but not exported"
    assert set(syms) == set(_lib.EXPORTS)
    assert L.rl4co_b200_abi_version() == _lib.ABI_VERSION


def test_struct_layouts_match_header():
    from eam_rl4co_b200 import _lib
    # 8 pointers + int64 + 3 int32 (+pad) ; see include/rl4co_b200.h
    assert C.sizeof(_lib.Cache) == 8 * 8 + 8 + 16
    assert _lib.RolloutArgs.seed.offset == 24 and _lib.RolloutArgs.inst_offset.offset == 32
    assert C.sizeof(_lib.RolloutArgs) == 176 and _lib.RolloutArgs.err.offset == 152
    assert _lib.RolloutArgs.impl.offset == 160 and _lib.RolloutArgs.start_mod.offset == 164


def test_argument_errors_return_codes_without_launching():
    from eam_rl4co_b200 import _lib
    L = _lib.lib()
    rc = L.rl4co_rollout(None, None, None)
    assert rc == -1 and b"null args" in L.rl4co_last_error()
    cache = _lib.Cache(None, None, None, None, None, None, None, None, 1, 300, 128, 8)
    rc = L.rl4co_decode_step(0, C.byref(cache), None, None, None, None, None, None, None, None,
                             C.c_int64(1), None)
    assert rc == -1 and b"n_nodes=300" in L.rl4co_last_error()
    cache = _lib.Cache(None, None, None, None, None, None, None, None, 1, 20, 64, 8)
    rc = L.rl4co_decode_step(0, C.byref(cache), None, None, None, None, None, None, None, None,
                             C.c_int64(1), None)
    assert rc == -1 and b"embed_dim=128" in L.rl4co_last_error()
    # the evaluate kernels check their arguments before any CUDA call
    assert L.rl4co_eval_ll_max_nodes() == 128
    rc = L.rl4co_eval_ll_fwd(None, None, None, None, None, None, None, None, None, None, None, C.c_int64(1), C.c_int64(1),
                             C.c_int32(20), C.c_float(10.0), C.c_float(1.0), None)
    assert rc == -1 and b"null pointer" in L.rl4co_last_error()
    one = C.c_void_p(16)                                           # never dereferenced: the size checks come first
    rc = L.rl4co_eval_ll_bwd(*([one] * 18), C.c_int64(2), C.c_int64(3), C.c_int32(200), C.c_float(10.0), C.c_float(1.0), None)
    assert rc == -1 and b"n_nodes=200" in L.rl4co_last_error()
    rc = L.rl4co_eval_ll_bwd(*([one] * 18), C.c_int64(2), C.c_int64(3), C.c_int32(50), C.c_float(10.0), C.c_float(0.0), None)
    assert rc == -1 and b"temperature" in L.rl4co_last_error()
    sz = (C.c_int64(2), C.c_int64(3), C.c_int64(4))
    rc = L.rl4co_eval_query_fwd(one, None, None, None, one, None, None, one, *sz, C.c_int32(129), None)
    assert rc == -1 and b"n_nodes=129" in L.rl4co_last_error()
    rc = L.rl4co_eval_query_fwd(one, one, None, None, one, None, None, one, *sz, C.c_int32(20), None)
    assert rc == -1 and b"pa and first" in L.rl4co_last_error()
    rc = L.rl4co_eval_query_bwd(one, one, None, one, one, None, None, *sz, C.c_int32(20), None)
    assert rc == -1 and b"dwcap and rem" in L.rl4co_last_error()
    rc = L.rl4co_eval_query_bwd(one, one, None, None, None, None, None, *sz, C.c_int32(20), None)
    assert rc == -1 and b"null dq / cur / dpb" in L.rl4co_last_error()
    assert L.rl4co_instnorm_max_nodes() == 128
    rc = L.rl4co_instnorm_fwd(one, one, one, one, one, one, C.c_int64(2), C.c_int32(129), C.c_float(1e-5), None)
    assert rc == -1 and b"N=129" in L.rl4co_last_error()
    rc = L.rl4co_instnorm_bwd(one, one, one, one, one, one, one, None, C.c_int64(2), C.c_int32(20), None)
    assert rc == -1 and b"null pointer" in L.rl4co_last_error()


def test_no_fallback_when_library_missing(monkeypatch):
    from eam_rl4co_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/librl4co_b200.so")
    with pytest.raises(RuntimeError, match="no CPU or PyTorch fallback"):
        _lib.lib()


def test_cpu_tensors_are_refused():
    import torch
    from eam_rl4co_b200 import _lib
    with pytest.raises(RuntimeError, match="CUDA tensors"):
        _lib.ptr(torch.zeros(3))
