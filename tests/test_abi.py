"""CPU tests: the C-ABI library loads without a GPU and exports every symbol include/b200ops.h
declares; argument validation fails loudly; nothing silently falls back to the CPU."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from b200ops import _lib
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    return _lib.load()


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200ops.h")).read()
    return sorted(set(re.findall(r"B200_API\s+[\w\s\*]+?\b(b200_\w+)\s*\(", text)))


def test_header_declares_entry_points():
    syms = declared_symbols()
    assert "b200_paged_attention_decode" in syms and "b200_reshape_and_cache_flash" in syms
    assert len(syms) >= 10


def test_every_declared_symbol_is_exported_and_bound(lib):
    from b200ops import _lib
    raw = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared_symbols():
        assert hasattr(raw, name), f"{name} declared in include/b200ops.h but not exported"
        assert name in _lib.SIGNATURES, f"{name} has no ctypes signature in b200ops/_lib.py"
    for name in _lib.SIGNATURES:
        assert name in declared_symbols(), f"{name} bound in python but not declared in the header"


def test_version_and_workspace_query(lib):
    assert lib.b200_version() >= 1000
    n = lib.b200_paged_attention_workspace_bytes(256, 32, 8, 128)
    assert n > 256 * 4 and n < 64 << 20


def test_no_cpu_fallback():
    from b200ops import ops
    x = torch.randn(4, 64)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.scaled_fp8_quant(x)
    kc = torch.zeros(2, 16, 2, 64, dtype=torch.bfloat16)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.reshape_and_cache_flash(torch.zeros(1, 2, 64, dtype=torch.bfloat16),
                                    torch.zeros(1, 2, 64, dtype=torch.bfloat16), kc, kc,
                                    torch.zeros(1, dtype=torch.int64), "auto", torch.ones(1),
                                    torch.ones(1))


@pytest.mark.skipif(torch.cuda.is_available(), reason="CPU-only check")
def test_compute_entry_point_fails_without_device(lib):
    from b200ops import _lib
    assert lib.b200_sm_count() == _lib.ERR_CUDA
    assert "no CUDA device" in _lib.last_error()


def test_argument_validation_before_launch(lib):
    from b200ops import _lib
    a = _lib.PagedAttnArgs()
    a.num_seqs, a.num_heads, a.num_kv_heads, a.head_size = 4, 30, 8, 128
    rc = lib.b200_paged_attention_decode(ctypes.byref(a), None)
    assert rc == _lib.ERR_INVALID and "divisible" in _lib.last_error()
    rc = lib.b200_reshape_and_cache_flash(None, None, None, None, None, 4, 2, 64, 16, 128, 128, 1, 1,
                                          1, _lib.BF16, _lib.F16, None, None, 0, None)
    assert rc == _lib.ERR_INVALID
    with pytest.raises(RuntimeError):
        _lib.check(rc, "reshape_and_cache_flash")


def test_no_cpu_fallback_round1_next_rows():
    """The operators added for the SURVEY 8f rows refuse CPU tensors like the rest (no CPU fallback), and their
    argument checks run before any launch (validation through the C ABI works without a device)."""
    from b200ops import _lib, ops
    lib = _lib.load()
    x = torch.randn(4, 256)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.per_token_group_quant_fp8(x.to(torch.bfloat16), 128)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.apply_top_k_top_p(x, torch.ones(4, dtype=torch.int32), None)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.greedy_sample(x)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.apply_repetition_penalties(x, x > 0, x < 0, torch.ones(4))
    a8 = torch.zeros(4, 256, dtype=torch.float8_e4m3fn)
    w8 = torch.zeros(128, 256, dtype=torch.float8_e4m3fn)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.cutlass_scaled_mm(a8, w8.t(), torch.ones(4, 2), torch.ones(2, 1), torch.bfloat16)
    # C-ABI validation before launch
    rc = lib.b200_fp8_quant_token_group(None, None, None, 4, 100, 64, 100, 100, 1, 1, 1e-10, -448.0, 448.0, 0,
                                        _lib.BF16, None)
    assert rc == _lib.ERR_INVALID and "divisible" in _lib.last_error()
    rc = lib.b200_fp8_gemm_blockscaled(None, None, None, None, None, 4, 128, 256, 256, 256, 128, 2, 1, 1, 2,
                                       _lib.BF16, 0, None)
    assert rc == _lib.ERR_INVALID and "null pointer" in _lib.last_error()
    rc = lib.b200_top_k_top_p(None, None, 0, None, 4, 0, 0, None)
    assert rc == _lib.ERR_INVALID
    rc = lib.b200_argmax(None, None, 4, 100, 100, _lib.I32, None)
    assert rc == _lib.ERR_INVALID and "bf16 / f16 / f32" in _lib.last_error()
