"""TEST INFRASTRUCTURE ONLY -- loader for oracle/_ref/refcpu.so, the reference's own CPU paged
attention (csrc/cpu/cpu_attn.cpp) compiled UNMODIFIED by oracle/build_ref_cpu_attn.sh.

Used (a) to pin oracle/ref_ops.py (tests/golden/make_golden.py), (b) as the timed CPU baseline /
`bench.py --impl reference`. Never imported by the product package.
"""
from __future__ import annotations

import os

import torch

_SO = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "refcpu.so")
_loaded = False


def available() -> bool:
    return os.path.exists(_SO)


def load() -> None:
    global _loaded
    if not _loaded:
        if not available():
            raise RuntimeError(f"{_SO} missing: run oracle/build_ref_cpu_attn.sh (needs /root/reference)")
        torch.ops.load_library(_SO)
        _loaded = True


def pick_isa(block_size: int, head_size: int, kv_cache_dtype: str = "auto") -> str:
    """vllm/v1/attention/backends/cpu_attn.py:461-500 (_get_attn_isa), x86 subset."""
    if kv_cache_dtype != "auto":
        return "vec"
    if block_size % 32 != 0:
        return "vec16"
    return "vec"


def make_packed_cache(num_blocks, num_kv_heads, block_size, head_size, dtype):
    """CPU backend cache layout (tests/kernels/attention/test_cpu_attn.py:488-494)."""
    packed = torch.empty(num_blocks, num_kv_heads, block_size, head_size * 2, dtype=dtype)
    packed = packed.view(num_blocks, num_kv_heads, block_size * 2, -1)
    return packed.chunk(2, dim=2)


def reshape_and_cache(key, value, key_cache, value_cache, slot_mapping, isa, k_scale=1.0,
                      v_scale=1.0, kv_cache_dtype="auto"):
    load()
    torch.ops._b200_refcpu.cpu_attn_reshape_and_cache(key, value, key_cache, value_cache,
                                                      slot_mapping, isa, k_scale, v_scale,
                                                      kv_cache_dtype)


def scheduler_metadata(num_reqs, num_heads, num_kv_heads, head_dim, seq_lens, dtype,
                       query_start_loc, isa, enable_kv_split=True, sliding_window=-1,
                       kv_cache_dtype="auto"):
    load()
    return torch.ops._b200_refcpu.get_scheduler_metadata(
        num_reqs, num_heads, num_kv_heads, head_dim, seq_lens, dtype, query_start_loc, True,
        sliding_window, isa, enable_kv_split, None, kv_cache_dtype)


def attention(query, key_cache, value_cache, output, query_start_loc, seq_lens, scale, block_table,
              metadata, alibi_slopes=None, sliding_window=-1, softcap=0.0, k_scale=1.0, v_scale=1.0,
              kv_cache_dtype="auto"):
    load()
    torch.ops._b200_refcpu.cpu_attention_with_kv_cache(
        query, key_cache, value_cache, output, query_start_loc, seq_lens, scale, True, alibi_slopes,
        sliding_window, block_table, softcap, metadata, None, None, k_scale, v_scale, kv_cache_dtype)
