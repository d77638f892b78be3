"""ctypes binding of libb200ops.so (the C ABI declared in include/b200ops.h).

The library is built in-tree (``make -C vllm-fork_b200`` or ``__graft_entry__.build()``) and lives
next to this file.  There is NO CPU fallback: if the library is missing, or a compute entry point is
called without a CUDA device, a RuntimeError is raised.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200ops.so")

# element types (include/b200ops.h)
F32, F16, BF16, FP8_E4M3, U8, I32, I64 = range(7)

OK = 0
ERR_INVALID, ERR_UNSUPPORTED, ERR_CUDA, ERR_WORKSPACE = -1, -2, -3, -4

c_void_p, c_int, c_int64, c_float, c_size_t = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_size_t


class PagedAttnArgs(C.Structure):
    """struct b200_paged_attn_args (include/b200ops.h)."""

    _fields_ = [
        ("out", c_void_p),
        ("lse", c_void_p),
        ("q", c_void_p),
        ("k_cache", c_void_p),
        ("v_cache", c_void_p),
        ("block_table", c_void_p),
        ("seq_lens", c_void_p),
        ("k_scale", c_void_p),
        ("v_scale", c_void_p),
        ("alibi_slopes", c_void_p),
        ("workspace", c_void_p),
        ("workspace_bytes", c_size_t),
        ("num_blocks", c_int64),
        ("q_stride_s", c_int64),
        ("q_stride_h", c_int64),
        ("out_stride_s", c_int64),
        ("out_stride_h", c_int64),
        ("kv_block_stride", c_int64),
        ("kv_token_stride", c_int64),
        ("kv_head_stride", c_int64),
        ("bt_stride", c_int64),
        ("num_seqs", C.c_int32),
        ("num_heads", C.c_int32),
        ("num_kv_heads", C.c_int32),
        ("head_size", C.c_int32),
        ("block_size", C.c_int32),
        ("max_seq_len", C.c_int32),
        ("q_dtype", C.c_int32),
        ("kv_dtype", C.c_int32),
        ("sliding_window", C.c_int32),
        ("scale", c_float),
        ("softcap", c_float),
        ("output_scale", c_void_p),
    ]


class VarlenAttnArgs(C.Structure):
    """struct b200_varlen_attn_args (include/b200ops.h)."""

    _fields_ = [
        ("out", c_void_p),
        ("lse", c_void_p),
        ("q", c_void_p),
        ("k_cache", c_void_p),
        ("v_cache", c_void_p),
        ("cu_seqlens_q", c_void_p),
        ("seq_lens", c_void_p),
        ("block_table", c_void_p),
        ("k_scale", c_void_p),
        ("v_scale", c_void_p),
        ("alibi_slopes", c_void_p),
        ("output_scale", c_void_p),
        ("q_stride_t", c_int64),
        ("q_stride_h", c_int64),
        ("out_stride_t", c_int64),
        ("out_stride_h", c_int64),
        ("kv_block_stride", c_int64),
        ("kv_token_stride", c_int64),
        ("kv_head_stride", c_int64),
        ("bt_stride", c_int64),
        ("num_blocks", c_int64),
        ("num_seqs", C.c_int32),
        ("num_tokens", C.c_int32),
        ("num_heads", C.c_int32),
        ("num_kv_heads", C.c_int32),
        ("head_size", C.c_int32),
        ("block_size", C.c_int32),
        ("q_dtype", C.c_int32),
        ("kv_dtype", C.c_int32),
        ("causal", C.c_int32),
        ("sliding_window", C.c_int32),
        ("scale", c_float),
        ("softcap", c_float),
    ]


# name -> (restype, argtypes); every symbol include/b200ops.h declares must be listed here
# (tests/test_abi.py cross-checks this table against the header).
SIGNATURES = {
    "b200_last_error": (C.c_char_p, []),
    "b200_version": (c_int, []),
    "b200_sm_count": (c_int, []),
    "b200_set_pdl": (c_int, [c_int]),
    "b200_reshape_and_cache_flash": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_int64,
         c_int64, c_int64, c_int64, c_int64, c_int, c_int, c_void_p, c_void_p, c_int, c_void_p],
    ),
    "b200_reshape_and_cache": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_int,
         c_int64, c_int64, c_int, c_int, c_void_p, c_void_p, c_void_p],
    ),
    "b200_fp8_quant_static": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int64, c_int64, c_int, c_int, c_int64,
         c_int64, c_int, c_void_p],
    ),
    "b200_fp8_quant_dynamic_tensor": (
        c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int64, c_int64, c_int, c_void_p]),
    "b200_fp8_quant_dynamic_token": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int64, c_int64, c_int, c_void_p],
    ),
    "b200_fp8_quant_token_group": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int64, c_int64, c_int64, c_int64,
         c_float, c_float, c_float, c_int, c_int, c_void_p],
    ),
    "b200_fp8_gemm_blockscaled": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int64, c_int64, c_int64,
         c_int64, c_int64, c_int64, c_int64, c_int, c_int, c_void_p],
    ),
    "b200_argmax": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_int64, c_int, c_void_p]),
    "b200_top_k_top_p": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int64, c_int, c_int64, c_void_p]),
    "b200_apply_repetition_penalties": (
        c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p]),
    "b200_fp8_gemm": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int64,
         c_int64, c_int64, c_int, c_int, c_int, c_void_p],
    ),
    "b200_fp8_gemm_pair": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int64,
         c_int64, c_int64, c_int, c_int, c_int, c_int, c_void_p],
    ),
    "b200_fp8_gemm_splitk": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int64,
         c_int64, c_int64, c_int, c_int, c_int, c_int, c_void_p],
    ),
    "b200_fp8_gemm_silu_mul_fp8": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int64,
         c_int64, c_int64, c_int, c_int, c_int, c_void_p],
    ),
    "b200_fp8_gemm_v3_workspace_bytes": (c_size_t, []),
    "b200_fp8_gemm_v3": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int64,
         c_int64, c_int64, c_int, c_int, c_int, c_void_p, c_size_t, c_void_p],
    ),
    "b200_silu_and_mul": (
        c_int, [c_void_p, c_void_p, c_int64, c_int, c_int64, c_int64, c_int, c_void_p]),
    "b200_silu_and_mul_fp8": (
        c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int64, c_int64, c_int, c_void_p]),
    "b200_rms_norm": (
        c_int, [c_void_p, c_void_p, c_void_p, c_float, c_int64, c_int, c_int64, c_int, c_void_p]),
    "b200_fused_add_rms_norm": (
        c_int, [c_void_p, c_void_p, c_void_p, c_float, c_int64, c_int, c_int64, c_int, c_void_p]),
    "b200_rms_norm_static_fp8_quant": (
        c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_int64, c_int, c_int64, c_int,
                c_void_p]),
    "b200_fused_add_rms_norm_static_fp8_quant": (
        c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_int64, c_int, c_int64,
                c_int, c_void_p]),
    "b200_rotary_embedding": (
        c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int64, c_int64, c_int64,
                c_int, c_int, c_int, c_int, c_void_p]),
    "b200_rope_and_cache_flash": (
        c_int, [c_void_p] * 8 + [c_int64, c_int, c_int, c_int, c_int, c_int] + [c_int64] * 6
        + [c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_void_p]),
    "b200_moe_align_block_size": (
        c_int, [c_void_p, c_int, c_int64, c_int, c_int, c_void_p, c_int64, c_void_p, c_int64, c_void_p,
                c_void_p, c_void_p, c_void_p]),
    "b200_moe_topk_softmax": (
        c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_int,
                c_void_p]),
    "b200_moe_sum": (
        c_int, [c_void_p, c_void_p, c_int64, c_int, c_int, c_int64, c_int64, c_void_p, c_int, c_void_p,
                c_int, c_void_p]),
    "b200_moe_permute_fp8": (
        c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int64, c_int,
                c_int64, c_int, c_void_p]),
    "b200_moe_unpermute_reduce": (
        c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int64, c_int,
                c_void_p]),
    "b200_moe_grouped_gemm_fp8": (
        c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                c_int64, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "b200_moe_permute_rows": (
        c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int64, c_int64, c_int,
                c_int64, c_int64, c_int64, c_void_p]),
    "b200_moe_grouped_gemm_fp8_blockscaled": (
        c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int,
                c_int, c_int, c_int, c_void_p]),
    "b200_nvfp4_quant": (
        c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int64, c_int, c_int, c_void_p]),
    "b200_nvfp4_gemm": (
        c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int64,
                c_int64, c_int64, c_int, c_void_p]),
    "b200_moe_mm_data": (
        c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int64, c_int,
                c_int, c_int, c_int, c_int, c_void_p]),
    "b200_shuffle_rows": (
        c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64, c_void_p]),
    "b200_nvfp4_experts_quant": (
        c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int64, c_int, c_int64,
                c_int, c_void_p]),
    "b200_silu_and_mul_nvfp4_experts_quant": (
        c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int64, c_int, c_int64,
                c_int, c_void_p]),
    "b200_nvfp4_moe_gemm": (
        c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                c_int64, c_int, c_int, c_int, c_int, c_void_p]),
    "b200_paged_attention_v1": (
        c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_float, c_void_p,
                c_void_p, c_int, c_int, c_void_p, c_int, c_int, c_void_p, c_void_p, c_int64, c_int64,
                c_int64, c_int64, c_int, c_void_p]),
    "b200_paged_attention_v2": (
        c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                c_int, c_float, c_void_p, c_void_p, c_int, c_int, c_void_p, c_int, c_int, c_void_p,
                c_void_p, c_int64, c_int64, c_int64, c_int64, c_int, c_int, c_void_p]),
    "b200_silu_and_mul_nvfp4": (
        c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int64, c_int, c_void_p]),
    "b200_merge_attn_states": (
        c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int]
        + [c_int64] * 12 + [c_int, c_int, c_void_p]),
    "b200_ipc_alloc": (c_int, [c_size_t, C.POINTER(c_void_p), c_void_p]),
    "b200_ipc_open": (c_int, [c_void_p, C.POINTER(c_void_p)]),
    "b200_ipc_close": (c_int, [c_void_p]),
    "b200_ipc_free": (c_int, [c_void_p]),
    "b200_allreduce_meta_bytes": (c_size_t, []),
    "b200_allreduce_init": (c_int, [C.POINTER(c_void_p), c_int, c_int, C.POINTER(c_void_p), c_size_t, c_size_t]),
    "b200_allreduce_heap": (c_int, [c_void_p, C.POINTER(c_size_t), C.POINTER(c_size_t)]),
    "b200_allreduce_register": (c_int, [c_void_p, C.POINTER(c_void_p), c_size_t]),
    "b200_allreduce_destroy": (c_int, [c_void_p]),
    "b200_allreduce_set_multicast": (c_int, [c_void_p, c_void_p]),
    "b200_allreduce_has_multicast": (c_int, [c_void_p]),
    "b200_allreduce_run": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p]),
    "b200_allreduce_fused_add_rms_norm": (
        c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_int64, c_int, c_int,
                c_void_p]),
    "b200_debug_set_trace": (c_int, [c_void_p]),
    "b200_dcp_lse_merge": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int,
                                   c_int, c_int, c_void_p]),
    "b200_paged_attention_workspace_bytes": (c_size_t, [c_int, c_int, c_int, c_int]),
    "b200_paged_attention_decode": (c_int, [C.POINTER(PagedAttnArgs), c_void_p]),
    "b200_paged_attention_decode_static": (c_int, [C.POINTER(PagedAttnArgs), c_void_p]),
    "b200_paged_attention_varlen": (c_int, [C.POINTER(VarlenAttnArgs), c_void_p]),
}

_lib = None


def load() -> C.CDLL:
    """Load libb200ops.so (once). Raises RuntimeError if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"b200ops: {LIB_PATH} not found - build it with `make -C vllm-fork_b200` "
            "(or __graft_entry__.build()). There is no CPU / PyTorch fallback."
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError => ABI mismatch, fail loudly
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error() -> str:
    return load().b200_last_error().decode("utf-8", "replace")


def check(rc: int, op: str) -> None:
    """Turn a negative return code into the exception the reference op would raise."""
    if rc == OK:
        return
    msg = last_error() or f"{op} failed with code {rc}"
    if rc == ERR_UNSUPPORTED:
        raise NotImplementedError(msg)
    raise RuntimeError(msg)
