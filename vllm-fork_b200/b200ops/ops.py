"""Host-side mirror of the reference operator surface (vllm/_custom_ops.py) over libb200ops.so.

Same function names, argument meaning and error behaviour as the reference's Python wrappers, so that
parity tests read like the reference's own kernel tests and a maintainer can rebind
``vllm._custom_ops.<name> = b200ops.ops.<name>``.  PyTorch is used only for
device memory and the current CUDA stream.  There is no CPU fallback.
"""
from __future__ import annotations

import os

import torch

from . import _lib
from ._lib import BF16, F16, F32, FP8_E4M3

FP8_DTYPE = torch.float8_e4m3fn
_FP8_KV_NAMES = ("fp8", "fp8_e4m3")


def _dt(t: torch.Tensor) -> int:
    if t.dtype == torch.bfloat16:
        return BF16
    if t.dtype == torch.float16:
        return F16
    if t.dtype == torch.float32:
        return F32
    if t.dtype in (torch.float8_e4m3fn, torch.uint8):
        return FP8_E4M3
    raise RuntimeError(f"b200ops: unsupported dtype {t.dtype}")


def _cuda(*ts: torch.Tensor | None) -> None:
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError(
                "b200ops: tensors must live on a CUDA device (there is no CPU fallback)")


def _p(t: torch.Tensor | None) -> int | None:
    return None if t is None else t.data_ptr()


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


# ------------------------------------------------------------------------------------------------
# KV-cache write                                      reference: vllm/_custom_ops.py:2614-2633
# ------------------------------------------------------------------------------------------------
def reshape_and_cache_flash(
    key: torch.Tensor,
    value: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    slot_mapping: torch.Tensor,
    kv_cache_dtype: str,
    k_scale: torch.Tensor,
    v_scale: torch.Tensor,
) -> None:
    """key/value [T, H, D]; key_cache/value_cache [num_blocks, block_size, H, D] views (NHD or HND
    strides); slot_mapping int64 [T'] (T' <= T, -1 = skip); kv_cache_dtype "auto" | "fp8" | "fp8_e4m3";
    k_scale/v_scale float32 [1] or [H]."""
    _cuda(key, value, key_cache, value_cache, slot_mapping, k_scale, v_scale)
    lib = _lib.load()
    if key.dim() != 3 or value.dim() != 3 or key_cache.dim() != 4:
        raise RuntimeError("reshape_and_cache_flash: key/value must be 3-D and the caches 4-D")
    if key.stride(2) != 1 or value.stride(2) != 1 or key_cache.stride(3) != 1:
        raise RuntimeError("reshape_and_cache_flash: last dimension must be contiguous")
    if key_cache.stride(0) != value_cache.stride(0):
        raise RuntimeError("reshape_and_cache_flash: key_cache/value_cache block strides differ")
    if slot_mapping.dtype != torch.int64:
        raise RuntimeError("reshape_and_cache_flash: slot_mapping must be int64")
    if k_scale.shape != v_scale.shape:
        raise RuntimeError("k_scale and v_scale must have the same shape")
    num_heads, head_size = key.shape[1], key.shape[2]
    if k_scale.numel() not in (1, num_heads):
        raise RuntimeError("k_scale and v_scale must be of shape [1] or [num_heads]")
    if kv_cache_dtype == "auto":
        cache_dt = _dt(key)
        if key_cache.dtype != key.dtype:
            raise RuntimeError("reshape_and_cache_flash: kv_cache_dtype 'auto' needs cache dtype == key dtype")
    elif kv_cache_dtype in _FP8_KV_NAMES:
        cache_dt = FP8_E4M3
        if key_cache.element_size() != 1:
            raise RuntimeError("reshape_and_cache_flash: fp8 kv cache must have 1-byte elements")
    else:
        raise RuntimeError(f"Unsupported data type of kv cache: {kv_cache_dtype}")
    rc = lib.b200_reshape_and_cache_flash(
        _p(key), _p(value), _p(key_cache), _p(value_cache), _p(slot_mapping),
        slot_mapping.shape[0], num_heads, head_size, key_cache.shape[1],
        key.stride(0), value.stride(0), key_cache.stride(0), key_cache.stride(1),
        key_cache.stride(2), _dt(key), cache_dt, _p(k_scale), _p(v_scale),
        1 if k_scale.numel() > 1 else 0, _stream())
    _lib.check(rc, "reshape_and_cache_flash")


def reshape_and_cache(
    key: torch.Tensor,
    value: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    slot_mapping: torch.Tensor,
    kv_cache_dtype: str,
    k_scale: torch.Tensor,
    v_scale: torch.Tensor,
) -> None:
    """Legacy layout: key_cache [nb, H, D/x, bs, x], value_cache [nb, H, D, bs] (contiguous)."""
    _cuda(key, value, key_cache, value_cache, slot_mapping, k_scale, v_scale)
    lib = _lib.load()
    if key_cache.dim() != 5 or value_cache.dim() != 4:
        raise RuntimeError("reshape_and_cache: key_cache must be 5-D and value_cache 4-D")
    if not (key_cache.is_contiguous() and value_cache.is_contiguous()):
        raise RuntimeError("reshape_and_cache: caches must be contiguous")
    if kv_cache_dtype == "auto":
        cache_dt = _dt(key)
    elif kv_cache_dtype in _FP8_KV_NAMES:
        cache_dt = FP8_E4M3
    else:
        raise RuntimeError(f"Unsupported data type of kv cache: {kv_cache_dtype}")
    rc = lib.b200_reshape_and_cache(
        _p(key), _p(value), _p(key_cache), _p(value_cache), _p(slot_mapping),
        slot_mapping.shape[0], key.shape[1], key.shape[2], key_cache.shape[3], key_cache.shape[4],
        key.stride(0), value.stride(0), _dt(key), cache_dt, _p(k_scale), _p(v_scale), _stream())
    _lib.check(rc, "reshape_and_cache")


# ------------------------------------------------------------------------------------------------
# FP8 activation quantisation                         reference: vllm/_custom_ops.py:1832-1898
# ------------------------------------------------------------------------------------------------
def static_scaled_fp8_quant(result: torch.Tensor, input: torch.Tensor, scale: torch.Tensor,
                            group_shape: tuple[int, int] | None = None) -> None:
    """torch.ops._C.static_scaled_fp8_quant (csrc/libtorch_stable/.../fp8/common.cu:183-335)."""
    _cuda(result, input, scale)
    lib = _lib.load()
    if input.stride(-1) != 1:
        raise RuntimeError("last dimension of input must be contiguous")
    if result.stride(-1) != 1:
        raise RuntimeError("last dimension of output must be contiguous")
    hidden = input.shape[-1]
    num_tokens = input.numel() // hidden if hidden else 0
    if scale.dim() == 0 or scale.numel() == 1:
        gm, gn, si, sj = max(num_tokens, 1), hidden, 0, 0
    elif scale.dim() == 1:
        if group_shape is None:
            raise RuntimeError(
                "1D scale requires explicit group_shape to disambiguate per-channel vs per-token "
                "quantization. Use group_shape=(-1, 1) for per-channel or group_shape=(1, -1) for "
                "per-token.")
        if len(group_shape) != 2:
            raise RuntimeError(f"group_shape must have exactly 2 elements, got {len(group_shape)}")
        gm = num_tokens if group_shape[0] == -1 else int(group_shape[0])
        gn = hidden if group_shape[1] == -1 else int(group_shape[1])
        em, en = num_tokens // gm, hidden // gn
        if scale.numel() != em * en:
            raise RuntimeError(
                f"1D scale length ({scale.numel()}) does not match expected size ({em * en}) for "
                f"group_shape {tuple(group_shape)} with input shape ({num_tokens}, {hidden})")
        if em == 1:
            si, sj = 0, scale.stride(0)
        elif en == 1:
            si, sj = scale.stride(0), 0
        else:
            raise RuntimeError(
                "1D scale can only be used when one of the scale dimensions is 1. For 2D group "
                "scaling, use a 2D scale tensor.")
    elif scale.dim() == 2:
        s0, s1 = scale.shape
        if num_tokens % s0 != 0:
            raise RuntimeError(f"num_tokens ({num_tokens}) must be divisible by scale.size(0) ({s0})")
        if hidden % s1 != 0:
            raise RuntimeError(f"hidden_size ({hidden}) must be divisible by scale.size(1) ({s1})")
        gm, gn = num_tokens // s0, hidden // s1
        if group_shape is not None:
            if len(group_shape) != 2:
                raise RuntimeError(
                    f"group_shape must have exactly 2 elements, got {len(group_shape)}")
            egm = num_tokens if group_shape[0] == -1 else int(group_shape[0])
            egn = hidden if group_shape[1] == -1 else int(group_shape[1])
            if (egm, egn) != (gm, gn):
                raise RuntimeError(
                    f"Explicit group_shape {tuple(group_shape)} does not match inferred group shape "
                    f"({gm}, {gn}) from 2D scale tensor shape ({s0}, {s1})")
        si, sj = scale.stride(0), scale.stride(1)
    else:
        raise RuntimeError(f"scale must be 0D, 1D, or 2D tensor, but got {scale.dim()}D")
    rc = lib.b200_fp8_quant_static(
        _p(result), _p(input), _p(scale), num_tokens, hidden,
        input.stride(-2) if input.dim() > 1 else hidden,
        result.stride(-2) if result.dim() > 1 else hidden,
        gm, gn, si, sj, _dt(input), _stream())
    _lib.check(rc, "static_scaled_fp8_quant")


def dynamic_scaled_fp8_quant(result: torch.Tensor, input: torch.Tensor, scale: torch.Tensor) -> None:
    """torch.ops._C.dynamic_scaled_fp8_quant (common.cu:337-381): scale[0] is written."""
    _cuda(result, input, scale)
    lib = _lib.load()
    if input.stride(-1) != 1:
        raise RuntimeError("last dimension of input must be contiguous")
    if result.stride(-1) != 1:
        raise RuntimeError("last dimension of output must be contiguous")
    hidden = input.shape[-1]
    num_tokens = input.numel() // hidden if hidden else 0
    rc = lib.b200_fp8_quant_dynamic_tensor(
        _p(result), _p(input), _p(scale), num_tokens, hidden,
        input.stride(-2) if input.dim() > 1 else hidden,
        result.stride(-2) if result.dim() > 1 else hidden, _dt(input), _stream())
    _lib.check(rc, "dynamic_scaled_fp8_quant")


def dynamic_per_token_scaled_fp8_quant(result: torch.Tensor, input: torch.Tensor,
                                       scale: torch.Tensor, scale_ub: torch.Tensor | None) -> None:
    """torch.ops._C.dynamic_per_token_scaled_fp8_quant (common.cu:383-422): scale[t] is written."""
    _cuda(result, input, scale, scale_ub)
    lib = _lib.load()
    if input.stride(-1) != 1:
        raise RuntimeError("last dimension of input must be contiguous")
    if result.stride(-1) != 1:
        raise RuntimeError("last dimension of output must be contiguous")
    hidden = input.shape[-1]
    num_tokens = input.numel() // hidden if hidden else 0
    rc = lib.b200_fp8_quant_dynamic_token(
        _p(result), _p(input), _p(scale), _p(scale_ub), num_tokens, hidden,
        input.stride(-2) if input.dim() > 1 else hidden,
        result.stride(-2) if result.dim() > 1 else hidden, _dt(input), _stream())
    _lib.check(rc, "dynamic_per_token_scaled_fp8_quant")


def scaled_fp8_quant(
    input: torch.Tensor,
    scale: torch.Tensor | None = None,
    num_token_padding: int | None = None,
    scale_ub: torch.Tensor | None = None,
    use_per_token_if_dynamic: bool = False,
    output: torch.Tensor | None = None,
    group_shape: tuple[int, int] | None = None,
) -> tuple[torch.Tensor, torch.Tensor]:
    """Same contract as vllm._custom_ops.scaled_fp8_quant (vllm/_custom_ops.py:1832-1898)."""
    assert input.ndim == 2
    shape = tuple(input.shape)
    if num_token_padding:
        shape = (max(num_token_padding, input.shape[0]), shape[1])
    if output is None:
        output = torch.empty(shape, device=input.device, dtype=FP8_DTYPE)
    else:
        assert num_token_padding is None, "padding not supported if output passed in"
        assert output.dtype == FP8_DTYPE
    if scale is None:
        if use_per_token_if_dynamic:
            scale = torch.empty((shape[0], 1), device=input.device, dtype=torch.float32)
            dynamic_per_token_scaled_fp8_quant(output, input, scale, scale_ub)
        else:
            scale = torch.empty(1, device=input.device, dtype=torch.float32)
            dynamic_scaled_fp8_quant(output, input, scale)
    else:
        static_scaled_fp8_quant(output, input, scale, group_shape)
    return output, scale


def per_token_group_fp8_quant(input: torch.Tensor, output_q: torch.Tensor, output_s: torch.Tensor,
                              group_size: int, eps: float, fp8_min: float, fp8_max: float,
                              scale_ue8m0: bool, dummy_is_scale_transposed: bool = False,
                              dummy_is_tma_aligned: bool = False) -> None:
    """torch.ops._C.per_token_group_fp8_quant (per_token_group_quant.cu:195-300): the scale layout is
    taken from output_s' strides, exactly what the two dummy flags describe in the reference."""
    _cuda(input, output_q, output_s)
    if not input.is_contiguous() or not output_q.is_contiguous():
        raise RuntimeError("per_token_group_fp8_quant: input and output_q must be contiguous")
    if output_s.dtype != torch.float32:
        raise RuntimeError("per_token_group_fp8_quant: output_s must be float32")
    hidden = input.shape[-1]
    if hidden % group_size != 0:
        raise RuntimeError(f"per_token_group_fp8_quant: the last dimension of `x` {hidden} must be "
                           f"divisible by `group_size` {group_size}")
    tokens = input.numel() // hidden if hidden else 0
    s2 = output_s if output_s.ndim == 2 else output_s.view(tokens, -1) if output_s.is_contiguous() else None
    if s2 is None or s2.ndim != 2 or tuple(s2.shape) != (tokens, hidden // group_size):
        raise RuntimeError("per_token_group_fp8_quant: output_s must be a 2-D [tokens, groups] view")
    rc = _lib.load().b200_fp8_quant_token_group(
        _p(output_q), _p(s2), _p(input), tokens, hidden, group_size, hidden, hidden,
        s2.stride(0), s2.stride(1), float(eps), float(fp8_min), float(fp8_max), int(bool(scale_ue8m0)),
        _dt(input), _stream())
    _lib.check(rc, "per_token_group_fp8_quant")


def per_token_group_quant_fp8(x: torch.Tensor, group_size: int, eps: float = 1e-10,
                              dtype: torch.dtype | None = None, column_major_scales: bool = False,
                              tma_aligned_scales: bool = False, out_q: torch.Tensor | None = None,
                              use_ue8m0: bool = False) -> tuple[torch.Tensor, torch.Tensor]:
    """Same contract as fp8_utils.per_token_group_quant_fp8 (fp8_utils.py:533-660): allocates the
    scale tensor row-major, column-major or TMA-aligned column-major and runs the CUDA kernel."""
    dtype = FP8_DTYPE if dtype is None else dtype
    assert dtype == FP8_DTYPE, "only torch.float8_e4m3fn is supported"
    assert x.shape[-1] % group_size == 0, (
        f"the last dimension of `x` {x.shape[-1]} must be divisible by `group_size` {group_size}")
    assert x.stride(-1) == 1, "`x` groups must be contiguous"
    assert out_q is None or out_q.shape == x.shape
    x_q = out_q if out_q is not None else torch.empty(x.shape, device=x.device, dtype=dtype)
    if column_major_scales:
        if tma_aligned_scales:
            assert x.dim() == 2, "TMA-aligned scales: 2-D activations"
            m, sf_k = x.shape[-2], x.shape[-1] // group_size
            tma_m = (m + 3) // 4 * 4  # get_tma_aligned_size(m, 4): 16-byte aligned columns of floats
            x_s = torch.empty_strided((m, sf_k), (1, tma_m), device=x.device, dtype=torch.float32)
        else:
            shape = x.shape[:-2] + (x.shape[-1] // group_size, x.shape[-2])
            x_s = torch.empty(shape, device=x.device, dtype=torch.float32).permute(-1, -2)
    else:
        x_s = torch.empty(x.shape[:-1] + (x.shape[-1] // group_size,), device=x.device,
                          dtype=torch.float32)
    fi = torch.finfo(dtype)
    xc = x if x.is_contiguous() else x.contiguous()
    per_token_group_fp8_quant(xc, x_q, x_s, group_size, eps, fi.min, fi.max, use_ue8m0,
                              column_major_scales, tma_aligned_scales)
    return x_q, x_s


# ------------------------------------------------------------------------------------------------
# FP8 W8A8 GEMM                                        reference: vllm/_custom_ops.py:756-804
# ------------------------------------------------------------------------------------------------
def set_pdl(level: int) -> int:
    """Programmatic dependent launch between consecutive b200 kernels of a stream (include/b200ops.h
    b200_set_pdl): 0 off, 1 prologue overlap (default), 2 + GEMM weight tiles requested before the wait (the
    caller promises that GEMM weights are not written by the kernel preceding the GEMM). Returns the previous
    level. Results are identical at every level."""
    return int(_lib.load().b200_set_pdl(int(level)))


def cutlass_scaled_mm_supports_fp8(cuda_device_capability: int) -> bool:
    """torch.ops._C.cutlass_scaled_mm_supports_fp8: this build targets sm_100a only."""
    return cuda_device_capability // 10 == 10


# which FP8 GEMM kernel cutlass_scaled_mm launches: "v2" = one CTA per tile (csrc/gemm_fp8.cu),
# "v3" = persistent CTA pairs (cta_group::2, UMMA M = 256), whole-tile raster or stream-K (csrc/gemm_fp8_v3.cu),
# "v4" = split-K over the CTA pairs of a cluster (csrc/gemm_fp8_v4.cu), "v5" = CTA pairs of UMMA M = 128 sharing the
# batch tile (csrc/gemm_fp8_v5.cu). All are tcgen05 kernels with the same arithmetic and all are parity-green; this
# is a tuning switch (env B200_GEMM_IMPL), not a fallback. "auto" (default) follows measurements on B200.
GEMM_IMPL = os.environ.get("B200_GEMM_IMPL", "auto")
LAST_GEMM_IMPL = ""
_gemm_ws: dict[tuple, torch.Tensor] = {}
_sms: dict[str, int] = {}


def _sm_count(device: torch.device) -> int:
    key = str(device)
    if key not in _sms:
        _sms[key] = torch.cuda.get_device_properties(device).multi_processor_count
    return _sms[key]


def pick_gemm_impl(M: int, N: int, K: int, out_tma_addressable: bool, sms: int) -> str:
    """Which FP8 GEMM kernel `cutlass_scaled_mm` launches for an [M, K] x [K, N] problem on a GPU with `sms` SMs.
    Measured on B200 (profiles/r2t_gemm_impls.md, r3d_vs_wheel_gemm.md; CUDA-graph timed, weights rotated beyond L2):
     * M > 256, or at least as many 256-row weight tiles as CTA pairs (gate_up) -> "v3" (persistent pairs);
     * batches <= 64 with few weight tiles -> "v4" (split-K over a cluster: the fp32 partial exchange is small);
     * every other decode batch -> "v5" (CTA pairs of UMMA M = 128: half the bytes pulled per SM; 10-25 % faster
       than the one-CTA kernel on the qkv / o / down shapes at batch 128-256);
     * "v2" remains for outputs that a TMA store cannot address (unaligned base or row stride).
    K does not enter the rule today (it did in round 1: long-K shapes went to v4 at every batch)."""
    pairs = sms // 2
    tiles = (N + 255) // 256
    if M > 256 or tiles >= pairs:
        return "v3"
    if M <= 64 and tiles * 2 <= pairs and out_tma_addressable:
        return "v4"
    if out_tma_addressable:
        return "v5"
    return "v2"


def _gemm_workspace(device: torch.device) -> torch.Tensor:
    """Zero-initialised stream-K workspace of the v3 kernel. The kernel leaves its flags at zero, so one buffer
    serves every call that runs on ONE stream; two v3 GEMMs running concurrently on different streams would
    corrupt each other's flags, hence one buffer per (device, stream) for eager launches. Under stream capture
    the per-device buffer is used (no allocation or memset is recorded into the graph): graphs captured from
    this module must not be replayed concurrently with each other on one device."""
    if torch.cuda.is_current_stream_capturing():
        key = (str(device), None)
    else:
        key = (str(device), torch.cuda.current_stream(device).cuda_stream)
    ws = _gemm_ws.get(key)
    if ws is None:
        # (a first call under capture allocates + zeroes inside the graph's pool, as before: legal, one memset node)
        ws = torch.zeros(_lib.load().b200_fp8_gemm_v3_workspace_bytes(), dtype=torch.uint8, device=device)
        _gemm_ws[key] = ws
        if key[1] is not None:   # warm-up call outside capture: the capture-time buffer is made now as well
            _gemm_ws.setdefault((str(device), None), torch.zeros_like(ws))
    return ws


def _cutlass_scaled_mm_blockwise(a, b, scale_a, scale_b, out_dtype, bias, out, target_shape, split=0):
    """The blockwise branch of torch.ops._C.cutlass_scaled_mm (c3x/scaled_mm_helper.hpp:40-60; error text kept)."""
    M, K = a.shape
    N = b.shape[1]
    if scale_a.dim() != 2:
        raise RuntimeError("a scale must be 2d tensor.")
    if scale_b.dim() != 2:
        raise RuntimeError("b scale must be 2d tensor.")
    kb, nb = (K + 127) // 128, (N + 127) // 128
    if tuple(scale_a.shape) != (M, kb):
        raise RuntimeError("a_scale_group_shape must be [1, 128].")
    if tuple(scale_b.shape) != (kb, nb):
        raise RuntimeError("b_scale_group_shape must be [128, 128].")
    if bias is not None:
        raise RuntimeError("Bias not yet supported blockwise scaled_mm")
    if out is None:
        out = torch.empty((M, N), dtype=out_dtype, device=a.device)
    elif out.dtype != out_dtype or out.stride(-1) != 1:
        raise RuntimeError("cutlass_scaled_mm: bad out tensor")
    rc = _lib.load().b200_fp8_gemm_blockscaled(
        _p(out), _p(a), _p(b), _p(scale_a), _p(scale_b), M, N, K, a.stride(0) if M > 1 else K, b.stride(1),
        out.stride(0) if out.dim() > 1 else N, scale_a.stride(0), scale_a.stride(1), scale_b.stride(0),
        scale_b.stride(1), _dt(out), split, _stream())
    _lib.check(rc, "cutlass_scaled_mm")
    return out.view(*target_shape)


def cutlass_scaled_mm_supports_block_fp8(cuda_device_capability: int) -> bool:
    """torch.ops._C.cutlass_scaled_mm_supports_block_fp8 (scaled_mm_entry.cu:161-180): sm_100a only here."""
    return cuda_device_capability // 10 == 10


def w8a8_block_fp8_cutlass_mm(A: torch.Tensor, B: torch.Tensor, As: torch.Tensor, Bs: torch.Tensor,
                              block_size: list[int], output_dtype: torch.dtype = torch.float16) -> torch.Tensor:
    """kernels/linear/scaled_mm/cutlass.py:329-343 (`cutlass_scaled_mm(A, B, As, Bs, block_size, dtype)`):
    B is the weight [N, K], Bs its [N/128, K/128] scales."""
    assert list(block_size) == [128, 128], "only 128 x 128 weight blocks"
    return cutlass_scaled_mm(A, B.T, out_dtype=output_dtype, scale_a=As, scale_b=Bs.T)


def cutlass_scaled_mm(
    a: torch.Tensor,
    b: torch.Tensor,
    scale_a: torch.Tensor,
    scale_b: torch.Tensor,
    out_dtype: torch.dtype,
    bias: torch.Tensor | None = None,
    out: torch.Tensor | None = None,
    impl: str | None = None,
) -> torch.Tensor:
    """output = torch.mm(scale_a * a, scale_b * b).to(out_dtype) (+ bias).
    a: e4m3 [M, K] row-major; b: e4m3 [K, N] column-major (weight.t()); scale_a [1] / [M,1];
    scale_b [1] / [N,1] (or [1,N]); fp32 accumulate."""
    assert out_dtype is torch.bfloat16 or out_dtype is torch.float16
    assert bias is None or (bias.numel() == b.shape[1] and bias.dtype == out_dtype)
    _cuda(a, b, scale_a, scale_b, bias, out)
    lib = _lib.load()
    target_shape = (*a.shape[:-1], b.shape[1])
    a = a.view(-1, a.shape[-1])
    M, K = a.shape
    N = b.shape[1]
    if a.element_size() != 1 or b.element_size() != 1:
        raise RuntimeError("cutlass_scaled_mm: a and b must be fp8_e4m3")
    if b.shape[0] != K:
        raise RuntimeError("cutlass_scaled_mm: a.shape[1] must equal b.shape[0]")
    if b.stride(0) != 1 or (M > 1 and a.stride(1) != 1) or (M == 1 and K > 1 and a.stride(1) != 1):
        raise RuntimeError("cutlass_scaled_mm: a must be row-major and b column-major")
    if K % 16 != 0 or N % 16 != 0:
        raise RuntimeError("cutlass_scaled_mm: b.shape[0] and b.shape[1] must be multiples of 16")
    if scale_a.dtype != torch.float32 or scale_b.dtype != torch.float32:
        raise RuntimeError("cutlass_scaled_mm: scales must be float32")
    if scale_a.numel() not in (1, M) or scale_b.numel() not in (1, N):
        # blockwise scaling (scaled_mm_helper.hpp:40-60): a_s [M, ceil(K/128)], b_s [ceil(K/128), ceil(N/128)]
        # impl="split<n>" bounds the K split of the blockwise kernel (tests); default auto
        split = int(impl[5:]) if impl and impl.startswith("split") else 0
        return _cutlass_scaled_mm_blockwise(a, b, scale_a, scale_b, out_dtype, bias, out, target_shape, split)
    if not (scale_a.is_contiguous() and scale_b.is_contiguous()):
        raise RuntimeError("cutlass_scaled_mm: scales must be contiguous")
    if out is None:
        out = torch.empty((M, N), dtype=out_dtype, device=a.device)
    elif out.dtype != out_dtype or out.stride(-1) != 1:
        raise RuntimeError("cutlass_scaled_mm: bad out tensor")
    common = (_p(out), _p(a), _p(b), _p(scale_a), _p(scale_b), _p(bias), M, N, K,
              a.stride(0) if M > 1 else K, b.stride(1), out.stride(0) if out.dim() > 1 else N,
              1 if scale_a.numel() > 1 else 0, 1 if scale_b.numel() > 1 else 0, _dt(out))
    which = impl or GEMM_IMPL
    if which == "auto":
        which = pick_gemm_impl(M, N, K, out.data_ptr() % 16 == 0 and (out.stride(0) * 2) % 16 == 0,
                               _sm_count(a.device))
    global LAST_GEMM_IMPL
    LAST_GEMM_IMPL = which      # for benchmark tables: which kernel the dispatch picked
    if which.startswith("v5"):   # "v5" (auto batch tile) or "v5:<batch tile>": CTA pairs of UMMA M = 128
        bt = int(which[3:]) if len(which) > 3 else 0
        rc = lib.b200_fp8_gemm_pair(*common, bt, _stream())
    elif which.startswith("v4"):   # "v4" (auto split) or "v4:<split>"
        split = int(which[3:]) if len(which) > 3 else 0
        rc = lib.b200_fp8_gemm_splitk(*common, split, _stream())
    elif which == "v3":
        ws = _gemm_workspace(a.device)
        rc = lib.b200_fp8_gemm_v3(*common, _p(ws), ws.numel(), _stream())
    else:
        rc = lib.b200_fp8_gemm(*common, _stream())
    _lib.check(rc, "cutlass_scaled_mm")
    return out.view(*target_shape)


def scaled_mm_silu_and_mul_quant(result: torch.Tensor, a: torch.Tensor, w_gate_up: torch.Tensor,
                                 scale_a: torch.Tensor, scale_w: torch.Tensor, quant_scale: torch.Tensor,
                                 act_dtype: torch.dtype = torch.bfloat16, bias: torch.Tensor | None = None) -> None:
    """result (e4m3 [M, I]) = silu_and_mul_quant(cutlass_scaled_mm(a, w_gate_up.t(), ...).to(act_dtype), quant_scale)
    in one kernel (decode batches, M <= 256; bit-identical to the unfused pair).
    a e4m3 [M, K]; w_gate_up e4m3 [2I, K] row-major (gate rows, then up rows)."""
    _cuda(result, a, w_gate_up, scale_a, scale_w, quant_scale, bias)
    M, K = a.shape
    I2, Kw = w_gate_up.shape
    if Kw != K or I2 % 2 or a.element_size() != 1 or w_gate_up.element_size() != 1:
        raise RuntimeError("scaled_mm_silu_and_mul_quant: a [M,K] and w_gate_up [2I,K] must be fp8 with equal K")
    I = I2 // 2
    if tuple(result.shape) != (M, I) or result.element_size() != 1 or result.stride(1) != 1:
        raise RuntimeError("scaled_mm_silu_and_mul_quant: result must be fp8 [M, I]")
    if a.stride(1) != 1 or w_gate_up.stride(1) != 1:
        raise RuntimeError("scaled_mm_silu_and_mul_quant: a and w_gate_up must be row-major")
    if scale_a.numel() not in (1, M) or scale_w.numel() not in (1, I2):
        raise RuntimeError("scaled_mm_silu_and_mul_quant: scale_a must have 1 or M elements, scale_w 1 or 2I")
    rc = _lib.load().b200_fp8_gemm_silu_mul_fp8(
        _p(result), _p(a), _p(w_gate_up), _p(scale_a), _p(scale_w), _p(bias), _p(quant_scale), M, I, K,
        a.stride(0), w_gate_up.stride(0), result.stride(0), 1 if scale_a.numel() > 1 else 0,
        1 if scale_w.numel() > 1 else 0, BF16 if act_dtype is torch.bfloat16 else F16, _stream())
    _lib.check(rc, "scaled_mm_silu_and_mul_quant")


# ------------------------------------------------------------------------------------------------
# Gated activation                      reference: torch.ops._C.silu_and_mul / silu_and_mul_quant
# ------------------------------------------------------------------------------------------------
def silu_and_mul(result: torch.Tensor, input: torch.Tensor) -> None:
    """result [..., d] = silu(input[..., :d]) * input[..., d:]."""
    _cuda(result, input)
    lib = _lib.load()
    d = input.shape[-1] // 2
    if result.shape[-1] != d or input.stride(-1) != 1 or result.stride(-1) != 1:
        raise RuntimeError("silu_and_mul: input [..., 2d] and result [..., d], last dim contiguous")
    n = input.numel() // (2 * d) if d else 0
    i2, r2 = input.view(-1, 2 * d), result.view(-1, d)
    rc = lib.b200_silu_and_mul(_p(r2), _p(i2), n, d, i2.stride(0), r2.stride(0), _dt(input),
                               _stream())
    _lib.check(rc, "silu_and_mul")


def silu_and_mul_quant(result: torch.Tensor, input: torch.Tensor, scale: torch.Tensor) -> None:
    """result (e4m3) [..., d] = static_scaled_fp8_quant(silu_and_mul(input), scale)."""
    _cuda(result, input, scale)
    lib = _lib.load()
    d = input.shape[-1] // 2
    if result.shape[-1] != d or result.element_size() != 1:
        raise RuntimeError("silu_and_mul_quant: result must be fp8 [..., d]")
    n = input.numel() // (2 * d) if d else 0
    i2, r2 = input.view(-1, 2 * d), result.view(-1, d)
    rc = lib.b200_silu_and_mul_fp8(_p(r2), _p(i2), _p(scale), n, d, i2.stride(0), r2.stride(0),
                                   _dt(input), _stream())
    _lib.check(rc, "silu_and_mul_quant")


# ------------------------------------------------------------------------------------------------
# NVFP4 activation quantisation                         reference: vllm/_custom_ops.py:1523-1598
# ------------------------------------------------------------------------------------------------
def create_fp4_scale_tensor(m: int, n: int, device, is_sf_swizzled_layout: bool) -> torch.Tensor:
    """int32-packed e4m3 scales: [round_up(m,128), round_up(n/16,4)/4] swizzled, else uint8 [m, n/16]."""
    if is_sf_swizzled_layout:
        rm = (m + 127) // 128 * 128
        rn = (n // 16 + 3) // 4 * 4
        return torch.empty((rm, rn // 4), device=device, dtype=torch.int32)
    return torch.empty((m, n // 16), device=device, dtype=torch.uint8)


def scaled_fp4_quant(input: torch.Tensor, input_global_scale: torch.Tensor,
                     is_sf_swizzled_layout: bool = True, output: torch.Tensor | None = None,
                     output_scale: torch.Tensor | None = None) -> tuple[torch.Tensor, torch.Tensor]:
    """Returns (uint8 [m, n/2] packed e2m1, float8_e4m3fn scales in the swizzled / row-major layout).
    output / output_scale: caller-allocated results (the `scaled_fp4_quant.out` overload, torch_bindings.cpp:218)."""
    _cuda(input, input_global_scale, output, output_scale)
    assert input.ndim >= 1, f"input.ndim needs to be >= 1, but got {input.ndim}."
    other_dims = 1 if input.ndim == 1 else -1
    input = input.reshape(other_dims, input.shape[-1])
    m, n = input.shape
    assert n % 16 == 0, f"last dim has to be multiple of 16, but got {n}."
    assert input.dtype in (torch.float16, torch.bfloat16), (
        f"input.dtype needs to be fp16 or bf16 but got {input.dtype}.")
    if input.stride(1) != 1:
        input = input.contiguous()
    if output is None:
        output = torch.empty((m, n // 2), device=input.device, dtype=torch.uint8)
    elif tuple(output.shape) != (m, n // 2) or output.element_size() != 1 or not output.is_contiguous():
        raise RuntimeError(f"scaled_fp4_quant: output must be a contiguous 1-byte tensor [{m}, {n // 2}]")
    if output_scale is None:
        output_scale = create_fp4_scale_tensor(m, n, input.device, is_sf_swizzled_layout)
    else:
        need = create_fp4_scale_tensor(m, n, "meta", is_sf_swizzled_layout)
        if output_scale.numel() * output_scale.element_size() != need.numel() * need.element_size() or \
                not output_scale.is_contiguous():
            raise RuntimeError("scaled_fp4_quant: output_scale has the wrong size for this layout")
    rc = _lib.load().b200_nvfp4_quant(_p(output), _p(output_scale), _p(input), _p(input_global_scale),
                                      m, n, input.stride(0), 1 if is_sf_swizzled_layout else 0,
                                      _dt(input), _stream())
    _lib.check(rc, "scaled_fp4_quant")
    return output, output_scale.view(torch.float8_e4m3fn)


def silu_and_mul_nvfp4_quant(result: torch.Tensor, result_block_scale: torch.Tensor, input: torch.Tensor,
                             input_global_scale: torch.Tensor) -> None:
    """torch.ops._C.silu_and_mul_nvfp4_quant(result, result_block_scale, input, input_global_scale):
    result uint8 [m, d/2], result_block_scale swizzled e4m3 (as scaled_fp4_quant), input [m, 2d]."""
    _cuda(result, result_block_scale, input, input_global_scale)
    if input.dim() != 2 or input.stride(1) != 1:
        raise RuntimeError("silu_and_mul_nvfp4_quant: input must be 2-D with a contiguous last dim")
    m, d = input.shape[0], input.shape[1] // 2
    if d % 16 != 0:
        raise RuntimeError("silu_and_mul_nvfp4_quant: The N dimension must be multiple of 16.")
    if input.dtype not in (torch.float16, torch.bfloat16):
        raise RuntimeError("silu_and_mul_nvfp4_quant: Unsupported input data type for quantize_to_fp4.")
    rm, rn = (m + 127) // 128 * 128, (d // 16 + 3) // 4 * 4
    if (tuple(result.shape) != (m, d // 2) or result.dtype != torch.uint8 or not result.is_contiguous()
            or result_block_scale.numel() * result_block_scale.element_size() != rm * rn
            or not result_block_scale.is_contiguous()):
        raise RuntimeError("silu_and_mul_nvfp4_quant: result [m, d/2] uint8 and a swizzled block-scale buffer")
    rc = _lib.load().b200_silu_and_mul_nvfp4(_p(result), _p(result_block_scale), _p(input),
                                             _p(input_global_scale), m, d, input.stride(0), _dt(input),
                                             _stream())
    _lib.check(rc, "silu_and_mul_nvfp4_quant")


def cutlass_scaled_fp4_mm(a: torch.Tensor, b: torch.Tensor, block_scale_a: torch.Tensor,
                          block_scale_b: torch.Tensor, alpha: torch.Tensor, out_dtype: torch.dtype,
                          out: torch.Tensor | None = None) -> torch.Tensor:
    """out = alpha * (a (.) block_scale_a) @ (b (.) block_scale_b).T  (vllm/_custom_ops.py:733-745; checks of
    nvfp4_scaled_mm_kernels.cu:279-348). a uint8 [m, k/2], b uint8 [n, k/2] packed e2m1; block scales
    float8_e4m3fn in the swizzled layout scaled_fp4_quant returns; alpha f32 [1]."""
    _cuda(a, b, block_scale_a, block_scale_b, alpha, out)
    assert a.ndim == 2 and b.ndim == 2
    if a.dtype != torch.uint8 or b.dtype != torch.uint8:
        raise RuntimeError("cutlass_scaled_fp4_mm: a and b must be uint8 (packed fp4)")
    if block_scale_a.element_size() not in (1, 4) or block_scale_b.element_size() not in (1, 4):
        raise RuntimeError("cutlass_scaled_fp4_mm: scale_a / scale_b must be float8_e4m3fn")
    if alpha.dtype != torch.float32:
        raise RuntimeError("cutlass_scaled_fp4_mm: alpha must be float32")
    m, n = a.shape[0], b.shape[0]
    if a.shape[1] != b.shape[1]:
        raise RuntimeError(f"cutlass_scaled_fp4_mm: a and b shapes cannot be multiplied ({tuple(a.shape)} and {tuple(b.shape)})")
    k = a.shape[1] * 2
    if k % 32 != 0:
        raise RuntimeError(f"cutlass_scaled_fp4_mm: Expected k to be divisible by 32, but got a shape {tuple(a.shape)}")
    if n % 32 != 0:
        raise RuntimeError(f"cutlass_scaled_fp4_mm: Expected n to be divisible by 32, but got b shape {tuple(b.shape)}")
    if a.stride(1) != 1 or b.stride(1) != 1:
        raise RuntimeError("cutlass_scaled_fp4_mm: a and b must be row-major")
    ru = lambda x, y: (x + y - 1) // y * y
    sf_k = ru(k // 16, 4)
    for name, sf, rows in (("scale_a", block_scale_a, m), ("scale_b", block_scale_b, n)):
        if not sf.is_contiguous() or sf.numel() * sf.element_size() != ru(rows, 128) * sf_k:
            raise RuntimeError(
                f"cutlass_scaled_fp4_mm: {name} must be padded and swizzled to a shape ({ru(rows, 128)}x{sf_k}), "
                f"but got a shape {tuple(sf.shape)}")
    assert out_dtype is torch.bfloat16 or out_dtype is torch.float16
    if out is None:
        out = torch.empty((m, n), dtype=out_dtype, device=a.device)
    elif out.dtype != out_dtype or out.stride(-1) != 1 or tuple(out.shape) != (m, n):
        raise RuntimeError("cutlass_scaled_fp4_mm: bad out tensor")
    rc = _lib.load().b200_nvfp4_gemm(_p(out), _p(a), _p(b), _p(block_scale_a), _p(block_scale_b), _p(alpha),
                                     m, n, k, a.stride(0), b.stride(0), out.stride(0), _dt(out), _stream())
    _lib.check(rc, "cutlass_scaled_fp4_mm")
    return out


# ------------------------------------------------------------------------------------------------
# NVFP4 fused MoE operators (packed expert-sorted rows)     reference: vllm/_custom_ops.py:847-931,
# 1010-1046, 1601-1725; caller run_cutlass_moe_fp4 (fused_moe/experts/cutlass_moe.py:498-688)
# ------------------------------------------------------------------------------------------------
MAX_TOKENS_PER_EXPERT_FP4_MOE = int(os.environ.get("VLLM_MAX_TOKENS_PER_EXPERT_FP4_MOE", "163840"))
SWAP_AB_THRESHOLD = 64   # moe_data.cu:14


def get_cutlass_moe_mm_data(topk_ids: torch.Tensor, expert_offsets: torch.Tensor, problem_sizes1: torch.Tensor,
                            problem_sizes2: torch.Tensor, input_permutation: torch.Tensor,
                            output_permutation: torch.Tensor, num_experts: int, n: int, k: int,
                            blockscale_offsets: torch.Tensor | None = None, is_gated: bool = True) -> None:
    """torch.ops._C.get_cutlass_moe_mm_data (vllm/_custom_ops.py:847-899, moe_data.cu:236-289): all outputs
    int32, caller-allocated. Swap-AB problem sizes exactly when the reference would use them (FP8 path with
    at most 64 (token, k) pairs, moe_data.cu:254-255)."""
    _cuda(topk_ids, expert_offsets, problem_sizes1, problem_sizes2, input_permutation, output_permutation,
          blockscale_offsets)
    for name, t in (("topk_ids", topk_ids), ("expert_offsets", expert_offsets), ("problem_sizes1", problem_sizes1),
                    ("problem_sizes2", problem_sizes2), ("input_permutation", input_permutation),
                    ("output_permutation", output_permutation), ("blockscale_offsets", blockscale_offsets)):
        if t is not None and (t.dtype != torch.int32 or not t.is_contiguous()):
            raise RuntimeError(f"get_cutlass_moe_mm_data: {name} must be a contiguous int32 tensor")
    numel = topk_ids.numel()
    if expert_offsets.numel() < num_experts + 1 or problem_sizes1.numel() < 3 * num_experts or \
            problem_sizes2.numel() < 3 * num_experts or input_permutation.numel() < numel or \
            output_permutation.numel() < numel or \
            (blockscale_offsets is not None and blockscale_offsets.numel() < num_experts + 1):
        raise RuntimeError("get_cutlass_moe_mm_data: an output tensor is too small")
    swap_ab = blockscale_offsets is None and numel <= SWAP_AB_THRESHOLD
    topk = topk_ids.shape[-1] if topk_ids.dim() > 1 else 1
    rc = _lib.load().b200_moe_mm_data(_p(topk_ids), _p(expert_offsets), _p(problem_sizes1), _p(problem_sizes2),
                                      _p(input_permutation), _p(output_permutation), _p(blockscale_offsets),
                                      num_experts, numel, topk, n, k, 1 if is_gated else 0, 1 if swap_ab else 0,
                                      _stream())
    _lib.check(rc, "get_cutlass_moe_mm_data")


def shuffle_rows(input_tensor: torch.Tensor, dst2src_map: torch.Tensor,
                 output_tensor: torch.Tensor | None = None) -> torch.Tensor:
    """output[i, :] = input[dst2src_map[i], :] (vllm/_custom_ops.py:919-931, torch.ops._moe_C.shuffle_rows)."""
    _cuda(input_tensor, dst2src_map, output_tensor)
    if input_tensor.dim() != 2 or input_tensor.stride(1) != 1:
        raise RuntimeError("shuffle_rows: input must be 2-D with a contiguous last dim")
    if dst2src_map.dtype != torch.int32 or not dst2src_map.is_contiguous():
        raise RuntimeError("shuffle_rows: dst2src_map must be a contiguous int32 tensor")
    rows = dst2src_map.shape[0]
    if output_tensor is None:
        output_tensor = torch.empty((rows, input_tensor.shape[1]), device=input_tensor.device,
                                    dtype=input_tensor.dtype)
    elif tuple(output_tensor.shape) != (rows, input_tensor.shape[1]) or not output_tensor.is_contiguous() or \
            output_tensor.dtype != input_tensor.dtype:
        raise RuntimeError("shuffle_rows: bad output tensor")
    es = input_tensor.element_size()
    rc = _lib.load().b200_shuffle_rows(_p(output_tensor), _p(input_tensor), _p(dst2src_map), rows,
                                       input_tensor.shape[0], input_tensor.shape[1] * es,
                                       input_tensor.stride(0) * es, _stream())
    _lib.check(rc, "shuffle_rows")
    return output_tensor


def _experts_quant_out(op: str, gated: bool, output, output_scale, input, input_global_scale,
                       input_offset_by_experts, output_scale_offset_by_experts) -> None:
    _cuda(output, output_scale, input, input_global_scale, input_offset_by_experts, output_scale_offset_by_experts)
    if input.dim() != 2 or input.stride(1) != 1:
        raise RuntimeError(f"{op}: input must be 2-D with a contiguous last dim")
    if input.dtype not in (torch.float16, torch.bfloat16):
        raise RuntimeError(f"{op}: input.dtype needs to be fp16 or bf16 but got {input.dtype}.")
    m_topk, k = input.shape[0], input.shape[1] // (2 if gated else 1)
    if k % 16 != 0:
        raise RuntimeError(f"{op}: k must be a multiple of 16, but got {k}")
    for name, t in (("input_offset_by_experts", input_offset_by_experts),
                    ("output_scale_offset_by_experts", output_scale_offset_by_experts)):
        if t.dtype != torch.int32 or not t.is_contiguous():
            raise RuntimeError(f"{op}: {name} must be a contiguous int32 tensor")
    n_experts = input_offset_by_experts.numel() - 1
    if output_scale_offset_by_experts.numel() != n_experts + 1 or n_experts < 1:
        raise RuntimeError(f"{op}: the two offset tensors must both have num_experts + 1 entries")
    if input_global_scale is not None and (input_global_scale.dtype != torch.float32 or
                                           input_global_scale.numel() < n_experts):
        raise RuntimeError(f"{op}: input_global_scale must be float32 [num_experts]")
    if output.element_size() != 1 or not output.is_contiguous() or tuple(output.shape) != (m_topk, k // 2):
        raise RuntimeError(f"{op}: output must be a contiguous 1-byte tensor [{m_topk}, {k // 2}]")
    padded_k = (k // 16 + 3) // 4 * 4
    if not output_scale.is_contiguous() or (output_scale.numel() * output_scale.element_size()) % padded_k != 0:
        raise RuntimeError(f"{op}: output_scale must be contiguous with rows of {padded_k} bytes")
    fn = _lib.load().b200_silu_and_mul_nvfp4_experts_quant if gated else _lib.load().b200_nvfp4_experts_quant
    rc = fn(_p(output), _p(output_scale), _p(input), _p(input_global_scale), _p(input_offset_by_experts),
            _p(output_scale_offset_by_experts), n_experts, m_topk, k, input.stride(0), _dt(input), _stream())
    _lib.check(rc, op)


def scaled_fp4_experts_quant_out(output, output_scale, input, input_global_scale, input_offset_by_experts,
                                 output_scale_offset_by_experts) -> None:
    """torch.ops._C.scaled_fp4_experts_quant (torch_bindings.cpp:225-229): caller-allocated outputs."""
    _experts_quant_out("scaled_fp4_experts_quant", False, output, output_scale, input, input_global_scale,
                       input_offset_by_experts, output_scale_offset_by_experts)


def silu_and_mul_scaled_fp4_experts_quant_out(output, output_scale, input, input_global_scale,
                                              input_offset_by_experts, output_scale_offset_by_experts) -> None:
    """torch.ops._C.silu_and_mul_scaled_fp4_experts_quant (torch_bindings.cpp:231-236)."""
    _experts_quant_out("silu_and_mul_scaled_fp4_experts_quant", True, output, output_scale, input,
                       input_global_scale, input_offset_by_experts, output_scale_offset_by_experts)


def _experts_quant(gated: bool, input_tensor, input_global_scale, expert_offsets, blockscale_offsets, topk,
                   max_scale_rows):
    assert input_tensor.ndim == 2, f"input.ndim needs to be == 2, but got {input_tensor.ndim}."
    m_numtopk = input_tensor.shape[0]
    k = input_tensor.shape[1] // (2 if gated else 1)
    assert m_numtopk <= MAX_TOKENS_PER_EXPERT_FP4_MOE * topk, (
        f"m_numtopk must be less than MAX_TOKENS_PER_EXPERT({MAX_TOKENS_PER_EXPERT_FP4_MOE}) for cutlass_moe_fp4, "
        f"observed m_numtopk = {m_numtopk}. Use VLLM_MAX_TOKENS_PER_EXPERT_FP4_MOE to set this value.")
    padded_k = (k // 16 + 3) // 4
    n_experts = expert_offsets.numel() - 1
    # The reference allocates MAX_TOKENS_PER_EXPERT * topk scale rows (160 k rows: ~tens of MB of address space it
    # never touches); the rows an expert-sorted batch can reach are bounded by m_numtopk + 127 per expert.
    rows = max_scale_rows if max_scale_rows is not None else \
        min(MAX_TOKENS_PER_EXPERT_FP4_MOE * topk, (m_numtopk + 127 * n_experts + 127) // 128 * 128)
    output = torch.empty(m_numtopk, k // 2, device=input_tensor.device, dtype=torch.uint8)
    output_scales = torch.empty(rows, padded_k, dtype=torch.int32, device=input_tensor.device)
    _experts_quant_out("silu_and_mul_scaled_fp4_experts_quant" if gated else "scaled_fp4_experts_quant", gated,
                       output, output_scales, input_tensor, input_global_scale, expert_offsets, blockscale_offsets)
    return output, output_scales.view(torch.float8_e4m3fn)


def scaled_fp4_experts_quant(input_tensor: torch.Tensor, input_global_scale: torch.Tensor,
                             expert_offsets: torch.Tensor, blockscale_offsets: torch.Tensor, topk: int,
                             max_scale_rows: int | None = None) -> tuple[torch.Tensor, torch.Tensor]:
    """vllm/_custom_ops.py:1601-1660: NVFP4 quantisation of expert-sorted rows with one global scale and one
    swizzled scale-factor matrix per expert. Returns (uint8 [m_topk, k/2], float8_e4m3fn scales)."""
    return _experts_quant(False, input_tensor, input_global_scale, expert_offsets, blockscale_offsets, topk,
                          max_scale_rows)


def silu_and_mul_scaled_fp4_experts_quant(input_tensor: torch.Tensor, input_global_scale: torch.Tensor,
                                          expert_offsets: torch.Tensor, blockscale_offsets: torch.Tensor,
                                          topk: int, max_scale_rows: int | None = None):
    """vllm/_custom_ops.py:1663-1725: input [m_topk, 2k] gate || up; quantises silu(gate) * up."""
    return _experts_quant(True, input_tensor, input_global_scale, expert_offsets, blockscale_offsets, topk,
                          max_scale_rows)


def cutlass_fp4_group_mm(out: torch.Tensor, a: torch.Tensor, b: torch.Tensor, a_blockscale: torch.Tensor,
                         b_blockscales: torch.Tensor, alphas: torch.Tensor, problem_sizes: torch.Tensor,
                         expert_offsets: torch.Tensor, sf_offsets: torch.Tensor) -> None:
    """torch.ops._C.cutlass_fp4_group_mm (torch_bindings.cpp:193-197; checks of nvfp4_blockwise_moe_kernel.cu:
    676-746): a uint8 [m_topk, k/2] expert-sorted, b uint8 [E, n, k/2], a_blockscale / b_blockscales
    float8_e4m3fn swizzled per expert, alphas f32 [E], problem_sizes int32 [E, 3] (m_e, n, k),
    expert_offsets / sf_offsets int32 [E], out bf16 / f16 [m_topk, n]."""
    _cuda(out, a, b, a_blockscale, b_blockscales, alphas, problem_sizes, expert_offsets, sf_offsets)
    if a.dtype != torch.uint8 or b.dtype != torch.uint8:
        raise RuntimeError("cutlass_fp4_group_mm: a and b must be uint8 (packed fp4)")
    if a.dim() != 2 or b.dim() != 3 or not a.is_contiguous() or not b.is_contiguous():
        raise RuntimeError("cutlass_fp4_group_mm: a must be a contiguous 2-D and b a contiguous 3-D tensor")
    if a_blockscale.element_size() not in (1, 4) or b_blockscales.element_size() not in (1, 4):
        raise RuntimeError("cutlass_fp4_group_mm: block scales must be float8_e4m3fn")
    if alphas.dtype != torch.float32:
        raise RuntimeError("cutlass_fp4_group_mm: alphas must be float32")
    for name, t in (("problem_sizes", problem_sizes), ("expert_offsets", expert_offsets), ("sf_offsets", sf_offsets)):
        if t.dtype != torch.int32 or not t.is_contiguous():
            raise RuntimeError(f"cutlass_fp4_group_mm: {name} must be a contiguous int32 tensor")
    E, n, half_k = b.shape
    m_topk = a.shape[0]
    if a.shape[1] != half_k:
        raise RuntimeError("cutlass_fp4_group_mm: a and b disagree on k")
    if problem_sizes.dim() != 2 or tuple(problem_sizes.shape) != (E, 3):
        raise RuntimeError("cutlass_fp4_group_mm: problem_sizes must be [num_experts, 3]")
    if expert_offsets.numel() < E or sf_offsets.numel() < E or alphas.numel() < E:
        raise RuntimeError("cutlass_fp4_group_mm: expert_offsets / sf_offsets / alphas need num_experts entries")
    k = half_k * 2
    ru = lambda x, y: (x + y - 1) // y * y
    if not b_blockscales.is_contiguous() or \
            b_blockscales.numel() * b_blockscales.element_size() != E * ru(n, 128) * ru(k // 16, 4):
        raise RuntimeError(f"cutlass_fp4_group_mm: b_blockscales must be padded and swizzled per expert to "
                           f"({E}x{ru(n, 128)}x{ru(k // 16, 4)})")
    if out.dtype not in (torch.bfloat16, torch.float16) or tuple(out.shape) != (m_topk, n) or not out.is_contiguous():
        raise RuntimeError("cutlass_fp4_group_mm: out must be a contiguous bf16 / f16 [m_topk, n] tensor")
    rc = _lib.load().b200_nvfp4_moe_gemm(_p(out), _p(a), _p(b), _p(a_blockscale), _p(b_blockscales), _p(alphas),
                                         _p(problem_sizes), _p(expert_offsets), _p(sf_offsets), m_topk, E, n, k,
                                         _dt(out), _stream())
    _lib.check(rc, "cutlass_fp4_group_mm")


def cutlass_fp4_moe_mm(out_tensors, a_tensors, b_tensors, a_scales, b_scales, alphas, problem_sizes,
                       expert_offsets, sf_offsets) -> None:
    """vllm/_custom_ops.py:1010-1046 (same argument order as the reference wrapper)."""
    cutlass_fp4_group_mm(out_tensors, a_tensors, b_tensors, a_scales, b_scales, alphas, problem_sizes,
                         expert_offsets, sf_offsets)


# ------------------------------------------------------------------------------------------------
# RMSNorm (+ residual, + static FP8 quant) and RoPE     reference: vllm/_custom_ops.py rms_norm,
# fused_add_rms_norm, rms_norm_static_fp8_quant, fused_add_rms_norm_static_fp8_quant, rotary_embedding
# ------------------------------------------------------------------------------------------------
def _rows(t: torch.Tensor):
    h = t.shape[-1]
    t2 = t.view(-1, h)
    return t2, t2.shape[0], h


def rms_norm(out: torch.Tensor, input: torch.Tensor, weight: torch.Tensor, epsilon: float) -> None:
    _cuda(out, input, weight)
    i2, n, h = _rows(input)
    if not out.is_contiguous():
        raise RuntimeError("rms_norm: out must be contiguous")
    rc = _lib.load().b200_rms_norm(_p(out), _p(i2), _p(weight), float(epsilon), n, h, i2.stride(0),
                                   _dt(input), _stream())
    _lib.check(rc, "rms_norm")


def fused_add_rms_norm(input: torch.Tensor, residual: torch.Tensor, weight: torch.Tensor,
                       epsilon: float) -> None:
    _cuda(input, residual, weight)
    i2, n, h = _rows(input)
    if not residual.is_contiguous():
        raise RuntimeError("fused_add_rms_norm: residual must be contiguous")
    rc = _lib.load().b200_fused_add_rms_norm(_p(i2), _p(residual), _p(weight), float(epsilon), n, h,
                                             i2.stride(0), _dt(input), _stream())
    _lib.check(rc, "fused_add_rms_norm")


def rms_norm_static_fp8_quant(out: torch.Tensor, input: torch.Tensor, weight: torch.Tensor,
                              scale: torch.Tensor, epsilon: float) -> None:
    _cuda(out, input, weight, scale)
    i2, n, h = _rows(input)
    if not out.is_contiguous():
        raise RuntimeError("rms_norm_static_fp8_quant: out must be contiguous")
    rc = _lib.load().b200_rms_norm_static_fp8_quant(_p(out), _p(i2), _p(weight), _p(scale),
                                                    float(epsilon), n, h, i2.stride(0), _dt(input),
                                                    _stream())
    _lib.check(rc, "rms_norm_static_fp8_quant")


def fused_add_rms_norm_static_fp8_quant(out: torch.Tensor, input: torch.Tensor,
                                        residual: torch.Tensor, weight: torch.Tensor,
                                        scale: torch.Tensor, epsilon: float) -> None:
    _cuda(out, input, residual, weight, scale)
    i2, n, h = _rows(input)
    if not (out.is_contiguous() and residual.is_contiguous()):
        raise RuntimeError("fused_add_rms_norm_static_fp8_quant: out / residual must be contiguous")
    rc = _lib.load().b200_fused_add_rms_norm_static_fp8_quant(
        _p(out), _p(i2), _p(residual), _p(weight), _p(scale), float(epsilon), n, h, i2.stride(0),
        _dt(input), _stream())
    _lib.check(rc, "fused_add_rms_norm_static_fp8_quant")


def rotary_embedding(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor | None,
                     head_size: int, cos_sin_cache: torch.Tensor, is_neox: bool) -> None:
    """In place. query [T, Hq*D] or [T, Hq, D]; key likewise or None; positions int64 [T]."""
    _cuda(positions, query, key, cos_sin_cache)
    T = positions.numel()
    if positions.dtype != torch.int64:
        raise RuntimeError("rotary_embedding: positions must be int64")
    if cos_sin_cache.dtype != query.dtype:
        raise RuntimeError("rotary_embedding: cos_sin_cache must have the query dtype")
    q2 = query.view(T, -1, head_size) if query.dim() == 2 else query
    k2 = None if key is None else (key.view(T, -1, head_size) if key.dim() == 2 else key)
    rc = _lib.load().b200_rotary_embedding(
        _p(positions), _p(q2), _p(k2), _p(cos_sin_cache), T, cos_sin_cache.shape[1], q2.stride(0),
        0 if k2 is None else k2.stride(0), q2.stride(1), q2.shape[1],
        0 if k2 is None else k2.shape[1], 1 if is_neox else 0, _dt(query), _stream())
    _lib.check(rc, "rotary_embedding")


def rope_and_reshape_and_cache_flash(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor,
                                     value: torch.Tensor, head_size: int, cos_sin_cache: torch.Tensor,
                                     is_neox: bool, key_cache: torch.Tensor, value_cache: torch.Tensor,
                                     slot_mapping: torch.Tensor, kv_cache_dtype: str, k_scale: torch.Tensor,
                                     v_scale: torch.Tensor) -> None:
    """rotary_embedding(positions, query, key, ...) followed by reshape_and_cache_flash(key, value, ...) in one
    launch (AttentionImpl.do_rope_and_kv_cache_update, backend.py:950-1006); bit-identical to the pair.
    query [T, Hq, D] / key, value [T, Hkv, D] (views of the fused QKV output are fine), rotated in place."""
    _cuda(positions, query, key, value, cos_sin_cache, key_cache, value_cache, slot_mapping, k_scale, v_scale)
    T = positions.numel()
    if positions.dtype != torch.int64 or slot_mapping.dtype != torch.int64:
        raise RuntimeError("rope_and_reshape_and_cache_flash: positions and slot_mapping must be int64")
    if slot_mapping.shape[0] != T:
        raise RuntimeError("rope_and_reshape_and_cache_flash: one slot per token is required")
    if cos_sin_cache.dtype != query.dtype or key.dtype != query.dtype or value.dtype != query.dtype:
        raise RuntimeError("rope_and_reshape_and_cache_flash: query / key / value / cos_sin_cache dtypes differ")
    q3 = query.view(T, -1, head_size) if query.dim() == 2 else query
    k3 = key.view(T, -1, head_size) if key.dim() == 2 else key
    v3 = value.view(T, -1, head_size) if value.dim() == 2 else value
    for t in (q3, k3, v3):
        if t.stride(2) != 1 or t.stride(1) != head_size:
            raise RuntimeError("rope_and_reshape_and_cache_flash: heads must be contiguous")
    if key_cache.dim() != 4 or key_cache.stride(3) != 1 or key_cache.stride() != value_cache.stride():
        raise RuntimeError("rope_and_reshape_and_cache_flash: caches must be 4-D views with equal strides")
    hkv = k3.shape[1]
    if k_scale.numel() not in (1, hkv) or k_scale.shape != v_scale.shape:
        raise RuntimeError("k_scale and v_scale must be of shape [1] or [num_heads]")
    if kv_cache_dtype == "auto":
        cache_dt = _dt(query)
        if key_cache.dtype != query.dtype:
            raise RuntimeError("rope_and_reshape_and_cache_flash: kv_cache_dtype 'auto' needs cache dtype == key dtype")
    elif kv_cache_dtype in _FP8_KV_NAMES:
        cache_dt = FP8_E4M3
        if key_cache.element_size() != 1:
            raise RuntimeError("rope_and_reshape_and_cache_flash: fp8 kv cache must have 1-byte elements")
    else:
        raise RuntimeError(f"Unsupported data type of kv cache: {kv_cache_dtype}")
    rc = _lib.load().b200_rope_and_cache_flash(
        _p(positions), _p(q3), _p(k3), _p(v3), _p(cos_sin_cache), _p(key_cache), _p(value_cache),
        _p(slot_mapping), T, cos_sin_cache.shape[1], head_size, q3.shape[1], hkv, key_cache.shape[1],
        q3.stride(0), k3.stride(0), v3.stride(0), key_cache.stride(0), key_cache.stride(1), key_cache.stride(2),
        1 if is_neox else 0, _dt(query), cache_dt, _p(k_scale), _p(v_scale), 1 if k_scale.numel() > 1 else 0,
        _stream())
    _lib.check(rc, "rope_and_reshape_and_cache_flash")


# ------------------------------------------------------------------------------------------------
# Paged-attention decode
# ------------------------------------------------------------------------------------------------
_workspaces: dict[tuple, torch.Tensor] = {}


def paged_attention_workspace(num_seqs: int, num_heads: int, num_kv_heads: int, head_size: int,
                              device: torch.device) -> torch.Tensor:
    """Persistent split-KV workspace (the metadata builder owns it in the backend; cached here for
    direct op calls). Sized once for the largest batch seen, like triton_attn.py:154-175."""
    lib = _lib.load()
    n = lib.b200_paged_attention_workspace_bytes(num_seqs, num_heads, num_kv_heads, head_size)
    key = (str(device), num_heads, head_size)
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < n:
        ws = torch.zeros(n, dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


def unified_attention(
    q: torch.Tensor,
    k: torch.Tensor,
    v: torch.Tensor,
    out: torch.Tensor,
    cu_seqlens_q: torch.Tensor,
    max_seqlen_q: int,
    seqused_k: torch.Tensor,
    max_seqlen_k: int,
    softmax_scale: float,
    causal: bool,
    window_size,
    block_table: torch.Tensor,
    softcap: float,
    q_descale,
    k_descale: torch.Tensor | None,
    v_descale: torch.Tensor | None,
    alibi_slopes: torch.Tensor | None = None,
    output_scale: torch.Tensor | None = None,
    sinks: torch.Tensor | None = None,
    lse: torch.Tensor | None = None,
) -> torch.Tensor:
    """Variable-length (prefill / chunked-prefill / mixed) paged attention; same positional arguments as
    the reference's unified_attention (vllm/v1/attention/ops/triton_unified_attention.py:802-826).

    q / out [T, Hq, D] packed over sequences (cu_seqlens_q int32 [S+1]); k / v = key_cache / value_cache
    [num_blocks, block_size, Hkv, D] views (strides forwarded); seqused_k int32 [S] = context + query
    length; block_table int32 [S, max_blocks]; window_size = (left, right) as FlashAttention, (-1, -1) =
    none (the kernel's SLIDING_WINDOW is 1 + left, :1128); k_descale / v_descale: device f32 scalars for
    FP8 caches; output_scale: `out` is float8_e4m3fn and receives e4m3(attention / output_scale).
    Semantics: ref_paged_attn, tests/kernels/attention/test_triton_unified_attention.py:33-88.
    max_seqlen_q / max_seqlen_k are accepted for signature parity (the kernel reads lengths on device).
    lse (extension): optional float32 [T, Hq] natural-log LSE."""
    _cuda(q, k, v, out, cu_seqlens_q, seqused_k, block_table, k_descale, v_descale, alibi_slopes, output_scale, lse)
    lib = _lib.load()
    if not causal:
        raise RuntimeError("unified_attention: only causal attention is supported")
    if q_descale is not None:
        raise RuntimeError("unified_attention: q_descale (FP8 query) is not supported")
    if sinks is not None:
        raise RuntimeError("unified_attention: attention sinks are not supported")
    if (output_scale is not None) != (out.element_size() == 1):
        raise RuntimeError("unified_attention: out must be fp8 exactly when output_scale is given")
    if q.dim() != 3 or k.dim() != 4 or v.dim() != 4:
        raise RuntimeError("unified_attention: q must be 3-D and the caches 4-D")
    if q.stride(2) != 1 or out.stride(2) != 1 or k.stride(3) != 1:
        raise RuntimeError("unified_attention: last dimension must be contiguous")
    if k.stride() != v.stride() or k.shape != v.shape:
        raise RuntimeError("unified_attention: key_cache and value_cache must share shape and strides")
    if block_table.dtype != torch.int32 or seqused_k.dtype != torch.int32 or cu_seqlens_q.dtype != torch.int32:
        raise RuntimeError("unified_attention: block_table, seqused_k and cu_seqlens_q must be int32")
    if block_table.stride(1) != 1 or not cu_seqlens_q.is_contiguous() or not seqused_k.is_contiguous():
        raise RuntimeError("unified_attention: index tensors must be contiguous")
    T, Hq, D = q.shape
    nb, bs, Hkv, Dk = k.shape
    S = seqused_k.numel()
    if Dk != D or out.shape != q.shape:
        raise RuntimeError("unified_attention: shape mismatch between q, out and the cache")
    if cu_seqlens_q.numel() != S + 1 or block_table.shape[0] < S:
        raise RuntimeError("unified_attention: cu_seqlens_q must have num_seqs + 1 entries")
    kv_fp8 = k.element_size() == 1
    if kv_fp8 and (k_descale is None or v_descale is None):
        raise RuntimeError("unified_attention: fp8 kv cache needs k_descale and v_descale")
    if lse is not None and (lse.dtype != torch.float32 or lse.shape != (T, Hq) or not lse.is_contiguous()):
        raise RuntimeError("unified_attention: lse must be contiguous float32 [num_tokens, num_heads]")
    left = -1 if window_size is None else int(window_size[0])
    a = _lib.VarlenAttnArgs()
    a.out, a.lse, a.q = _p(out), _p(lse), _p(q)
    a.k_cache, a.v_cache = _p(k), _p(v)
    a.cu_seqlens_q, a.seq_lens, a.block_table = _p(cu_seqlens_q), _p(seqused_k), _p(block_table)
    a.k_scale, a.v_scale = _p(k_descale), _p(v_descale)
    a.alibi_slopes, a.output_scale = _p(alibi_slopes), _p(output_scale)
    a.q_stride_t, a.q_stride_h = q.stride(0), q.stride(1)
    a.out_stride_t, a.out_stride_h = out.stride(0), out.stride(1)
    a.kv_block_stride, a.kv_token_stride, a.kv_head_stride = k.stride(0), k.stride(1), k.stride(2)
    a.bt_stride = block_table.stride(0)
    a.num_blocks = nb
    a.num_seqs, a.num_tokens, a.num_heads, a.num_kv_heads, a.head_size, a.block_size = S, T, Hq, Hkv, D, bs
    a.q_dtype = _dt(q)
    a.kv_dtype = FP8_E4M3 if kv_fp8 else _dt(k)
    a.causal = 1
    a.sliding_window = left + 1 if left >= 0 else 0
    a.scale = float(softmax_scale)
    a.softcap = float(softcap) if softcap else 0.0
    import ctypes
    rc = lib.b200_paged_attention_varlen(ctypes.byref(a), _stream())
    _lib.check(rc, "unified_attention")
    return out


def paged_attention_decode(
    out: torch.Tensor,
    query: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    block_table: torch.Tensor,
    seq_lens: torch.Tensor,
    scale: float,
    *,
    k_scale: torch.Tensor | None = None,
    v_scale: torch.Tensor | None = None,
    alibi_slopes: torch.Tensor | None = None,
    sliding_window: int | None = None,
    softcap: float | None = None,
    lse: torch.Tensor | None = None,
    max_seq_len: int = 0,
    workspace: torch.Tensor | None = None,
    output_scale: torch.Tensor | None = None,
) -> torch.Tensor:
    """Decode attention over a paged KV cache.
    output_scale (device f32 scalar): `out` is float8_e4m3fn and receives the statically quantised result
    e4m3(attention / output_scale) - the fused FP8 output of AttentionImpl.forward (backend.py:895-934).

    query/out [S, Hq, D]; key_cache/value_cache [num_blocks, block_size, Hkv, D] views (any of the
    NHD / HND / split layouts - strides are forwarded); block_table int32 [S, max_blocks];
    seq_lens int32 [S]. fp8 caches (float8_e4m3fn or uint8) need k_scale / v_scale device scalars.
    Semantics: ref_paged_attn in tests/kernels/attention/test_triton_unified_attention.py:33-88
    with query_len == 1.
    """
    _cuda(out, query, key_cache, value_cache, block_table, seq_lens, k_scale, v_scale,
          alibi_slopes, lse, workspace, output_scale)
    lib = _lib.load()
    if (output_scale is not None) != (out.element_size() == 1):
        raise RuntimeError("paged_attention_decode: out must be fp8 exactly when output_scale is given")
    if query.dim() != 3 or key_cache.dim() != 4 or value_cache.dim() != 4:
        raise RuntimeError("paged_attention_decode: query must be 3-D and the caches 4-D")
    if query.stride(2) != 1 or out.stride(2) != 1 or key_cache.stride(3) != 1:
        raise RuntimeError("paged_attention_decode: last dimension must be contiguous")
    if key_cache.stride() != value_cache.stride() or key_cache.shape != value_cache.shape:
        raise RuntimeError("paged_attention_decode: key_cache and value_cache must share shape and strides")
    if block_table.dtype != torch.int32 or seq_lens.dtype != torch.int32:
        raise RuntimeError("paged_attention_decode: block_table and seq_lens must be int32")
    if block_table.stride(1) != 1:
        raise RuntimeError("paged_attention_decode: block_table rows must be contiguous")
    S, Hq, D = query.shape
    nb, bs, Hkv, Dk = key_cache.shape
    if Dk != D:
        raise RuntimeError("paged_attention_decode: head_size mismatch between query and cache")
    kv_fp8 = key_cache.element_size() == 1
    if kv_fp8 and (k_scale is None or v_scale is None):
        raise RuntimeError("paged_attention_decode: fp8 kv cache needs k_scale and v_scale")
    if workspace is None:
        workspace = paged_attention_workspace(S, Hq, Hkv, D, query.device)
    a = _lib.PagedAttnArgs()
    a.out, a.lse, a.q = _p(out), _p(lse), _p(query)
    a.k_cache, a.v_cache = _p(key_cache), _p(value_cache)
    a.block_table, a.seq_lens = _p(block_table), _p(seq_lens)
    a.k_scale, a.v_scale = _p(k_scale), _p(v_scale)
    a.alibi_slopes = _p(alibi_slopes)
    a.workspace, a.workspace_bytes = _p(workspace), workspace.numel() * workspace.element_size()
    a.num_blocks = nb
    a.q_stride_s, a.q_stride_h = query.stride(0), query.stride(1)
    a.out_stride_s, a.out_stride_h = out.stride(0), out.stride(1)
    a.kv_block_stride, a.kv_token_stride, a.kv_head_stride = (
        key_cache.stride(0), key_cache.stride(1), key_cache.stride(2))
    a.bt_stride = block_table.stride(0)
    a.num_seqs, a.num_heads, a.num_kv_heads, a.head_size, a.block_size = S, Hq, Hkv, D, bs
    a.max_seq_len = int(max_seq_len)
    a.q_dtype = _dt(query)
    a.kv_dtype = FP8_E4M3 if kv_fp8 else _dt(key_cache)
    a.sliding_window = int(sliding_window) if sliding_window else 0
    a.scale = float(scale)
    a.softcap = float(softcap) if softcap else 0.0
    a.output_scale = _p(output_scale)
    import ctypes
    rc = lib.b200_paged_attention_decode(ctypes.byref(a), _stream())
    _lib.check(rc, "paged_attention_decode")
    return out


# ------------------------------------------------------------------------------------------------
# Legacy paged attention     reference: torch.ops._C.paged_attention_v1 / v2 (wheel-only)
# ------------------------------------------------------------------------------------------------
def _legacy_common(op, out, query, key_cache, value_cache, num_kv_heads, block_tables, seq_lens,
                   block_size, kv_cache_dtype, k_scale, v_scale, alibi_slopes):
    _cuda(out, query, key_cache, value_cache, block_tables, seq_lens, k_scale, v_scale, alibi_slopes)
    if query.dim() != 3 or key_cache.dim() != 5 or value_cache.dim() != 4:
        raise RuntimeError(f"{op}: query [S,H,D], key_cache [nb,Hkv,D/x,bs,x], value_cache [nb,Hkv,D,bs]")
    if query.stride(2) != 1 or query.stride(1) != query.shape[2] or not out.is_contiguous():
        raise RuntimeError(f"{op}: query heads and out must be contiguous")
    if not (key_cache[0].is_contiguous() and value_cache[0].is_contiguous()):
        raise RuntimeError(f"{op}: each cache block must be contiguous")
    if block_tables.dtype != torch.int32 or seq_lens.dtype != torch.int32:
        raise RuntimeError(f"{op}: block_tables and seq_lens must be int32")
    if kv_cache_dtype not in ("auto", "fp8", "fp8_e4m3"):
        raise RuntimeError(f"{op}: Unsupported data type of kv cache: {kv_cache_dtype}")
    fp8 = kv_cache_dtype != "auto"
    if fp8 and key_cache.element_size() != 1:
        raise RuntimeError(f"{op}: fp8 kv_cache_dtype needs a 1-byte cache tensor")
    if value_cache.shape[3] != block_size or value_cache.shape[1] != num_kv_heads:
        raise RuntimeError(f"{op}: value_cache shape does not match num_kv_heads / block_size")
    if key_cache.stride(0) != value_cache.stride(0) or key_cache.stride(1) != value_cache.stride(1):
        raise RuntimeError(f"{op}: key_cache and value_cache must share block / head strides")
    return _lib.FP8_E4M3 if fp8 else _dt(query)


def paged_attention_v1(out: torch.Tensor, query: torch.Tensor, key_cache: torch.Tensor,
                       value_cache: torch.Tensor, num_kv_heads: int, scale: float,
                       block_tables: torch.Tensor, seq_lens: torch.Tensor, block_size: int,
                       max_seq_len: int, alibi_slopes: torch.Tensor | None, kv_cache_dtype: str,
                       k_scale: torch.Tensor, v_scale: torch.Tensor, tp_rank: int = 0,
                       blocksparse_local_blocks: int = 0, blocksparse_vert_stride: int = 0,
                       blocksparse_block_size: int = 64, blocksparse_head_sliding_step: int = 0) -> None:
    """torch.ops._C.paged_attention_v1: single-pass decode attention over the x-packed cache layout."""
    kvdt = _legacy_common("paged_attention_v1", out, query, key_cache, value_cache, num_kv_heads,
                          block_tables, seq_lens, block_size, kv_cache_dtype, k_scale, v_scale, alibi_slopes)
    S, H, D = query.shape
    rc = _lib.load().b200_paged_attention_v1(
        _p(out), _p(query), _p(key_cache), _p(value_cache), S, H, num_kv_heads, D, float(scale),
        _p(block_tables), _p(seq_lens), block_size, int(max_seq_len), _p(alibi_slopes), _dt(query), kvdt,
        _p(k_scale), _p(v_scale), query.stride(0), key_cache.stride(0), key_cache.stride(1),
        block_tables.stride(0), int(blocksparse_vert_stride), _stream())
    _lib.check(rc, "paged_attention_v1")


def paged_attention_v2(out: torch.Tensor, exp_sums: torch.Tensor, max_logits: torch.Tensor,
                       tmp_out: torch.Tensor, query: torch.Tensor, key_cache: torch.Tensor,
                       value_cache: torch.Tensor, num_kv_heads: int, scale: float,
                       block_tables: torch.Tensor, seq_lens: torch.Tensor, block_size: int,
                       max_seq_len: int, alibi_slopes: torch.Tensor | None, kv_cache_dtype: str,
                       k_scale: torch.Tensor, v_scale: torch.Tensor, tp_rank: int = 0,
                       blocksparse_local_blocks: int = 0, blocksparse_vert_stride: int = 0,
                       blocksparse_block_size: int = 64, blocksparse_head_sliding_step: int = 0) -> None:
    """torch.ops._C.paged_attention_v2: 512-token partitions (exp_sums / max_logits / tmp_out) + reduce."""
    kvdt = _legacy_common("paged_attention_v2", out, query, key_cache, value_cache, num_kv_heads,
                          block_tables, seq_lens, block_size, kv_cache_dtype, k_scale, v_scale, alibi_slopes)
    _cuda(exp_sums, max_logits, tmp_out)
    S, H, D = query.shape
    P = exp_sums.shape[2]
    if (exp_sums.dtype != torch.float32 or max_logits.dtype != torch.float32 or tmp_out.dtype != query.dtype
            or tuple(exp_sums.shape) != (S, H, P) or tuple(max_logits.shape) != (S, H, P)
            or tuple(tmp_out.shape) != (S, H, P, D)
            or not (exp_sums.is_contiguous() and max_logits.is_contiguous() and tmp_out.is_contiguous())):
        raise RuntimeError("paged_attention_v2: exp_sums / max_logits [S,H,P] f32 and tmp_out [S,H,P,D] contiguous")
    rc = _lib.load().b200_paged_attention_v2(
        _p(out), _p(exp_sums), _p(max_logits), _p(tmp_out), _p(query), _p(key_cache), _p(value_cache),
        S, H, num_kv_heads, D, float(scale), _p(block_tables), _p(seq_lens), block_size, int(max_seq_len),
        _p(alibi_slopes), _dt(query), kvdt, _p(k_scale), _p(v_scale), query.stride(0),
        key_cache.stride(0), key_cache.stride(1), block_tables.stride(0), P,
        int(blocksparse_vert_stride), _stream())
    _lib.check(rc, "paged_attention_v2")


# ------------------------------------------------------------------------------------------------
# merge_attn_states                     reference: vllm/_custom_ops.py:178-197, ops/merge_attn_states.py
# ------------------------------------------------------------------------------------------------
def merge_attn_states(output: torch.Tensor, prefix_output: torch.Tensor, prefix_lse: torch.Tensor,
                      suffix_output: torch.Tensor, suffix_lse: torch.Tensor,
                      output_lse: torch.Tensor | None = None,
                      prefill_tokens_with_context: int | None = None,
                      output_scale: torch.Tensor | None = None) -> None:
    """Merge prefix (KV-cache context) and suffix (new tokens) partial attention by their LSEs.
    outputs [T, H, D]; LSEs float32 [H, T]."""
    _cuda(output, prefix_output, prefix_lse, suffix_output, suffix_lse, output_lse, output_scale)
    assert prefix_output.stride(1) == suffix_output.stride(1), (
        "merge_attn_states requires prefix_output and suffix_output to have matching head strides, got "
        f"{prefix_output.stride(1)} and {suffix_output.stride(1)}")
    if prefix_output.dim() != 3 or prefix_output.shape != suffix_output.shape or output.shape != prefix_output.shape:
        raise RuntimeError("merge_attn_states: outputs must be [T, H, D] of one shape")
    if prefix_output.dtype != suffix_output.dtype or prefix_output.dtype not in (
            torch.float32, torch.float16, torch.bfloat16):
        raise RuntimeError("merge_attn_states: inputs must be float32 / float16 / bfloat16")
    if prefix_lse.dtype != torch.float32 or suffix_lse.dtype != torch.float32:
        raise RuntimeError("merge_attn_states: LSEs must be float32")
    if prefix_output.stride(2) != 1 or suffix_output.stride(2) != 1 or output.stride(2) != 1:
        raise RuntimeError("merge_attn_states: last dim must be contiguous")
    T, H, D = prefix_output.shape
    if tuple(prefix_lse.shape) != (H, T) or tuple(suffix_lse.shape) != (H, T):
        raise RuntimeError("merge_attn_states: LSEs must be [num_heads, num_tokens]")
    fp8_out = output.dtype not in (torch.float32, torch.float16, torch.bfloat16)
    if fp8_out:
        assert output_scale is not None, f"output_scale is required when output is {output.dtype}"
    elif output.dtype != prefix_output.dtype:
        raise RuntimeError("merge_attn_states: output dtype must match the inputs or be FP8")
    ol = output_lse
    rc = _lib.load().b200_merge_attn_states(
        _p(output), _p(ol), _p(prefix_output), _p(prefix_lse), _p(suffix_output), _p(suffix_lse),
        _p(output_scale), T, H, D, prefix_output.stride(0), suffix_output.stride(0), prefix_output.stride(1),
        output.stride(0), output.stride(1), prefix_lse.stride(0), prefix_lse.stride(1),
        suffix_lse.stride(0), suffix_lse.stride(1), 0 if ol is None else ol.stride(0),
        0 if ol is None else ol.stride(1),
        -1 if prefill_tokens_with_context is None else int(prefill_tokens_with_context),
        _dt(prefix_output), _lib.FP8_E4M3 if fp8_out else _dt(output), _stream())
    _lib.check(rc, "merge_attn_states")


# ------------------------------------------------------------------------------------------------
# Sampler-side ops      reference: vllm/v1/sample/sampler.py (greedy_sample),
# vllm/v1/sample/ops/topk_topp_sampler.py:349-408 (apply_top_k_top_p), vllm/_custom_ops.py:309-357
# ------------------------------------------------------------------------------------------------
def greedy_sample(logits: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """logits.argmax(dim=-1).view(-1) (Sampler.greedy_sample): int64 [rows]; ties -> lowest index."""
    _cuda(logits, out)
    if logits.dim() != 2 or logits.stride(1) != 1:
        raise RuntimeError("greedy_sample: logits must be [rows, vocab] with a contiguous last dim")
    rows, vocab = logits.shape
    if out is None:
        out = torch.empty(rows, dtype=torch.int64, device=logits.device)
    elif out.dtype != torch.int64 or out.numel() != rows or not out.is_contiguous():
        raise RuntimeError("greedy_sample: out must be a contiguous int64 [rows] tensor")
    rc = _lib.load().b200_argmax(_p(out), _p(logits), rows, vocab, logits.stride(0) if rows > 1 else vocab,
                                 _dt(logits), _stream())
    _lib.check(rc, "greedy_sample")
    return out


def apply_top_k_top_p(logits: torch.Tensor, k: torch.Tensor | None, p: torch.Tensor | None) -> torch.Tensor:
    """Same contract as topk_topp_sampler.apply_top_k_top_p: fp32 logits [rows, vocab] masked in place with -inf
    (top-k keeps ties of the k-th value, top-p drops the ascending prefix whose softmax mass is <= 1 - p and always
    keeps the largest token) and returned. k int32 / int64 [rows], p float32 [rows]."""
    if p is None and k is None:
        return logits
    _cuda(logits, k, p)
    if logits.dim() != 2 or logits.dtype != torch.float32 or logits.stride(1) != 1:
        raise RuntimeError("apply_top_k_top_p: logits must be float32 [rows, vocab] with a contiguous last dim")
    rows, vocab = logits.shape
    kd = 0
    if k is not None:
        if k.dtype not in (torch.int32, torch.int64) or k.numel() != rows or not k.is_contiguous():
            raise RuntimeError("apply_top_k_top_p: k must be a contiguous int32 / int64 [rows] tensor")
        kd = _lib.I32 if k.dtype == torch.int32 else _lib.I64
    if p is not None and (p.dtype != torch.float32 or p.numel() != rows or not p.is_contiguous()):
        raise RuntimeError("apply_top_k_top_p: p must be a contiguous float32 [rows] tensor")
    rc = _lib.load().b200_top_k_top_p(_p(logits), _p(k), kd, _p(p), rows, vocab,
                                      logits.stride(0) if rows > 1 else vocab, _stream())
    _lib.check(rc, "apply_top_k_top_p")
    return logits


def apply_repetition_penalties(logits: torch.Tensor, prompt_mask: torch.Tensor, output_mask: torch.Tensor,
                               repetition_penalties: torch.Tensor) -> None:
    """torch.ops._C.apply_repetition_penalties_ (vllm/_custom_ops.py:325-357), in place."""
    _cuda(logits, prompt_mask, output_mask, repetition_penalties)
    if not (logits.is_contiguous() and prompt_mask.is_contiguous() and output_mask.is_contiguous()
            and repetition_penalties.is_contiguous()):
        raise RuntimeError("apply_repetition_penalties_: tensors must be contiguous")
    if prompt_mask.dtype != torch.bool or output_mask.dtype != torch.bool:
        raise RuntimeError("apply_repetition_penalties_: masks must be bool")
    if prompt_mask.shape != logits.shape or output_mask.shape != logits.shape:
        raise RuntimeError("apply_repetition_penalties_: masks must have the shape of logits")
    rows, vocab = logits.shape
    pen = repetition_penalties.to(logits.dtype)
    rc = _lib.load().b200_apply_repetition_penalties(_p(logits), _p(prompt_mask), _p(output_mask), _p(pen),
                                                     rows, vocab, _dt(logits), _stream())
    _lib.check(rc, "apply_repetition_penalties_")
