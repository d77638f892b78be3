"""Fused MoE on libb200ops.so: routing ops with the reference's names and `fused_experts`.

Reference surface: vllm/_custom_ops.py (moe_align_block_size :2237-2254, moe_sum :2228-2234,
topk_softmax :2385-2402), vllm/model_executor/layers/fused_moe/moe_align_block_size.py:11-103 and
fused_moe.py:1448-1535, 1587-1853 (fused_experts / fused_experts_impl).
"""
from __future__ import annotations

import torch

from . import _lib, ops
from ._lib import I32, I64
from .ops import _cuda, _dt, _p, _stream


def _idt(t: torch.Tensor) -> int:
    if t.dtype == torch.int32:
        return I32
    if t.dtype == torch.int64:
        return I64
    raise RuntimeError(f"b200ops: index tensor must be int32 or int64, got {t.dtype}")


def moe_align_block_size_op(topk_ids: torch.Tensor, num_experts: int, block_size: int,
                            sorted_token_ids: torch.Tensor, experts_ids: torch.Tensor,
                            num_tokens_post_pad: torch.Tensor,
                            maybe_expert_map: torch.Tensor | None = None,
                            inv_perm: torch.Tensor | None = None) -> None:
    """torch.ops._moe_C.moe_align_block_size (csrc/libtorch_stable/moe/torch_bindings.cpp:27-32)."""
    _cuda(topk_ids, sorted_token_ids, experts_ids, num_tokens_post_pad, maybe_expert_map, inv_perm)
    if not topk_ids.is_contiguous():
        raise RuntimeError("moe_align_block_size: topk_ids must be contiguous")
    if maybe_expert_map is not None and maybe_expert_map.dtype != torch.int32:
        raise RuntimeError("moe_align_block_size: expert_map must be int32")
    rc = _lib.load().b200_moe_align_block_size(
        _p(topk_ids), _idt(topk_ids), topk_ids.numel(), num_experts, block_size, _p(sorted_token_ids),
        sorted_token_ids.numel(), _p(experts_ids), experts_ids.numel(), _p(num_tokens_post_pad),
        _p(maybe_expert_map), _p(inv_perm), _stream())
    _lib.check(rc, "moe_align_block_size")


def moe_align_block_size(topk_ids: torch.Tensor, block_size: int, num_experts: int,
                         expert_map: torch.Tensor | None = None, pad_sorted_ids: bool = False,
                         ignore_invalid_experts: bool = False, return_inv_perm: bool = False):
    """Same contract as fused_moe/moe_align_block_size.py:11-103."""
    numel = topk_ids.numel()
    max_num_tokens_padded = numel + num_experts * (block_size - 1)
    if pad_sorted_ids:
        max_num_tokens_padded = (max_num_tokens_padded + block_size - 1) // block_size * block_size
    if numel < num_experts:
        max_num_tokens_padded = min(numel * block_size, max_num_tokens_padded)
    dev = topk_ids.device
    sorted_ids = torch.empty((max_num_tokens_padded,), dtype=torch.int32, device=dev)
    max_num_m_blocks = (max_num_tokens_padded + block_size - 1) // block_size
    expert_ids = torch.empty((max_num_m_blocks,), dtype=torch.int32, device=dev)
    num_tokens_post_pad = torch.empty((1,), dtype=torch.int32, device=dev)
    inv = torch.empty((numel,), dtype=torch.int32, device=dev) if return_inv_perm else None
    moe_align_block_size_op(topk_ids, num_experts, block_size, sorted_ids, expert_ids,
                            num_tokens_post_pad, expert_map if ignore_invalid_experts else None, inv)
    if expert_map is not None and not ignore_invalid_experts:
        expert_ids = expert_map[expert_ids]
    if return_inv_perm:
        return sorted_ids, expert_ids, num_tokens_post_pad, inv
    return sorted_ids, expert_ids, num_tokens_post_pad


def topk_softmax(topk_weights: torch.Tensor, topk_ids: torch.Tensor,
                 token_expert_indices: torch.Tensor, gating_output: torch.Tensor,
                 renormalize: bool = False, bias: torch.Tensor | None = None) -> None:
    """torch.ops._moe_C.topk_softmax: weights f32 [M,k], ids i32 [M,k], token_expert_indices i32 [M,k]."""
    _cuda(topk_weights, topk_ids, token_expert_indices, gating_output, bias)
    if topk_weights.dtype != torch.float32 or topk_ids.dtype != torch.int32 or \
            token_expert_indices.dtype != torch.int32:
        raise RuntimeError("topk_softmax: weights must be float32 and indices int32")
    if not gating_output.is_contiguous():
        raise RuntimeError("topk_softmax: gating_output must be contiguous")
    M, E = gating_output.shape
    rc = _lib.load().b200_moe_topk_softmax(
        _p(topk_weights), _p(topk_ids), _p(token_expert_indices), _p(gating_output), _p(bias), M, E,
        topk_weights.shape[-1], 1 if renormalize else 0, _dt(gating_output), _stream())
    _lib.check(rc, "topk_softmax")


def moe_sum(input: torch.Tensor, output: torch.Tensor, topk_ids: torch.Tensor | None = None,
            expert_map: torch.Tensor | None = None) -> None:
    """torch.ops._moe_C.moe_sum: output[t,:] = sum_k input[t,k,:] (fp32 accumulate)."""
    _cuda(input, output, topk_ids, expert_map)
    if input.dim() != 3 or input.stride(2) != 1 or not output.is_contiguous():
        raise RuntimeError("moe_sum: input [tokens, topk, d] (d contiguous), output contiguous")
    T, K, d = input.shape
    rc = _lib.load().b200_moe_sum(
        _p(output), _p(input), T, K, d, input.stride(0), input.stride(1), _p(topk_ids),
        0 if topk_ids is None else _idt(topk_ids), _p(expert_map), _dt(input), _stream())
    _lib.check(rc, "moe_sum")


def _block_size_for(tokens_per_expert: float) -> int:
    for b in (16, 32, 64):
        if tokens_per_expert <= b:
            return b
    return 128


def fused_experts(
    hidden_states: torch.Tensor,
    w1: torch.Tensor,
    w2: torch.Tensor,
    topk_weights: torch.Tensor,
    topk_ids: torch.Tensor,
    activation: str = "silu",
    apply_router_weight_on_input: bool = False,
    use_fp8_w8a8: bool = False,
    per_channel_quant: bool = False,
    global_num_experts: int = -1,
    expert_map: torch.Tensor | None = None,
    w1_scale: torch.Tensor | None = None,
    w2_scale: torch.Tensor | None = None,
    a1_scale: torch.Tensor | None = None,
    a2_scale: torch.Tensor | None = None,
    block_shape=None,
    **unsupported,
) -> torch.Tensor:
    """out[t] = sum_k w[t,k] * W2[e] . silu_and_mul(W1[e] . x_t), FP8 W8A8 (fused_moe.py:1448-1535, 1587-1853).

    hidden_states [M,H] bf16/f16; w1 [E, 2I, H], w2 [E, H, I] e4m3; topk_weights [M,k]; topk_ids int [M,k];
    expert_map int32 [E_global] -> local id or -1 (expert parallel). Quantisation schemes, as the reference's
    moe_kernel_quantize_input selects them (fused_moe/utils.py:138-161, call sites fused_moe.py:1764-1770,1813-1819):
      * per-tensor: a*_scale given -> static, else dynamic over the whole batch; w*_scale [E] (or [E, N] with
        per_channel_quant);
      * per_channel_quant=True and a*_scale None: dynamic PER-TOKEN activation scales, w*_scale [E, N];
      * block_shape=[128, 128]: per-token-group (1 x 128) activation scales, w*_scale [E, N/128, K/128].
    apply_router_weight_on_input (topk == 1): the router weight multiplies the input, not the output
    (fused_moe.py:1757-1762)."""
    if not use_fp8_w8a8:
        raise NotImplementedError("b200ops.fused_experts: only use_fp8_w8a8=True is built (NVFP4: cutlass_moe_fp4)")
    if activation != "silu" or any(v not in (None, False) for v in unsupported.values()):
        raise NotImplementedError("b200ops.fused_experts: unsupported option")
    if block_shape is not None and (list(block_shape) != [128, 128] or per_channel_quant):
        raise NotImplementedError("b200ops.fused_experts: block_shape must be [128, 128] without per_channel_quant")
    _cuda(hidden_states, w1, w2, topk_weights, topk_ids, expert_map, w1_scale, w2_scale, a1_scale, a2_scale)
    lib = _lib.load()
    M, H = hidden_states.shape
    E, N1, K1 = w1.shape
    I = N1 // 2
    if K1 != H or tuple(w2.shape) != (E, H, I):
        raise RuntimeError("fused_experts: weight shapes do not match hidden_states")
    if not (w1.is_contiguous() and w2.is_contiguous() and hidden_states.stride(1) == 1):
        raise RuntimeError("fused_experts: weights must be contiguous")
    topk = topk_ids.shape[1]
    numel = M * topk
    Eg = global_num_experts if global_num_experts > 0 else E
    dev, dt = hidden_states.device, hidden_states.dtype
    if apply_router_weight_on_input:
        if topk != 1:
            raise RuntimeError("apply_router_weight_on_input is only implemented for topk=1")
        hidden_states = hidden_states * topk_weights.to(dt)
    bs = _block_size_for(numel / max(E, 1))
    sorted_ids, expert_ids, npp, inv = moe_align_block_size(
        topk_ids, bs, Eg, expert_map, pad_sorted_ids=True, ignore_invalid_experts=True,
        return_inv_perm=True)
    P = sorted_ids.numel()
    blockwise = block_shape is not None
    per_token = per_channel_quant and not blockwise

    def permute_rows(q, scales):
        """token-order quantised rows [M, K] (+ scale rows [M, C]) -> sorted order [P, K] (+ [P, C])"""
        K = q.shape[1]
        dst = torch.empty(P, K, dtype=torch.uint8, device=dev)
        C = scales.shape[1]
        sdst = torch.empty(P, C, dtype=torch.float32, device=dev)
        rc = lib.b200_moe_permute_rows(_p(dst), _p(q), _p(sorted_ids), _p(npp), _p(sdst), _p(scales), C,
                                       scales.stride(0), P, topk, numel, K, q.stride(0), _stream())
        _lib.check(rc, "moe_permute_rows")
        return dst, sdst

    def grouped(out, a8, w, a_s, w_s, N, K, a_rows):
        if blockwise:
            rc = lib.b200_moe_grouped_gemm_fp8_blockscaled(_p(out), _p(a8), _p(w), _p(expert_ids), _p(npp), _p(a_s),
                                                           _p(w_s), P, E, N, K, bs, _dt(out), _stream())
        else:
            rc = lib.b200_moe_grouped_gemm_fp8(_p(out), _p(a8), _p(w), _p(expert_ids), _p(npp), _p(a_s), _p(w_s),
                                               None, P, E, N, K, bs, 1 if per_channel_quant else 0,
                                               1 if a_rows else 0, _dt(out), _stream())
        _lib.check(rc, "moe_grouped_gemm_fp8")

    if blockwise:
        if w1_scale is None or w2_scale is None or not w1_scale.is_contiguous() or not w2_scale.is_contiguous():
            raise RuntimeError("fused_experts: block_shape needs contiguous w1_scale / w2_scale")
        q, s = ops.per_token_group_quant_fp8(hidden_states, 128)
        a8, a1_rows = permute_rows(q.view(torch.uint8), s)
        a1_s = a1_rows
    elif per_token and a1_scale is None:
        q, s = ops.scaled_fp8_quant(hidden_states, None, use_per_token_if_dynamic=True)
        a8, a1_rows = permute_rows(q.view(torch.uint8), s.view(M, 1))
        a1_s = a1_rows
    else:
        if a1_scale is None:  # dynamic per-tensor: scale = max|x| / 448 (same kernel as scaled_fp8_quant)
            _, a1_scale = ops.scaled_fp8_quant(hidden_states)
        a8 = torch.empty(P, H, dtype=torch.uint8, device=dev)
        rc = lib.b200_moe_permute_fp8(_p(a8), _p(hidden_states), _p(sorted_ids), _p(npp), _p(a1_scale), P,
                                      topk, numel, H, hidden_states.stride(0), _dt(hidden_states), _stream())
        _lib.check(rc, "moe_permute")
        a1_s = a1_scale
    # blocks past num_tokens_post_pad are skipped by the GEMMs and never gathered: zero-fill only when a
    # dynamic activation scale will reduce over the whole buffer
    dynamic2 = a2_scale is None
    inter = (torch.zeros if dynamic2 else torch.empty)(P, N1, dtype=dt, device=dev)
    grouped(inter, a8, w1, a1_s, w1_scale, N1, H, a_rows=(per_token and a1_scale is None))
    h8 = torch.empty(P, I, dtype=torch.float8_e4m3fn, device=dev)
    a2_rows = False
    if blockwise:
        act = torch.empty(P, I, dtype=dt, device=dev)
        ops.silu_and_mul(act, inter)
        h8, a2_s = ops.per_token_group_quant_fp8(act, 128)
    elif dynamic2:
        act = torch.empty(P, I, dtype=dt, device=dev)
        ops.silu_and_mul(act, inter)
        if per_token:
            _, a2_s = ops.scaled_fp8_quant(act, None, use_per_token_if_dynamic=True, output=h8)
            a2_rows = True
        else:
            _, a2_s = ops.scaled_fp8_quant(act, output=h8)
    else:
        ops.silu_and_mul_quant(h8, inter, a2_scale)
        a2_s = a2_scale
    out2 = torch.empty(P, H, dtype=dt, device=dev)
    grouped(out2, h8, w2, a2_s, w2_scale, H, I, a_rows=a2_rows)
    out = torch.empty(M, H, dtype=dt, device=dev)
    tw = None if apply_router_weight_on_input else topk_weights.to(torch.float32).contiguous()
    rc = lib.b200_moe_unpermute_reduce(_p(out), _p(out2), _p(inv), _p(tw), None, M, topk, H, out.stride(0),
                                       _dt(out), _stream())
    _lib.check(rc, "moe_unpermute")
    return out


def cutlass_moe_fp4(
    a: torch.Tensor,
    a1_gscale: torch.Tensor,
    w1_fp4: torch.Tensor,
    w1_blockscale: torch.Tensor,
    w1_alphas: torch.Tensor,
    a2_gscale: torch.Tensor,
    w2_fp4: torch.Tensor,
    w2_blockscale: torch.Tensor,
    w2_alphas: torch.Tensor,
    topk_weights: torch.Tensor,
    topk_ids: torch.Tensor,
    apply_router_weight_on_input: bool = False,
    output: torch.Tensor | None = None,
) -> torch.Tensor:
    """NVFP4 W4A4 fused MoE, the operator sequence of run_cutlass_moe_fp4 (fused_moe/experts/cutlass_moe.py:
    498-688) with SiLU gating: get_cutlass_moe_mm_data -> shuffle_rows -> scaled_fp4_experts_quant ->
    cutlass_fp4_group_mm (w1) -> silu_and_mul_scaled_fp4_experts_quant -> cutlass_fp4_group_mm (w2) ->
    top-k weighted un-permute.

    a [m, k] bf16/f16; a1_gscale / a2_gscale f32 [e] per-expert activation global scales; w1_fp4 uint8
    [e, 2n, k/2], w1_blockscale float8_e4m3fn [e, 2n, k/16] swizzled per expert, w1_alphas f32 [e];
    w2_fp4 uint8 [e, k, n/2], w2_blockscale [e, k, n/16], w2_alphas f32 [e]; topk_weights [m, topk];
    topk_ids int32 [m, topk] (-1 = dropped)."""
    _cuda(a, a1_gscale, w1_fp4, w1_blockscale, w1_alphas, a2_gscale, w2_fp4, w2_blockscale, w2_alphas,
          topk_weights, topk_ids, output)
    assert topk_weights.shape == topk_ids.shape, "topk shape mismatch"
    assert w1_fp4.dtype == torch.uint8, "weight 1 must be uint8"
    assert w2_fp4.dtype == torch.uint8, "weight 2 must be uint8"
    assert w1_fp4.ndim == 3 and w2_fp4.ndim == 3 and w1_blockscale.ndim == 3 and w2_blockscale.ndim == 3, \
        "All Weights must be of rank 3 for cutlass_moe_fp4"
    m, k = a.shape
    e, w1_n, half_k = w1_fp4.shape
    e2, k_w2, half_n = w2_fp4.shape
    n = half_n * 2
    assert e == e2, f"Number of experts must match between weights. {e}, {e2}"
    assert k == half_k * 2 and k == k_w2, "Hidden size mismatch between a, w1 and w2"
    assert w1_n == 2 * n, "mismatch in expected `n`"
    assert a.dtype in (torch.half, torch.bfloat16), "Invalid input dtype"
    assert topk_weights.size(0) == m and topk_ids.size(0) == m, "topk must be provided for each row of a"
    if topk_ids.dtype != torch.int32:
        topk_ids = topk_ids.to(torch.int32)
    topk_ids = topk_ids.contiguous()
    topk = topk_ids.size(1)
    dev, dt = a.device, a.dtype
    numel = m * topk
    expert_offsets = torch.empty(e + 1, dtype=torch.int32, device=dev)
    blockscale_offsets = torch.empty(e + 1, dtype=torch.int32, device=dev)
    problem_sizes1 = torch.empty((e, 3), dtype=torch.int32, device=dev)
    problem_sizes2 = torch.empty((e, 3), dtype=torch.int32, device=dev)
    a_map = torch.empty(numel, dtype=torch.int32, device=dev)
    c_map = torch.empty(numel, dtype=torch.int32, device=dev)
    if apply_router_weight_on_input:
        assert topk == 1, "apply_router_weight_on_input is only implemented for topk=1"
        a = a * topk_weights.to(dt)
    ops.get_cutlass_moe_mm_data(topk_ids, expert_offsets, problem_sizes1, problem_sizes2, a_map, c_map, e, n, k,
                                blockscale_offsets, is_gated=True)
    # dropped pairs (topk_ids == -1) leave -1 in the tail of a_map: those packed rows are zero-filled
    rep_a = ops.shuffle_rows(a, a_map)
    rep_a_fp4, rep_a_blockscale = ops.scaled_fp4_experts_quant(rep_a, a1_gscale, expert_offsets,
                                                               blockscale_offsets, topk)
    c1 = torch.empty((numel, w1_n), dtype=dt, device=dev)
    ops.cutlass_fp4_moe_mm(c1, rep_a_fp4, w1_fp4, rep_a_blockscale, w1_blockscale, w1_alphas, problem_sizes1,
                           expert_offsets[:-1], blockscale_offsets[:-1])
    int_fp4, int_blockscale = ops.silu_and_mul_scaled_fp4_experts_quant(c1, a2_gscale, expert_offsets,
                                                                        blockscale_offsets, topk)
    c3 = torch.empty((numel, k), dtype=dt, device=dev)
    ops.cutlass_fp4_moe_mm(c3, int_fp4, w2_fp4, int_blockscale, w2_blockscale, w2_alphas, problem_sizes2,
                           expert_offsets[:-1], blockscale_offsets[:-1])
    if output is None:
        output = torch.empty((m, k), dtype=dt, device=dev)
    # c_map is exactly the inverse permutation b200_moe_unpermute_reduce consumes (packed row of each flat
    # (token, j) pair); dropped pairs carry the packed row count expert_offsets[e] and are skipped
    tw = None if apply_router_weight_on_input else topk_weights.to(torch.float32).contiguous()
    rc = _lib.load().b200_moe_unpermute_reduce(_p(output), _p(c3), _p(c_map), _p(tw), _p(expert_offsets[e:]), m,
                                               topk, k, output.stride(0), _dt(output), _stream())
    _lib.check(rc, "moe_unpermute")
    return output
