"""torch.library registration of the hot-path operators in namespace ``b200``.

The reference reaches its kernels through ``torch.ops._C*`` operators whose schemas live in
csrc/libtorch_stable/torch_bindings.cpp (:116-119, :187-190, :218-222, :353-362, :369-390, :417-420, :571-584,
:634-648, :889-903) and csrc/libtorch_stable/moe/torch_bindings.cpp (:8-10, :30-41); every one of them also
carries a fake (meta) implementation for torch.compile (vllm/_custom_ops.py:24-33 ``register_fake``).
This module re-registers the SAME schemas, verbatim, as ``torch.ops.b200.<name>``:

* the CUDA implementation of each op is the matching wrapper of b200ops/ops.py / b200ops/fused_moe.py (one
  ctypes call into libb200ops.so - there is no CPU kernel registered: a CPU tensor raises
  NotImplementedError from the dispatcher, i.e. no CPU fallback);
* every op has a fake implementation (all ops mutate caller-allocated outputs and return ``()``, so the
  fake is a no-op; ``scaled_fp4_quant`` / ``cutlass_scaled_mm_supports_fp8`` return the right metadata).

``paged_attention_v1/v2`` use the schema recovered from the installed wheel: the op is
absent from the reference tree. ``fused_experts`` mirrors the ``vllm::fused_experts`` custom op
(fused_moe.py:1448-1535, registered with direct_register_custom_op).
"""
from __future__ import annotations

import torch

from . import fused_moe, ops

NAMESPACE = "b200"
_lib = torch.library.Library(NAMESPACE, "DEF")

# name -> (schema tail exactly as the reference declares it, reference location)
SCHEMAS: dict[str, tuple[str, str]] = {
    "reshape_and_cache_flash": (
        "(Tensor key, Tensor value, Tensor! key_cache, Tensor! value_cache, Tensor slot_mapping, "
        "str kv_cache_dtype, Tensor k_scale, Tensor v_scale) -> ()", "torch_bindings.cpp:897-903"),
    "reshape_and_cache": (
        "(Tensor key, Tensor value, Tensor! key_cache, Tensor! value_cache, Tensor slot_mapping, "
        "str kv_cache_dtype, Tensor k_scale, Tensor v_scale) -> ()", "torch_bindings.cpp:889-894"),
    "static_scaled_fp8_quant": (
        "(Tensor! result, Tensor input, Tensor scale, int[]? group_shape=None) -> ()", "torch_bindings.cpp:634-636"),
    "dynamic_scaled_fp8_quant": (
        "(Tensor! result, Tensor input, Tensor! scale) -> ()", "torch_bindings.cpp:639-642"),
    "dynamic_per_token_scaled_fp8_quant": (
        "(Tensor! result, Tensor input, Tensor! scale, Tensor? scale_ub) -> ()", "torch_bindings.cpp:645-648"),
    "per_token_group_fp8_quant": (
        "(Tensor input, Tensor! output_q, Tensor! output_s, int group_size, float eps, float fp8_min, "
        "float fp8_max, bool scale_ue8m0, bool dummy_is_scale_transposed, bool dummy_is_tma_aligned) -> ()",
        "torch_bindings.cpp:13-19"),
    "cutlass_scaled_mm": (
        "(Tensor! out, Tensor a, Tensor b, Tensor a_scales, Tensor b_scales, Tensor? bias) -> ()",
        "torch_bindings.cpp:116-119"),
    "cutlass_scaled_mm_supports_fp8": ("(int cuda_device_capability) -> bool", "torch_bindings.cpp:131"),
    "cutlass_scaled_fp4_mm": (
        "(Tensor! out, Tensor a, Tensor b, Tensor block_scale_a, Tensor block_scale_b, Tensor alpha) -> ()",
        "torch_bindings.cpp:187-190"),
    "scaled_fp4_quant.out": (
        "(Tensor input, Tensor input_scale, bool is_sf_swizzled_layout, *, Tensor(a!) output, "
        "Tensor(b!) output_scale) -> ()", "torch_bindings.cpp:218-222"),
    "silu_and_mul_nvfp4_quant": (
        "(Tensor! result, Tensor! result_block_scale, Tensor input, Tensor input_global_scale) -> ()",
        "torch_bindings.cpp:251-254"),
    "scaled_fp4_experts_quant": (
        "(Tensor! output, Tensor! output_scale, Tensor input, Tensor input_global_scale, "
        "Tensor input_offset_by_experts, Tensor output_scale_offset_by_experts) -> ()", "torch_bindings.cpp:225-229"),
    "cutlass_fp4_group_mm": (
        "(Tensor! out, Tensor a, Tensor b, Tensor a_blockscale, Tensor b_blockscales, Tensor alphas, "
        "Tensor problem_sizes, Tensor expert_offsets, Tensor sf_offsets) -> ()", "torch_bindings.cpp:193-197"),
    "silu_and_mul_scaled_fp4_experts_quant": (
        "(Tensor! output, Tensor! output_scale, Tensor input, Tensor input_global_scale, "
        "Tensor input_offset_by_experts, Tensor output_scale_offset_by_experts) -> ()", "torch_bindings.cpp:231-236"),
    "get_cutlass_moe_mm_data": (
        "(Tensor topk_ids, Tensor! expert_offsets, Tensor! problem_sizes1, Tensor! problem_sizes2, "
        "Tensor! input_permutation, Tensor! output_permutation, int num_experts, int n, int k, "
        "Tensor? blockscale_offsets, bool is_gated) -> ()", "torch_bindings.cpp:151-157"),
    "shuffle_rows": (
        "(Tensor input_tensor, Tensor dst2src_map, Tensor! output_tensor) -> ()", "moe/torch_bindings.cpp:119-121"),
    "silu_and_mul": ("(Tensor! result, Tensor input) -> ()", "torch_bindings.cpp:571"),
    "silu_and_mul_quant": ("(Tensor! result, Tensor input, Tensor scale) -> ()", "torch_bindings.cpp:583-584"),
    "merge_attn_states": (
        "(Tensor! output, Tensor!? output_lse, Tensor prefix_output, Tensor prefix_lse, Tensor suffix_output, "
        "Tensor suffix_lse, int? prefill_tokens_with_context, Tensor? output_scale=None) -> ()",
        "torch_bindings.cpp:353-362 (the reference writes `int!?`; torch.library's Python parser has no "
        "mutable-int spelling, the argument is read-only here as well)"),
    "rms_norm": ("(Tensor! result, Tensor input, Tensor? weight, float epsilon) -> ()", "torch_bindings.cpp:369"),
    "fused_add_rms_norm": (
        "(Tensor! input, Tensor! residual, Tensor? weight, float epsilon) -> ()", "torch_bindings.cpp:375"),
    "rms_norm_static_fp8_quant": (
        "(Tensor! result, Tensor input, Tensor weight, Tensor scale, float epsilon) -> ()", "torch_bindings.cpp:381"),
    "fused_add_rms_norm_static_fp8_quant": (
        "(Tensor! result, Tensor input, Tensor! residual, Tensor weight, Tensor scale, float epsilon) -> ()",
        "torch_bindings.cpp:387"),
    "rotary_embedding": (
        "(Tensor positions, Tensor! query, Tensor!? key, int head_size, Tensor cos_sin_cache, bool is_neox, "
        "int rope_dim_offset=0, bool inverse=False) -> ()", "torch_bindings.cpp:417-420"),
    "apply_repetition_penalties_": (
        "(Tensor! logits, Tensor prompt_mask, Tensor output_mask, Tensor repetition_penalties) -> ()",
        "torch_bindings.cpp:539-541"),
    "moe_align_block_size": (
        "(Tensor topk_ids, int num_experts, int block_size, Tensor! sorted_token_ids, Tensor! experts_ids, "
        "Tensor! num_tokens_post_pad, Tensor? maybe_expert_map) -> ()", "moe/torch_bindings.cpp:36-41"),
    "moe_sum": (
        "(Tensor input, Tensor! output, Tensor? topk_ids=None, Tensor? expert_map=None) -> ()",
        "moe/torch_bindings.cpp:30-32"),
    "topk_softmax": (
        "(Tensor! topk_weights, Tensor! topk_indices, Tensor! token_expert_indices, Tensor gating_output, "
        "bool renormalize, Tensor? bias, Tensor? is_padding) -> ()", "moe/torch_bindings.cpp:8-10"),
    "paged_attention_v1": (
        "(Tensor! out, Tensor query, Tensor key_cache, Tensor value_cache, int num_kv_heads, float scale, "
        "Tensor block_tables, Tensor seq_lens, int block_size, int max_seq_len, Tensor? alibi_slopes, "
        "str kv_cache_dtype, Tensor k_scale, Tensor v_scale, int tp_rank, int blocksparse_local_blocks, "
        "int blocksparse_vert_stride, int blocksparse_block_size, int blocksparse_head_sliding_step) -> ()",
        "wheel _C.abi3.so schema (SURVEY.md 8b; absent from the reference tree)"),
    "paged_attention_v2": (
        "(Tensor! out, Tensor! exp_sums, Tensor! max_logits, Tensor! tmp_out, Tensor query, Tensor key_cache, "
        "Tensor value_cache, int num_kv_heads, float scale, Tensor block_tables, Tensor seq_lens, int block_size, "
        "int max_seq_len, Tensor? alibi_slopes, str kv_cache_dtype, Tensor k_scale, Tensor v_scale, int tp_rank, "
        "int blocksparse_local_blocks, int blocksparse_vert_stride, int blocksparse_block_size, "
        "int blocksparse_head_sliding_step) -> ()", "wheel _C.abi3.so schema (SURVEY.md 8b)"),
    # the fork's flash-layout decode attention has no _C schema in the reference (its backends call Triton /
    # FlashAttention functions); this op is what B200AttentionImpl.forward lowers to
    "paged_attention_decode": (
        "(Tensor! out, Tensor query, Tensor key_cache, Tensor value_cache, Tensor block_table, Tensor seq_lens, "
        "float scale, Tensor? k_scale, Tensor? v_scale, Tensor? alibi_slopes, int? sliding_window, "
        "float? softcap, Tensor!? lse, int max_seq_len, Tensor!? workspace, Tensor? output_scale) -> ()",
        "B200AttentionImpl.forward (vllm/v1/attention/backend.py:895-1006 contract)"),
    "fused_experts": (
        "(Tensor hidden_states, Tensor w1, Tensor w2, Tensor topk_weights, Tensor topk_ids, "
        "str activation='silu', bool apply_router_weight_on_input=False, bool use_fp8_w8a8=False, "
        "bool per_channel_quant=False, int global_num_experts=-1, Tensor? expert_map=None, "
        "Tensor? w1_scale=None, Tensor? w2_scale=None, Tensor? a1_scale=None, Tensor? a2_scale=None, "
        "int[]? block_shape=None) -> Tensor", "fused_moe.py:1448-1535 (vllm::fused_experts)"),
}


# ---------------------------------------------------------------------------------------------- CUDA impls
def _reshape_and_cache_flash(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype, k_scale, v_scale):
    ops.reshape_and_cache_flash(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype, k_scale, v_scale)


def _reshape_and_cache(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype, k_scale, v_scale):
    ops.reshape_and_cache(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype, k_scale, v_scale)


def _static_scaled_fp8_quant(result, input, scale, group_shape=None):
    ops.static_scaled_fp8_quant(result, input, scale, None if group_shape is None else tuple(group_shape))


def _dynamic_scaled_fp8_quant(result, input, scale):
    ops.dynamic_scaled_fp8_quant(result, input, scale)


def _dynamic_per_token_scaled_fp8_quant(result, input, scale, scale_ub):
    ops.dynamic_per_token_scaled_fp8_quant(result, input, scale, scale_ub)


def _per_token_group_fp8_quant(input, output_q, output_s, group_size, eps, fp8_min, fp8_max, scale_ue8m0,
                               dummy_is_scale_transposed, dummy_is_tma_aligned):
    ops.per_token_group_fp8_quant(input, output_q, output_s, group_size, eps, fp8_min, fp8_max, scale_ue8m0,
                                  dummy_is_scale_transposed, dummy_is_tma_aligned)


def _cutlass_scaled_mm(out, a, b, a_scales, b_scales, bias):
    ops.cutlass_scaled_mm(a, b, a_scales, b_scales, out.dtype, bias, out=out)


def _cutlass_scaled_mm_supports_fp8(cuda_device_capability):
    return ops.cutlass_scaled_mm_supports_fp8(cuda_device_capability)


def _cutlass_scaled_fp4_mm(out, a, b, block_scale_a, block_scale_b, alpha):
    ops.cutlass_scaled_fp4_mm(a, b, block_scale_a, block_scale_b, alpha, out.dtype, out=out)


def _scaled_fp4_quant_out(input, input_scale, is_sf_swizzled_layout, *, output, output_scale):
    ops.scaled_fp4_quant(input, input_scale, is_sf_swizzled_layout, output=output, output_scale=output_scale)


def _silu_and_mul_nvfp4_quant(result, result_block_scale, input, input_global_scale):
    ops.silu_and_mul_nvfp4_quant(result, result_block_scale, input, input_global_scale)


def _scaled_fp4_experts_quant(output, output_scale, input, input_global_scale, input_offset_by_experts,
                              output_scale_offset_by_experts):
    ops.scaled_fp4_experts_quant_out(output, output_scale, input, input_global_scale, input_offset_by_experts,
                                     output_scale_offset_by_experts)


def _cutlass_fp4_group_mm(out, a, b, a_blockscale, b_blockscales, alphas, problem_sizes, expert_offsets,
                          sf_offsets):
    ops.cutlass_fp4_group_mm(out, a, b, a_blockscale, b_blockscales, alphas, problem_sizes, expert_offsets,
                             sf_offsets)


def _silu_and_mul_scaled_fp4_experts_quant(output, output_scale, input, input_global_scale,
                                           input_offset_by_experts, output_scale_offset_by_experts):
    ops.silu_and_mul_scaled_fp4_experts_quant_out(output, output_scale, input, input_global_scale,
                                                  input_offset_by_experts, output_scale_offset_by_experts)


def _get_cutlass_moe_mm_data(topk_ids, expert_offsets, problem_sizes1, problem_sizes2, input_permutation,
                             output_permutation, num_experts, n, k, blockscale_offsets, is_gated):
    ops.get_cutlass_moe_mm_data(topk_ids, expert_offsets, problem_sizes1, problem_sizes2, input_permutation,
                                output_permutation, num_experts, n, k, blockscale_offsets, is_gated)


def _shuffle_rows(input_tensor, dst2src_map, output_tensor):
    ops.shuffle_rows(input_tensor, dst2src_map, output_tensor)


def _silu_and_mul(result, input):
    ops.silu_and_mul(result, input)


def _silu_and_mul_quant(result, input, scale):
    ops.silu_and_mul_quant(result, input, scale)


def _merge_attn_states(output, output_lse, prefix_output, prefix_lse, suffix_output, suffix_lse,
                       prefill_tokens_with_context, output_scale=None):
    ops.merge_attn_states(output, prefix_output, prefix_lse, suffix_output, suffix_lse, output_lse,
                          prefill_tokens_with_context, output_scale)


def _rms_norm(result, input, weight, epsilon):
    if weight is None:
        weight = torch.ones(input.shape[-1], dtype=input.dtype, device=input.device)
    ops.rms_norm(result, input, weight, epsilon)


def _fused_add_rms_norm(input, residual, weight, epsilon):
    if weight is None:
        weight = torch.ones(input.shape[-1], dtype=input.dtype, device=input.device)
    ops.fused_add_rms_norm(input, residual, weight, epsilon)


def _rms_norm_static_fp8_quant(result, input, weight, scale, epsilon):
    ops.rms_norm_static_fp8_quant(result, input, weight, scale, epsilon)


def _fused_add_rms_norm_static_fp8_quant(result, input, residual, weight, scale, epsilon):
    ops.fused_add_rms_norm_static_fp8_quant(result, input, residual, weight, scale, epsilon)


def _rotary_embedding(positions, query, key, head_size, cos_sin_cache, is_neox, rope_dim_offset=0, inverse=False):
    if rope_dim_offset != 0 or inverse:
        raise NotImplementedError("b200::rotary_embedding: rope_dim_offset / inverse are not on the decode path")
    ops.rotary_embedding(positions, query, key, head_size, cos_sin_cache, is_neox)


def _apply_repetition_penalties_(logits, prompt_mask, output_mask, repetition_penalties):
    ops.apply_repetition_penalties(logits, prompt_mask, output_mask, repetition_penalties)


def _moe_align_block_size(topk_ids, num_experts, block_size, sorted_token_ids, experts_ids, num_tokens_post_pad,
                          maybe_expert_map):
    fused_moe.moe_align_block_size_op(topk_ids, num_experts, block_size, sorted_token_ids, experts_ids,
                                      num_tokens_post_pad, maybe_expert_map)


def _moe_sum(input, output, topk_ids=None, expert_map=None):
    fused_moe.moe_sum(input, output, topk_ids, expert_map)


def _topk_softmax(topk_weights, topk_indices, token_expert_indices, gating_output, renormalize, bias, is_padding):
    if is_padding is not None:
        raise NotImplementedError("b200::topk_softmax: is_padding is not supported")
    fused_moe.topk_softmax(topk_weights, topk_indices, token_expert_indices, gating_output, renormalize, bias)


def _paged_attention_v1(out, query, key_cache, value_cache, num_kv_heads, scale, block_tables, seq_lens,
                        block_size, max_seq_len, alibi_slopes, kv_cache_dtype, k_scale, v_scale, tp_rank,
                        blocksparse_local_blocks, blocksparse_vert_stride, blocksparse_block_size,
                        blocksparse_head_sliding_step):
    ops.paged_attention_v1(out, query, key_cache, value_cache, num_kv_heads, scale, block_tables, seq_lens,
                           block_size, max_seq_len, alibi_slopes, kv_cache_dtype, k_scale, v_scale, tp_rank,
                           blocksparse_local_blocks, blocksparse_vert_stride, blocksparse_block_size,
                           blocksparse_head_sliding_step)


def _paged_attention_v2(out, exp_sums, max_logits, tmp_out, query, key_cache, value_cache, num_kv_heads, scale,
                        block_tables, seq_lens, block_size, max_seq_len, alibi_slopes, kv_cache_dtype, k_scale,
                        v_scale, tp_rank, blocksparse_local_blocks, blocksparse_vert_stride,
                        blocksparse_block_size, blocksparse_head_sliding_step):
    ops.paged_attention_v2(out, exp_sums, max_logits, tmp_out, query, key_cache, value_cache, num_kv_heads,
                           scale, block_tables, seq_lens, block_size, max_seq_len, alibi_slopes, kv_cache_dtype,
                           k_scale, v_scale, tp_rank, blocksparse_local_blocks, blocksparse_vert_stride,
                           blocksparse_block_size, blocksparse_head_sliding_step)


def _paged_attention_decode(out, query, key_cache, value_cache, block_table, seq_lens, scale, k_scale, v_scale,
                            alibi_slopes, sliding_window, softcap, lse, max_seq_len, workspace, output_scale):
    ops.paged_attention_decode(out, query, key_cache, value_cache, block_table, seq_lens, scale,
                               k_scale=k_scale, v_scale=v_scale, alibi_slopes=alibi_slopes,
                               sliding_window=sliding_window, softcap=softcap, lse=lse,
                               max_seq_len=max_seq_len, workspace=workspace, output_scale=output_scale)


def _fused_experts(hidden_states, w1, w2, topk_weights, topk_ids, activation="silu",
                   apply_router_weight_on_input=False, use_fp8_w8a8=False, per_channel_quant=False,
                   global_num_experts=-1, expert_map=None, w1_scale=None, w2_scale=None, a1_scale=None,
                   a2_scale=None, block_shape=None):
    return fused_moe.fused_experts(hidden_states, w1, w2, topk_weights, topk_ids, activation=activation,
                                   apply_router_weight_on_input=apply_router_weight_on_input,
                                   use_fp8_w8a8=use_fp8_w8a8, per_channel_quant=per_channel_quant,
                                   global_num_experts=global_num_experts, expert_map=expert_map,
                                   w1_scale=w1_scale, w2_scale=w2_scale, a1_scale=a1_scale, a2_scale=a2_scale,
                                   block_shape=block_shape)


_IMPLS = {
    "reshape_and_cache_flash": _reshape_and_cache_flash, "reshape_and_cache": _reshape_and_cache,
    "static_scaled_fp8_quant": _static_scaled_fp8_quant, "dynamic_scaled_fp8_quant": _dynamic_scaled_fp8_quant,
    "dynamic_per_token_scaled_fp8_quant": _dynamic_per_token_scaled_fp8_quant,
    "per_token_group_fp8_quant": _per_token_group_fp8_quant, "cutlass_scaled_mm": _cutlass_scaled_mm,
    "cutlass_scaled_mm_supports_fp8": _cutlass_scaled_mm_supports_fp8,
    "cutlass_scaled_fp4_mm": _cutlass_scaled_fp4_mm, "scaled_fp4_quant.out": _scaled_fp4_quant_out,
    "silu_and_mul_nvfp4_quant": _silu_and_mul_nvfp4_quant, "scaled_fp4_experts_quant": _scaled_fp4_experts_quant,
    "cutlass_fp4_group_mm": _cutlass_fp4_group_mm, "silu_and_mul": _silu_and_mul,
    "silu_and_mul_scaled_fp4_experts_quant": _silu_and_mul_scaled_fp4_experts_quant,
    "get_cutlass_moe_mm_data": _get_cutlass_moe_mm_data, "shuffle_rows": _shuffle_rows,
    "silu_and_mul_quant": _silu_and_mul_quant, "merge_attn_states": _merge_attn_states, "rms_norm": _rms_norm,
    "fused_add_rms_norm": _fused_add_rms_norm, "rms_norm_static_fp8_quant": _rms_norm_static_fp8_quant,
    "fused_add_rms_norm_static_fp8_quant": _fused_add_rms_norm_static_fp8_quant,
    "rotary_embedding": _rotary_embedding, "apply_repetition_penalties_": _apply_repetition_penalties_,
    "moe_align_block_size": _moe_align_block_size, "moe_sum": _moe_sum, "topk_softmax": _topk_softmax,
    "paged_attention_v1": _paged_attention_v1, "paged_attention_v2": _paged_attention_v2,
    "paged_attention_decode": _paged_attention_decode, "fused_experts": _fused_experts,
}


def _fake_none(*args, **kwargs):
    """all these ops write caller-allocated outputs and return (): nothing to compute under FakeTensor"""
    return None


def _fake_fused_experts(hidden_states, *args, **kwargs):
    return torch.empty_like(hidden_states)   # fused_moe.py:1537-1560 fused_experts_fake


for _name, (_schema, _where) in SCHEMAS.items():
    _lib.define(_name + _schema)
    if _name == "cutlass_scaled_mm_supports_fp8":
        # no tensor argument: not dispatched by device (the reference registers it the same way, :131-132)
        _lib.impl(_name, _IMPLS[_name], "CompositeExplicitAutograd")
        continue
    _lib.impl(_name, _IMPLS[_name], "CUDA")
    torch.library.register_fake(f"{NAMESPACE}::{_name}", _fake_fused_experts if _name == "fused_experts" else _fake_none,
                                lib=_lib)


def registered_ops() -> list[str]:
    return sorted(SCHEMAS)
