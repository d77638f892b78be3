"""B200 attention backend for the reference's backend registry (AttentionBackendEnum.CUSTOM).

Mirrors the plug point of vllm/v1/attention/backend.py (AttentionBackend :56-401, AttentionImpl
:895-1006, AttentionMetadataBuilder :623-772) the way TritonAttentionBackend
(vllm/v1/attention/backends/triton_attn.py:322-876) and FlashAttentionBackend
(backends/flash_attn.py:139-173, 779-1175) implement it:

* KV cache: logical shape (num_blocks, num_kv_heads, block_size, 2*head_size), K|V packed in the last
  dim, physical order NHD (stride order (0, 2, 1, 3)); views
  ``key_cache, value_cache = kv_cache.transpose(1, 2).split(head_size, dim=-1)``.
* ``forward_includes_kv_cache_update = False``: the KV write is ``do_kv_cache_update`` ->
  reshape_and_cache_flash; ``forward`` is the paged attention: the split-KV decode kernel when every
  request has one query token (max_query_len == 1), the variable-length kernel (unified_attention:
  prefill / chunked prefill / mixed batches) otherwise - the same dispatch triton_attn.py makes between
  its 2-D and 3-D kernels.
* both kernels read lengths / tables on the device (no host sync) => CUDA-graph support level ALWAYS;
  the metadata builder owns the persistent split-KV workspace.

With B200OPS_VLLM_BASES=1 (set by the plugin before importing this module) the classes derive from
vLLM's ABCs and `register()` installs the backend in the CUSTOM slot
(vllm/v1/attention/backends/registry.py:242-294). Otherwise the same classes work stand-alone (tests/test_gpu_backend.py drives them with a duck-typed CommonAttentionMetadata).
"""
from __future__ import annotations

import dataclasses

import torch

from . import ops

import os

_HAVE_VLLM = False
_CG_ALWAYS = "ALWAYS"
if os.environ.get("B200OPS_VLLM_BASES", "0") == "1":  # set by b200ops.vllm_plugin.register()
    # derive from the reference's ABCs only when asked to (importing vLLM is slow and its ABCs
    # differ between vLLM generations)
    from vllm.v1.attention.backend import (AttentionBackend as _BackendBase,
                                           AttentionCGSupport as _CGSupport,
                                           AttentionImpl as _ImplBase,
                                           AttentionMetadataBuilder as _BuilderBase)
    _HAVE_VLLM = True
    _CG_ALWAYS = _CGSupport.ALWAYS   # the enum member the reference compares against (backend.py:606-626)
if not _HAVE_VLLM:

    class _BackendBase:  # type: ignore[no-redef]
        pass

    class _ImplBase:  # type: ignore[no-redef]
        pass

    class _BuilderBase:  # type: ignore[no-redef]
        def __init__(self, *a, **k):
            pass


@dataclasses.dataclass
class B200AttentionMetadata:
    """Per-step metadata (the subset of TritonAttentionMetadata, triton_attn.py:59-131, the two kernels
    need)."""
    num_actual_tokens: int
    max_query_len: int
    max_seq_len: int
    query_start_loc: torch.Tensor
    seq_lens: torch.Tensor
    block_table: torch.Tensor
    slot_mapping: torch.Tensor
    workspace: torch.Tensor


class B200AttentionMetadataBuilder(_BuilderBase):
    """Builds B200AttentionMetadata from CommonAttentionMetadata (backend.py:404-600)."""

    # AttentionCGSupport.ALWAYS: decode and variable-length kernels are both graph-safe (the enum member when
    # the vLLM ABCs are in use, its name otherwise)
    _cudagraph_support = _CG_ALWAYS

    def __init__(self, kv_cache_spec=None, layer_names=None, vllm_config=None, device=None,
                 num_heads: int | None = None, num_kv_heads: int | None = None,
                 head_size: int | None = None, max_num_seqs: int = 1024):
        # the vLLM base keeps kv_cache_spec / layer_names / vllm_config / device (backend.py:640-651)
        self.kv_cache_spec, self.layer_names = kv_cache_spec, layer_names
        self.vllm_config = vllm_config
        self.device = torch.device(device) if device is not None else torch.device("cuda")
        if vllm_config is not None and num_heads is None:
            mc, pc = vllm_config.model_config, vllm_config.parallel_config
            num_heads = mc.get_num_attention_heads(pc)
            num_kv_heads = mc.get_num_kv_heads(pc)
            head_size = mc.get_head_size()
            max_num_seqs = vllm_config.scheduler_config.max_num_seqs
        self.num_heads, self.num_kv_heads, self.head_size = num_heads, num_kv_heads, head_size
        # persistent workspace sized once for the largest batch (triton_attn.py:154-175)
        self.workspace = ops.paged_attention_workspace(max_num_seqs, num_heads, num_kv_heads,
                                                       head_size, self.device)

    def build(self, common_prefix_len: int, common_attn_metadata, fast_build: bool = False):
        m = common_attn_metadata
        return B200AttentionMetadata(
            num_actual_tokens=m.num_actual_tokens, max_query_len=m.max_query_len,
            max_seq_len=m.max_seq_len, query_start_loc=m.query_start_loc, seq_lens=m.seq_lens,
            block_table=m.block_table_tensor, slot_mapping=m.slot_mapping, workspace=self.workspace)

    def build_for_cudagraph_capture(self, common_attn_metadata):
        return self.build(0, common_attn_metadata)


class B200AttentionImpl(_ImplBase):
    """AttentionImpl (backend.py:895-1006) on libb200ops.so."""

    can_return_lse_for_decode = True
    supports_quant_query_input = False  # the kernel takes bf16/f16 queries (fp8 KV is expanded exactly)

    def __init__(self, num_heads: int, head_size: int, scale: float, num_kv_heads: int | None = None,
                 alibi_slopes: list[float] | None = None, sliding_window: int | None = None,
                 kv_cache_dtype: str = "auto", logits_soft_cap: float | None = None,
                 attn_type: str = "decoder", kv_sharing_target_layer_name: str | None = None,
                 **extra) -> None:
        if attn_type not in ("decoder", "DECODER") and "DECODER" not in str(attn_type).upper():
            raise NotImplementedError("B200AttentionImpl supports decoder self-attention only")
        if kv_cache_dtype not in B200AttentionBackend.supported_kv_cache_dtypes:
            raise NotImplementedError(f"kv_cache_dtype {kv_cache_dtype} not supported")
        self.num_heads, self.head_size, self.scale = num_heads, head_size, float(scale)
        self.num_kv_heads = num_kv_heads or num_heads
        self.alibi_slopes = None if alibi_slopes is None else torch.tensor(alibi_slopes, dtype=torch.float32)
        self.sliding_window = sliding_window
        self.kv_cache_dtype = kv_cache_dtype
        self.logits_soft_cap = logits_soft_cap
        self.kv_sharing_target_layer_name = kv_sharing_target_layer_name

    def _views(self, kv_cache: torch.Tensor):
        # flash_attn.py:941 / triton_attn.py:651-653
        key_cache, value_cache = kv_cache.transpose(1, 2).split(self.head_size, dim=-1)
        return key_cache, value_cache

    def do_kv_cache_update(self, layer, key: torch.Tensor, value: torch.Tensor, kv_cache: torch.Tensor,
                           slot_mapping: torch.Tensor) -> None:
        """flash_attn.py:1141-1175: scatter the new K/V rows into the paged cache."""
        key_cache, value_cache = self._views(kv_cache)
        fp8 = self.kv_cache_dtype.startswith("fp8")
        if fp8:
            key_cache, value_cache = key_cache.view(torch.uint8), value_cache.view(torch.uint8)
        ops.reshape_and_cache_flash(key, value, key_cache, value_cache, slot_mapping,
                                    self.kv_cache_dtype, layer._k_scale, layer._v_scale)

    # ---- RoPE + KV-cache fusion hooks (AttentionImpl, vllm/v1/attention/backend.py:958-1006)
    def fused_rope_kvcache_supported(self) -> bool:
        return True

    def do_rope_and_kv_cache_update(self, layer, query: torch.Tensor, key: torch.Tensor, value: torch.Tensor,
                                    positions: torch.Tensor, cos_sin_cache: torch.Tensor, is_neox: bool,
                                    kv_cache: torch.Tensor, layer_slot_mapping: torch.Tensor) -> None:
        """in-place RoPE of query / key and the KV-cache write, one launch."""
        key_cache, value_cache = self._views(kv_cache)
        if self.kv_cache_dtype.startswith("fp8"):
            key_cache, value_cache = key_cache.view(torch.uint8), value_cache.view(torch.uint8)
        ops.rope_and_reshape_and_cache_flash(positions, query, key, value, self.head_size, cos_sin_cache, is_neox,
                                             key_cache, value_cache, layer_slot_mapping, self.kv_cache_dtype,
                                             layer._k_scale, layer._v_scale)

    def forward(self, layer, query: torch.Tensor, key, value, kv_cache: torch.Tensor, attn_metadata,
                output: torch.Tensor | None = None, output_scale=None, output_block_scale=None,
                lse: torch.Tensor | None = None) -> torch.Tensor:
        assert output is not None, "Output tensor must be provided."
        if output_block_scale is not None:
            raise NotImplementedError("fused block-scaled (NVFP4) output quantization is not supported by B200AttentionImpl")
        if attn_metadata is None:  # profiling run (triton_attn.py:610-612)
            return output.fill_(0)
        n = attn_metadata.num_actual_tokens
        key_cache, value_cache = self._views(kv_cache)
        fp8 = self.kv_cache_dtype.startswith("fp8")
        if fp8:
            key_cache, value_cache = key_cache.view(torch.uint8), value_cache.view(torch.uint8)
        if self.alibi_slopes is not None and self.alibi_slopes.device != query.device:
            self.alibi_slopes = self.alibi_slopes.to(query.device)
        q = query[:n].view(n, self.num_heads, self.head_size)
        o = output[:n].view(n, self.num_heads, self.head_size)
        if attn_metadata.max_query_len > 1:
            # prefill / chunked prefill / mixed batch (triton_attn.py:735-770 -> unified_attention)
            S = attn_metadata.seq_lens.numel()
            win = (self.sliding_window - 1, 0) if self.sliding_window else (-1, -1)
            ops.unified_attention(
                q, key_cache, value_cache, o, attn_metadata.query_start_loc[: S + 1], attn_metadata.max_query_len,
                attn_metadata.seq_lens, attn_metadata.max_seq_len, self.scale, True, win,
                attn_metadata.block_table, self.logits_soft_cap or 0.0, None,
                layer._k_scale if fp8 else None, layer._v_scale if fp8 else None,
                alibi_slopes=self.alibi_slopes, output_scale=output_scale, lse=lse)
            return output
        ops.paged_attention_decode(
            o, q, key_cache, value_cache, attn_metadata.block_table[:n], attn_metadata.seq_lens[:n],
            self.scale, k_scale=layer._k_scale if fp8 else None, v_scale=layer._v_scale if fp8 else None,
            alibi_slopes=self.alibi_slopes, sliding_window=self.sliding_window,
            softcap=self.logits_soft_cap, lse=lse, max_seq_len=attn_metadata.max_seq_len,
            workspace=attn_metadata.workspace, output_scale=output_scale)
        return output


class B200AttentionBackend(_BackendBase):
    accept_output_buffer: bool = True
    supported_dtypes = [torch.float16, torch.bfloat16]
    supported_kv_cache_dtypes = ["auto", "float16", "bfloat16", "fp8", "fp8_e4m3"]
    forward_includes_kv_cache_update: bool = False

    @staticmethod
    def get_name() -> str:
        return "CUSTOM"  # must be a member name of AttentionBackendEnum (attention.py:423)

    @staticmethod
    def get_impl_cls():
        return B200AttentionImpl

    @staticmethod
    def get_builder_cls():
        return B200AttentionMetadataBuilder

    @staticmethod
    def get_supported_kernel_block_sizes():
        try:
            from vllm.v1.attention.backend import MultipleOf
            return [MultipleOf(16)]
        except Exception:
            return [16, 32, 64, 128]

    @staticmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int, head_size: int,
                           cache_dtype_str: str = "auto") -> tuple[int, ...]:
        if block_size % 16 != 0:
            raise ValueError("Block size must be a multiple of 16.")
        return (num_blocks, num_kv_heads, block_size, 2 * head_size)

    @staticmethod
    def get_kv_cache_stride_order(include_num_layers_dimension: bool = False) -> tuple[int, ...]:
        # NHD physical order: one 16-token page of all heads is one contiguous TMA box
        if include_num_layers_dimension:
            return (1, 0, 3, 2, 4)
        return (0, 2, 1, 3)

    # head sizes BOTH kernels serve: the decode kernel also takes 256, but the variable-length (prefill) kernel
    # that forward() routes every max_query_len > 1 batch to handles 64 and 128 only (attention_prefill.cu),
    # so 256 must not be accepted at backend selection
    @classmethod
    def supports_head_size(cls, head_size: int) -> bool:
        return head_size in (64, 128)

    @classmethod
    def get_supported_head_sizes(cls) -> list[int]:
        return [64, 128]

    @classmethod
    def supports_compute_capability(cls, capability) -> bool:
        major = getattr(capability, "major", None)
        if major is None:
            major = capability[0] if isinstance(capability, (tuple, list)) else int(capability) // 10
        return major == 10

    @classmethod
    def supports_kv_cache_dtype(cls, kv_cache_dtype) -> bool:
        return kv_cache_dtype is None or kv_cache_dtype in cls.supported_kv_cache_dtypes

    @classmethod
    def get_required_kv_cache_layout(cls):
        return "NHD"


def register() -> None:
    """Entry point for the `vllm.general_plugins` group (vllm/plugins/__init__.py:18-87): puts this
    backend in the CUSTOM slot; select it with --attention-backend CUSTOM."""
    from vllm.v1.attention.backends.registry import AttentionBackendEnum, register_backend
    register_backend(AttentionBackendEnum.CUSTOM, "b200ops.backend.B200AttentionBackend")
