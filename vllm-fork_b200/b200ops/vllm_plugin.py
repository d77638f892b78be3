"""vLLM general plugin: makes the reference use libb200ops.so for the decode hot path.

Loaded in every worker process through the entry-point group ``vllm.general_plugins``
(vllm/plugins/__init__.py:18,38-42,77-87; allow-listed by VLLM_PLUGINS), declared in
vllm-fork_b200/pyproject.toml as ``b200ops = "b200ops.vllm_plugin:register"``. It

1. registers the ``torch.ops.b200.*`` operators (b200ops/torch_ops.py: the reference's schemas + fakes);
2. rebinds the Python wrappers of ``vllm/_custom_ops.py`` that the hot path calls to their b200ops twins
   (same names, argument meaning and error behaviour, b200ops/ops.py), and ``fused_experts``;
3. installs ``B200AttentionBackend`` in the CUSTOM slot of the backend registry
   (vllm/v1/attention/backends/registry.py:242-294) - select it with ``--attention-backend CUSTOM``.
"""
from __future__ import annotations

import os

# wrappers of vllm/_custom_ops.py replaced 1:1 (name in vllm._custom_ops == name in b200ops.ops)
CUSTOM_OPS_REBOUND = (
    "reshape_and_cache_flash", "reshape_and_cache", "scaled_fp8_quant", "cutlass_scaled_mm",
    "cutlass_scaled_mm_supports_fp8", "scaled_fp4_quant", "cutlass_scaled_fp4_mm", "merge_attn_states",
    "paged_attention_v1", "paged_attention_v2", "rms_norm", "fused_add_rms_norm",
    "rms_norm_static_fp8_quant", "fused_add_rms_norm_static_fp8_quant", "rotary_embedding",
)
MOE_OPS_REBOUND = ("moe_sum", "topk_softmax")

_registered = False


def register() -> None:
    """Entry point. Idempotent (vLLM may load plugins more than once per process)."""
    global _registered
    if _registered:
        return
    os.environ["B200OPS_VLLM_BASES"] = "1"   # b200ops.backend derives from vLLM's ABCs
    from . import fused_moe, ops, torch_ops  # noqa: F401  (torch_ops registers torch.ops.b200.*)
    import vllm._custom_ops as ref
    for name in CUSTOM_OPS_REBOUND:
        if hasattr(ops, name):
            setattr(ref, name, getattr(ops, name))
    for name in MOE_OPS_REBOUND:
        setattr(ref, name, getattr(fused_moe, name))
    try:  # the function object model code imports (fused_moe/__init__.py re-exports it)
        import vllm.model_executor.layers.fused_moe.fused_moe as ref_moe
        ref_moe.b200_fused_experts = fused_moe.fused_experts
    except Exception:  # pragma: no cover - layout differs between vLLM generations
        pass
    from . import backend
    backend.register()
    _registered = True
