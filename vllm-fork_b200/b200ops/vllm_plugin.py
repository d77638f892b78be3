"""vLLM general plugin: makes the reference use libb200ops.so for the decode hot path.

Loaded in every worker process through the entry-point group ``vllm.general_plugins``
(vllm/plugins/__init__.py:18,38-42,77-87; allow-listed by VLLM_PLUGINS), declared in
vllm-fork_b200/pyproject.toml as ``b200ops = "b200ops.vllm_plugin:register"``. It

1. registers the ``torch.ops.b200.*`` operators (b200ops/torch_ops.py: the reference's schemas + fakes);
2. rebinds the Python wrappers of ``vllm/_custom_ops.py`` that the hot path calls to their b200ops twins
   (same names, argument meaning and error behaviour, b200ops/ops.py), and replaces ``fused_experts``
   (fused_moe/fused_moe.py:1587-1628 and the package re-export) by a dispatcher: FP8 W8A8 + SiLU calls on CUDA
   tensors run ``b200ops.fused_moe.fused_experts``; everything this library does not cover (unquantised, int8 /
   int4 / mx schemes, zero points, biases, other activations) is handed back to the function that was there
   before - delegation to the host framework, not a CPU path of ours.
   Limitation: the rebound wrappers are ctypes calls, which Dynamo cannot trace - run with ``enforce_eager`` (CUDA
   graphs of the decode step still work: every entry point is capturable). The ``torch.ops.b200.*`` operators with
   fakes exist for a compiled graph (torch_ops.py) but the wrappers are not routed through them yet;
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


def make_fused_experts_dispatcher(original, b200_impl):
    """Stand-in for the reference's ``fused_experts(hidden_states, w1, w2, topk_weights, topk_ids, [inplace,]
    activation, apply_router_weight_on_input, global_num_experts, expert_map, quant_config)``
    (fused_moe.py:1587-1628; the installed wheel generation also has ``inplace``). Arguments are bound with the
    ORIGINAL function's signature, so positional and keyword call styles both work."""
    import inspect
    sig = inspect.signature(original)

    def covered(args) -> bool:
        qc = args.get("quant_config")
        act = args.get("activation", "silu")
        if qc is None or not getattr(qc, "use_fp8_w8a8", False):
            return False
        if getattr(act, "value", act) != "silu" or not args["hidden_states"].is_cuda:
            return False
        for name in ("use_int8_w8a8", "use_int8_w8a16", "use_int4_w4a16"):
            if getattr(qc, name, False):
                return False
        for name in ("ocp_mx_scheme", "w1_zp", "w2_zp", "w1_bias", "w2_bias"):
            if getattr(qc, name, None) is not None:
                return False
        return True

    def fused_experts(*a, **kw):
        bound = sig.bind(*a, **kw)
        bound.apply_defaults()
        args = bound.arguments
        if not covered(args):
            return original(*a, **kw)
        qc = args["quant_config"]
        out = b200_impl(
            args["hidden_states"], args["w1"], args["w2"], args["topk_weights"], args["topk_ids"],
            activation="silu", apply_router_weight_on_input=args.get("apply_router_weight_on_input", False),
            use_fp8_w8a8=True, per_channel_quant=bool(getattr(qc, "per_act_token_quant", False)),
            global_num_experts=args.get("global_num_experts", -1), expert_map=args.get("expert_map"),
            w1_scale=qc.w1_scale, w2_scale=qc.w2_scale, a1_scale=getattr(qc, "a1_scale", None),
            a2_scale=getattr(qc, "a2_scale", None), block_shape=getattr(qc, "block_shape", None))
        if args.get("inplace", False):
            args["hidden_states"].copy_(out)
            return args["hidden_states"]
        return out

    fused_experts._b200_dispatcher = True
    fused_experts._b200_original = original
    fused_experts.__doc__ = original.__doc__
    return fused_experts


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
        import vllm.model_executor.layers.fused_moe as ref_moe_pkg
        import vllm.model_executor.layers.fused_moe.fused_moe as ref_moe
        if not getattr(ref_moe.fused_experts, "_b200_dispatcher", False):
            disp = make_fused_experts_dispatcher(ref_moe.fused_experts, fused_moe.fused_experts)
            ref_moe.fused_experts = disp
            if hasattr(ref_moe_pkg, "fused_experts"):
                ref_moe_pkg.fused_experts = disp
    except ImportError:  # pragma: no cover - layout differs between vLLM generations
        pass
    from . import backend
    backend.register()
    _registered = True
