"""Print the schemas of the installed vllm wheel's ops that overlap our C ABI (GPU box only)."""
import os
import torch
import vllm
d = os.path.dirname(vllm.__file__)
for lib in ("_C.abi3.so", "_C_stable_libtorch.abi3.so", "_moe_C.abi3.so"):
    try:
        torch.ops.load_library(os.path.join(d, lib)); print("loaded", lib)
    except Exception as e:
        print("fail", lib, str(e)[:300])
ops = ["_C::paged_attention_v1", "_C::paged_attention_v2", "_C::cutlass_scaled_mm", "_C::scaled_fp4_quant",
       "_C_cache_ops::reshape_and_cache_flash", "_C_cache_ops::reshape_and_cache", "_moe_C::moe_align_block_size",
       "_C::merge_attn_states", "_C::static_scaled_fp8_quant", "_C::dynamic_scaled_fp8_quant",
       "_C::silu_and_mul", "_C::silu_and_mul_quant", "_C::rms_norm", "_C::fused_add_rms_norm",
       "_C::rms_norm_static_fp8_quant", "_C::rotary_embedding", "_moe_C::topk_softmax", "_C::cutlass_scaled_fp4_mm",
       "_C::silu_and_mul_nvfp4_quant", "_C::dynamic_per_token_scaled_fp8_quant",
       "_C::fused_add_rms_norm_static_fp8_quant", "_moe_C::moe_sum"]
for op in ops:
    found = False
    for ov in ("", "out"):
        try:
            print(torch._C._get_schema(op, ov)); found = True
        except Exception:
            pass
    if not found:
        print(op, "MISSING")
