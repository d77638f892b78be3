"""CPU: the operator boundary as the reference specifies it.

(1) b200ops/torch_ops.py registers every hot-path operator in namespace `b200` with the reference's schema
    (compared argument by argument - name, type, mutability, default - with the schema text of
    csrc/libtorch_stable/torch_bindings.cpp restated in torch_ops.SCHEMAS), each with a fake implementation
    (runs under FakeTensorMode without a GPU) and NO CPU kernel (a CPU tensor raises: no CPU fallback);
(2) the vLLM plugin entry point (vllm-fork_b200/pyproject.toml, group vllm.general_plugins) rebinds
    vllm._custom_ops and installs the backend in AttentionBackendEnum.CUSTOM; the classes derive from the
    installed vLLM's real ABCs and the builder reports AttentionCGSupport.ALWAYS (the enum, not a string).
Runs in a subprocess: the plugin sets B200OPS_VLLM_BASES before b200ops.backend is imported."""
import os
import subprocess
import sys
import textwrap

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "vllm-fork_b200")


def _run(code: str, timeout=900):
    env = dict(os.environ, PYTHONPATH=PKG + os.pathsep + os.environ.get("PYTHONPATH", ""))
    env.pop("B200OPS_VLLM_BASES", None)
    r = subprocess.run([sys.executable, "-c", textwrap.dedent(code)], capture_output=True, text=True, env=env,
                       timeout=timeout, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + "\n" + r.stderr[-3000:]
    return r.stdout


def test_torch_library_ops_registered_with_reference_schemas():
    out = _run("""
        import torch
        from b200ops import torch_ops
        from torch._subclasses.fake_tensor import FakeTensorMode
        names = torch_ops.registered_ops()
        assert len(names) >= 28, names
        for name in names:
            base, _, overload = name.partition(".")
            op = getattr(getattr(torch.ops.b200, base), overload or "default")
            want = torch._C.parse_schema("b200::" + name + torch_ops.SCHEMAS[name][0])
            got = op._schema
            assert len(got.arguments) == len(want.arguments), name
            for a, b in zip(got.arguments, want.arguments):
                assert a.name == b.name and str(a.type) == str(b.type) and a.kwarg_only == b.kwarg_only, (name, a, b)
                assert (a.alias_info is not None and a.alias_info.is_write) == \\
                       (b.alias_info is not None and b.alias_info.is_write), (name, a.name)
                assert a.default_value == b.default_value, (name, a.name)
            assert str(got.returns) == str(want.returns), name
        # fake implementations: trace-able without a GPU
        with FakeTensorMode():
            x = torch.empty(4, 256, dtype=torch.bfloat16, device="cuda")
            o = torch.empty(4, 128, dtype=torch.bfloat16, device="cuda")
            assert torch.ops.b200.silu_and_mul(o, x) is None
            o8 = torch.empty(4, 256, dtype=torch.float8_e4m3fn, device="cuda")
            sc = torch.empty(1, dtype=torch.float32, device="cuda")
            assert torch.ops.b200.static_scaled_fp8_quant(o8, x, sc) is None
            w = torch.empty(512, 256, dtype=torch.float8_e4m3fn, device="cuda")
            out = torch.empty(4, 512, dtype=torch.bfloat16, device="cuda")
            assert torch.ops.b200.cutlass_scaled_mm(out, o8, w.t(), sc, sc, None) is None
            r = torch.ops.b200.fused_experts(x, w, w, x, x, use_fp8_w8a8=True)
            assert r.shape == x.shape and r.dtype == x.dtype
        # no CPU kernel behind any op: the dispatcher refuses CPU tensors
        try:
            torch.ops.b200.silu_and_mul(torch.empty(4, 128), torch.empty(4, 256))
        except NotImplementedError:
            pass
        else:
            raise AssertionError("a CPU call must raise (no CPU fallback)")
        assert torch.ops.b200.cutlass_scaled_mm_supports_fp8(100) is True
        assert torch.ops.b200.cutlass_scaled_mm_supports_fp8(90) is False
        print("OK", len(names))
    """)
    assert "OK" in out


def test_pyproject_declares_vllm_plugin_entry_point():
    import tomllib
    with open(os.path.join(PKG, "pyproject.toml"), "rb") as f:
        cfg = tomllib.load(f)
    eps = cfg["project"]["entry-points"]["vllm.general_plugins"]
    assert eps == {"b200ops": "b200ops.vllm_plugin:register"}
    mod, fn = eps["b200ops"].split(":")
    assert os.path.exists(os.path.join(PKG, *mod.split(".")) + ".py") and fn == "register"


def test_plugin_registers_backend_through_vllm_registry():
    pytest.importorskip("vllm")
    out = _run("""
        import importlib.metadata, sys
        from b200ops import vllm_plugin
        vllm_plugin.register()
        vllm_plugin.register()   # idempotent
        import vllm._custom_ops as ref
        from b200ops import backend, fused_moe, ops
        for name in vllm_plugin.CUSTOM_OPS_REBOUND:
            assert getattr(ref, name) is getattr(ops, name), name
        for name in vllm_plugin.MOE_OPS_REBOUND:
            assert getattr(ref, name) is getattr(fused_moe, name), name
        from vllm.v1.attention.backend import (AttentionBackend, AttentionCGSupport, AttentionImpl,
                                               AttentionMetadataBuilder)
        from vllm.v1.attention.backends.registry import AttentionBackendEnum
        cls = AttentionBackendEnum.CUSTOM.get_class()          # registry.py: resolves the registered path
        assert cls is backend.B200AttentionBackend and issubclass(cls, AttentionBackend)
        assert AttentionBackendEnum[cls.get_name()] is AttentionBackendEnum.CUSTOM
        assert issubclass(cls.get_impl_cls(), AttentionImpl)
        assert issubclass(cls.get_builder_cls(), AttentionMetadataBuilder)
        assert not getattr(cls.get_impl_cls(), "__abstractmethods__", None)
        assert not getattr(cls.get_builder_cls(), "__abstractmethods__", None)
        cg = cls.get_builder_cls().get_cudagraph_support(None, None)
        assert cg is AttentionCGSupport.ALWAYS, cg
        impl = cls.get_impl_cls()(32, 128, 128 ** -0.5, 8, None, None, "fp8", None, "decoder", None)
        assert isinstance(impl, AttentionImpl) and impl.num_kv_heads == 8
        assert cls.get_kv_cache_shape(10, 16, 8, 128, "auto") == (10, 8, 16, 256)
        assert cls.get_kv_cache_stride_order() == (0, 2, 1, 3)
        assert cls.supports_head_size(128) and not cls.supports_head_size(256)
        assert cls.get_supported_head_sizes() == [64, 128]
        import torch
        assert hasattr(torch.ops.b200, "reshape_and_cache_flash")
        print("OK")
    """)
    assert "OK" in out


def test_fused_experts_dispatcher_routes_fp8_silu_to_b200_and_delegates_the_rest():
    """The stand-in the plugin installs for the reference's fused_experts (fused_moe.py:1587-1628): bound with the
    original's signature (positional or keyword), FP8 W8A8 + SiLU on CUDA tensors -> b200 implementation, anything
    else -> the function that was bound before. Pure host logic: both sides are recorded stand-ins."""
    import types

    import torch
    from b200ops.vllm_plugin import make_fused_experts_dispatcher
    calls = []

    def original(hidden_states, w1, w2, topk_weights, topk_ids, inplace=False, activation="silu",
                 apply_router_weight_on_input=False, global_num_experts=-1, expert_map=None, quant_config=None):
        calls.append(("original", inplace))
        return hidden_states

    def ours(hidden_states, w1, w2, topk_weights, topk_ids, **kw):
        calls.append(("b200", kw))
        return hidden_states + 1

    class CudaLike(torch.Tensor):   # a CPU tensor that claims to live on the GPU (host logic only)
        @property
        def is_cuda(self):
            return True

    disp = make_fused_experts_dispatcher(original, ours)
    x = torch.zeros(2, 4).as_subclass(CudaLike)
    w = torch.zeros(1)
    qc = types.SimpleNamespace(use_fp8_w8a8=True, use_int8_w8a8=False, use_int8_w8a16=False, use_int4_w4a16=False,
                               ocp_mx_scheme=None, per_act_token_quant=True, w1_scale=w, w2_scale=w, w1_zp=None,
                               w2_zp=None, a1_scale=None, a2_scale=None, block_shape=[128, 128], w1_bias=None,
                               w2_bias=None)
    silu = types.SimpleNamespace(value="silu")
    r = disp(x, w, w, w, w, False, silu, True, 8, None, qc)          # positional, enum-like activation
    assert calls[-1][0] == "b200" and float(r[0, 0]) == 1.0
    kw = calls[-1][1]
    assert kw["use_fp8_w8a8"] and kw["per_channel_quant"] and kw["apply_router_weight_on_input"]
    assert kw["global_num_experts"] == 8 and kw["block_shape"] == [128, 128] and kw["w1_scale"] is w
    r = disp(x, w, w, w, w, inplace=True, quant_config=qc)            # keyword style, inplace writes the input
    assert calls[-1][0] == "b200" and r is x and float(x[0, 0]) == 1.0
    # everything the library does not cover goes back to the host framework's own function
    n = len(calls)
    disp(x, w, w, w, w, quant_config=None)                                              # unquantised
    disp(x, w, w, w, w, activation=types.SimpleNamespace(value="gelu"), quant_config=qc)  # other activation
    qc.w1_bias = w
    disp(x, w, w, w, w, quant_config=qc)                                                # biases
    qc.w1_bias = None
    disp(torch.zeros(2, 4), w, w, w, w, quant_config=qc)                                # CPU tensor
    assert [c[0] for c in calls[n:]] == ["original"] * 4
    assert disp._b200_dispatcher and disp._b200_original is original


def test_plugin_replaces_fused_experts_in_the_installed_vllm():
    pytest.importorskip("vllm")
    out = _run("""
        import vllm.model_executor.layers.fused_moe as pkg
        import vllm.model_executor.layers.fused_moe.fused_moe as mod
        before = mod.fused_experts
        from b200ops import vllm_plugin
        vllm_plugin.register()
        vllm_plugin.register()   # idempotent: no dispatcher around a dispatcher
        assert mod.fused_experts is not before and mod.fused_experts._b200_dispatcher
        assert mod.fused_experts._b200_original is before
        assert pkg.fused_experts is mod.fused_experts
        print("OK")
    """)
    assert "OK" in out
