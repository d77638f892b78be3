"""GPU: the plugin path end to end on the box's installed vLLM (the reference's generation of the ABCs):
`register()` -> AttentionBackendEnum.CUSTOM resolved through the registry -> backend / builder / impl
instantiated THROUGH the real ABCs -> KV write + decode + prefill on the GPU, equal to the direct operator
calls (those are oracle-checked in test_gpu_attention*.py); torch.ops.b200.* run through the dispatcher,
`torch.library.opcheck` on three of them; a prefill through the backend for every advertised head size."""
import os
import subprocess
import sys
import textwrap

import numpy as np
import pytest
import torch

from oracle import ref_ops

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "vllm-fork_b200")
DEV = "cuda:0"


def _run(code: str, timeout=1200):
    env = dict(os.environ, PYTHONPATH=PKG + os.pathsep + ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    env.pop("B200OPS_VLLM_BASES", None)
    r = subprocess.run([sys.executable, "-c", textwrap.dedent(code)], capture_output=True, text=True, env=env,
                       timeout=timeout, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-4000:] + "\n" + r.stderr[-4000:]
    return r.stdout


def test_registered_backend_through_real_abcs():
    pytest.importorskip("vllm")
    out = _run("""
        import dataclasses, torch
        from b200ops import vllm_plugin
        vllm_plugin.register()
        from b200ops import ops
        from vllm.v1.attention.backend import (AttentionBackend, AttentionCGSupport, AttentionImpl,
                                               AttentionMetadataBuilder, CommonAttentionMetadata)
        from vllm.v1.attention.backends.registry import AttentionBackendEnum
        B = AttentionBackendEnum.CUSTOM.get_class()
        assert issubclass(B, AttentionBackend) and B.get_name() == "CUSTOM"
        dev = torch.device("cuda:0")
        Hq, Hkv, D, bs = 32, 8, 128, 16
        lens = [300, 17, 1024, 65, 1]
        S = len(lens)
        g = torch.Generator().manual_seed(0)
        max_blocks = (max(lens) + bs - 1) // bs
        nb = S * max_blocks + 5
        shape, order = B.get_kv_cache_shape(nb, bs, Hkv, D, "fp8"), B.get_kv_cache_stride_order()
        phys = torch.zeros(tuple(shape[i] for i in order), dtype=torch.uint8, device=dev)
        kv_cache = phys.permute(*[order.index(i) for i in range(4)])
        impl = B.get_impl_cls()(Hq, D, D ** -0.5, Hkv, None, None, "fp8", None, "decoder", None)
        assert isinstance(impl, AttentionImpl)
        builder = B.get_builder_cls()(None, ["layer.0"], None, dev, num_heads=Hq, num_kv_heads=Hkv, head_size=D,
                                      max_num_seqs=64)
        assert isinstance(builder, AttentionMetadataBuilder)
        assert type(builder).get_cudagraph_support(None, None) is AttentionCGSupport.ALWAYS

        class Layer:
            _k_scale = torch.tensor([0.05], device=dev)
            _v_scale = torch.tensor([0.04], device=dev)
            _q_scale = torch.tensor([1.0], device=dev)

        bt = torch.randperm(nb, generator=g)[: S * max_blocks].to(torch.int32).view(S, max_blocks).to(dev)
        T = sum(lens)
        key = (torch.rand(T, Hkv, D, generator=g) * 2 - 1).to(torch.bfloat16).to(dev)
        value = (torch.rand(T, Hkv, D, generator=g) * 2 - 1).to(torch.bfloat16).to(dev)
        slots = torch.cat([bt[s, torch.arange(L, device=dev) // bs].long() * bs + torch.arange(L, device=dev) % bs
                           for s, L in enumerate(lens)])
        impl.do_kv_cache_update(Layer, key, value, kv_cache, slots)
        # the real CommonAttentionMetadata dataclass of this vLLM generation, filled field by field
        fields = {f.name for f in dataclasses.fields(CommonAttentionMetadata)}
        vals = dict(query_start_loc=torch.arange(S + 1, dtype=torch.int32, device=dev),
                    query_start_loc_cpu=torch.arange(S + 1, dtype=torch.int32),
                    seq_lens=torch.tensor(lens, dtype=torch.int32, device=dev),
                    num_reqs=S, num_actual_tokens=S, max_query_len=1, max_seq_len=max(lens),
                    block_table_tensor=bt, slot_mapping=slots[:S])
        try:
            meta_in = CommonAttentionMetadata(**{k: v for k, v in vals.items() if k in fields})
            print("CommonAttentionMetadata: real dataclass")
        except TypeError as e:   # this generation needs more fields: say which, use a duck-typed stand-in
            print("CommonAttentionMetadata differs:", e)
            meta_in = type("M", (), vals)()
        meta = builder.build(0, meta_in)
        q = (torch.rand(S, Hq * D, generator=g) * 2 - 1).to(torch.bfloat16).to(dev)
        out = torch.empty(S, Hq * D, dtype=torch.bfloat16, device=dev)
        impl.forward(Layer, q, None, None, kv_cache, meta, output=out)
        kc, vc = kv_cache.transpose(1, 2).split(D, dim=-1)
        direct = torch.empty(S, Hq, D, dtype=torch.bfloat16, device=dev)
        torch.ops.b200.paged_attention_decode(direct, q.view(S, Hq, D), kc.view(torch.uint8), vc.view(torch.uint8), bt,
                                              vals["seq_lens"], D ** -0.5, Layer._k_scale, Layer._v_scale, None, None,
                                              None, None, max(lens), None, None)
        torch.cuda.synchronize()
        assert torch.equal(out.view(S, Hq, D), direct)
        assert out.float().abs().sum().item() > 0
        # dispatcher path of a few more ops == the Python wrappers
        x = torch.randn(64, 512, device=dev, dtype=torch.bfloat16)
        o1, o2 = torch.empty(64, 256, device=dev, dtype=torch.bfloat16), torch.empty(64, 256, device=dev, dtype=torch.bfloat16)
        torch.ops.b200.silu_and_mul(o1, x); ops.silu_and_mul(o2, x)
        assert torch.equal(o1, o2)
        print("OK")
    """)
    assert "OK" in out
    print(out)


def test_opcheck_three_ops():
    sys.path.insert(0, PKG)
    from b200ops import torch_ops  # noqa: F401
    x = torch.randn(32, 512, device=DEV, dtype=torch.bfloat16)
    o = torch.empty(32, 256, device=DEV, dtype=torch.bfloat16)
    tests = ("test_schema", "test_faketensor")
    torch.library.opcheck(torch.ops.b200.silu_and_mul.default, (o, x), test_utils=tests)
    # (ops with FP8 outputs cannot go through opcheck's test_schema: it clones and compares every argument with
    # aten ops that have no Float8 kernels)
    w0 = torch.ones(512, device=DEV, dtype=torch.bfloat16)
    torch.library.opcheck(torch.ops.b200.rms_norm.default, (torch.empty_like(x), x, w0, 1e-5), test_utils=tests)
    res = torch.randn(32, 512, device=DEV, dtype=torch.bfloat16)
    w = torch.ones(512, device=DEV, dtype=torch.bfloat16)
    torch.library.opcheck(torch.ops.b200.fused_add_rms_norm.default, (x.clone(), res, w, 1e-5), test_utils=tests)
    key = torch.randn(8, 2, 128, device=DEV, dtype=torch.bfloat16)
    kc = torch.zeros(4, 16, 2, 128, device=DEV, dtype=torch.bfloat16)
    vc = torch.zeros(4, 16, 2, 128, device=DEV, dtype=torch.bfloat16)
    slots = torch.arange(8, device=DEV, dtype=torch.int64)
    one = torch.ones(1, device=DEV)
    torch.library.opcheck(torch.ops.b200.reshape_and_cache_flash.default,
                          (key, key.clone(), kc, vc, slots, "auto", one, one), test_utils=tests)


def test_backend_prefill_for_every_advertised_head_size():
    """a head size accepted at backend selection must survive its first prefill."""
    from b200ops.backend import B200AttentionBackend as B
    import dataclasses

    @dataclasses.dataclass
    class CommonMeta:
        query_start_loc: torch.Tensor
        seq_lens: torch.Tensor
        num_reqs: int
        num_actual_tokens: int
        max_query_len: int
        max_seq_len: int
        block_table_tensor: torch.Tensor
        slot_mapping: torch.Tensor

    class Layer:
        _k_scale = torch.tensor([0.05], device=DEV)
        _v_scale = torch.tensor([0.04], device=DEV)

    assert not B.supports_head_size(256)
    for D in B.get_supported_head_sizes():
        Hq, Hkv, bs = 8, 2, 16
        qlens = [70, 1, 33]
        S = len(qlens)
        g = torch.Generator().manual_seed(D)
        max_blocks = 8
        nb = S * max_blocks + 2
        shape, order = B.get_kv_cache_shape(nb, bs, Hkv, D, "auto"), B.get_kv_cache_stride_order()
        phys = torch.zeros(tuple(shape[i] for i in order), dtype=torch.bfloat16, device=DEV)
        kv_cache = phys.permute(*[order.index(i) for i in range(4)])
        bt = torch.randperm(nb, generator=g)[: S * max_blocks].to(torch.int32).view(S, max_blocks)
        impl = B.get_impl_cls()(Hq, D, D ** -0.5, Hkv, None, None, "auto", None)
        builder = B.get_builder_cls()(device=DEV, num_heads=Hq, num_kv_heads=Hkv, head_size=D, max_num_seqs=16)
        T = sum(qlens)
        key = (torch.rand(T, Hkv, D, generator=g) * 2 - 1).to(torch.bfloat16)
        value = (torch.rand(T, Hkv, D, generator=g) * 2 - 1).to(torch.bfloat16)
        slots = torch.cat([bt[s, torch.arange(L) // bs].long() * bs + torch.arange(L) % bs
                           for s, L in enumerate(qlens)])
        impl.do_kv_cache_update(Layer, key.to(DEV), value.to(DEV), kv_cache, slots.to(DEV))
        kv_cpu = torch.zeros(nb, bs, Hkv, 2 * D, dtype=torch.bfloat16)
        ref_ops.reshape_and_cache_flash(key, value, kv_cpu[..., :D], kv_cpu[..., D:], slots, "auto",
                                        torch.ones(1), torch.ones(1))
        cu = torch.zeros(S + 1, dtype=torch.int32)
        cu[1:] = torch.tensor(qlens).cumsum(0)
        q = (torch.rand(T, Hq * D, generator=g) * 2 - 1).to(torch.bfloat16)
        meta = builder.build(0, CommonMeta(
            query_start_loc=cu.to(DEV), seq_lens=torch.tensor(qlens, dtype=torch.int32, device=DEV), num_reqs=S,
            num_actual_tokens=T, max_query_len=max(qlens), max_seq_len=max(qlens),
            block_table_tensor=bt.to(DEV), slot_mapping=slots.to(DEV)))
        out = torch.empty(T, Hq * D, dtype=torch.bfloat16, device=DEV)
        impl.forward(Layer, q.to(DEV), None, None, kv_cache, meta, output=out)
        ref = ref_ops.paged_attention_varlen(q.view(T, Hq, D), kv_cpu[..., :D], kv_cpu[..., D:], qlens, qlens, bt,
                                             D ** -0.5)
        np.testing.assert_allclose(out.float().cpu().numpy().reshape(T, Hq, D), ref, atol=1e-2, rtol=1e-2)
