"""GPU parity at the backend level: B200AttentionBackend driven like the reference's
tests/v1/attention/test_attention_backends.py (KV cache allocated through get_kv_cache_shape +
get_kv_cache_stride_order, a mock layer with _k_scale/_v_scale, CommonAttentionMetadata-like input),
checked against the oracle (itself pinned to the reference CPU kernels)."""
import dataclasses

import numpy as np
import pytest
import torch

from oracle import ref_ops

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@dataclasses.dataclass
class CommonMeta:  # the fields of CommonAttentionMetadata (backend.py:404-470) the builder reads
    query_start_loc: torch.Tensor
    seq_lens: torch.Tensor
    num_reqs: int
    num_actual_tokens: int
    max_query_len: int
    max_seq_len: int
    block_table_tensor: torch.Tensor
    slot_mapping: torch.Tensor


class MockLayer:  # tests/v1/attention/utils.py MockAttentionLayer
    def __init__(self, k, v):
        self._k_scale = torch.tensor([k], device=DEV)
        self._v_scale = torch.tensor([v], device=DEV)
        self._q_scale = torch.tensor([1.0], device=DEV)


@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8"])
@pytest.mark.parametrize("heads", [(32, 8), (14, 2), (8, 8)])
def test_backend_decode_roundtrip(kv_cache_dtype, heads):
    from b200ops.backend import B200AttentionBackend as B
    Hq, Hkv = heads
    D, bs = (64 if Hq == 14 else 128), 16
    lens = [300, 17, 1024, 65, 1]
    S = len(lens)
    assert B.get_name() == "CUSTOM" and not B.forward_includes_kv_cache_update
    assert B.supports_head_size(D) and B.supports_compute_capability((10, 0))
    g = torch.Generator().manual_seed(0)
    max_blocks = (max(lens) + bs - 1) // bs
    nb = S * max_blocks + 5
    shape = B.get_kv_cache_shape(nb, bs, Hkv, D, kv_cache_dtype)
    order = B.get_kv_cache_stride_order()
    assert shape == (nb, Hkv, bs, 2 * D) and order == (0, 2, 1, 3)
    cdt = torch.uint8 if kv_cache_dtype == "fp8" else torch.bfloat16
    # allocate physically in stride order, expose the logical view (gpu_model_runner.py:7384)
    phys = torch.zeros(tuple(shape[i] for i in order), dtype=cdt, device=DEV)
    inv = [order.index(i) for i in range(4)]
    kv_cache = phys.permute(*inv)
    assert kv_cache.shape == shape
    bt = torch.randperm(nb, generator=g)[: S * max_blocks].to(torch.int32).view(S, max_blocks)
    impl = B.get_impl_cls()(Hq, D, D ** -0.5, Hkv, None, None, kv_cache_dtype, None)
    builder = B.get_builder_cls()(device=DEV, num_heads=Hq, num_kv_heads=Hkv, head_size=D, max_num_seqs=64)
    layer = MockLayer(0.05, 0.04)
    # CPU mirror for the oracle
    kv_cpu = torch.zeros(nb, bs, Hkv, 2 * D, dtype=cdt)
    kc_c, vc_c = kv_cpu[..., :D], kv_cpu[..., D:]
    # fill the context through do_kv_cache_update (many tokens per call)
    T = sum(lens)
    key = (torch.rand(T, Hkv, D, generator=g) * 2 - 1).to(torch.bfloat16)
    value = (torch.rand(T, Hkv, D, generator=g) * 2 - 1).to(torch.bfloat16)
    slots = torch.cat([bt[s, torch.arange(L) // bs].long() * bs + torch.arange(L) % bs
                       for s, L in enumerate(lens)])
    impl.do_kv_cache_update(layer, key.to(DEV), value.to(DEV), kv_cache, slots.to(DEV))
    ref_ops.reshape_and_cache_flash(key, value, kc_c, vc_c, slots, kv_cache_dtype,
                                    layer._k_scale.cpu(), layer._v_scale.cpu())
    assert torch.equal(phys.cpu().view(torch.uint8), kv_cpu.view(torch.uint8))
    q = (torch.rand(S, Hq * D, generator=g) * 2 - 1).to(torch.bfloat16)
    meta = builder.build(0, CommonMeta(
        query_start_loc=torch.arange(S + 1, dtype=torch.int32, device=DEV),
        seq_lens=torch.tensor(lens, dtype=torch.int32, device=DEV), num_reqs=S, num_actual_tokens=S,
        max_query_len=1, max_seq_len=max(lens), block_table_tensor=bt.to(DEV),
        slot_mapping=slots[:S].to(DEV)))
    out = torch.empty(S + 3, Hq * D, dtype=torch.bfloat16, device=DEV)  # padded for CUDA graphs
    qg = torch.zeros(S + 3, Hq * D, dtype=torch.bfloat16, device=DEV)
    qg[:S] = q.to(DEV)
    impl.forward(layer, qg, None, None, kv_cache, meta, output=out)
    fp8 = kv_cache_dtype == "fp8"
    ref = ref_ops.paged_attention_decode(q.view(S, Hq, D), kc_c, vc_c, bt, torch.tensor(lens, dtype=torch.int32),
                                         D ** -0.5, k_scale=layer._k_scale.cpu() if fp8 else None,
                                         v_scale=layer._v_scale.cpu() if fp8 else None)
    np.testing.assert_allclose(out[:S].float().cpu().numpy().reshape(S, Hq, D), ref, atol=1e-2, rtol=1e-2)
    # profiling run: metadata None -> zeros
    assert impl.forward(layer, qg, None, None, kv_cache, None, output=out).abs().sum().item() == 0


@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8"])
def test_backend_prefill_then_decode(kv_cache_dtype):
    """A whole request life cycle through the backend, the way the model runner drives it: chunked prefill
    (mixed batch: prompts of different lengths, one request already decoding) -> do_kv_cache_update ->
    forward (variable-length kernel), then a decode step for everybody (decode kernel). Both checked
    against the oracle on the CPU mirror of the cache."""
    from b200ops.backend import B200AttentionBackend as B
    Hq, Hkv, D, bs = 32, 8, 128, 16
    fp8 = kv_cache_dtype == "fp8"
    ctx = [0, 0, 0, 200]            # tokens already in the cache
    qlens = [130, 64, 7, 1]         # new tokens this step
    S = len(ctx)
    g = torch.Generator().manual_seed(1)
    max_blocks = 24
    nb = S * max_blocks + 3
    cdt = torch.uint8 if fp8 else torch.bfloat16
    shape, order = B.get_kv_cache_shape(nb, bs, Hkv, D, kv_cache_dtype), B.get_kv_cache_stride_order()
    phys = torch.zeros(tuple(shape[i] for i in order), dtype=cdt, device=DEV)
    kv_cache = phys.permute(*[order.index(i) for i in range(4)])
    bt = torch.randperm(nb, generator=g)[: S * max_blocks].to(torch.int32).view(S, max_blocks)
    impl = B.get_impl_cls()(Hq, D, D ** -0.5, Hkv, None, None, kv_cache_dtype, None)
    builder = B.get_builder_cls()(device=DEV, num_heads=Hq, num_kv_heads=Hkv, head_size=D, max_num_seqs=64)
    assert str(builder._cudagraph_support).endswith("ALWAYS")
    layer = MockLayer(0.05, 0.04)
    ks, vs = (layer._k_scale.cpu(), layer._v_scale.cpu()) if fp8 else (None, None)
    kv_cpu = torch.zeros(nb, bs, Hkv, 2 * D, dtype=cdt)
    kc_c, vc_c = kv_cpu[..., :D], kv_cpu[..., D:]

    def slots_of(s, lo, hi):
        pos = torch.arange(lo, hi)
        return bt[s, pos // bs].long() * bs + pos % bs

    def write(slots, n):
        key = (torch.rand(n, Hkv, D, generator=g) * 2 - 1).to(torch.bfloat16)
        value = (torch.rand(n, Hkv, D, generator=g) * 2 - 1).to(torch.bfloat16)
        impl.do_kv_cache_update(layer, key.to(DEV), value.to(DEV), kv_cache, slots.to(DEV))
        ref_ops.reshape_and_cache_flash(key, value, kc_c, vc_c, slots, kv_cache_dtype,
                                        layer._k_scale.cpu(), layer._v_scale.cpu())

    write(slots_of(3, 0, ctx[3]), ctx[3])                       # request 3 is past its prefill
    # ---- step 1: mixed batch
    slots = torch.cat([slots_of(s, ctx[s], ctx[s] + qlens[s]) for s in range(S)])
    T = sum(qlens)
    write(slots, T)
    lens = [c + q for c, q in zip(ctx, qlens)]
    cu = torch.zeros(S + 1, dtype=torch.int32)
    cu[1:] = torch.tensor(qlens).cumsum(0)
    q = (torch.rand(T, Hq * D, generator=g) * 2 - 1).to(torch.bfloat16)
    meta = builder.build(0, CommonMeta(
        query_start_loc=cu.to(DEV), seq_lens=torch.tensor(lens, dtype=torch.int32, device=DEV), num_reqs=S,
        num_actual_tokens=T, max_query_len=max(qlens), max_seq_len=max(lens), block_table_tensor=bt.to(DEV),
        slot_mapping=slots.to(DEV)))
    out = torch.empty(T + 5, Hq * D, dtype=torch.bfloat16, device=DEV)
    qg = torch.zeros(T + 5, Hq * D, dtype=torch.bfloat16, device=DEV)
    qg[:T] = q.to(DEV)
    impl.forward(layer, qg, None, None, kv_cache, meta, output=out)
    ref = ref_ops.paged_attention_varlen(q.view(T, Hq, D), kc_c, vc_c, qlens, lens, bt, D ** -0.5,
                                         k_scale=ks, v_scale=vs)
    np.testing.assert_allclose(out[:T].float().cpu().numpy().reshape(T, Hq, D), ref, atol=1e-2, rtol=1e-2)
    # ---- step 2: everybody decodes one token
    slots = torch.cat([slots_of(s, lens[s], lens[s] + 1) for s in range(S)])
    write(slots, S)
    lens = [n + 1 for n in lens]
    q = (torch.rand(S, Hq * D, generator=g) * 2 - 1).to(torch.bfloat16)
    meta = builder.build(0, CommonMeta(
        query_start_loc=torch.arange(S + 1, dtype=torch.int32, device=DEV),
        seq_lens=torch.tensor(lens, dtype=torch.int32, device=DEV), num_reqs=S, num_actual_tokens=S,
        max_query_len=1, max_seq_len=max(lens), block_table_tensor=bt.to(DEV), slot_mapping=slots.to(DEV)))
    out = torch.empty(S, Hq * D, dtype=torch.bfloat16, device=DEV)
    impl.forward(layer, q.to(DEV), None, None, kv_cache, meta, output=out)
    ref = ref_ops.paged_attention_decode(q.view(S, Hq, D), kc_c, vc_c, bt, torch.tensor(lens, dtype=torch.int32),
                                         D ** -0.5, k_scale=ks, v_scale=vs)
    np.testing.assert_allclose(out.float().cpu().numpy().reshape(S, Hq, D), ref, atol=1e-2, rtol=1e-2)
