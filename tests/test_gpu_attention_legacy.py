"""GPU parity: paged_attention_v1 / v2 (legacy x-packed cache layout) vs the oracle, plus
reshape_and_cache -> paged_attention round trip. Mirrors tests/kernels/attention/test_attention.py:112-330
of the reference (its parameter grid: head sizes 32..256, block sizes 16/32, GQA (40,40)/(64,8), alibi,
kv_cache_dtype auto/fp8; tolerance there atol = rtol = 1e-3 for 16-bit KV, 1e-2 for fp8)."""
import numpy as np
import pytest
import torch

from oracle import ref_ops

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
PARTITION = 512


def _ops():
    from b200ops import ops
    return ops


def _make(S, Hq, Hkv, D, bs, lens, dtype, fp8, seed=0):
    g = torch.Generator().manual_seed(seed)
    max_blocks = max(1, (max(lens) + bs - 1) // bs)
    nb = S * max_blocks + 3
    bt = torch.randperm(nb, generator=g)[: S * max_blocks].to(torch.int32).view(S, max_blocks)
    s = D ** -0.5
    q = ((torch.rand(S, Hq, D, generator=g) * 2 - 1) * s * 4).to(dtype)
    k = ((torch.rand(nb, bs, Hkv, D, generator=g) * 2 - 1)).to(dtype)
    v = ((torch.rand(nb, bs, Hkv, D, generator=g) * 2 - 1)).to(dtype)
    ks = vs = None
    if fp8:
        ks, vs = torch.tensor([1.0 / 128]), torch.tensor([1.0 / 200])
        k = torch.from_numpy(ref_ops.f32_to_e4m3_bits(k.float().numpy() / np.float32(ks.item())))
        v = torch.from_numpy(ref_ops.f32_to_e4m3_bits(v.float().numpy() / np.float32(vs.item())))
    kc, vc = ref_ops.flash_cache_to_legacy(k, v)
    return q, kc, vc, bt, torch.tensor(lens, dtype=torch.int32), ks, vs


def _tol(fp8):
    return (1e-2, 1e-2) if fp8 else (5e-3, 1e-2)


@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("heads", [(40, 40), (32, 8), (14, 2), (16, 1)])
@pytest.mark.parametrize("D", [32, 64, 80, 96, 112, 128, 192, 256])
@pytest.mark.parametrize("bs", [16, 32])
@pytest.mark.parametrize("kv", ["auto", "fp8"])
def test_paged_attention_legacy(version, heads, D, bs, kv):
    Hq, Hkv = heads
    if (D not in (64, 128)) and heads != (14, 2):
        pytest.skip("grid thinned: odd head sizes only on one head config")
    if bs == 32 and heads == (40, 40):
        pytest.skip("grid thinned")
    fp8 = kv == "fp8"
    if fp8 and D % 16 != 0:
        pytest.skip("fp8 cache needs head_size % 16 == 0")
    lens = [1, 17, 513, 1100, 333]
    S = len(lens)
    dtype = torch.bfloat16 if D != 96 else torch.float16
    q, kc, vc, bt, sl, ks, vs = _make(S, Hq, Hkv, D, bs, lens, dtype, fp8)
    scale = D ** -0.5
    ref = ref_ops.paged_attention_legacy(q, kc, vc, bt, sl, scale, k_scale=ks, v_scale=vs)
    ops = _ops()
    out = torch.full((S, Hq, D), float("nan"), dtype=dtype, device=DEV)
    one = torch.ones(1, device=DEV)
    args = (q.to(DEV), kc.to(DEV), vc.to(DEV), Hkv, scale, bt.to(DEV), sl.to(DEV), bs, max(lens), None, kv,
            ks.to(DEV) if fp8 else one, vs.to(DEV) if fp8 else one)
    if version == "v1":
        ops.paged_attention_v1(out, *args)
    else:
        P = (max(lens) + PARTITION - 1) // PARTITION
        exp_sums = torch.empty(S, Hq, P, dtype=torch.float32, device=DEV)
        max_logits = torch.empty_like(exp_sums)
        tmp_out = torch.empty(S, Hq, P, D, dtype=dtype, device=DEV)
        ops.paged_attention_v2(out, exp_sums, max_logits, tmp_out, *args)
    atol, rtol = _tol(fp8)
    np.testing.assert_allclose(out.float().cpu().numpy(), ref, atol=atol, rtol=rtol)


@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("bs", [8, 16, 32])
def test_paged_attention_legacy_alibi_and_partials(version, bs):
    """ALiBi bias (test_attention.py:96-101) and the v2 side outputs: max_logits must be the true
    per-partition max of scale*q.k + bias (natural-log domain), exp_sums the sum of exp(logit - max)."""
    S, Hq, Hkv, D = 3, 8, 2, 64
    lens = [700, 40, 1030]
    q, kc, vc, bt, sl, _, _ = _make(S, Hq, Hkv, D, bs, lens, torch.float16, False, seed=3)
    slopes = torch.tensor([2.0 ** -(i + 1) for i in range(Hq)], dtype=torch.float32)
    scale = D ** -0.5
    ref = ref_ops.paged_attention_legacy(q, kc, vc, bt, sl, scale, alibi_slopes=slopes)
    ops = _ops()
    one = torch.ones(1, device=DEV)
    out = torch.empty(S, Hq, D, dtype=torch.float16, device=DEV)
    args = (q.to(DEV), kc.to(DEV), vc.to(DEV), Hkv, scale, bt.to(DEV), sl.to(DEV), bs, max(lens), slopes.to(DEV),
            "auto", one, one)
    if version == "v1":
        ops.paged_attention_v1(out, *args)
    else:
        P = 3
        exp_sums = torch.zeros(S, Hq, P, dtype=torch.float32, device=DEV)
        max_logits = torch.zeros_like(exp_sums)
        tmp_out = torch.zeros(S, Hq, P, D, dtype=torch.float16, device=DEV)
        ops.paged_attention_v2(out, exp_sums, max_logits, tmp_out, *args)
        # recompute partition statistics on the host for sequence 2 (3 partitions: 512, 512, 6)
        k, _v = ref_ops.legacy_cache_to_flash(kc, vc)
        s = 2
        L = lens[s]
        idx = bt[s, : (L + bs - 1) // bs].long()
        kk = k[idx].reshape(-1, Hkv, D)[:L].float().numpy().astype(np.float64)
        kk = np.repeat(kk, Hq // Hkv, axis=1)
        att = np.einsum("hd,khd->hk", q[s].float().numpy().astype(np.float64) * scale, kk)
        att = att + slopes.numpy().astype(np.float64)[:, None] * (np.arange(L) - L + 1)[None, :]
        for part in range(3):
            seg = att[:, part * 512: min(L, (part + 1) * 512)]
            mx = seg.max(axis=1)
            np.testing.assert_allclose(max_logits[s, :, part].cpu().numpy(), mx, atol=2e-3, rtol=1e-3)
            np.testing.assert_allclose(exp_sums[s, :, part].cpu().numpy(), np.exp(seg - mx[:, None]).sum(1),
                                       rtol=2e-3)
    np.testing.assert_allclose(out.float().cpu().numpy(), ref, atol=5e-3, rtol=1e-2)


@pytest.mark.parametrize("kv", ["auto", "fp8"])
def test_reshape_and_cache_then_paged_attention_v1(kv):
    """write path + read path of the legacy layout together (PagedAttention.write_to_paged_cache ->
    forward_decode, vllm/v1/attention/ops/paged_attn.py:33-51)."""
    ops = _ops()
    S, Hq, Hkv, D, bs = 4, 32, 8, 128, 16
    lens = [100, 257, 16, 31]
    g = torch.Generator().manual_seed(7)
    max_blocks = (max(lens) + bs - 1) // bs
    nb = S * max_blocks + 1
    bt = torch.randperm(nb, generator=g)[: S * max_blocks].to(torch.int32).view(S, max_blocks)
    T = sum(lens)
    key = (torch.rand(T, Hkv, D, generator=g) * 2 - 1).to(torch.bfloat16)
    value = (torch.rand(T, Hkv, D, generator=g) * 2 - 1).to(torch.bfloat16)
    q = (torch.rand(S, Hq, D, generator=g) * 2 - 1).to(torch.bfloat16)
    slots = torch.cat([bt[s, torch.arange(L) // bs].long() * bs + torch.arange(L) % bs
                       for s, L in enumerate(lens)])
    fp8 = kv == "fp8"
    x = 16 if fp8 else 8
    cdt = torch.uint8 if fp8 else torch.bfloat16
    kc = torch.zeros(nb, Hkv, D // x, bs, x, dtype=cdt, device=DEV)
    vc = torch.zeros(nb, Hkv, D, bs, dtype=cdt, device=DEV)
    ks = torch.tensor([1.0 / 64], device=DEV)
    vs = torch.tensor([1.0 / 32], device=DEV)
    ops.reshape_and_cache(key.to(DEV), value.to(DEV), kc, vc, slots.to(DEV), kv, ks, vs)
    out = torch.empty(S, Hq, D, dtype=torch.bfloat16, device=DEV)
    sl = torch.tensor(lens, dtype=torch.int32)
    ops.paged_attention_v1(out, q.to(DEV), kc, vc, Hkv, D ** -0.5, bt.to(DEV), sl.to(DEV), bs, max(lens), None,
                           kv, ks, vs)
    ref = ref_ops.paged_attention_legacy(q, kc.cpu(), vc.cpu(), bt, sl, D ** -0.5,
                                         k_scale=ks.cpu() if fp8 else None, v_scale=vs.cpu() if fp8 else None)
    atol, rtol = _tol(fp8)
    np.testing.assert_allclose(out.float().cpu().numpy(), ref, atol=atol, rtol=rtol)


def test_paged_attention_legacy_validation():
    ops = _ops()
    q = torch.zeros(1, 8, 64, dtype=torch.bfloat16, device=DEV)
    kc = torch.zeros(2, 2, 8, 16, 8, dtype=torch.bfloat16, device=DEV)
    vc = torch.zeros(2, 2, 64, 16, dtype=torch.bfloat16, device=DEV)
    bt = torch.zeros(1, 1, dtype=torch.int32, device=DEV)
    sl = torch.ones(1, dtype=torch.int32, device=DEV)
    one = torch.ones(1, device=DEV)
    out = torch.empty_like(q)
    with pytest.raises(NotImplementedError, match="block-sparse"):
        ops.paged_attention_v1(out, q, kc, vc, 2, 0.125, bt, sl, 16, 1, None, "auto", one, one,
                               blocksparse_vert_stride=2)
    with pytest.raises(RuntimeError, match="Unsupported data type of kv cache"):
        ops.paged_attention_v1(out, q, kc, vc, 2, 0.125, bt, sl, 16, 1, None, "int8", one, one)
    with pytest.raises(RuntimeError, match="int32"):
        ops.paged_attention_v1(out, q, kc, vc, 2, 0.125, bt.long(), sl, 16, 1, None, "auto", one, one)
