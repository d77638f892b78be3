"""GPU parity: RMSNorm family (+static FP8 quant) and rotary_embedding vs the oracle
(SURVEY.md section 8f rank-1 rows). Mirrors tests/kernels/core/test_layernorm.py and
test_pos_encoding.py of the reference."""
import numpy as np
import pytest
import torch

from oracle import ref_ops

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _ops():
    from b200ops import ops
    return ops


def close_in_ulps(got, ref, dtype, frac=0.998):
    """block-reduction order may move rsqrt by 1 ulp(fp32); with the reference's two roundings
    (T(x*s), then *w) that is at most 2 ulp of `dtype`, on a handful of elements."""
    ulp = 2 ** -7 if dtype == torch.bfloat16 else 2 ** -10
    np.testing.assert_allclose(got.float().cpu().numpy(), ref.float().numpy(), rtol=2.1 * ulp, atol=1e-6)
    assert (got.cpu() == ref).float().mean() >= frac


@pytest.mark.parametrize("tokens,hidden", [(1, 896), (7, 4096), (256, 4096), (83, 8192), (5, 16384)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("add_residual", [False, True])
def test_rms_norm(tokens, hidden, dtype, add_residual):
    g = torch.Generator().manual_seed(0)
    x = torch.randn(tokens, hidden, generator=g).to(dtype)
    w = (torch.randn(hidden, generator=g) * 0.1 + 1).to(dtype)
    res = torch.randn(tokens, hidden, generator=g).to(dtype) if add_residual else None
    ref, ref_res = ref_ops.rms_norm(x, w, 1e-5, res)
    if add_residual:
        xg, rg = x.to(DEV), res.to(DEV)
        _ops().fused_add_rms_norm(xg, rg, w.to(DEV), 1e-5)
        assert torch.equal(rg.cpu(), ref_res)  # the residual stream is bit-exact
        close_in_ulps(xg, ref, dtype)
    else:
        out = torch.empty_like(x, device=DEV)
        _ops().rms_norm(out, x.to(DEV), w.to(DEV), 1e-5)
        close_in_ulps(out, ref, dtype)


@pytest.mark.parametrize("tokens,hidden", [(3, 1024), (256, 4096), (17, 8192)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("add_residual", [False, True])
def test_rms_norm_static_fp8_quant(tokens, hidden, dtype, add_residual):
    g = torch.Generator().manual_seed(1)
    x = torch.randn(tokens, hidden, generator=g).to(dtype)
    w = (torch.randn(hidden, generator=g) * 0.1 + 1).to(dtype)
    res = torch.randn(tokens, hidden, generator=g).to(dtype) if add_residual else None
    scale = torch.tensor([0.01])
    ref, ref_res = ref_ops.rms_norm_static_fp8_quant(x, w, scale, 1e-6, res)
    out = torch.empty(tokens, hidden, dtype=torch.float8_e4m3fn, device=DEV)
    if add_residual:
        rg = res.to(DEV)
        _ops().fused_add_rms_norm_static_fp8_quant(out, x.to(DEV), rg, w.to(DEV), scale.to(DEV), 1e-6)
        assert torch.equal(rg.cpu(), ref_res)
    else:
        _ops().rms_norm_static_fp8_quant(out, x.to(DEV), w.to(DEV), scale.to(DEV), 1e-6)
    got = out.view(torch.uint8).cpu().numpy()
    assert (got == ref).mean() > 0.995
    a, b = ref_ops.e4m3_bits_to_f32(got), ref_ops.e4m3_bits_to_f32(ref)
    assert np.abs(a - b).max() <= 32.0 and np.all(np.abs(a - b) <= np.abs(b) * 0.13 + 2 ** -9)


@pytest.mark.parametrize("is_neox", [True, False])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("heads,kv_heads,D,rot", [(32, 8, 128, 128), (14, 2, 64, 64), (8, 8, 128, 64)])
def test_rotary_embedding(is_neox, dtype, heads, kv_heads, D, rot):
    g = torch.Generator().manual_seed(2)
    T, max_pos = 37, 8192
    inv = 1.0 / (10000 ** (torch.arange(0, rot, 2).float() / rot))
    fr = torch.outer(torch.arange(max_pos).float(), inv)
    cache = torch.cat([fr.cos(), fr.sin()], dim=-1).to(dtype)
    pos = torch.randint(0, max_pos, (T,), generator=g)
    qkv = torch.randn(T, (heads + 2 * kv_heads) * D, generator=g).to(dtype)
    q = qkv[:, : heads * D].view(T, heads, D)
    k = qkv[:, heads * D:(heads + kv_heads) * D].view(T, kv_heads, D)
    rq, rk = ref_ops.rotary_embedding(pos, q, k, D, cache, is_neox)
    qkv_g = qkv.to(DEV)
    qg = qkv_g[:, : heads * D].view(T, heads, D)
    kg = qkv_g[:, heads * D:(heads + kv_heads) * D].view(T, kv_heads, D)
    _ops().rotary_embedding(pos.to(DEV), qg, kg, D, cache.to(DEV), is_neox)
    # x*c - y*s in fp32: fma contraction may differ by 1 ulp of the output dtype on rare elements
    close_in_ulps(qg, rq, dtype, frac=0.99)
    close_in_ulps(kg, rk, dtype, frac=0.99)
    # v part of the fused tensor untouched
    assert torch.equal(qkv_g[:, (heads + kv_heads) * D:].cpu(), qkv[:, (heads + kv_heads) * D:])


@pytest.mark.parametrize("kv", ["auto", "fp8"])
@pytest.mark.parametrize("neox", [True, False])
@pytest.mark.parametrize("heads,D,rot", [((32, 8), 128, 128), ((14, 2), 64, 64), ((8, 1), 128, 64), ((4, 4), 80, 40)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_rope_and_cache_flash_equals_unfused_pair(kv, neox, heads, D, rot, dtype):
    """fused rotary_embedding + reshape_and_cache_flash (backend hook do_rope_and_kv_cache_update) must be
    bit-identical to the two oracle-checked launches: rotated q / k and every cache byte."""
    Hq, Hkv = heads
    T, bs, nb = 57, 16, 9
    g = torch.Generator().manual_seed(21)
    qkv = torch.randn(T, (Hq + 2 * Hkv) * D, generator=g).to(dtype).to(DEV)
    pos = torch.randint(0, 500, (T,), generator=g).to(DEV)
    cache = torch.randn(500, rot, generator=g).to(dtype).to(DEV)
    slots = torch.randperm(nb * bs, generator=g)[:T].long()
    slots[3] = -1
    slots = slots.to(DEV)
    per_head = Hkv > 1
    ks = (torch.rand(Hkv if per_head else 1, generator=g) * 0.05 + 0.01).to(DEV)
    vs = (torch.rand(Hkv if per_head else 1, generator=g) * 0.05 + 0.01).to(DEV)
    cdt = torch.uint8 if kv == "fp8" else dtype
    ops = _ops()

    def views(t):
        return (t[:, :Hq * D].view(T, Hq, D), t[:, Hq * D:(Hq + Hkv) * D].view(T, Hkv, D),
                t[:, (Hq + Hkv) * D:].view(T, Hkv, D))

    a, b = qkv.clone(), qkv.clone()
    c1 = torch.zeros(nb, bs, Hkv, 2 * D, dtype=cdt, device=DEV)
    c2 = torch.zeros_like(c1)
    q1, k1, v1 = views(a)
    ops.rotary_embedding(pos, q1, k1, D, cache, neox)
    ops.reshape_and_cache_flash(k1, v1, c1[..., :D], c1[..., D:], slots, kv, ks, vs)
    q2, k2, v2 = views(b)
    ops.rope_and_reshape_and_cache_flash(pos, q2, k2, v2, D, cache, neox, c2[..., :D], c2[..., D:], slots, kv, ks, vs)
    torch.cuda.synchronize()
    assert torch.equal(a.view(torch.int16), b.view(torch.int16))
    assert torch.equal(c1.view(torch.uint8), c2.view(torch.uint8))
