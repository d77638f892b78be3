"""GPU parity: reshape_and_cache_flash / reshape_and_cache vs the oracle, through the C ABI.
Mirrors tests/kernels/attention/test_cache.py:167-463 of the reference (bit-exact for "auto",
byte-exact for fp8 -- stricter than the reference's dequantised atol=1e-3/rtol=0.1)."""
import pytest
import torch

from helpers import make_kv_cache
from oracle import ref_ops

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _ops():
    from b200ops import ops
    return ops


@pytest.mark.parametrize("layout", ["NHD", "HND"])
@pytest.mark.parametrize("kv_dtype", ["auto", "fp8"])
@pytest.mark.parametrize("head_size", [64, 80, 256])
@pytest.mark.parametrize("block_size", [8, 16, 32])
@pytest.mark.parametrize("per_head_scale", [False, True])
@pytest.mark.parametrize("in_dtype", [torch.bfloat16, torch.float16])
def test_reshape_and_cache_flash(layout, kv_dtype, head_size, block_size, per_head_scale, in_dtype):
    if kv_dtype == "auto" and per_head_scale:
        pytest.skip("scales unused for auto")
    torch.manual_seed(0)
    T, H, nb = 42, 8, 64
    qkv = torch.randn(T, 3, H, head_size).to(in_dtype)  # key/value are strided views (fused QKV)
    key, value = qkv[:, 1], qkv[:, 2]
    slots = torch.randperm(nb * block_size)[:T].to(torch.int64)
    slots[3] = -1
    slots[17] = -1
    cdt = torch.uint8 if kv_dtype == "fp8" else in_dtype
    raw_c, kc_c, vc_c = make_kv_cache(nb, block_size, H, head_size, cdt, layout)
    raw_g, kc_g, vc_g = make_kv_cache(nb, block_size, H, head_size, cdt, layout)
    raw_g = raw_g.to(DEV)
    kv = raw_g if layout == "NHD" else raw_g.transpose(1, 2)
    kc_g, vc_g = kv[..., :head_size], kv[..., head_size:]
    n = H if per_head_scale else 1
    k_scale = (key.float().abs().amax() / 64.0).repeat(n) * (torch.arange(n) * 0.1 + 1)
    v_scale = (value.float().abs().amax() / 64.0).repeat(n) * (torch.arange(n) * 0.2 + 1)
    ref_ops.reshape_and_cache_flash(key, value, kc_c, vc_c, slots, kv_dtype, k_scale, v_scale)
    _ops().reshape_and_cache_flash(key.to(DEV), value.to(DEV), kc_g, vc_g, slots.to(DEV), kv_dtype,
                                   k_scale.to(DEV), v_scale.to(DEV))
    torch.cuda.synchronize()
    assert torch.equal(raw_g.cpu().view(torch.uint8), raw_c.view(torch.uint8))


def test_reshape_and_cache_flash_unaligned_and_short_slot_mapping():
    """num_heads=13, head_size=46 (test_cache.py:429-463) and slot_mapping shorter than key."""
    torch.manual_seed(1)
    T, H, D, bs, nb = 19, 13, 46, 16, 8
    key = torch.randn(T + 5, H, D).to(torch.bfloat16)
    value = torch.randn(T + 5, H, D).to(torch.bfloat16)
    slots = torch.randperm(nb * bs)[:T].to(torch.int64)
    for kv_dtype, cdt in (("auto", torch.bfloat16), ("fp8", torch.uint8)):
        raw_c, kc_c, vc_c = make_kv_cache(nb, bs, H, D, cdt, "NHD")
        raw_g = raw_c.clone().to(DEV)
        sc = torch.tensor([0.07])
        ref_ops.reshape_and_cache_flash(key, value, kc_c, vc_c, slots, kv_dtype, sc, sc)
        _ops().reshape_and_cache_flash(key.to(DEV), value.to(DEV), raw_g[..., :D], raw_g[..., D:],
                                       slots.to(DEV), kv_dtype, sc.to(DEV), sc.to(DEV))
        assert torch.equal(raw_g.cpu().view(torch.uint8), raw_c.view(torch.uint8))


def test_reshape_and_cache_flash_split_layout_f32_and_empty():
    torch.manual_seed(2)
    T, H, D, bs, nb = 11, 4, 128, 16, 4
    key, value = torch.randn(T, H, D), torch.randn(T, H, D)
    slots = torch.randperm(nb * bs)[:T].to(torch.int64)
    raw_c, kc_c, vc_c = make_kv_cache(nb, bs, H, D, torch.float32, packed=False)
    raw_g = raw_c.clone().to(DEV)
    one = torch.ones(1)
    ref_ops.reshape_and_cache_flash(key, value, kc_c, vc_c, slots, "auto", one, one)
    _ops().reshape_and_cache_flash(key.to(DEV), value.to(DEV), raw_g[0], raw_g[1], slots.to(DEV),
                                   "auto", one.to(DEV), one.to(DEV))
    assert torch.equal(raw_g.cpu(), raw_c)
    # empty slot_mapping: no launch, no error
    _ops().reshape_and_cache_flash(key.to(DEV), value.to(DEV), raw_g[0], raw_g[1],
                                   torch.zeros(0, dtype=torch.int64, device=DEV), "auto",
                                   one.to(DEV), one.to(DEV))
    with pytest.raises(RuntimeError, match="Unsupported data type of kv cache"):
        _ops().reshape_and_cache_flash(key.to(DEV), value.to(DEV), raw_g[0], raw_g[1],
                                       slots.to(DEV), "int3", one.to(DEV), one.to(DEV))


@pytest.mark.parametrize("kv_dtype", ["auto", "fp8"])
@pytest.mark.parametrize("head_size", [64, 128])
def test_reshape_and_cache_legacy(kv_dtype, head_size):
    torch.manual_seed(3)
    T, H, bs, nb = 23, 6, 16, 8
    cdt = torch.uint8 if kv_dtype == "fp8" else torch.bfloat16
    x = 16 // (1 if kv_dtype == "fp8" else 2)
    key = torch.randn(T, H, head_size).to(torch.bfloat16)
    value = torch.randn(T, H, head_size).to(torch.bfloat16)
    slots = torch.randperm(nb * bs)[:T].to(torch.int64)
    slots[5] = -1
    kc = torch.zeros(nb, H, head_size // x, bs, x, dtype=cdt)
    vc = torch.zeros(nb, H, head_size, bs, dtype=cdt)
    kg, vg = kc.clone().to(DEV), vc.clone().to(DEV)
    ks, vs = torch.tensor([0.03]), torch.tensor([0.05])
    ref_ops.reshape_and_cache(key, value, kc, vc, slots, kv_dtype, ks, vs)
    _ops().reshape_and_cache(key.to(DEV), value.to(DEV), kg, vg, slots.to(DEV), kv_dtype,
                             ks.to(DEV), vs.to(DEV))
    assert torch.equal(kg.cpu().view(torch.uint8), kc.view(torch.uint8))
    assert torch.equal(vg.cpu().view(torch.uint8), vc.view(torch.uint8))
