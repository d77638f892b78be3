"""GPU parity against the reference's OWN CUDA kernels, on identical inputs in one process.

The image ships the vllm 0.22 wheel (a build of upstream vLLM with sm_100 cubins); its
torch.ops._C / _C_cache_ops / _moe_C operators have exactly the schemas our ops mirror. Here every
shared operator is run twice - the wheel's kernel and ours through the C ABI - and compared:
bit-exact for byte / integer / index work (KV-cache bytes, FP8 codes and scales, NVFP4 codes and block
scales, MoE routing tables), within output rounding for the GEMMs / attention. The wheel is LIBRARY code
used only as a checker here; nothing under b200ops imports it. Skipped where the wheel cannot be loaded.
"""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
_state = {}


def wheel():
    """torch.ops with the wheel's libraries loaded, or skip."""
    if "ok" not in _state:
        try:
            import importlib.util
            spec = importlib.util.find_spec("vllm")
            d = os.path.dirname(spec.origin)
            for lib in ("_C.abi3.so", "_C_stable_libtorch.abi3.so", "_moe_C.abi3.so"):
                torch.ops.load_library(os.path.join(d, lib))
            torch.ops._C.cutlass_scaled_mm  # noqa: B018  (raises if the namespace is empty)
            _state["ok"] = True
        except Exception as e:  # noqa: BLE001
            _state["ok"] = False
            _state["why"] = repr(e)[:200]
    if not _state["ok"]:
        pytest.skip(f"vllm wheel ops unavailable: {_state.get('why')}")
    return torch.ops


def ours():
    from b200ops import ops
    return ops


def moe():
    from b200ops import fused_moe
    return fused_moe


def gen(seed):
    return torch.Generator(device=DEV).manual_seed(seed)


def bits(t):
    return t.contiguous().view(torch.uint8)


# ------------------------------------------------------------------------------------------------ KV cache
@pytest.mark.parametrize("kv", ["auto", "fp8"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_reshape_and_cache_flash_bytes(kv, dtype):
    W = wheel()
    T, H, D, bs, nb = 77, 8, 128, 16, 40
    g = gen(1)
    qkv = torch.randn(T, (32 + 2 * H) * D, device=DEV, generator=g).to(dtype)
    key = qkv[:, 32 * D:(32 + H) * D].view(T, H, D)          # strided views of a fused QKV output
    value = qkv[:, (32 + H) * D:].view(T, H, D)
    slots = torch.randperm(nb * bs, device=DEV, generator=g)[:T].long()
    slots[5] = -1
    cdt = torch.uint8 if kv == "fp8" else dtype
    ks, vs = torch.tensor([0.03], device=DEV), torch.tensor([0.05], device=DEV)
    c1 = torch.zeros(nb, bs, H, 2 * D, dtype=cdt, device=DEV)
    c2 = torch.zeros_like(c1)
    W._C_cache_ops.reshape_and_cache_flash(key, value, c1[..., :D], c1[..., D:], slots, kv, ks, vs)
    ours().reshape_and_cache_flash(key, value, c2[..., :D], c2[..., D:], slots, kv, ks, vs)
    assert torch.equal(bits(c1), bits(c2))


@pytest.mark.parametrize("kv", ["auto", "fp8"])
def test_reshape_and_cache_legacy_bytes(kv):
    W = wheel()
    T, H, D, bs, nb = 50, 4, 64, 16, 9
    g = gen(2)
    key = torch.randn(T, H, D, device=DEV, generator=g).to(torch.bfloat16)
    value = torch.randn(T, H, D, device=DEV, generator=g).to(torch.bfloat16)
    slots = torch.randperm(nb * bs, device=DEV, generator=g)[:T].long()
    x = 16 if kv == "fp8" else 8
    cdt = torch.uint8 if kv == "fp8" else torch.bfloat16
    ks, vs = torch.tensor([0.04], device=DEV), torch.tensor([0.02], device=DEV)
    k1 = torch.zeros(nb, H, D // x, bs, x, dtype=cdt, device=DEV)
    v1 = torch.zeros(nb, H, D, bs, dtype=cdt, device=DEV)
    k2, v2 = torch.zeros_like(k1), torch.zeros_like(v1)
    W._C_cache_ops.reshape_and_cache(key, value, k1, v1, slots, kv, ks, vs)
    ours().reshape_and_cache(key, value, k2, v2, slots, kv, ks, vs)
    assert torch.equal(bits(k1), bits(k2)) and torch.equal(bits(v1), bits(v2))


# ------------------------------------------------------------------------------------------------ FP8 quantisers
@pytest.mark.parametrize("shape", [(1, 4096), (256, 4096), (83, 14336), (7, 1000)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
def test_fp8_quantisers_bytes_and_scales(shape, dtype):
    W = wheel()
    x = (torch.randn(*shape, device=DEV, generator=gen(3)) * 3).to(dtype)
    f8 = torch.float8_e4m3fn
    # static
    sc = torch.tensor([0.021], device=DEV)
    a, b = torch.empty(shape, dtype=f8, device=DEV), torch.empty(shape, dtype=f8, device=DEV)
    W._C.static_scaled_fp8_quant(a, x, sc)
    ours().static_scaled_fp8_quant(b, x, sc)
    assert torch.equal(bits(a), bits(b))
    # dynamic per tensor (the wheel accumulates its scale with atomicMax: zero it first)
    s1, s2 = torch.zeros(1, device=DEV), torch.zeros(1, device=DEV)
    W._C.dynamic_scaled_fp8_quant(a, x, s1)
    ours().dynamic_scaled_fp8_quant(b, x, s2)
    assert torch.equal(s1, s2) and torch.equal(bits(a), bits(b))
    # dynamic per token (+ upper bound)
    for ub in (None, torch.tensor([2.0], device=DEV)):
        t1, t2 = torch.empty(shape[0], 1, device=DEV), torch.empty(shape[0], 1, device=DEV)
        W._C.dynamic_per_token_scaled_fp8_quant(a, x, t1, ub)
        ours().dynamic_per_token_scaled_fp8_quant(b, x, t2, ub)
        assert torch.equal(t1, t2) and torch.equal(bits(a), bits(b))


def _wheel_group_quant(W, x, q, s, g, ue, col, tma):
    fi = torch.finfo(torch.float8_e4m3fn)
    try:
        W._C.per_token_group_fp8_quant(x, q, s, g, 1e-10, fi.min, fi.max, ue, col, tma)
    except (RuntimeError, TypeError) as e:
        if "argument" not in str(e) and "schema" not in str(e):
            raise
        try:
            W._C.per_token_group_fp8_quant(x, q, s, g, 1e-10, fi.min, fi.max, ue)
        except (RuntimeError, TypeError, AttributeError) as e2:
            pytest.skip(f"wheel per_token_group_fp8_quant has another schema: {e2!r}"[:200])
    except AttributeError as e:
        pytest.skip(f"wheel has no per_token_group_fp8_quant: {e!r}"[:200])


@pytest.mark.parametrize("shape", [(7, 512), (256, 4096), (83, 14336), (2050, 7168)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("group", [64, 128])
@pytest.mark.parametrize("ue8m0", [False, True])
@pytest.mark.parametrize("col", [False, True])
def test_per_token_group_fp8_quant_bytes_and_scales(shape, dtype, group, ue8m0, col):
    W = wheel()
    x = (torch.randn(*shape, device=DEV, generator=gen(5)) * 3).to(dtype)
    f8 = torch.float8_e4m3fn
    a, b = torch.empty(shape, dtype=f8, device=DEV), torch.empty(shape, dtype=f8, device=DEV)
    G = shape[1] // group
    if col:
        s1 = torch.empty(G, shape[0], device=DEV).t()
        s2 = torch.empty(G, shape[0], device=DEV).t()
    else:
        s1, s2 = torch.empty(shape[0], G, device=DEV), torch.empty(shape[0], G, device=DEV)
    _wheel_group_quant(W, x, a, s1, group, ue8m0, col, False)
    fi = torch.finfo(f8)
    ours().per_token_group_fp8_quant(x, b, s2, group, 1e-10, fi.min, fi.max, ue8m0, col, False)
    assert torch.equal(s1, s2) and torch.equal(bits(a), bits(b))


# ------------------------------------------------------------------------------------------------ activation / norm / rope
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_silu_and_mul_and_quant(dtype):
    W = wheel()
    x = (torch.randn(256, 2 * 14336, device=DEV, generator=gen(4)) * 2).to(dtype)
    a, b = torch.empty(256, 14336, dtype=dtype, device=DEV), torch.empty(256, 14336, dtype=dtype, device=DEV)
    W._C.silu_and_mul(a, x)
    ours().silu_and_mul(b, x)
    # both evaluate expf in fp32; the division differs by <= 1 ulp(fp32) -> a handful of 16-bit roundings flip
    assert (a == b).float().mean().item() > 0.999
    torch.testing.assert_close(a.float(), b.float(), rtol=2 ** -7, atol=1e-6)
    sc = torch.tensor([0.02], device=DEV)
    qa = torch.empty(256, 14336, dtype=torch.float8_e4m3fn, device=DEV)
    qb = torch.empty_like(qa)
    W._C.silu_and_mul_quant(qa, x, sc)
    ours().silu_and_mul_quant(qb, x, sc)
    assert (bits(qa) == bits(qb)).float().mean().item() > 0.999


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("hidden", [4096, 896, 8192])
def test_rms_norm_family(dtype, hidden):
    W = wheel()
    g = gen(5)
    x = torch.randn(67, hidden, device=DEV, generator=g).to(dtype)
    res = torch.randn(67, hidden, device=DEV, generator=g).to(dtype)
    w = (torch.rand(hidden, device=DEV, generator=g) + 0.5).to(dtype)
    sc = torch.tensor([0.01], device=DEV)
    a, b = torch.empty_like(x), torch.empty_like(x)
    W._C.rms_norm(a, x, w, 1e-5)
    ours().rms_norm(b, x, w, 1e-5)
    # the two kernels reduce the sum of squares in different orders: <= 1 ulp of the 16-bit output
    torch.testing.assert_close(a.float(), b.float(), rtol=2 ** -7, atol=1e-5)
    assert (a == b).float().mean().item() > 0.99
    x1, r1, x2, r2 = x.clone(), res.clone(), x.clone(), res.clone()
    W._C.fused_add_rms_norm(x1, r1, w, 1e-5)
    ours().fused_add_rms_norm(x2, r2, w, 1e-5)
    assert torch.equal(r1, r2)                                  # the residual add is exact arithmetic
    torch.testing.assert_close(x1.float(), x2.float(), rtol=2 ** -7, atol=1e-5)
    q1 = torch.empty(67, hidden, dtype=torch.float8_e4m3fn, device=DEV)
    q2 = torch.empty_like(q1)
    r1, r2 = res.clone(), res.clone()
    W._C.fused_add_rms_norm_static_fp8_quant(q1, x, r1, w, sc, 1e-5)
    ours().fused_add_rms_norm_static_fp8_quant(q2, x, r2, w, sc, 1e-5)
    assert torch.equal(r1, r2)
    assert (bits(q1) == bits(q2)).float().mean().item() > 0.99
    d = (q1.float() - q2.float()).abs()
    assert (d <= 32.0).all()                                    # never more than one e4m3 step
    W._C.rms_norm_static_fp8_quant(q1, x, w, sc, 1e-5)
    ours().rms_norm_static_fp8_quant(q2, x, w, sc, 1e-5)
    assert (bits(q1) == bits(q2)).float().mean().item() > 0.99


@pytest.mark.parametrize("neox", [True, False])
def test_rotary_embedding(neox):
    W = wheel()
    g = gen(6)
    T, Hq, Hkv, D = 99, 32, 8, 128
    qkv = torch.randn(T, (Hq + 2 * Hkv) * D, device=DEV, generator=g).to(torch.bfloat16)
    pos = torch.randint(0, 4096, (T,), device=DEV, generator=g)
    cache = torch.randn(4096, D, device=DEV, generator=g).to(torch.bfloat16)
    a, b = qkv.clone(), qkv.clone()
    W._C.rotary_embedding(pos, a[:, :Hq * D], a[:, Hq * D:(Hq + Hkv) * D], D, cache, neox)
    ours().rotary_embedding(pos, b[:, :Hq * D], b[:, Hq * D:(Hq + Hkv) * D], D, cache, neox)
    assert torch.equal(a, b)


# ------------------------------------------------------------------------------------------------ GEMMs
def _fp8(shape, seed):
    return torch.randn(*shape, device=DEV, generator=gen(seed)).clamp(-448, 448).to(torch.float8_e4m3fn)


@pytest.mark.parametrize("impl", ["v2", "v3"])
@pytest.mark.parametrize("mnk", [(256, 6144, 4096), (256, 4096, 14336), (33, 1024, 512), (1, 4096, 4096), (1000, 1280, 2048)])
@pytest.mark.parametrize("per_token,per_channel,use_bias", [(False, False, False), (True, True, True)])
def test_cutlass_scaled_mm(mnk, per_token, per_channel, use_bias, impl):
    W = wheel()
    M, N, K = mnk
    a, w = _fp8((M, K), 7), _fp8((N, K), 8)
    g = gen(9)
    sa = (torch.rand(M if per_token else 1, 1, device=DEV, generator=g) + 0.5) * 0.05
    sb = (torch.rand(N if per_channel else 1, 1, device=DEV, generator=g) + 0.5) * 0.05
    bias = (torch.rand(N, device=DEV, generator=g) - 0.5).to(torch.bfloat16) if use_bias else None
    ref = torch.empty(M, N, dtype=torch.bfloat16, device=DEV)
    W._C.cutlass_scaled_mm(ref, a, w.t(), sa, sb, bias)
    out = ours().cutlass_scaled_mm(a, w.t(), sa, sb, torch.bfloat16, bias, impl=impl)
    # same fp32-accumulate contract, different accumulation order + the epilogue's multiply order:
    # one bf16 ulp, plus cancellation noise relative to the row's magnitude
    torch.testing.assert_close(out.float(), ref.float(), rtol=2 ** -7, atol=2e-3 * ref.float().abs().max().item())
    assert (out == ref).float().mean().item() > 0.95


@pytest.mark.parametrize("shape", [(128, 64), (150, 160), (256, 14336), (1, 4096), (33, 2048)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("swizzled", [True, False])
def test_scaled_fp4_quant_bits(shape, dtype, swizzled):
    """same instruction sequence (rcp.approx.ftz, cvt.rn.satfinite.e2m1x2) -> codes AND scales bit-exact."""
    W = wheel()
    x = torch.randn(*shape, device=DEV, generator=gen(10)).to(dtype)
    gs = ((448.0 * 6.0) / x.float().abs().max()).reshape(1)
    c1, s1 = W._C.scaled_fp4_quant(x, gs, swizzled)
    c2, s2 = ours().scaled_fp4_quant(x, gs, swizzled)
    assert torch.equal(c1, c2)
    assert torch.equal(bits(s1).reshape(-1), bits(s2).reshape(-1))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_silu_and_mul_nvfp4_quant_bits(dtype):
    W = wheel()
    m, d = 150, 1792
    x = torch.randn(m, 2 * d, device=DEV, generator=gen(11)).to(dtype)
    gs = torch.tensor([500.0], device=DEV)
    o1 = torch.empty(m, d // 2, dtype=torch.uint8, device=DEV)
    s1 = ours().create_fp4_scale_tensor(m, d, DEV, True)
    o2, s2 = torch.empty_like(o1), torch.empty_like(s1)
    s1.zero_(); s2.zero_()
    W._C.silu_and_mul_nvfp4_quant(o1, s1, x, gs)
    ours().silu_and_mul_nvfp4_quant(o2, s2, x, gs)
    assert torch.equal(o1, o2) and torch.equal(bits(s1), bits(s2))


@pytest.mark.parametrize("mnk", [(128, 256, 256), (150, 128, 128), (33, 4096, 2048), (256, 6144, 4096), (512, 512, 1024)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_cutlass_scaled_fp4_mm(mnk, dtype):
    W = wheel()
    m, n, k = mnk
    a = torch.randn(m, k, device=DEV, generator=gen(12)).to(dtype)
    b = torch.randn(n, k, device=DEV, generator=gen(13)).to(dtype)
    gsa = ((448.0 * 6.0) / a.float().abs().max()).reshape(1)
    gsb = ((448.0 * 6.0) / b.float().abs().max()).reshape(1)
    a4, sfa = ours().scaled_fp4_quant(a, gsa)
    b4, sfb = ours().scaled_fp4_quant(b, gsb)
    alpha = 1.0 / (gsa * gsb)
    ref = torch.empty(m, n, dtype=dtype, device=DEV)
    W._C.cutlass_scaled_fp4_mm(ref, a4, b4, sfa, sfb, alpha)
    out = ours().cutlass_scaled_fp4_mm(a4, b4, sfa, sfb, alpha, dtype)
    torch.testing.assert_close(out.float(), ref.float(), rtol=2 ** -7, atol=2e-3 * ref.float().abs().max().item())
    assert (out == ref).float().mean().item() > 0.95


# ------------------------------------------------------------------------------------------------ MoE routing
@pytest.mark.parametrize("M,k,E,bs", [(256, 2, 8, 16), (1, 2, 8, 64), (1000, 4, 64, 128), (33, 8, 256, 32), (4096, 2, 8, 128)])
def test_moe_align_block_size_tables(M, k, E, bs):
    W = wheel()
    g = gen(14)
    ids = torch.stack([torch.randperm(E, device=DEV, generator=g)[:k] for _ in range(M)]).to(torch.int32)
    numel = ids.numel()
    maxp = numel + E * (bs - 1)
    if numel < E:
        maxp = min(numel * bs, maxp)
    s1 = torch.empty(maxp, dtype=torch.int32, device=DEV)
    e1 = torch.empty((maxp + bs - 1) // bs, dtype=torch.int32, device=DEV)
    n1 = torch.empty(1, dtype=torch.int32, device=DEV)
    W._moe_C.moe_align_block_size(ids, E, bs, s1, e1, n1, None)
    s2, e2, n2 = moe().moe_align_block_size(ids, bs, E)
    npad = int(n1.item())
    assert int(n2.item()) == npad
    nblk = npad // bs
    assert torch.equal(e1[:nblk], e2[:nblk])
    # within an expert's block range the wheel's order depends on atomics; compare as sets per block row
    a = s1[:npad].view(nblk, bs).sort(dim=1).values
    b = s2[:npad].view(nblk, bs).sort(dim=1).values
    # tokens of one expert may be distributed over its blocks in any order: compare per expert
    for e in range(E):
        rows = (e1[:nblk] == e)
        assert torch.equal(a[rows].reshape(-1).sort().values, b[rows].reshape(-1).sort().values)


@pytest.mark.parametrize("M,E,k,renorm", [(256, 8, 2, True), (77, 64, 6, False), (1, 128, 8, True), (1000, 256, 8, False)])
def test_topk_softmax(M, E, k, renorm):
    W = wheel()
    gate = torch.randn(M, E, device=DEV, generator=gen(15))
    w1 = torch.empty(M, k, device=DEV); i1 = torch.empty(M, k, dtype=torch.int32, device=DEV)
    t1 = torch.empty(M, k, dtype=torch.int32, device=DEV)
    w2, i2, t2 = torch.empty_like(w1), torch.empty_like(i1), torch.empty_like(t1)
    W._moe_C.topk_softmax(w1, i1, t1, gate, renorm, None)
    moe().topk_softmax(w2, i2, t2, gate, renorm)
    assert torch.equal(i1, i2) and torch.equal(t1, t2)
    torch.testing.assert_close(w1, w2, rtol=1e-5, atol=1e-7)


def test_moe_sum():
    W = wheel()
    x = torch.randn(300, 4, 4096, device=DEV, generator=gen(16)).to(torch.bfloat16)
    a, b = torch.empty(300, 4096, dtype=torch.bfloat16, device=DEV), torch.empty(300, 4096, dtype=torch.bfloat16, device=DEV)
    W._moe_C.moe_sum(x, a)
    moe().moe_sum(x, b)
    # the reference TREE accumulates the top-k rows in fp32 and rounds once (moe_align_sum_kernels.cu:395-413);
    # ours must equal that exactly. The wheel is an older upstream build that accumulates in the 16-bit
    # dtype (torch.sum-like), so it only agrees to a couple of bf16 ulps.
    assert torch.equal(b, x.float().sum(dim=1).to(torch.bfloat16))
    torch.testing.assert_close(a.float(), b.float(), rtol=2 ** -6, atol=0.13)


# ------------------------------------------------------------------------------------------------ attention
def test_merge_attn_states():
    W = wheel()
    g = gen(17)
    T, H, D = 300, 16, 128
    po = torch.randn(T, H, D, device=DEV, generator=g).to(torch.bfloat16)
    so = torch.randn(T, H, D, device=DEV, generator=g).to(torch.bfloat16)
    pl = torch.randn(H, T, device=DEV, generator=g) * 3
    sl = torch.randn(H, T, device=DEV, generator=g) * 3
    pl[0, :10] = float("inf"); sl[1, 5:9] = float("-inf"); pl[1, 5:7] = float("-inf")
    o1, o2 = torch.empty_like(po), torch.empty_like(po)
    l1, l2 = torch.empty(H, T, device=DEV), torch.empty(H, T, device=DEV)
    W._C.merge_attn_states(o1, l1, po, pl, so, sl, 200, None)
    ours().merge_attn_states(o2, po, pl, so, sl, l2, 200)
    assert torch.equal(o1, o2) and torch.equal(l1, l2)          # same expf / division sequence
    q1 = torch.empty(T, H, D, dtype=torch.float8_e4m3fn, device=DEV)
    q2 = torch.empty_like(q1)
    sc = torch.tensor([0.03], device=DEV)
    W._C.merge_attn_states(q1, None, po, pl, so, sl, None, sc)
    ours().merge_attn_states(q2, po, pl, so, sl, None, None, sc)
    assert torch.equal(bits(q1), bits(q2))


@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("kv", ["auto", "fp8"])
@pytest.mark.parametrize("heads,D,bs", [((32, 8), 128, 16), ((14, 2), 64, 16), ((8, 8), 256, 32), ((16, 1), 96, 16)])
def test_paged_attention_legacy(version, kv, heads, D, bs):
    W = wheel()
    Hq, Hkv = heads
    if kv == "fp8" and D % 16:
        pytest.skip("fp8 cache needs head_size % 16 == 0")
    g = gen(18)
    lens = [1, 17, 513, 1100, 2049, 333]
    S = len(lens)
    mb = (max(lens) + bs - 1) // bs
    nb = S * mb + 3
    bt = torch.randperm(nb, device=DEV, generator=g)[: S * mb].to(torch.int32).view(S, mb)
    sl = torch.tensor(lens, dtype=torch.int32, device=DEV)
    dt = torch.float16 if D == 96 else torch.bfloat16
    q = (torch.randn(S, Hq, D, device=DEV, generator=g) * 0.5).to(dt)
    fp8 = kv == "fp8"
    x = 16 if fp8 else 8
    if fp8:
        kc = torch.randint(0, 0x78, (nb, Hkv, D // x, bs, x), dtype=torch.uint8, device=DEV, generator=g)
        vc = torch.randint(0, 0x78, (nb, Hkv, D, bs), dtype=torch.uint8, device=DEV, generator=g)
        kc |= (torch.randint(0, 2, kc.shape, dtype=torch.uint8, device=DEV, generator=g) << 7)
        vc |= (torch.randint(0, 2, vc.shape, dtype=torch.uint8, device=DEV, generator=g) << 7)
    else:
        kc = torch.randn(nb, Hkv, D // x, bs, x, device=DEV, generator=g).to(dt)
        vc = torch.randn(nb, Hkv, D, bs, device=DEV, generator=g).to(dt)
    ks, vs = torch.tensor([0.01], device=DEV), torch.tensor([0.004], device=DEV)
    slopes = torch.tensor([2.0 ** -(i % 8 + 1) for i in range(Hq)], device=DEV)
    o1 = torch.empty(S, Hq, D, dtype=dt, device=DEV)
    o2 = torch.empty_like(o1)
    tail = (Hkv, D ** -0.5, bt, sl, bs, max(lens), slopes, kv, ks, vs, 0, 0, 0, 64, 0)
    if version == "v1":
        W._C.paged_attention_v1(o1, q, kc, vc, *tail)
        ours().paged_attention_v1(o2, q, kc, vc, *tail)
    else:
        P = (max(lens) + 511) // 512
        es = [torch.empty(S, Hq, P, device=DEV) for _ in range(2)]
        ml = [torch.empty(S, Hq, P, device=DEV) for _ in range(2)]
        tmp = [torch.empty(S, Hq, P, D, dtype=dt, device=DEV) for _ in range(2)]
        W._C.paged_attention_v2(o1, es[0], ml[0], tmp[0], q, kc, vc, *tail)
        ours().paged_attention_v2(o2, es[1], ml[1], tmp[1], q, kc, vc, *tail)
        for s, L in enumerate(lens):   # side outputs of the partitions that exist
            p = (L + 511) // 512
            torch.testing.assert_close(ml[0][s, :, :p], ml[1][s, :, :p], rtol=1e-3, atol=2e-3)
            torch.testing.assert_close(es[0][s, :, :p], es[1][s, :, :p], rtol=5e-3, atol=1e-4)
    tol = 2e-2 if fp8 else 1e-2   # north_star: bf16 attention output <= 2e-2 abs
    torch.testing.assert_close(o1.float(), o2.float(), atol=tol, rtol=2e-2)


@pytest.mark.parametrize("kv", ["auto", "fp8"])
@pytest.mark.parametrize("heads,D", [((32, 8), 128), ((14, 2), 64), ((8, 1), 128)])
def test_flash_layout_decode_vs_wheel_paged_attention(kv, heads, D):
    """the headline kernel (TMA, flash layout [nb, bs, Hkv, D]) against the wheel's paged_attention_v2 on the
    same KV bytes re-laid out in the legacy x-packed format."""
    W = wheel()
    Hq, Hkv = heads
    bs = 16
    g = gen(19)
    lens = [4096, 1, 1000, 2500, 16, 4095]
    S = len(lens)
    mb = (max(lens) + bs - 1) // bs
    nb = S * mb + 5
    bt = torch.randperm(nb, device=DEV, generator=g)[: S * mb].to(torch.int32).view(S, mb)
    sl = torch.tensor(lens, dtype=torch.int32, device=DEV)
    q = (torch.randn(S, Hq, D, device=DEV, generator=g) * 0.5).to(torch.bfloat16)
    fp8 = kv == "fp8"
    if fp8:
        cache = torch.randint(0, 0x78, (nb, bs, Hkv, 2 * D), dtype=torch.uint8, device=DEV, generator=g)
        cache |= (torch.randint(0, 2, cache.shape, dtype=torch.uint8, device=DEV, generator=g) << 7)
    else:
        cache = torch.randn(nb, bs, Hkv, 2 * D, device=DEV, generator=g).to(torch.bfloat16)
    kf, vf = cache[..., :D], cache[..., D:]
    x = 16 if fp8 else 8
    kc = kf.reshape(nb, bs, Hkv, D // x, x).permute(0, 2, 3, 1, 4).contiguous()
    vc = vf.permute(0, 2, 3, 1).contiguous()
    ks, vs = torch.tensor([0.01], device=DEV), torch.tensor([0.004], device=DEV)
    P = (max(lens) + 511) // 512
    ref = torch.empty(S, Hq, D, dtype=torch.bfloat16, device=DEV)
    W._C.paged_attention_v2(ref, torch.empty(S, Hq, P, device=DEV), torch.empty(S, Hq, P, device=DEV),
                            torch.empty(S, Hq, P, D, dtype=torch.bfloat16, device=DEV), q, kc, vc, Hkv, D ** -0.5,
                            bt, sl, bs, max(lens), None, kv, ks, vs, 0, 0, 0, 64, 0)
    out = torch.empty_like(ref)
    ours().paged_attention_decode(out, q, kf, vf, bt, sl, D ** -0.5, k_scale=ks if fp8 else None,
                                  v_scale=vs if fp8 else None, max_seq_len=max(lens))
    torch.testing.assert_close(out.float(), ref.float(), atol=2e-2, rtol=2e-2)


# ------------------------------------------------------------------------------------------------ NVFP4 MoE
def _wheel_fp4_moe_or_skip(W):
    for name in ("get_cutlass_moe_mm_data", "scaled_fp4_experts_quant", "cutlass_fp4_group_mm"):
        if not hasattr(W._C, name):
            pytest.skip(f"wheel has no _C.{name}")


@pytest.mark.parametrize("m,topk,E", [(64, 8, 64), (224, 6, 40), (256, 2, 8)])
def test_get_cutlass_moe_mm_data_vs_wheel(m, topk, E):
    """offsets and problem sizes exact; the permutations per-expert set-equal (the wheel orders by atomicAdd)."""
    W = wheel()
    _wheel_fp4_moe_or_skip(W)
    ids = torch.stack([torch.randperm(E, device=DEV, generator=gen(m + i))[:topk] for i in range(m)]).to(torch.int32)
    n, k = 1024, 1536

    def run(fn):
        eo = torch.zeros(E + 1, dtype=torch.int32, device=DEV)
        bso = torch.zeros(E + 1, dtype=torch.int32, device=DEV)
        ps1 = torch.zeros(E, 3, dtype=torch.int32, device=DEV)
        ps2 = torch.zeros(E, 3, dtype=torch.int32, device=DEV)
        am = torch.zeros(m * topk, dtype=torch.int32, device=DEV)
        cm = torch.zeros(m * topk, dtype=torch.int32, device=DEV)
        fn(ids, eo, ps1, ps2, am, cm, E, n, k, bso, True)
        return [t.cpu().numpy() for t in (eo, bso, ps1, ps2, am, cm)]

    a = run(W._C.get_cutlass_moe_mm_data)
    b = run(ours().get_cutlass_moe_mm_data)
    for x, y in zip(a[:4], b[:4]):
        np.testing.assert_array_equal(x, y)
    eo = a[0]
    for e in range(E):
        assert sorted(a[4][eo[e]:eo[e + 1]].tolist()) == sorted(b[4][eo[e]:eo[e + 1]].tolist())
    flat = ids.cpu().numpy().reshape(-1)
    for cm in (a[5], b[5]):
        assert ((cm >= eo[flat]) & (cm < eo[flat + 1])).all()


@pytest.mark.parametrize("gated", [False, True])
@pytest.mark.parametrize("m,topk,E,k", [(64, 8, 64, 1024), (224, 6, 40, 1536), (256, 2, 8, 4096)])
def test_scaled_fp4_experts_quant_bits_vs_wheel(m, topk, E, k, gated):
    """codes of every packed row and scale bytes of every valid (row, column) bit-identical to the wheel's kernel."""
    W = wheel()
    _wheel_fp4_moe_or_skip(W)
    name = "silu_and_mul_scaled_fp4_experts_quant" if gated else "scaled_fp4_experts_quant"
    if not hasattr(W._C, name):
        pytest.skip(f"wheel has no _C.{name}")
    from oracle import ref_ops
    ids = np.stack([np.random.default_rng(i).choice(E, topk, replace=False) for i in range(m)]).astype(np.int32)
    d = ref_ops.cutlass_moe_mm_data(ids, E, 64, k)
    eo = torch.from_numpy(d["expert_offsets"]).to(DEV)
    bso = torch.from_numpy(d["blockscale_offsets"]).to(DEV)
    rows = m * topk
    x = torch.randn(rows, k * (2 if gated else 1), device=DEV, generator=gen(20)).to(torch.bfloat16)
    gs = (torch.rand(E, device=DEV, generator=gen(21)) * 300 + 50).float()
    rc = (k // 16 + 3) // 4 * 4
    sf_rows = int(d["blockscale_offsets"][E])
    o1 = torch.zeros(rows, k // 2, dtype=torch.uint8, device=DEV)
    s1 = torch.zeros(sf_rows, rc // 4, dtype=torch.int32, device=DEV)
    o2, s2 = torch.zeros_like(o1), torch.zeros_like(s1)
    getattr(W._C, name)(o1, s1, x, gs, eo, bso)
    fn = ours().silu_and_mul_scaled_fp4_experts_quant_out if gated else ours().scaled_fp4_experts_quant_out
    fn(o2, s2, x, gs, eo, bso)
    assert torch.equal(o1, o2)
    f1, f2 = bits(s1).cpu().numpy().reshape(-1), bits(s2).cpu().numpy().reshape(-1)
    for e in range(E):
        lo, hi = int(d["expert_offsets"][e]), int(d["expert_offsets"][e + 1])
        if hi == lo:
            continue
        mask = ref_ops.swizzle_sf(np.ones((hi - lo, k // 16), dtype=np.uint8)) != 0
        base = int(d["blockscale_offsets"][e]) * rc
        np.testing.assert_array_equal(f1[base:base + mask.size][mask], f2[base:base + mask.size][mask])


@pytest.mark.parametrize("m,topk,E,n,k", [(64, 8, 64, 512, 1024), (224, 6, 40, 1024, 1536), (256, 2, 8, 2048, 1024)])
def test_cutlass_fp4_group_mm_vs_wheel(m, topk, E, n, k):
    W = wheel()
    _wheel_fp4_moe_or_skip(W)
    from oracle import ref_ops
    ids = np.stack([np.random.default_rng(i).choice(E, topk, replace=False) for i in range(m)]).astype(np.int32)
    d = ref_ops.cutlass_moe_mm_data(ids, E, n, k, is_gated=False)
    eo = torch.from_numpy(d["expert_offsets"]).to(DEV)
    bso = torch.from_numpy(d["blockscale_offsets"]).to(DEV)
    ps = torch.from_numpy(d["problem_sizes1"]).to(DEV)
    rows = m * topk
    x = torch.randn(rows, k, device=DEV, generator=gen(30)).to(torch.bfloat16)
    a_gs = (torch.rand(E, device=DEV, generator=gen(31)) * 300 + 50).float()
    a4, a_sf = ours().scaled_fp4_experts_quant(x, a_gs, eo, bso, topk)
    rn, rc = (n + 127) // 128 * 128, (k // 16 + 3) // 4 * 4
    w4 = torch.empty(E, n, k // 2, dtype=torch.uint8, device=DEV)
    w_sf = torch.empty(E, rn, rc, dtype=torch.uint8, device=DEV)
    w_gs = torch.empty(E, device=DEV)
    for e in range(E):
        w = (torch.randn(n, k, device=DEV, generator=gen(40 + e)) / 8).to(torch.bfloat16)
        w_gs[e] = 448.0 * 6.0 / w.float().abs().max()
        q, s = ours().scaled_fp4_quant(w, w_gs[e:e + 1])
        w4[e], w_sf[e] = q, bits(s).view(rn, rc)
    w_sf = w_sf.view(torch.float8_e4m3fn)
    alphas = 1.0 / (a_gs * w_gs)
    ref = torch.zeros(rows, n, dtype=torch.bfloat16, device=DEV)
    out = torch.zeros_like(ref)
    try:
        W._C.cutlass_fp4_group_mm(ref, a4, w4, a_sf, w_sf, alphas, ps, eo[:-1].contiguous(), bso[:-1].contiguous())
        torch.cuda.synchronize()
    except RuntimeError as e:   # the wheel's grouped kernel may be compiled out for this arch
        pytest.skip(f"wheel cutlass_fp4_group_mm unavailable: {repr(e)[:120]}")
    ours().cutlass_fp4_group_mm(out, a4, w4, a_sf, w_sf, alphas, ps, eo[:-1].contiguous(), bso[:-1].contiguous())
    torch.testing.assert_close(out.float(), ref.float(), rtol=2 ** -7, atol=2e-3 * ref.float().abs().max().item())
    assert (out == ref).float().mean().item() > 0.95
