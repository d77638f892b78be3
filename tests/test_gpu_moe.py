"""GPU parity: moe_align_block_size (bit-exact), topk_softmax, moe_sum, grouped FP8 GEMM and
fused_experts vs the oracle. Mirrors tests/kernels/moe/test_moe_align_block_size.py:189-327,
test_moe.py:304-420 and test_cutlass_moe.py:395-488 of the reference."""
import numpy as np
import pytest
import torch

from oracle import ref_ops

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _moe():
    from b200ops import fused_moe
    return fused_moe


@pytest.mark.parametrize("m", [1, 3, 256, 2256, 4096])
@pytest.mark.parametrize("topk", [1, 2, 16, 32])
@pytest.mark.parametrize("E", [32, 160, 256, 257])
@pytest.mark.parametrize("bs", [32, 128])
@pytest.mark.parametrize("pad_sorted", [False, True])
def test_moe_align_block_size(m, topk, E, bs, pad_sorted):
    g = torch.Generator().manual_seed(0)
    ids = torch.rand(m, E, generator=g).argsort(dim=1)[:, :topk].to(torch.int32)  # per-row randperm(E)[:k]
    s, e, n = ref_ops.moe_align_block_size(ids.numpy(), bs, E, pad_sorted_ids=pad_sorted)
    gs, ge, gn = _moe().moe_align_block_size(ids.to(DEV), bs, E, pad_sorted_ids=pad_sorted)
    assert int(gn.item()) == n
    # bit-exact including the order inside each expert segment (our kernel is deterministic/stable)
    np.testing.assert_array_equal(gs.cpu().numpy(), s)
    np.testing.assert_array_equal(ge.cpu().numpy(), e)


def test_moe_align_expert_map_invalid_ids_int64_and_determinism():
    g = torch.Generator().manual_seed(1)
    E = 64
    ids = torch.randint(0, E, (333, 4), generator=g).to(torch.int64)
    ids[5, 1] = -1
    ids[7, 0] = E + 3
    emap = torch.full((E,), -1, dtype=torch.int32)
    emap[16:32] = torch.arange(16, dtype=torch.int32)
    s, e, n = ref_ops.moe_align_block_size(ids.numpy(), 16, E, emap.numpy(), ignore_invalid_experts=True)
    outs = []
    for _ in range(3):
        gs, ge, gn, inv = _moe().moe_align_block_size(ids.to(DEV), 16, E, emap.to(DEV),
                                                      ignore_invalid_experts=True, return_inv_perm=True)
        outs.append((gs.cpu(), ge.cpu(), int(gn.item())))
    np.testing.assert_array_equal(outs[0][0].numpy(), s)
    np.testing.assert_array_equal(outs[0][1].numpy(), e)
    assert outs[0][2] == n
    assert all(torch.equal(outs[0][0], o[0]) for o in outs[1:])
    inv = inv.cpu().numpy()
    kept = np.nonzero(inv >= 0)[0]
    np.testing.assert_array_equal(s[inv[kept]], kept)
    # without ignore_invalid_experts: expert_ids are mapped on the way out (python wrapper)
    s2, e2, n2 = ref_ops.moe_align_block_size(ids.numpy(), 16, E, emap.numpy())
    gs, ge, gn = _moe().moe_align_block_size(ids.to(DEV), 16, E, emap.to(DEV))
    np.testing.assert_array_equal(gs.cpu().numpy(), s2)
    np.testing.assert_array_equal(ge.cpu().numpy(), e2)


@pytest.mark.parametrize("M,E,topk", [(1, 8, 2), (256, 8, 2), (33, 64, 6), (77, 256, 8), (5, 160, 4), (3, 1000, 3)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("renorm", [False, True])
def test_topk_softmax(M, E, topk, dtype, renorm):
    g = torch.Generator().manual_seed(2)
    gating = torch.randn(M, E, generator=g).to(dtype)
    w, ids, tei = ref_ops.topk_softmax(gating, topk, renorm)
    gw = torch.empty(M, topk, dtype=torch.float32, device=DEV)
    gi = torch.empty(M, topk, dtype=torch.int32, device=DEV)
    gt = torch.empty(M, topk, dtype=torch.int32, device=DEV)
    _moe().topk_softmax(gw, gi, gt, gating.to(DEV), renorm)
    assert torch.equal(gi.cpu(), ids)
    assert torch.equal(gt.cpu(), tei)
    torch.testing.assert_close(gw.cpu(), w, rtol=2e-5, atol=1e-7)


@pytest.mark.parametrize("M,topk,d", [(1, 2, 128), (256, 2, 4096), (37, 8, 1024), (5, 3, 7168)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
def test_moe_sum(M, topk, d, dtype):
    g = torch.Generator().manual_seed(3)
    x = torch.randn(M, topk, d, generator=g).to(dtype)
    out = torch.empty(M, d, dtype=dtype, device=DEV)
    _moe().moe_sum(x.to(DEV), out)
    ref = ref_ops.moe_sum(x)
    if topk <= 2:
        assert torch.equal(out.cpu(), ref)  # a two-term fp32 sum has one rounding: exact
    else:
        torch.testing.assert_close(out.cpu(), ref, rtol=2 ** -7 if dtype == torch.bfloat16 else 1e-3, atol=1e-3)


def to_bits(t):
    return t.clamp(-448, 448).to(torch.float8_e4m3fn).view(torch.uint8)


@pytest.mark.parametrize("M,E,topk,H,I", [(1, 8, 2, 256, 512), (33, 8, 2, 512, 256), (256, 8, 2, 1024, 1408),
                                           (64, 32, 4, 256, 128), (7, 4, 1, 128, 128)])
@pytest.mark.parametrize("static_act", [True, False])
def test_fused_experts_fp8(M, E, topk, H, I, static_act):
    g = torch.Generator().manual_seed(4)
    x = (torch.randn(M, H, generator=g) * 0.5).to(torch.bfloat16)
    w1 = to_bits(torch.randn(E, 2 * I, H, generator=g))
    w2 = to_bits(torch.randn(E, H, I, generator=g))
    w1s = (torch.rand(E, generator=g) + 0.5) * (0.5 / H ** 0.5)
    w2s = (torch.rand(E, generator=g) + 0.5) * (0.5 / I ** 0.5)
    gating = torch.randn(M, E, generator=g)
    tw, ti, _ = ref_ops.topk_softmax(gating, topk, True)
    a1 = torch.tensor([x.float().abs().max().item() / 448.0])
    if static_act:
        a2 = torch.tensor([0.02])
    fm = _moe()
    out = fm.fused_experts(x.to(DEV), w1.to(DEV).view(torch.float8_e4m3fn), w2.to(DEV).view(torch.float8_e4m3fn),
                           tw.to(DEV), ti.to(DEV), use_fp8_w8a8=True, w1_scale=w1s.to(DEV), w2_scale=w2s.to(DEV),
                           a1_scale=a1.to(DEV) if static_act else None,
                           a2_scale=a2.to(DEV) if static_act else None)
    torch.cuda.synchronize()
    if not static_act:
        # dynamic: recover the a2 the GPU derived (max|act| / 448) through the same oracle ops
        M_, = (x.shape[0],)
        x8 = torch.from_numpy(ref_ops.e4m3_bits_to_f32(ref_ops.static_scaled_fp8_quant(x, a1)))
        acts = []
        for t in range(M_):
            for k in range(topk):
                e = int(ti[t, k])
                w = torch.from_numpy(ref_ops.e4m3_bits_to_f32(w1[e].numpy()))
                h = ((x8[t] @ w.t()) * (float(a1) * float(w1s[e]))).to(torch.bfloat16)
                acts.append(ref_ops.silu_and_mul(h[None]))
        a2 = torch.tensor([torch.cat(acts).float().abs().max().item() / 448.0])
    ref = ref_ops.fused_experts_fp8(x, w1, w2, tw, ti, w1s, w2s, a1, a2)
    got = out.float().cpu()
    scale = ref.abs().max().item()
    # reference tolerance for cutlass-vs-triton fp8 MoE: atol 5.5e-2 rtol 1e-2 (test_cutlass_moe.py);
    # ours: bf16 roundings of the intermediates only
    torch.testing.assert_close(got, ref, atol=2e-2 * scale + 1e-3, rtol=2e-2)
    assert (got - ref).abs().mean().item() <= 2e-3 * scale + 1e-4


def test_fused_experts_expert_parallel_shards_sum_to_full():
    """EP form of SURVEY 8e: every rank runs its local experts on all tokens; the sum of the shards
    equals the single-rank result (what the all-reduce after the MoE block computes)."""
    g = torch.Generator().manual_seed(5)
    M, E, topk, H, I = 64, 8, 2, 512, 256
    x = (torch.randn(M, H, generator=g) * 0.5).to(torch.bfloat16).to(DEV)
    w1 = to_bits(torch.randn(E, 2 * I, H, generator=g)).to(DEV)
    w2 = to_bits(torch.randn(E, H, I, generator=g)).to(DEV)
    w1s = ((torch.rand(E, generator=g) + 0.5) * 0.02).to(DEV)
    w2s = ((torch.rand(E, generator=g) + 0.5) * 0.03).to(DEV)
    tw, ti, _ = ref_ops.topk_softmax(torch.randn(M, E, generator=g), topk, True)
    a1, a2 = torch.tensor([0.01], device=DEV), torch.tensor([0.02], device=DEV)
    fm = _moe()
    f8 = torch.float8_e4m3fn
    full = fm.fused_experts(x, w1.view(f8), w2.view(f8), tw.to(DEV), ti.to(DEV), use_fp8_w8a8=True,
                            w1_scale=w1s, w2_scale=w2s, a1_scale=a1, a2_scale=a2).float()
    acc = torch.zeros_like(full)
    ep = 4
    per = E // ep
    for r in range(ep):
        emap = torch.full((E,), -1, dtype=torch.int32)
        emap[r * per:(r + 1) * per] = torch.arange(per, dtype=torch.int32)
        sl = slice(r * per, (r + 1) * per)
        acc += fm.fused_experts(x, w1[sl].contiguous().view(f8), w2[sl].contiguous().view(f8), tw.to(DEV),
                                ti.to(DEV), use_fp8_w8a8=True, global_num_experts=E,
                                expert_map=emap.to(DEV), w1_scale=w1s[sl].contiguous(),
                                w2_scale=w2s[sl].contiguous(), a1_scale=a1, a2_scale=a2).float()
    torch.testing.assert_close(acc, full, atol=2e-2 * full.abs().max().item(), rtol=2e-2)


# ------------------------------------------------------------------------------------------------
# The other quantisation schemes of fused_experts: per_channel_quant=True (dynamic
# per-token activation scales + per-channel weight scales), block_shape=[128,128], apply_router_weight_on_input.
# ------------------------------------------------------------------------------------------------
def _scheme_case(M, E, topk, H, I, scheme, seed=6):
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn(M, H, generator=g) * (torch.rand(M, 1, generator=g) + 0.2)).to(torch.bfloat16)
    w1 = to_bits(torch.randn(E, 2 * I, H, generator=g))
    w2 = to_bits(torch.randn(E, H, I, generator=g))
    if scheme == "channel":
        w1s = (torch.rand(E, 2 * I, generator=g) + 0.5) * (0.5 / H ** 0.5)
        w2s = (torch.rand(E, H, generator=g) + 0.5) * (0.5 / I ** 0.5)
    elif scheme == "block":
        w1s = (torch.rand(E, (2 * I + 127) // 128, (H + 127) // 128, generator=g) + 0.5) * (0.5 / H ** 0.5)
        w2s = (torch.rand(E, (H + 127) // 128, (I + 127) // 128, generator=g) + 0.5) * (0.5 / I ** 0.5)
    else:
        w1s = (torch.rand(E, generator=g) + 0.5) * (0.5 / H ** 0.5)
        w2s = (torch.rand(E, generator=g) + 0.5) * (0.5 / I ** 0.5)
    tw, ti, _ = ref_ops.topk_softmax(torch.randn(M, E, generator=g), topk, True)
    return x, w1, w2, w1s, w2s, tw, ti


def _run_scheme(x, w1, w2, w1s, w2s, tw, ti, scheme, **kw):
    f8 = torch.float8_e4m3fn
    return _moe().fused_experts(x.to(DEV), w1.to(DEV).view(f8), w2.to(DEV).view(f8), tw.to(DEV), ti.to(DEV),
                                use_fp8_w8a8=True, per_channel_quant=(scheme == "channel"),
                                block_shape=[128, 128] if scheme == "block" else None,
                                w1_scale=w1s.to(DEV), w2_scale=w2s.to(DEV), **kw).float().cpu()


@pytest.mark.parametrize("M,E,topk,H,I", [(1, 8, 2, 256, 512), (33, 8, 2, 512, 256), (256, 8, 2, 1024, 1408),
                                           (64, 32, 4, 256, 128), (200, 4, 1, 384, 640)])
@pytest.mark.parametrize("scheme", ["channel", "block"])
def test_fused_experts_fp8_per_token_and_block_schemes(M, E, topk, H, I, scheme):
    if scheme == "block" and I % 128:
        I = (I + 127) // 128 * 128      # per_token_group_quant_fp8 needs whole groups (as the reference asserts)
    x, w1, w2, w1s, w2s, tw, ti = _scheme_case(M, E, topk, H, I, scheme)
    got = _run_scheme(x, w1, w2, w1s, w2s, tw, ti, scheme)
    ref = ref_ops.fused_experts_fp8_scheme(x, w1, w2, tw, ti, w1s, w2s, scheme)
    scale = ref.abs().max().item()
    torch.testing.assert_close(got, ref, atol=2e-2 * scale + 1e-3, rtol=2e-2)
    assert (got - ref).abs().mean().item() <= 2e-3 * scale + 1e-4


def test_fused_experts_per_channel_static_activation_scales():
    """per_channel_quant with given a*_scale: static per-tensor activations (scaled_fp8_quant ignores
    use_per_token_if_dynamic when a scale is passed, _custom_ops.py:1881-1896) x per-channel weights."""
    x, w1, w2, w1s, w2s, tw, ti = _scheme_case(48, 8, 2, 256, 256, "channel")
    a1, a2 = torch.tensor([0.012]), torch.tensor([0.02])
    got = _run_scheme(x, w1, w2, w1s, w2s, tw, ti, "channel", a1_scale=a1.to(DEV), a2_scale=a2.to(DEV))
    ref = ref_ops.fused_experts_fp8_scheme(x, w1, w2, tw, ti, w1s, w2s, "channel", a1_scale=a1, a2_scale=a2)
    torch.testing.assert_close(got, ref, atol=2e-2 * ref.abs().max().item(), rtol=2e-2)


@pytest.mark.parametrize("scheme", ["tensor", "channel", "block"])
def test_fused_experts_apply_router_weight_on_input(scheme):
    x, w1, w2, w1s, w2s, tw, ti = _scheme_case(40, 8, 1, 256, 256, scheme)
    tw = torch.rand(40, 1) + 0.25
    got = _run_scheme(x, w1, w2, w1s, w2s, tw, ti, scheme, apply_router_weight_on_input=True)
    ref = ref_ops.fused_experts_fp8_scheme(x, w1, w2, tw, ti, w1s, w2s, scheme, apply_router_weight_on_input=True)
    torch.testing.assert_close(got, ref, atol=2e-2 * ref.abs().max().item(), rtol=2e-2)
    with pytest.raises(RuntimeError, match="topk=1"):
        x2, w1b, w2b, s1, s2, tw2, ti2 = _scheme_case(8, 8, 2, 256, 256, "tensor")
        _run_scheme(x2, w1b, w2b, s1, s2, tw2, ti2, "tensor", apply_router_weight_on_input=True)


def test_fused_experts_c4_shape_sampled_against_oracle():
    """BASELINE configs[3] full shape (M 256, E 8, top-2, H 4096, I 14336): the oracle is evaluated for a sample of
    the tokens only (each token's output depends on the others only through the static scales)."""
    M, E, topk, H, I = 256, 8, 2, 4096, 14336
    g = torch.Generator(device=DEV).manual_seed(8)
    f8 = torch.float8_e4m3fn
    x = (torch.randn(M, H, generator=g, device=DEV) * 0.5).to(torch.bfloat16)
    w1 = torch.randn(E, 2 * I, H, generator=g, device=DEV).clamp(-448, 448).to(f8)
    w2 = torch.randn(E, H, I, generator=g, device=DEV).clamp(-448, 448).to(f8)
    w1s = ((torch.rand(E, generator=g, device=DEV) + 0.5) * (0.5 / H ** 0.5))
    w2s = ((torch.rand(E, generator=g, device=DEV) + 0.5) * (0.5 / I ** 0.5))
    tw, ti, _ = ref_ops.topk_softmax(torch.randn(M, E, generator=torch.Generator().manual_seed(9)), topk, True)
    a1, a2 = torch.tensor([0.006], device=DEV), torch.tensor([0.03], device=DEV)
    out = _moe().fused_experts(x, w1, w2, tw.to(DEV), ti.to(DEV), use_fp8_w8a8=True, w1_scale=w1s, w2_scale=w2s,
                               a1_scale=a1, a2_scale=a2).float().cpu()
    sample = [0, 1, 77, 128, 255]
    ref = ref_ops.fused_experts_fp8_scheme(x[sample].cpu(), w1.view(torch.uint8).cpu(), w2.view(torch.uint8).cpu(),
                                           tw[sample], ti[sample], w1s.cpu(), w2s.cpu(), "tensor", a1_scale=a1.cpu(),
                                           a2_scale=a2.cpu())
    got = out[sample]
    scale = ref.abs().max().item()
    torch.testing.assert_close(got, ref, atol=2e-2 * scale + 1e-3, rtol=2e-2)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_moe_permute_fp8_and_unpermute_reduce_direct(dtype):
    """b200_moe_permute_fp8 / b200_moe_permute_rows / b200_moe_unpermute_reduce alone, against the definition
    (moe_permute_unpermute_op.cu semantics: gather rows by sorted_token_ids // topk; weighted scatter-sum back)."""
    from b200ops import _lib
    from b200ops.ops import _dt, _p, _stream
    lib = _lib.load()
    fm = _moe()
    M, E, topk, H, bs = 77, 16, 3, 512, 32
    g = torch.Generator().manual_seed(12)
    x = torch.randn(M, H, generator=g).to(dtype)
    ids = torch.stack([torch.randperm(E, generator=g)[:topk] for _ in range(M)]).to(torch.int32)
    emap = torch.arange(E, dtype=torch.int32)
    emap[3] = -1                                             # expert 3 lives on another rank
    sorted_ids, expert_ids, npp, inv = fm.moe_align_block_size(ids.to(DEV), bs, E, emap.to(DEV), pad_sorted_ids=True,
                                                               ignore_invalid_experts=True, return_inv_perm=True)
    P = sorted_ids.numel()
    scale = torch.tensor([0.02])
    a8 = torch.full((P, H), 0x55, dtype=torch.uint8, device=DEV)
    xd, sd = x.to(DEV), scale.to(DEV)       # keep the device tensors alive: _p() hands out raw pointers
    rc = lib.b200_moe_permute_fp8(_p(a8), _p(xd), _p(sorted_ids), _p(npp), _p(sd), P, topk, M * topk,
                                  H, H, _dt(x), _stream())
    assert rc == 0
    s_ids, n = sorted_ids.cpu().numpy(), int(npp.item())
    q = ref_ops.static_scaled_fp8_quant(x, scale)
    exp = np.full((P, H), 0x55, dtype=np.uint8)
    for i in range(n):
        exp[i] = q[s_ids[i] // topk] if s_ids[i] < M * topk else 0
    np.testing.assert_array_equal(a8.cpu().numpy(), exp)
    # permute_rows: bytes and scale rows
    qd = torch.from_numpy(q).to(DEV)
    srow = torch.rand(M, 4, generator=g).to(DEV)
    dst = torch.full((P, H), 0x55, dtype=torch.uint8, device=DEV)
    sdst = torch.full((P, 4), -1.0, device=DEV)
    rc = lib.b200_moe_permute_rows(_p(dst), _p(qd), _p(sorted_ids), _p(npp), _p(sdst), _p(srow), 4, 4, P, topk,
                                   M * topk, H, H, _stream())
    assert rc == 0
    np.testing.assert_array_equal(dst.cpu().numpy(), exp)
    exp_s = np.full((P, 4), -1.0, dtype=np.float32)
    for i in range(n):
        exp_s[i] = srow[s_ids[i] // topk].cpu().numpy() if s_ids[i] < M * topk else 0
    np.testing.assert_array_equal(sdst.cpu().numpy(), exp_s)
    # unpermute: out[t] = sum_k w[t,k] * src[inv[t,k]] (fp32 accumulate, one rounding), dropped pairs skipped
    src = torch.randn(P, H, generator=g).to(dtype)
    tw = torch.rand(M, topk, generator=g)
    out = torch.empty(M, H, dtype=dtype, device=DEV)
    srcd, twd = src.to(DEV), tw.to(DEV)
    rc = lib.b200_moe_unpermute_reduce(_p(out), _p(srcd), _p(inv), _p(twd), None, M, topk, H, H,
                                       _dt(out), _stream())
    assert rc == 0
    torch.cuda.synchronize()
    invc = inv.cpu().numpy().reshape(M, topk)
    assert ((invc < 0) == (ids.numpy() == 3)).all()
    ref = torch.zeros(M, H)
    for t in range(M):
        for k in range(topk):
            if invc[t, k] >= 0:
                ref[t] = torch.from_numpy((ref[t].numpy() + np.float32(tw[t, k].item()) * src[invc[t, k]].float().numpy()).astype(np.float32))
    # fp32 accumulate (FMA or mul + add: at most an fp32 ulp apart) and ONE rounding to the output dtype
    eps = 2.0 ** -8 if dtype == torch.bfloat16 else 2.0 ** -11
    torch.testing.assert_close(out.cpu().float(), ref, rtol=eps, atol=1e-6)
    assert (out.cpu() == ref.to(dtype)).float().mean().item() > 0.999
