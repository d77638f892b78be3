"""CPU tests: MoE routing oracle pinned to the reference's worked example and its golden properties."""
import numpy as np
import pytest
import torch

from oracle import ref_ops


def test_moe_align_worked_example_from_reference_docstring():
    # vllm/model_executor/layers/fused_moe/moe_align_block_size.py:59-72 (experts are 1-based in the
    # docstring's topk_ids; with num_experts = 5 the expert-0 segment is empty)
    topk_ids = np.array([[2, 3, 4], [1, 2, 4], [1, 3, 4], [1, 2, 3]])
    s, e, n = ref_ops.moe_align_block_size(topk_ids, 4, 5)
    assert n == 16
    assert s[:16].tolist() == [3, 6, 9, 12, 0, 4, 10, 12, 1, 7, 11, 12, 2, 5, 8, 12]
    assert e[:4].tolist() == [1, 2, 3, 4]
    assert (e[4:] == -1).all() and (s[16:] == 12).all()


def test_moe_align_properties_and_expert_map():
    g = np.random.default_rng(0)
    for m, topk, E, bs in [(1, 2, 32, 32), (3, 16, 160, 128), (256, 2, 8, 64), (2256, 8, 257, 32)]:
        ids = np.stack([g.permutation(E)[:topk] for _ in range(m)])
        s, e, n = ref_ops.moe_align_block_size(ids, bs, E)
        assert n % bs == 0 and n <= s.size
        flat = ids.reshape(-1)
        for b in range(n // bs):
            blk = s[b * bs:(b + 1) * bs]
            real = blk[blk < flat.size]
            assert (flat[real] == e[b]).all()
            assert (np.diff(real) > 0).all()  # stable order
        assert np.sort(s[s < flat.size]).tolist() == list(range(flat.size))
    # expert parallel: only experts 4..7 are local
    emap = np.full(8, -1, dtype=np.int32)
    emap[4:] = np.arange(4)
    ids = np.stack([g.permutation(8)[:2] for _ in range(64)])
    s, e, n = ref_ops.moe_align_block_size(ids, 16, 8, emap, ignore_invalid_experts=True)
    kept = s[s < 128]
    assert (ids.reshape(-1)[kept] >= 4).all() and kept.size == (ids >= 4).sum()
    assert e[: n // 16].max() <= 3


def test_topk_softmax_oracle():
    torch.manual_seed(0)
    g = torch.randn(17, 8)
    w, ids, tei = ref_ops.topk_softmax(g, 2, True)
    ref_w, ref_i = torch.topk(torch.softmax(g, -1), 2)
    assert torch.equal(ids.long(), ref_i)
    torch.testing.assert_close(w, ref_w / ref_w.sum(-1, keepdim=True))
    assert tei[3, 1] == 1 * 17 + 3


def test_nvfp4_oracle_matches_reference_test_helpers():
    """restated ref_nvfp4_quant / recover_swizzled_scales (test_nvfp4_quant.py:98-131)."""
    torch.manual_seed(42)
    x = torch.randn(150, 160).to(torch.bfloat16)
    gs = 448.0 * 6.0 / x.float().abs().max().item()
    codes, sf = ref_ops.nvfp4_quant(x, gs)
    # reference formulation, in torch
    xb = x.float().reshape(150, 10, 16)
    vec_max = xb.abs().max(-1, keepdim=True)[0]
    scale = (torch.tensor(gs, dtype=torch.float32) * (vec_max * (1.0 / 6.0))).to(torch.float8_e4m3fn).float()
    inv = torch.where(scale == 0, torch.zeros_like(scale), 1.0 / (scale * (1.0 / torch.tensor(gs, dtype=torch.float32))))
    ref_vals = (xb * inv).clamp(-6, 6)
    got_vals = torch.from_numpy(ref_ops.e2m1_code_to_f32(codes)).reshape(150, 10, 16)
    assert (got_vals - ref_vals).abs().max() <= 1.0  # nearest representable
    np.testing.assert_array_equal(ref_ops.e4m3_bits_to_f32(sf), scale.squeeze(-1).numpy())
    # swizzle round trip + reference recover_swizzled_scales
    flat = ref_ops.swizzle_sf(sf)
    np.testing.assert_array_equal(ref_ops.unswizzle_sf(flat, 150, 10), sf)
    t = torch.from_numpy(flat.copy()).reshape(1, 2, 3, 32, 4, 4).permute(0, 1, 4, 3, 2, 5).reshape(256, 12)
    np.testing.assert_array_equal(t[:150, :10].numpy(), sf)


# ------------------------------------------------------------------------------------------------
# NVFP4 fused MoE oracle (packed expert-sorted rows): pinned by definition-level properties and by the
# reference test's own model of the block (tests/kernels/moe/test_nvfp4_moe.py:121-160: dequantise, then a
# dense per-token MoE, atol = rtol = 1e-1).
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("m,topk,E", [(1, 1, 4), (7, 3, 5), (64, 8, 64), (100, 2, 8)])
def test_cutlass_moe_mm_data_oracle_definition(m, topk, E):
    rng = np.random.default_rng(m * 31 + E)
    ids = np.stack([rng.choice(E, size=topk, replace=False) for _ in range(m)]).astype(np.int32)
    ids[rng.random(ids.shape) < 0.1] = -1
    d = ref_ops.cutlass_moe_mm_data(ids, E, 96, 160)
    flat = ids.reshape(-1)
    off, bso = d["expert_offsets"], d["blockscale_offsets"]
    rows = int(off[E])
    assert rows == (flat >= 0).sum()
    for e in range(E):
        cnt = int((flat == e).sum())
        assert off[e + 1] - off[e] == cnt                                   # moe_data.cu:17-52, 55-66
        assert tuple(d["problem_sizes1"][e]) == (cnt, 192, 160)             # gated: 2n
        assert tuple(d["problem_sizes2"][e]) == (cnt, 160, 96)
        assert bso[e + 1] - bso[e] == (cnt + 127) // 128 * 128              # moe_data.cu:68-85
        # the packed rows of expert e are exactly the tokens routed to it (compute_arg_sorts, :87-117)
        toks = sorted(int(i) // topk for i in np.nonzero(flat == e)[0])
        assert sorted(d["a_map"][off[e]:off[e + 1]].tolist()) == toks
    for i, e in enumerate(flat):
        if e < 0:
            assert d["c_map"][i] == rows                                    # "invalid" pairs -> the packed row count
        else:
            assert off[e] <= d["c_map"][i] < off[e + 1]
            assert d["a_map"][d["c_map"][i]] == i // topk                   # c_map inverts a_map
    assert (d["a_map"][rows:] == -1).all()
    sw = ref_ops.cutlass_moe_mm_data(ids, E, 96, 160, swap_ab=True, with_blockscale=False)
    assert (sw["problem_sizes1"][:, [1, 0, 2]] == d["problem_sizes1"]).all()
    assert (sw["problem_sizes2"][:, [1, 0, 2]] == d["problem_sizes2"]).all()


def test_nvfp4_experts_quant_oracle_is_per_expert_quant():
    g = torch.Generator().manual_seed(0)
    off = np.array([0, 3, 3, 140, 141], dtype=np.int32)
    bso = np.array([0, 128, 128, 384, 512], dtype=np.int32)
    x = torch.randn(141, 96, generator=g).to(torch.bfloat16)
    gs = np.array([100.0, 7.0, 300.0, 50.0], dtype=np.float32)
    codes, sf, rows = ref_ops.nvfp4_experts_quant(x, gs, off, bso)
    rc = 8   # round_up(96 / 16, 4)
    assert sf.size == 512 * rc
    for e in range(4):
        lo, hi = off[e], off[e + 1]
        if hi == lo:
            continue
        c, s = ref_ops.nvfp4_quant(x[lo:hi], gs[e])
        np.testing.assert_array_equal(codes[lo:hi], c)
        np.testing.assert_array_equal(rows[lo:hi], s)
        n_pad = (hi - lo + 127) // 128 * 128
        got = ref_ops.unswizzle_sf(sf[bso[e] * rc:bso[e] * rc + n_pad * rc], hi - lo, 6)
        np.testing.assert_array_equal(got, s)


def test_cutlass_moe_fp4_oracle_vs_reference_test_model():
    m, n, k, E, topk = 24, 64, 128, 6, 2
    g = torch.Generator().manual_seed(7)
    a = (torch.randn(m, k, generator=g) / 10).to(torch.bfloat16)
    w1 = (torch.randn(E, 2 * n, k, generator=g) / 8).to(torch.bfloat16)
    w2 = (torch.randn(E, k, n, generator=g) / 8).to(torch.bfloat16)
    score = torch.randn(m, E, generator=g)
    tw, ti = torch.topk(torch.softmax(score, dim=-1), topk, dim=-1)

    def quant(w):
        gs = np.array([448.0 * 6.0 / w[e].float().abs().max().item() for e in range(E)], dtype=np.float32)
        cs = [ref_ops.nvfp4_quant(w[e], gs[e]) for e in range(E)]
        return np.stack([c for c, _ in cs]), np.stack([s for _, s in cs]), gs

    w1c, w1s, w1gs = quant(w1)
    w2c, w2s, w2gs = quant(w2)
    one = np.ones(E, dtype=np.float32)
    out = ref_ops.cutlass_moe_fp4(a, one, w1c, w1s, 1 / w1gs, one, w2c, w2s, 1 / w2gs, tw, ti.numpy().astype(np.int32))
    # reference test model: dequantised weights, activations as given, dense per-token MoE in fp32
    w1d = torch.from_numpy(np.stack([ref_ops.nvfp4_dequant(w1c[e], w1s[e], w1gs[e]) for e in range(E)]))
    w2d = torch.from_numpy(np.stack([ref_ops.nvfp4_dequant(w2c[e], w2s[e], w2gs[e]) for e in range(E)]))
    dense = torch.zeros(m, k)
    for t in range(m):
        for j in range(topk):
            e = int(ti[t, j])
            h = w1d[e] @ a[t].float()
            dense[t] += tw[t, j] * (w2d[e] @ (torch.nn.functional.silu(h[:n]) * h[n:]))
    torch.testing.assert_close(out.float(), dense, atol=1e-1, rtol=1e-1)
    # and dropping a pair removes exactly its contribution
    ti2 = ti.numpy().astype(np.int32).copy()
    ti2[:, 1] = -1
    out1 = ref_ops.cutlass_moe_fp4(a, one, w1c, w1s, 1 / w1gs, one, w2c, w2s, 1 / w2gs, tw, ti2)
    outk1 = ref_ops.cutlass_moe_fp4(a, one, w1c, w1s, 1 / w1gs, one, w2c, w2s, 1 / w2gs, tw[:, :1], ti2[:, :1])
    assert torch.equal(out1, outk1)


def _bits(t):
    return t.clamp(-448, 448).to(torch.float8_e4m3fn).view(torch.uint8)


def test_fused_experts_scheme_oracle_consistency():
    """the vectorised multi-scheme oracle equals the original per-token loop on the per-tensor scheme, and its
    block scheme equals w8a8_block_scaled_mm (native_w8a8_block_matmul restated) composed by hand."""
    g = torch.Generator().manual_seed(2)
    M, E, topk, H, I = 12, 4, 2, 256, 128
    x = (torch.randn(M, H, generator=g) * 0.5).to(torch.bfloat16)
    w1, w2 = _bits(torch.randn(E, 2 * I, H, generator=g)), _bits(torch.randn(E, H, I, generator=g))
    w1s, w2s = torch.rand(E, generator=g) * 0.05 + 0.01, torch.rand(E, generator=g) * 0.05 + 0.01
    tw, ti, _ = ref_ops.topk_softmax(torch.randn(M, E, generator=g), topk, True)
    a1, a2 = torch.tensor([0.01]), torch.tensor([0.02])
    old = ref_ops.fused_experts_fp8(x, w1, w2, tw, ti, w1s, w2s, a1, a2)
    new = ref_ops.fused_experts_fp8_scheme(x, w1, w2, tw, ti, w1s, w2s, "tensor", a1_scale=a1, a2_scale=a2)
    torch.testing.assert_close(new, old, rtol=1e-6, atol=1e-6)
    # block scheme, one token through one expert by hand
    w1b = torch.rand(E, 2 * I // 128, H // 128, generator=g) * 0.05 + 0.01
    w2b = torch.rand(E, H // 128, I // 128, generator=g) * 0.05 + 0.01
    ti1 = ti[:, :1].contiguous()
    tw1 = torch.ones(M, 1)
    got = ref_ops.fused_experts_fp8_scheme(x, w1, w2, tw1, ti1, w1b, w2b, "block")
    t = 5
    e = int(ti1[t, 0])
    q, s = ref_ops.per_token_group_quant_fp8(x[t:t + 1], 128)
    h = torch.from_numpy(ref_ops.w8a8_block_scaled_mm(q, w1[e].numpy(), s, w1b[e].numpy())).to(torch.bfloat16)
    act = ref_ops.silu_and_mul(h)
    q2, s2 = ref_ops.per_token_group_quant_fp8(act, 128)
    o = torch.from_numpy(ref_ops.w8a8_block_scaled_mm(q2, w2[e].numpy(), s2, w2b[e].numpy())).to(torch.bfloat16)
    torch.testing.assert_close(got[t], o[0].float(), rtol=1e-6, atol=1e-6)
