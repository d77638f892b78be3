"""CPU tests: MoE routing oracle pinned to the reference's worked example and its golden properties."""
import numpy as np
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
