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
