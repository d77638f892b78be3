"""GPU parity: sampler-side kernels (greedy argmax, top-k / top-p masking, repetition penalties) vs the oracle
restatement of the reference's sort-based definition (topk_topp_sampler.py:367-408; the reference's own test,
tests/v1/sample/test_topk_topp_sampler.py, compares its other implementations against that function).
Bit-exact for the kept logits and the argmax; the mask may differ only on the two boundary cases the sort-based
definition itself leaves to the sort order: ties of the top-p boundary value, and a cumulative mass within float
rounding of 1 - p."""
import numpy as np
import pytest
import torch

from oracle import ref_ops

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _ops():
    from b200ops import ops
    return ops


@pytest.mark.parametrize("rows,vocab", [(1, 17), (7, 1000), (256, 32000), (33, 128256), (5, 151936)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
def test_greedy_sample(rows, vocab, dtype):
    g = torch.Generator().manual_seed(rows + vocab)
    x = (torch.randn(rows, vocab, generator=g) * 4).to(dtype)   # 16-bit logits: plenty of exact ties
    x[0, vocab // 2] = x[0].max()                                  # forced tie: the lower index wins
    want = x.float().argmax(-1)
    got = _ops().greedy_sample(x.to(DEV))
    assert got.dtype == torch.int64
    assert torch.equal(got.cpu(), want)
    if rows > 2:
        x[2, 3] = float("nan")
        assert _ops().greedy_sample(x.to(DEV))[2].item() == 3


def _check_mask(got, ref, info, x, k, p):
    got, ref = got.cpu(), ref
    gk, rk = ~torch.isinf(got), ~torch.isinf(ref)
    assert torch.equal(got[gk], x[gk])                     # survivors are untouched, bit for bit
    diff = gk != rk
    if p is not None and diff.any():
        # legitimate differences: cumulative mass within rounding of 1 - p, or a tie of the boundary value
        cum = info["cum"]
        near = (cum - (1 - p.unsqueeze(1))).abs() <= 2e-6 + 1e-5 * (1 - p.unsqueeze(1))
        for r in diff.any(1).nonzero().flatten().tolist():
            vals = x[r][diff[r]]
            # a tie group (the same value several times in the row) is split by the sort order in the reference
            in_tie = torch.tensor([(x[r] == v).sum().item() >= 2 for v in vals.tolist()])
            assert (near[r][diff[r]] | in_tie).all(), (r, vals[:8])
        assert diff.float().mean() < 1e-3
    else:
        assert not diff.any()
    assert gk.any(1).all()                                   # at least one token survives in every row


@pytest.mark.parametrize("rows,vocab", [(3, 64), (16, 1000), (64, 32000), (9, 128256)])
@pytest.mark.parametrize("mode", ["k", "p", "kp"])
@pytest.mark.parametrize("kdtype", [torch.int32, torch.int64])
def test_apply_top_k_top_p(rows, vocab, mode, kdtype):
    g = torch.Generator().manual_seed(rows * 7 + vocab)
    x = (torch.randn(rows, vocab, generator=g) * 3).to(torch.bfloat16).float()   # bf16-valued: many ties
    k = p = None
    if "k" in mode:
        k = torch.randint(1, min(vocab, 200), (rows,), generator=g).to(kdtype)
        k[0] = vocab          # "no top-k" row
        k[-1] = 1
    if "p" in mode:
        p = torch.rand(rows, generator=g) * 0.98 + 0.01
        p[0] = 1.0
    ref, info = ref_ops.apply_top_k_top_p(x, k, p)
    out = _ops().apply_top_k_top_p(x.to(DEV).clone(), None if k is None else k.to(DEV), None if p is None else p.to(DEV))
    _check_mask(out, ref, info, x, k, p)
    if mode == "k":   # top-k alone is exact: everything >= the k-th largest value survives
        keep = x >= info["kth"].unsqueeze(1)
        keep[0] = True
        assert torch.equal(~torch.isinf(out.cpu()), keep)


def test_apply_top_k_top_p_deterministic_and_noop():
    ops = _ops()
    g = torch.Generator().manual_seed(5)
    x = torch.randn(32, 50000, generator=g).to(DEV)
    p = (torch.rand(32, generator=g) * 0.9 + 0.05).to(DEV)
    k = torch.randint(1, 100, (32,), generator=g).to(DEV)
    a = ops.apply_top_k_top_p(x.clone(), k, p)
    b = ops.apply_top_k_top_p(x.clone(), k, p)
    assert torch.equal(a, b)
    assert ops.apply_top_k_top_p(x, None, None) is x
    with pytest.raises(RuntimeError, match="float32"):
        ops.apply_top_k_top_p(x.to(torch.bfloat16), k, p)


@pytest.mark.parametrize("rows,vocab", [(1, 17), (8, 1000), (200, 32000), (3, 128256)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
def test_apply_repetition_penalties(rows, vocab, dtype):
    g = torch.Generator().manual_seed(vocab)
    x = (torch.randn(rows, vocab, generator=g) * 5).to(dtype)
    pm = torch.rand(rows, vocab, generator=g) < 0.1
    om = torch.rand(rows, vocab, generator=g) < 0.1
    pen = (torch.rand(rows, generator=g) + 0.6).to(dtype)
    want = ref_ops.apply_repetition_penalties(x, pm, om, pen)
    xg = x.to(DEV).clone()
    _ops().apply_repetition_penalties(xg, pm.to(DEV), om.to(DEV), pen.to(DEV))
    assert torch.equal(xg.cpu(), want)
