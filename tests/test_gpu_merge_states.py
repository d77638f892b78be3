"""GPU parity: merge_attn_states vs the oracle. Mirrors tests/kernels/attention/test_merge_attn_states.py
of the reference (NUM_TOKENS x HEAD sizes x dtypes, inf LSEs injected, prefill_tokens_with_context,
FP8 output with output_scale; tolerance there 1e-3 / 1e-2 by dtype)."""
import numpy as np
import pytest
import torch

from oracle import ref_ops

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _ops():
    from b200ops import ops
    return ops


@pytest.mark.parametrize("T", [1, 7, 256, 1031])
@pytest.mark.parametrize("H", [4, 32])
@pytest.mark.parametrize("D", [32, 128, 192])
@pytest.mark.parametrize("dtype", [torch.float32, torch.float16, torch.bfloat16])
@pytest.mark.parametrize("npre", [None, 0, "half"])
def test_merge_attn_states(T, H, D, dtype, npre):
    g = torch.Generator().manual_seed(T * 131 + H)
    po = torch.randn(T, H, D, generator=g).to(dtype)
    so = torch.randn(T, H, D, generator=g).to(dtype)
    pl = torch.randn(H, T, generator=g) * 3
    sl = torch.randn(H, T, generator=g) * 3
    # inject infinities like the reference test (+inf is treated as -inf; both -inf -> prefix copied)
    mask = torch.rand(H, T, generator=g)
    pl[mask < 0.1] = float("inf")
    sl[(mask > 0.05) & (mask < 0.15)] = float("-inf")
    n = None if npre is None else (0 if npre == 0 else T // 2)
    ref_out, ref_lse = ref_ops.merge_attn_states(po, pl, so, sl, n)
    out = torch.full((T, H, D), float("nan"), dtype=dtype, device=DEV)
    lse = torch.full((H, T), float("nan"), dtype=torch.float32, device=DEV)
    _ops().merge_attn_states(out, po.to(DEV), pl.to(DEV), so.to(DEV), sl.to(DEV), lse, n)
    tol = {torch.float32: 1e-5, torch.float16: 1e-3, torch.bfloat16: 8e-3}[dtype]
    np.testing.assert_allclose(out.float().cpu().numpy(), ref_out, atol=tol, rtol=tol)
    np.testing.assert_allclose(lse.cpu().numpy(), ref_lse, atol=1e-5, rtol=1e-5)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_merge_attn_states_fp8_output_strided_no_lse(dtype):
    T, H, D = 77, 8, 128
    g = torch.Generator().manual_seed(5)
    # inputs are head-slices of wider buffers: token stride != H * head stride
    pbuf = torch.randn(T, H, 2 * D, generator=g).to(dtype)
    sbuf = torch.randn(T, H + 2, 2 * D, generator=g).to(dtype)
    po, so = pbuf[:, :, :D], sbuf[:, 1:H + 1, :D]
    pl = torch.randn(H, T, generator=g)
    sl = torch.randn(H, T, generator=g)
    scale = torch.tensor([0.02])
    ref_bits, _ = ref_ops.merge_attn_states(po, pl, so, sl, 50, output_scale=scale)
    out = torch.zeros(T, H, D, dtype=torch.float8_e4m3fn, device=DEV)
    _ops().merge_attn_states(out, pbuf.to(DEV)[:, :, :D], pl.to(DEV), sbuf.to(DEV)[:, 1:H + 1, :D], sl.to(DEV),
                             None, 50, scale.to(DEV))
    got = out.view(torch.uint8).cpu().numpy()
    # expf / division differ by an ulp from numpy: allow a 1-code difference on < 0.5 % of the bytes
    diff = np.abs(got.astype(np.int16) - ref_bits.astype(np.int16))
    assert diff.max() <= 1 and (diff != 0).mean() < 5e-3
    with pytest.raises(AssertionError, match="output_scale is required"):
        _ops().merge_attn_states(out, pbuf.to(DEV)[:, :, :D], pl.to(DEV), sbuf.to(DEV)[:, 1:H + 1, :D],
                                 sl.to(DEV))
