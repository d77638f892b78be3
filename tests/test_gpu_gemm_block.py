"""GPU parity: block-scaled FP8 GEMM (1 x 128 activation scales, 128 x 128 weight scales) behind
cutlass_scaled_mm, vs the oracle restatement of native_w8a8_block_matmul.
Mirrors tests/kernels/quantization/test_block_fp8.py:161-201 (reference bound: mean rel diff < 1e-3; here the
per-element error is bounded by the output rounding, and an integer-valued case is bit-exact)."""
import numpy as np
import pytest
import torch

from oracle import ref_ops

pytestmark = pytest.mark.gpu
DEV = "cuda:0"

# (M, N, K): the reference's DSV3 kv_a_proj case (N % 128 != 0), its M/N/K grid, decode / prefill tiles,
# ragged M, K not a multiple of 128, several M tiles
MNK = [(32, 576, 7168), (1, 128, 128), (7, 512, 512), (16, 1536, 7168), (17, 2112, 7168), (33, 4096, 512),
       (64, 7168, 2048), (83, 24576, 1536), (128, 4096, 4096), (129, 2048, 7168), (300, 1280, 8192),
       (1024, 1024, 1024), (2050, 576, 512), (48, 640, 1040), (5, 144, 272)]


def _ops():
    from b200ops import ops
    return ops


def _inputs(M, N, K, seed):
    g = torch.Generator().manual_seed(seed)
    a8 = (torch.randn(M, K, generator=g)).clamp(-448, 448).to(torch.float8_e4m3fn)
    w8 = (torch.randn(N, K, generator=g)).clamp(-448, 448).to(torch.float8_e4m3fn)
    kb, nb = (K + 127) // 128, (N + 127) // 128
    a_s = (torch.rand(M, kb, generator=g) + 0.25) * 0.02
    b_s = (torch.rand(nb, kb, generator=g) + 0.25) * 0.02
    return a8, w8, a_s, b_s


@pytest.mark.parametrize("mnk", MNK)
@pytest.mark.parametrize("a_layout", ["col", "row"])
@pytest.mark.parametrize("out_dtype", [torch.bfloat16, torch.float16])
def test_block_scaled_mm(mnk, a_layout, out_dtype):
    M, N, K = mnk
    a8, w8, a_s, b_s = _inputs(M, N, K, seed=M + N + K)
    ref = ref_ops.w8a8_block_scaled_mm(a8.view(torch.uint8).numpy(), w8.view(torch.uint8).numpy(),
                                       a_s.numpy(), b_s.numpy())
    as_g = a_s.to(DEV)
    if a_layout == "col":   # CUTLASS wants column-major activation scales (test_block_fp8.py:190-193)
        as_g = as_g.t().contiguous().t()
    out = _ops().w8a8_block_fp8_cutlass_mm(a8.to(DEV), w8.to(DEV), as_g, b_s.to(DEV), [128, 128], out_dtype)
    torch.cuda.synchronize()
    got = out.float().cpu().numpy()
    assert got.shape == (M, N)
    rtol = 2 ** -8 if out_dtype == torch.bfloat16 else 2 ** -10
    np.testing.assert_allclose(got, ref, rtol=rtol, atol=1e-4)
    # the reference test's own bound: both sides rounded to the output dtype (test_block_fp8.py:195-201)
    ref_r = torch.from_numpy(ref).to(out_dtype).float().numpy()
    rel = np.abs(got - ref_r).mean() / np.abs(ref_r).mean()
    assert rel < 1e-3


@pytest.mark.parametrize("impl", [None, "split1", "split2", "split4", "split8"])
@pytest.mark.parametrize("mnk", [(16, 256, 512), (100, 384, 1024), (256, 1152, 768), (32, 512, 2048), (64, 1280, 3584)])
def test_block_scaled_mm_bit_exact_integers(mnk, impl):
    """Small-integer operands and power-of-two scales: every partial sum is exact in fp32, so the kernel
    must reproduce the float64 oracle bit for bit (catches any mis-indexed scale or k-block)."""
    M, N, K = mnk
    g = torch.Generator().manual_seed(7)
    a8 = torch.randint(-4, 5, (M, K), generator=g).float().to(torch.float8_e4m3fn)
    w8 = torch.randint(-3, 4, (N, K), generator=g).float().to(torch.float8_e4m3fn)
    kb, nb = K // 128, (N + 127) // 128
    a_s = torch.exp2(torch.randint(-3, 2, (M, kb), generator=g).float())
    b_s = torch.exp2(torch.randint(-2, 2, (nb, kb), generator=g).float())
    ref = ref_ops.w8a8_block_scaled_mm(a8.view(torch.uint8).numpy(), w8.view(torch.uint8).numpy(),
                                       a_s.numpy(), b_s.numpy())
    # impl bounds the split-K of the kernel (K slices over a thread-block cluster, DSMEM reduction): every split
    # must give the same bits
    out = _ops().cutlass_scaled_mm(a8.to(DEV), w8.to(DEV).t(), a_s.to(DEV), b_s.to(DEV).t(), torch.float16, impl=impl)
    assert np.abs(ref).max() < 2048 * 16   # representable range check for the exactness argument
    want = torch.from_numpy(ref).to(torch.float16)
    assert torch.equal(out.cpu(), want)


def test_block_scaled_mm_with_group_quant_pipeline():
    """per_token_group_quant_fp8 -> blockwise GEMM, the pair a block-FP8 linear runs (fp8_utils.py:apply_w8a8_block_fp8_linear)."""
    ops = _ops()
    M, N, K = 37, 768, 2048
    g = torch.Generator().manual_seed(11)
    x = (torch.randn(M, K, generator=g)).to(torch.bfloat16)
    w8 = (torch.randn(N, K, generator=g)).clamp(-448, 448).to(torch.float8_e4m3fn)
    b_s = (torch.rand(N // 128, K // 128, generator=g) + 0.25) * 0.02
    bits, a_s = ref_ops.per_token_group_quant_fp8(x, 128)
    ref = ref_ops.w8a8_block_scaled_mm(bits, w8.view(torch.uint8).numpy(), a_s, b_s.numpy())
    xq, xs = ops.per_token_group_quant_fp8(x.to(DEV), 128, column_major_scales=True)
    out = ops.w8a8_block_fp8_cutlass_mm(xq, w8.to(DEV), xs, b_s.to(DEV), [128, 128], torch.bfloat16)
    np.testing.assert_allclose(out.float().cpu().numpy(), ref, rtol=2 ** -8, atol=1e-4)


def test_block_scaled_mm_errors():
    ops = _ops()
    a8, w8, a_s, b_s = _inputs(16, 256, 256, 0)
    a8, w8, a_s, b_s = a8.to(DEV), w8.to(DEV), a_s.to(DEV), b_s.to(DEV)
    with pytest.raises(RuntimeError, match=r"a_scale_group_shape must be \[1, 128\]"):
        ops.cutlass_scaled_mm(a8, w8.t(), a_s[:, :1].repeat(1, 3), b_s.t(), torch.bfloat16)
    with pytest.raises(RuntimeError, match=r"b_scale_group_shape must be \[128, 128\]"):
        ops.cutlass_scaled_mm(a8, w8.t(), a_s, b_s.t()[:, :1].repeat(1, 3), torch.bfloat16)
    with pytest.raises(RuntimeError, match="Bias not yet supported blockwise"):
        ops.cutlass_scaled_mm(a8, w8.t(), a_s, b_s.t(), torch.bfloat16,
                              bias=torch.zeros(256, dtype=torch.bfloat16, device=DEV))
    out = ops.cutlass_scaled_mm(a8[:0], w8.t(), a_s[:0], b_s.t(), torch.bfloat16)
    assert out.shape == (0, 256)
