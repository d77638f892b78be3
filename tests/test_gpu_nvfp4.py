"""GPU parity: scaled_fp4_quant (NVFP4) vs the oracle. Mirrors
tests/kernels/quantization/test_nvfp4_quant.py:143-308 of the reference (exact there)."""
import numpy as np
import pytest
import torch

from oracle import ref_ops

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
SHAPES = [(128, 64), (128, 128), (256, 64), (256, 128), (90, 64), (150, 128), (150, 160), (1, 4096),
          (256, 14336), (33, 2048)]


def _ops():
    from b200ops import ops
    return ops


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("swizzled", [True, False])
def test_scaled_fp4_quant(shape, dtype, swizzled):
    M, K = shape
    g = torch.Generator().manual_seed(42)
    x = torch.randn(M, K, generator=g).to(dtype)
    gs = np.float32(448.0 * 6.0) / np.float32(x.float().abs().max().item())
    codes, sf, y = ref_ops.nvfp4_quant(x, gs, return_scaled=True)
    out, out_sf = _ops().scaled_fp4_quant(x.to(DEV), torch.tensor([gs], device=DEV), swizzled)
    assert out.shape == (M, K // 2) and out.dtype == torch.uint8
    got_sf_flat = out_sf.view(torch.uint8).cpu().numpy().reshape(-1)
    if swizzled:
        rm, rc = (M + 127) // 128 * 128, (K // 16 + 3) // 4 * 4
        assert got_sf_flat.size == rm * rc
        # layout + zero padding are exact properties: compare the whole swizzled buffer
        exp_flat = ref_ops.swizzle_sf(sf)
        got_sf = ref_ops.unswizzle_sf(got_sf_flat, M, K // 16)
        pad_mask = ref_ops.swizzle_sf(np.ones_like(sf)) == 0
        assert (got_sf_flat[pad_mask] == 0).all()
    else:
        got_sf = got_sf_flat.reshape(M, K // 16)
    sf_mismatch = (got_sf != sf).mean()
    got_codes = ref_ops.nvfp4_unpack(out.cpu().numpy())
    code_mismatch = (got_codes != codes).mean()
    # rcp.approx.ftz (the reference kernel's instruction, nvfp4_utils.cuh:158-176) vs the oracle's exact
    # reciprocal can only flip a value sitting on a rounding boundary. Such ties are COMMON for bf16 data
    # when amax has a factor 21 (gs = 2688/amax becomes a power of two over 7), so the bound is not a
    # fraction but a location: every differing code must be a near-tie of the oracle's scaled value.
    assert sf_mismatch <= 1e-3, f"scale mismatch fraction {sf_mismatch}"
    same = np.repeat(got_sf == sf, 16, axis=1)
    bad = (got_codes != codes) & same & ~ref_ops.nvfp4_near_tie(y)
    assert not bad.any(), f"{bad.sum()} fp4 codes differ away from a rounding tie"
    same_sf = np.repeat(got_sf == sf, 16, axis=1)
    d = np.abs(ref_ops.e2m1_code_to_f32(got_codes) - ref_ops.e2m1_code_to_f32(codes))
    assert (d[same_sf] <= 2.0).all()  # never more than one e2m1 step
    print(f"nvfp4 {shape} {dtype} swz={swizzled}: sf mismatch {sf_mismatch:.2e}, code mismatch {code_mismatch:.2e}")


def test_scaled_fp4_quant_zero_block_and_validation():
    x = torch.zeros(4, 64, dtype=torch.bfloat16)
    x[1, 16:32] = torch.arange(16).to(torch.bfloat16) - 8
    out, sf = _ops().scaled_fp4_quant(x.to(DEV), torch.tensor([1.0], device=DEV), False)
    codes, rsf = ref_ops.nvfp4_quant(x, 1.0)
    np.testing.assert_array_equal(sf.view(torch.uint8).cpu().numpy(), rsf)
    np.testing.assert_array_equal(ref_ops.nvfp4_unpack(out.cpu().numpy()), codes)
    with pytest.raises(AssertionError, match="multiple of 16"):
        _ops().scaled_fp4_quant(torch.zeros(2, 24, dtype=torch.bfloat16, device=DEV),
                                torch.tensor([1.0], device=DEV))


# ------------------------------------------------------------------------------------------------
# cutlass_scaled_fp4_mm. Mirrors tests/kernels/quantization/test_nvfp4_scaled_mm.py:51-108 (shapes
# there are (m, n, packed_k); tolerance there atol = rtol = 1e-1 against a bf16 matmul of the
# dequantised operands). Here the oracle is the exact block-scaled product (float64 accumulate), so
# the only admissible error is fp32 accumulation order + output rounding: rtol 2^-7 (bf16 has 8
# significand bits; f16 passes the same bound), atol 1e-3 * |out|_max for cancellation.
# ------------------------------------------------------------------------------------------------
MM_SHAPES = [(128, 128, 64), (128, 128, 128), (256, 128, 64), (128, 256, 128), (150, 128, 64), (128, 128, 96),
             # decode / swap-AB buckets, K tails (k % 256 != 0), N tails, large-M orientation
             (1, 4096, 2048), (16, 6144, 2048), (33, 4096, 2048), (64, 256, 80), (256, 14336 // 8, 2048),
             (100, 160, 48), (300, 384, 64), (512, 512, 1024), (1000, 4096 + 128, 160), (257, 96, 16)]


def _quant_pair(m, n, k, dtype, seed=42):
    g = torch.Generator().manual_seed(seed)
    a = torch.randn(m, k, generator=g).to(dtype)
    b = torch.randn(n, k, generator=g).to(dtype)
    gsa = (448.0 * 6.0) / a.float().abs().max()
    gsb = (448.0 * 6.0) / b.float().abs().max()
    return a, b, gsa.reshape(1).float(), gsb.reshape(1).float()


@pytest.mark.parametrize("shape", MM_SHAPES)
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_cutlass_scaled_fp4_mm(shape, dtype):
    m, n, packed_k = shape
    k = packed_k * 2
    ops = _ops()
    a, b, gsa, gsb = _quant_pair(m, n, k, dtype)
    a4, sfa = ops.scaled_fp4_quant(a.to(DEV), gsa.to(DEV))
    b4, sfb = ops.scaled_fp4_quant(b.to(DEV), gsb.to(DEV))
    alpha = (1.0 / (gsa * gsb)).to(DEV)
    out = ops.cutlass_scaled_fp4_mm(a4, b4, sfa, sfb, alpha, dtype)
    assert out.shape == (m, n) and out.dtype == dtype
    # oracle on the SAME quantised bits the GPU produced (the quantiser has its own test above)
    ac = ref_ops.nvfp4_unpack(a4.cpu().numpy())
    bc = ref_ops.nvfp4_unpack(b4.cpu().numpy())
    sa = ref_ops.unswizzle_sf(sfa.view(torch.uint8).cpu().numpy().reshape(-1), m, k // 16)
    sb = ref_ops.unswizzle_sf(sfb.view(torch.uint8).cpu().numpy().reshape(-1), n, k // 16)
    ref = ref_ops.nvfp4_scaled_mm(ac, sa, bc, sb, float(alpha.item()))
    got = out.float().cpu().numpy()
    tol = 2.0 ** -7 * np.abs(ref) + 1e-3 * np.abs(ref).max()
    bad = np.abs(got - ref) > tol
    assert not bad.any(), f"{bad.sum()} / {bad.size} outside tolerance, max err {np.abs(got - ref).max():.4g}"
    # the reference test's own criterion (vs the dequantised matmul of the ORIGINAL data)
    torch.testing.assert_close(out.float().cpu(), torch.from_numpy(ref), atol=1e-1, rtol=1e-1)


def test_cutlass_scaled_fp4_mm_exact_small_integers():
    """all block scales = 1 (e4m3 0x38), codes in {0, +-1, +-2}: every product and partial sum is a
    small integer, so the result must be BIT-exact whatever the accumulation order."""
    _exact_case(48, 256, 512)
    _exact_case(256, 512, 1024)    # several narrow batch tiles per weight slab (scale-column offset path)
    _exact_case(200, 384, 256)


def _exact_case(m, n, k):
    rng = np.random.default_rng(0)
    vals = np.array([0, 2, 4, 10, 12], dtype=np.uint8)  # e2m1 codes of 0, 1, 2, -1, -2
    ac = vals[rng.integers(0, 5, (m, k))]
    bc = vals[rng.integers(0, 5, (n, k))]
    one = np.full((m, k // 16), 0x38, dtype=np.uint8)
    oneb = np.full((n, k // 16), 0x38, dtype=np.uint8)
    a4 = torch.from_numpy(ref_ops.nvfp4_pack(ac)).to(DEV)
    b4 = torch.from_numpy(ref_ops.nvfp4_pack(bc)).to(DEV)
    sfa = torch.from_numpy(ref_ops.swizzle_sf(one)).to(DEV).view(torch.float8_e4m3fn)
    sfb = torch.from_numpy(ref_ops.swizzle_sf(oneb)).to(DEV).view(torch.float8_e4m3fn)
    out = _ops().cutlass_scaled_fp4_mm(a4, b4, sfa, sfb, torch.ones(1, device=DEV), torch.float16)
    ref = ref_ops.nvfp4_scaled_mm(ac, one, bc, oneb, 1.0)
    assert np.abs(ref).max() < 2048  # exactly representable in f16
    np.testing.assert_array_equal(out.float().cpu().numpy(), ref)


def test_cutlass_scaled_fp4_mm_validation():
    ops = _ops()
    a = torch.zeros(4, 24, dtype=torch.uint8, device=DEV)  # k = 48: not divisible by 32
    b = torch.zeros(32, 24, dtype=torch.uint8, device=DEV)
    sf = torch.zeros(128, 4, dtype=torch.uint8, device=DEV).view(torch.float8_e4m3fn)
    with pytest.raises(RuntimeError, match="divisible by 32"):
        ops.cutlass_scaled_fp4_mm(a, b, sf, sf, torch.ones(1, device=DEV), torch.bfloat16)
    a = torch.zeros(4, 32, dtype=torch.uint8, device=DEV)
    b = torch.zeros(32, 32, dtype=torch.uint8, device=DEV)
    with pytest.raises(RuntimeError, match="padded and swizzled"):
        ops.cutlass_scaled_fp4_mm(a, b, sf[:64], sf, torch.ones(1, device=DEV), torch.bfloat16)


# ------------------------------------------------------------------------------------------------
# silu_and_mul_nvfp4_quant. Mirrors tests/kernels/quantization/test_silu_mul_nvfp4_quant.py:31-80
# (there: dequantised fused vs unfused within atol = rtol = 3e-1).
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("shape", [(128, 256), (128, 128), (256, 256), (256, 128), (33, 14336 * 2 // 8), (1, 64)])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_silu_and_mul_nvfp4_quant(shape, dtype):
    m, two_d = shape
    d = two_d // 2
    g = torch.Generator().manual_seed(42)
    x = torch.randn(m, two_d, generator=g).to(dtype)
    gate, up = x[:, :d].float(), x[:, d:].float()
    act = (gate / (1.0 + torch.exp(-gate)) * up).to(dtype).float()
    gs = np.float32(448.0 * 6.0) / np.float32(act.abs().max().item())
    codes, sf, y = ref_ops.silu_and_mul_nvfp4_quant(x, gs, return_scaled=True)
    ops = _ops()
    out = torch.empty(m, d // 2, dtype=torch.uint8, device=DEV)
    out_sf = ops.create_fp4_scale_tensor(m, d, DEV, True)
    ops.silu_and_mul_nvfp4_quant(out, out_sf, x.to(DEV), torch.tensor([gs], device=DEV))
    got_sf_flat = out_sf.view(torch.uint8).cpu().numpy().reshape(-1)
    got_sf = ref_ops.unswizzle_sf(got_sf_flat, m, d // 16)
    pad_mask = ref_ops.swizzle_sf(np.ones_like(sf)) == 0
    assert (got_sf_flat[pad_mask] == 0).all()
    got_codes = ref_ops.nvfp4_unpack(out.cpu().numpy())
    # __expf/__fdividef (the reference's own silu, nvfp4_utils.cuh:304-307) vs exact exp may move the
    # activation by one ulp of the 16-bit dtype on a few elements; bound those, and everything else
    # must agree away from e2m1 rounding ties.
    assert (got_sf != sf).mean() <= 2e-3
    same = np.repeat(got_sf == sf, 16, axis=1)
    bad = (got_codes != codes) & same & ~ref_ops.nvfp4_near_tie(y)
    assert bad.mean() <= 1e-3, f"{bad.sum()} fp4 codes differ away from a rounding tie"
    # never more than one e2m1 step; and the reference test's criterion on dequantised values
    # (atol = rtol = 3e-1) for everything that is not a rounding tie
    stepd = np.abs(ref_ops.e2m1_code_to_f32(got_codes) - ref_ops.e2m1_code_to_f32(codes))
    assert (stepd[same] <= 2.0).all()
    deq_ref = ref_ops.nvfp4_dequant(codes, sf, gs)
    deq_got = ref_ops.nvfp4_dequant(got_codes, got_sf, gs)
    far = np.abs(deq_got - deq_ref) > 3e-1 + 3e-1 * np.abs(deq_ref)
    assert (far & ~ref_ops.nvfp4_near_tie(y)).mean() <= 1e-4
