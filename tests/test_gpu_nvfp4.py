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
    codes, sf = ref_ops.nvfp4_quant(x, gs)
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
    # rcp.approx vs exact reciprocal can only flip a value sitting exactly on a rounding boundary
    assert sf_mismatch <= 1e-3, f"scale mismatch fraction {sf_mismatch}"
    assert code_mismatch <= 2e-3, f"fp4 code mismatch fraction {code_mismatch}"
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
