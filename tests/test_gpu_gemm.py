"""GPU parity: cutlass_scaled_mm (FP8 W8A8 GEMM, tcgen05) and silu_and_mul(_quant) vs the oracle.
Mirrors tests/kernels/quantization/test_cutlass_scaled_mm.py:82-230 (reference tolerance
rtol=5e-1 atol=1.5e-1; ours is far tighter: the only error is the bf16/f16 output rounding)."""
import numpy as np
import pytest
import torch

from oracle import ref_ops

pytestmark = pytest.mark.gpu
DEV = "cuda:0"

MNK = [(1, 256, 128), (1, 16384, 1024), (1, 24576, 496), (16, 256, 496), (16, 16384, 128),
       (16, 24576, 4096), (32, 8192, 4096), (32, 16384, 4096), (33, 1024, 1024), (33, 8192, 128),
       (64, 2048, 496), (64, 16384, 1024), (100, 8192, 496), (128, 32768, 4096), (256, 4096, 4096),
       (512, 256, 1024), (512, 8192, 4096), (512, 16384, 128), (512, 24576, 128),
       (256, 6144, 4096), (256, 4096, 14336), (300, 1280, 8192), (2048, 2048, 2048)]


def _ops():
    from b200ops import ops
    return ops


def to_fp8_bits(t):
    return t.clamp(-448, 448).to(torch.float8_e4m3fn).view(torch.uint8)


IMPLS = ["v2", "v3"]
V3_MNK = [(1, 16384, 1024), (33, 1024, 1024), (128, 32768, 4096), (256, 4096, 4096), (256, 4096, 14336),
          (300, 1280, 8192), (2048, 2048, 2048), (512, 24576, 128)]


def run_case(M, N, K, per_token, per_channel, out_dtype, use_bias, seed=0, impl=None):
    g = torch.Generator().manual_seed(seed)
    a8 = to_fp8_bits(torch.randn(M, K, generator=g))
    w8 = to_fp8_bits(torch.randn(N, K, generator=g))
    sa = (torch.rand(M if per_token else 1, generator=g) + 0.5) * 0.05
    sb = (torch.rand(N if per_channel else 1, generator=g) + 0.5) * 0.05
    bias = (torch.rand(N, generator=g) * 5 - 2.5).to(out_dtype) if use_bias else None
    ref = ref_ops.scaled_mm(a8.numpy(), w8.numpy(), sa.numpy(), sb.numpy(),
                            None if bias is None else bias.float().numpy())
    a_g = a8.to(DEV).view(torch.float8_e4m3fn)
    b_g = w8.to(DEV).view(torch.float8_e4m3fn).t()  # [K, N] column-major
    out = _ops().cutlass_scaled_mm(a_g, b_g, sa.to(DEV).view(-1, 1) if per_token else sa.to(DEV),
                                   sb.to(DEV).view(-1, 1) if per_channel else sb.to(DEV), out_dtype,
                                   None if bias is None else bias.to(DEV), impl=impl)
    torch.cuda.synchronize()
    got = out.float().cpu().numpy()
    assert got.shape == (M, N)
    # output rounding: bf16 2^-9, f16 2^-12 relative (+ the bias rounding step)
    rtol = 2 ** -8 if out_dtype == torch.bfloat16 else 2 ** -10
    np.testing.assert_allclose(got, ref, rtol=rtol, atol=2e-2 if use_bias else 2e-3)


@pytest.mark.parametrize("impl", IMPLS)
@pytest.mark.parametrize("mnk", MNK)
@pytest.mark.parametrize("per_token,per_channel", [(False, False), (True, True)])
def test_scaled_mm_shapes(mnk, per_token, per_channel, impl):
    if impl == "v3" and (per_token or mnk not in V3_MNK):
        pytest.skip("v3 (non-default kernel) runs a thinned grid")
    run_case(*mnk, per_token, per_channel, torch.bfloat16, False, impl=impl)


@pytest.mark.parametrize("mnk", [(1, 256, 128), (33, 1024, 1024), (256, 4096, 4096), (512, 512, 512)])
@pytest.mark.parametrize("per_token,per_channel", [(True, False), (False, True)])
@pytest.mark.parametrize("out_dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("use_bias", [False, True])
@pytest.mark.parametrize("impl", IMPLS)
def test_scaled_mm_scales_bias_dtypes(mnk, per_token, per_channel, out_dtype, use_bias, impl):
    run_case(*mnk, per_token, per_channel, out_dtype, use_bias, seed=1, impl=impl)


@pytest.mark.parametrize("impl", IMPLS)
def test_scaled_mm_exact_small_integers(impl):
    """Integer-valued operands, power-of-two scales: every product and sum is exact in fp32 and in
    bf16 -> the GEMM must be BIT-exact (catches descriptor / swizzle / k-offset mistakes)."""
    g = torch.Generator().manual_seed(3)
    # the last shapes cut tiles across CTA pairs (stream-K partials) in the v3 kernel
    for M, N, K in [(7, 128, 128), (130, 384, 256), (256, 256, 1024), (513, 512, 384), (256, 4096, 4096),
                    (64, 1024, 14336), (1024, 1280, 2048)]:
        a = torch.randint(-2, 3, (M, K), generator=g).float()
        w = torch.randint(-2, 3, (N, K), generator=g).float()
        ref = (a.double() @ w.double().t()).float() * 0.25
        out = _ops().cutlass_scaled_mm(to_fp8_bits(a).to(DEV).view(torch.float8_e4m3fn),
                                       to_fp8_bits(w).to(DEV).view(torch.float8_e4m3fn).t(),
                                       torch.tensor([0.5], device=DEV), torch.tensor([0.5], device=DEV),
                                       torch.bfloat16, impl=impl)
        assert torch.equal(out.float().cpu(), ref.to(torch.bfloat16).float()), (M, N, K)


@pytest.mark.parametrize("impl", IMPLS)
def test_scaled_mm_strided_a_and_out_validation(impl):
    g = torch.Generator().manual_seed(4)
    M, N, K = 48, 256, 512
    wide = to_fp8_bits(torch.randn(M, 2 * K, generator=g))
    w8 = to_fp8_bits(torch.randn(N, K, generator=g))
    a_view = wide.to(DEV).view(torch.float8_e4m3fn)[:, K:]
    one = torch.ones(1, device=DEV)
    out = _ops().cutlass_scaled_mm(a_view, w8.to(DEV).view(torch.float8_e4m3fn).t(), one, one,
                                   torch.float16, impl=impl)
    ref = ref_ops.scaled_mm(wide[:, K:].numpy(), w8.numpy(), [1.0], [1.0])
    np.testing.assert_allclose(out.float().cpu().numpy(), ref, rtol=2 ** -10, atol=2e-3)
    with pytest.raises(RuntimeError, match="multiples of 16"):
        _ops().cutlass_scaled_mm(a_view[:, :500], w8.to(DEV).view(torch.float8_e4m3fn)[:, :500].t(),
                                 one, one, torch.float16, impl=impl)


@pytest.mark.parametrize("tokens,d", [(1, 128), (7, 17), (83, 14336), (256, 14336), (4096, 512)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_silu_and_mul(tokens, d, dtype):
    g = torch.Generator().manual_seed(5)
    x = (torch.randn(tokens, 2 * d, generator=g) * 2).to(dtype)
    ref = ref_ops.silu_and_mul(x)
    out = torch.empty(tokens, d, dtype=dtype, device=DEV)
    _ops().silu_and_mul(out, x.to(DEV))
    # expf differs by <= 1 ulp(fp32) between libm and CUDA -> at most 1 ulp of the output dtype
    ulp = 2 ** -7 if dtype == torch.bfloat16 else 2 ** -10
    np.testing.assert_allclose(out.float().cpu().numpy(), ref.float().numpy(), rtol=ulp, atol=1e-6)
    assert (out.cpu() == ref).float().mean() > 0.999


@pytest.mark.parametrize("tokens,d", [(5, 128), (256, 14336), (33, 200)])
def test_silu_and_mul_quant(tokens, d):
    g = torch.Generator().manual_seed(6)
    x = (torch.randn(tokens, 2 * d, generator=g) * 2).to(torch.bfloat16)
    scale = torch.tensor([0.02])
    ref = ref_ops.silu_and_mul_quant(x, scale)
    out = torch.empty(tokens, d, dtype=torch.float8_e4m3fn, device=DEV)
    _ops().silu_and_mul_quant(out, x.to(DEV), scale.to(DEV))
    got = out.view(torch.uint8).cpu().numpy()
    # identical except where the 1-ulp expf difference flips a rounding
    assert (got == ref).mean() > 0.999
    d1 = ref_ops.e4m3_bits_to_f32(got) - ref_ops.e4m3_bits_to_f32(ref)
    assert np.abs(d1).max() <= 32.0  # never more than one fp8 step (largest step = 32 at 256..448)


@pytest.mark.parametrize("M,I,K", [(256, 14336, 4096), (1, 256, 128), (33, 1792, 512), (128, 3584, 8192),
                                   (200, 128, 496), (64, 4864 // 128 * 128, 896)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("per_token,per_channel,use_bias", [(False, False, False), (True, True, True)])
def test_scaled_mm_silu_and_mul_quant_equals_unfused(M, I, K, dtype, per_token, per_channel, use_bias):
    """gate_up GEMM + silu_and_mul + FP8 quant in one kernel == cutlass_scaled_mm -> silu_and_mul_quant,
    BIT for bit (both halves of the pair are oracle-checked above)."""
    g = torch.Generator().manual_seed(M + I)
    a8 = to_fp8_bits(torch.randn(M, K, generator=g)).to(DEV).view(torch.float8_e4m3fn)
    w8 = to_fp8_bits(torch.randn(2 * I, K, generator=g)).to(DEV).view(torch.float8_e4m3fn)
    sa = ((torch.rand(M if per_token else 1, generator=g) + 0.5) * 0.03).to(DEV)
    sb = ((torch.rand(2 * I if per_channel else 1, generator=g) + 0.5) * 0.03).to(DEV)
    bias = (torch.rand(2 * I, generator=g) - 0.5).to(dtype).to(DEV) if use_bias else None
    qs = torch.tensor([0.05], device=DEV)
    ops = _ops()
    gu = ops.cutlass_scaled_mm(a8, w8.t(), sa.view(-1, 1) if per_token else sa, sb.view(-1, 1) if per_channel else sb,
                               dtype, bias, impl="v2")
    ref = torch.empty(M, I, dtype=torch.float8_e4m3fn, device=DEV)
    ops.silu_and_mul_quant(ref, gu, qs)
    out = torch.full((M, I), 0x7F, dtype=torch.uint8, device=DEV).view(torch.float8_e4m3fn)
    ops.scaled_mm_silu_and_mul_quant(out, a8, w8, sa, sb, qs, dtype, bias)
    torch.cuda.synchronize()
    assert torch.equal(out.view(torch.uint8).cpu(), ref.view(torch.uint8).cpu())


def test_scaled_mm_silu_and_mul_quant_validation():
    ops = _ops()
    one = torch.ones(1, device=DEV)
    a = torch.zeros(300, 128, dtype=torch.uint8, device=DEV).view(torch.float8_e4m3fn)
    w = torch.zeros(256, 128, dtype=torch.uint8, device=DEV).view(torch.float8_e4m3fn)
    with pytest.raises(NotImplementedError, match="M <= 256"):
        ops.scaled_mm_silu_and_mul_quant(torch.empty(300, 128, dtype=torch.float8_e4m3fn, device=DEV), a, w, one, one, one)
    w2 = torch.zeros(192, 128, dtype=torch.uint8, device=DEV).view(torch.float8_e4m3fn)
    with pytest.raises(NotImplementedError, match="multiple of 128"):
        ops.scaled_mm_silu_and_mul_quant(torch.empty(8, 96, dtype=torch.float8_e4m3fn, device=DEV), a[:8], w2, one, one, one)


# ------------------------------------------------------------------ v4: cluster split-K over DSMEM (decode batches)
V4_MNK = [(1, 256, 128), (16, 4096, 4096), (33, 1024, 1024), (64, 4096, 14336), (100, 6144, 4096),
          (128, 8192, 8192), (200, 272, 512), (256, 4096, 4096), (256, 6144, 4096), (256, 4096, 14336)]


@pytest.mark.parametrize("split", [0, 1, 2, 3, 4])
@pytest.mark.parametrize("mnk", V4_MNK)
def test_splitk_shapes(mnk, split):
    M, N, K = mnk
    if split > (K + 127) // 128:
        pytest.skip("more K slices than k-blocks")
    run_case(M, N, K, False, False, torch.bfloat16, False, impl="v4" if split == 0 else f"v4:{split}")


@pytest.mark.parametrize("mnk", [(33, 1024, 1024), (256, 4096, 4096)])
@pytest.mark.parametrize("per_token,per_channel", [(True, False), (False, True), (True, True)])
@pytest.mark.parametrize("out_dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("use_bias", [False, True])
def test_splitk_scales_bias_dtypes(mnk, per_token, per_channel, out_dtype, use_bias):
    run_case(*mnk, per_token, per_channel, out_dtype, use_bias, seed=1, impl="v4")


def test_splitk_exact_and_equal_to_v2():
    """(a) integer-valued operands: bit-exact whatever the split; (b) random operands: the K slices are summed
    in a different order than the unsplit kernel, so only output-rounding-level differences vs v2 are allowed
    and split = 1 must be bit-identical to v2."""
    g = torch.Generator().manual_seed(3)
    one = torch.tensor([0.5], device=DEV)
    for M, N, K in [(7, 256, 128), (130, 384, 512), (256, 256, 1024), (256, 4096, 4096), (64, 1024, 14336)]:
        a = torch.randint(-2, 3, (M, K), generator=g).float()
        w = torch.randint(-2, 3, (N, K), generator=g).float()
        ref = ((a.double() @ w.double().t()).float() * 0.25).to(torch.bfloat16).float()
        ag = to_fp8_bits(a).to(DEV).view(torch.float8_e4m3fn)
        wg = to_fp8_bits(w).to(DEV).view(torch.float8_e4m3fn).t()
        for split in (0, 1, 2, 3, 4):
            if split > K // 128:
                continue
            out = _ops().cutlass_scaled_mm(ag, wg, one, one, torch.bfloat16, impl="v4" if split == 0 else f"v4:{split}")
            assert torch.equal(out.float().cpu(), ref), (M, N, K, split)
    a8 = to_fp8_bits(torch.randn(256, 4096, generator=g)).to(DEV).view(torch.float8_e4m3fn)
    w8 = to_fp8_bits(torch.randn(6144, 4096, generator=g)).to(DEV).view(torch.float8_e4m3fn).t()
    sa, sb = torch.tensor([0.03], device=DEV), torch.tensor([0.02], device=DEV)
    o2 = _ops().cutlass_scaled_mm(a8, w8, sa, sb, torch.bfloat16, impl="v2")
    o1 = _ops().cutlass_scaled_mm(a8, w8, sa, sb, torch.bfloat16, impl="v4:1")
    o3 = _ops().cutlass_scaled_mm(a8, w8, sa, sb, torch.bfloat16, impl="v4:3")
    assert torch.equal(o1, o2)
    assert (o3.float() - o2.float()).abs().max().item() <= 2 ** -7 * o2.float().abs().max().item()
    assert (o3 == o2).float().mean().item() > 0.98


def test_splitk_rejects_large_batches():
    x = torch.zeros(300, 256, dtype=torch.uint8, device=DEV).view(torch.float8_e4m3fn)
    w = torch.zeros(256, 256, dtype=torch.uint8, device=DEV).view(torch.float8_e4m3fn).t()
    one = torch.ones(1, device=DEV)
    with pytest.raises(RuntimeError, match="decode batches only"):
        _ops().cutlass_scaled_mm(x, w, one, one, torch.bfloat16, impl="v4")


# ------------------------------------------------------------------ v5: CTA pairs of UMMA M = 128 (decode batches)
V5_MNK = [(1, 256, 128), (16, 4096, 4096), (33, 1024, 1024), (64, 4096, 14336), (100, 6144, 4096),
          (128, 8192, 1024), (128, 7168, 8192), (200, 272, 512), (256, 4096, 4096), (256, 6144, 4096),
          (256, 4096, 14336), (250, 1296, 496), (129, 64, 128)]


@pytest.mark.parametrize("bt", [0, 32, 64, 128, 256])
@pytest.mark.parametrize("mnk", V5_MNK)
def test_pair_shapes(mnk, bt):
    run_case(*mnk, False, False, torch.bfloat16, False, impl="v5" if bt == 0 else f"v5:{bt}")


@pytest.mark.parametrize("mnk", [(33, 1024, 1024), (256, 4096, 4096), (200, 272, 512)])
@pytest.mark.parametrize("per_token,per_channel", [(True, False), (False, True), (True, True)])
@pytest.mark.parametrize("out_dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("use_bias", [False, True])
def test_pair_scales_bias_dtypes(mnk, per_token, per_channel, out_dtype, use_bias):
    run_case(*mnk, per_token, per_channel, out_dtype, use_bias, seed=1, impl="v5")


def test_pair_exact_and_bit_identical_to_v2():
    """(a) integer-valued operands: bit-exact for every batch tile; (b) random operands: same k order and the same
    epilogue arithmetic as the one-CTA kernel, so the outputs are bit-identical to v2."""
    g = torch.Generator().manual_seed(4)
    one = torch.tensor([0.5], device=DEV)
    for M, N, K in [(7, 256, 128), (130, 384, 512), (256, 256, 1024), (256, 4096, 4096), (64, 1024, 14336),
                    (255, 208, 256)]:
        a = torch.randint(-2, 3, (M, K), generator=g).float()
        w = torch.randint(-2, 3, (N, K), generator=g).float()
        ref = ((a.double() @ w.double().t()).float() * 0.25).to(torch.bfloat16).float()
        ag = to_fp8_bits(a).to(DEV).view(torch.float8_e4m3fn)
        wg = to_fp8_bits(w).to(DEV).view(torch.float8_e4m3fn).t()
        for bt in (0, 32, 64, 128, 256):
            out = _ops().cutlass_scaled_mm(ag, wg, one, one, torch.bfloat16, impl="v5" if bt == 0 else f"v5:{bt}")
            assert torch.equal(out.float().cpu(), ref), (M, N, K, bt)
    for M, N, K in [(256, 6144, 4096), (100, 4096, 1024), (17, 1280, 8192)]:
        a8 = to_fp8_bits(torch.randn(M, K, generator=g)).to(DEV).view(torch.float8_e4m3fn)
        w8 = to_fp8_bits(torch.randn(N, K, generator=g)).to(DEV).view(torch.float8_e4m3fn).t()
        sa, sb = torch.tensor([0.03], device=DEV), torch.tensor([0.02], device=DEV)
        o2 = _ops().cutlass_scaled_mm(a8, w8, sa, sb, torch.bfloat16, impl="v2")
        for bt in (0, 64, 256):
            o5 = _ops().cutlass_scaled_mm(a8, w8, sa, sb, torch.bfloat16, impl="v5" if bt == 0 else f"v5:{bt}")
            assert torch.equal(o5, o2), (M, N, K, bt)


def test_pair_strided_and_rejections():
    g = torch.Generator().manual_seed(6)
    one = torch.ones(1, device=DEV)
    # a is a column slice of a wider buffer, out a column slice of a wider output (strides are part of the ABI)
    big = to_fp8_bits(torch.randn(40, 1024, generator=g)).to(DEV).view(torch.float8_e4m3fn)
    a = big[:, 256:768]
    w = to_fp8_bits(torch.randn(384, 512, generator=g)).to(DEV).view(torch.float8_e4m3fn)
    wide = torch.zeros(40, 1024, dtype=torch.bfloat16, device=DEV)
    o5 = _ops().cutlass_scaled_mm(a, w.t(), one, one, torch.bfloat16, out=wide[:, 128:512], impl="v5")
    o2 = _ops().cutlass_scaled_mm(a.contiguous(), w.t(), one, one, torch.bfloat16, impl="v2")
    assert torch.equal(o5, o2) and not wide[:, :128].any() and not wide[:, 512:].any()
    x = torch.zeros(300, 256, dtype=torch.uint8, device=DEV).view(torch.float8_e4m3fn)
    w0 = torch.zeros(256, 256, dtype=torch.uint8, device=DEV).view(torch.float8_e4m3fn).t()
    with pytest.raises(RuntimeError, match="decode batches only"):
        _ops().cutlass_scaled_mm(x, w0, one, one, torch.bfloat16, impl="v5")
