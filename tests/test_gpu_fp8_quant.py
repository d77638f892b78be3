"""GPU parity: scaled_fp8_quant (static / dynamic per-tensor / dynamic per-token) vs the oracle.
Mirrors tests/kernels/quantization/test_fp8_quant.py:63-221 (bit-exact bytes and scales here)."""
import numpy as np
import pytest
import torch

from oracle import ref_ops

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
HIDDEN = [17, 1024, 1025, 1026, 5137, 8193]
TOKENS = [1, 7, 4096]
DTYPES = [torch.bfloat16, torch.float16, torch.float32]


def _ops():
    from b200ops import ops
    return ops


def _x(tokens, hidden, dtype, seed=0):
    g = torch.Generator().manual_seed(seed)
    return (torch.rand(tokens, hidden, generator=g) * 2 - 1).mul(8).to(dtype) + 1e-6


@pytest.mark.parametrize("hidden", HIDDEN)
@pytest.mark.parametrize("tokens", TOKENS)
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("use_ub", [False, True])
def test_dynamic_per_token(hidden, tokens, dtype, use_ub):
    x = _x(tokens, hidden, dtype)
    ub = torch.tensor([6.0]) if use_ub else None
    bits, scales = ref_ops.dynamic_per_token_scaled_fp8_quant(x, ub)
    out, sc = _ops().scaled_fp8_quant(x.to(DEV), scale_ub=None if ub is None else ub.to(DEV),
                                      use_per_token_if_dynamic=True)
    assert sc.shape == (tokens, 1) and out.dtype == torch.float8_e4m3fn
    np.testing.assert_array_equal(sc.cpu().numpy(), scales)
    np.testing.assert_array_equal(out.view(torch.uint8).cpu().numpy(), bits)


@pytest.mark.parametrize("hidden", HIDDEN)
@pytest.mark.parametrize("tokens", TOKENS)
@pytest.mark.parametrize("dtype", DTYPES)
def test_dynamic_per_tensor_and_static(hidden, tokens, dtype):
    x = _x(tokens, hidden, dtype, seed=1)
    bits, scale = ref_ops.dynamic_scaled_fp8_quant(x)
    out, sc = _ops().scaled_fp8_quant(x.to(DEV))
    np.testing.assert_array_equal(sc.cpu().numpy(), scale)
    np.testing.assert_array_equal(out.view(torch.uint8).cpu().numpy(), bits)
    # static with the same scale gives the same bytes
    out2, _ = _ops().scaled_fp8_quant(x.to(DEV), scale=sc)
    assert torch.equal(out2.view(torch.uint8), out.view(torch.uint8))


def test_static_group_shapes_padding_and_strides():
    torch.manual_seed(2)
    M, K = 48, 512
    x = torch.randn(M, K).to(torch.bfloat16)
    o = _ops()
    # 2-D group scales (1x128) and (16x128)
    for sshape in [(M, K // 128), (M // 16, K // 128)]:
        s = torch.rand(*sshape) + 0.25
        bits = ref_ops.static_scaled_fp8_quant(x, s)
        out, _ = o.scaled_fp8_quant(x.to(DEV), scale=s.to(DEV))
        np.testing.assert_array_equal(out.view(torch.uint8).cpu().numpy(), bits)
    # 1-D per-channel / per-token need an explicit group_shape
    s = torch.rand(K) + 0.25
    bits = ref_ops.static_scaled_fp8_quant(x, s, group_shape=(-1, 1))
    out, _ = o.scaled_fp8_quant(x.to(DEV), scale=s.to(DEV), group_shape=(-1, 1))
    np.testing.assert_array_equal(out.view(torch.uint8).cpu().numpy(), bits)
    s = torch.rand(M) + 0.25
    bits = ref_ops.static_scaled_fp8_quant(x, s, group_shape=(1, -1))
    out, _ = o.scaled_fp8_quant(x.to(DEV), scale=s.to(DEV), group_shape=(1, -1))
    np.testing.assert_array_equal(out.view(torch.uint8).cpu().numpy(), bits)
    with pytest.raises(RuntimeError, match="1D scale requires explicit group_shape"):
        o.scaled_fp8_quant(x.to(DEV), scale=s.to(DEV))
    # num_token_padding pads the output rows
    out, sc = o.scaled_fp8_quant(x.to(DEV), num_token_padding=64)
    assert out.shape == (64, K)
    bits, _ = ref_ops.dynamic_scaled_fp8_quant(x)
    np.testing.assert_array_equal(out[:M].view(torch.uint8).cpu().numpy(), bits)
    # strided input rows (a column slice of a wider tensor)
    wide = torch.randn(M, 2 * K).to(torch.float16)
    xs = wide[:, K // 2: K // 2 + K]
    bits, scales = ref_ops.dynamic_per_token_scaled_fp8_quant(xs)
    out, sc = o.scaled_fp8_quant(wide.to(DEV)[:, K // 2: K // 2 + K], use_per_token_if_dynamic=True)
    np.testing.assert_array_equal(out.view(torch.uint8).cpu().numpy(), bits)
    np.testing.assert_array_equal(sc.cpu().numpy(), scales)


def test_saturation_zero_rows_and_min_scale():
    x = torch.zeros(3, 64, dtype=torch.bfloat16)
    x[1] = 1e-9
    x[2, 0] = 3e38
    bits, scales = ref_ops.dynamic_per_token_scaled_fp8_quant(x)
    out, sc = _ops().scaled_fp8_quant(x.to(DEV), use_per_token_if_dynamic=True)
    np.testing.assert_array_equal(sc.cpu().numpy(), scales)
    np.testing.assert_array_equal(out.view(torch.uint8).cpu().numpy(), bits)
    assert scales[0, 0] == np.float32(1.0) / (np.float32(448.0) * np.float32(512.0))
    # static scale far too small -> saturates to +-448, never NaN / inf
    big = torch.tensor([[1000.0, -1000.0, 0.5, float("inf")]], dtype=torch.float32)
    s = torch.tensor([1.0])
    out, _ = _ops().scaled_fp8_quant(big.to(DEV), scale=s.to(DEV))
    assert out.view(torch.uint8).cpu().tolist() == [[0x7E, 0xFE, 0x30, 0x7E]]
    np.testing.assert_array_equal(ref_ops.static_scaled_fp8_quant(big, s), [[0x7E, 0xFE, 0x30, 0x7E]])


# ------------------------------------------------------------------------------------------------
# per_token_group_fp8_quant: grid of tests/kernels/quantization/test_block_fp8.py:40-79 (+ f16 / f32,
# group sizes without a vector path, UE8M0 scales, the three scale layouts)
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("group_size", [64, 128, 512])
@pytest.mark.parametrize("tokens,hidden", [(7, 512), (2050, 4096), (7, 5120), (33, 13824)])
@pytest.mark.parametrize("layout", ["row", "col", "tma"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
def test_per_token_group_quant(tokens, hidden, group_size, layout, dtype):
    x = _x(tokens, hidden, dtype, seed=3)
    x[0, :group_size] = 0  # eps path
    bits, scales = ref_ops.per_token_group_quant_fp8(x, group_size)
    out, sc = _ops().per_token_group_quant_fp8(x.to(DEV), group_size, column_major_scales=layout != "row",
                                               tma_aligned_scales=layout == "tma")
    assert sc.shape == (tokens, hidden // group_size) and out.dtype == torch.float8_e4m3fn
    if layout == "col":
        assert sc.stride() == (1, tokens)
    if layout == "tma":
        assert sc.stride() == (1, (tokens + 3) // 4 * 4)
    np.testing.assert_array_equal(sc.cpu().numpy(), scales)
    np.testing.assert_array_equal(out.view(torch.uint8).cpu().numpy(), bits)


@pytest.mark.parametrize("group_size", [16, 24, 128, 1024, 2048])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_per_token_group_quant_ue8m0_and_odd_groups(group_size, dtype):
    tokens, hidden = 19, group_size * 6
    x = _x(tokens, hidden, dtype, seed=4) * 3
    x[1, 5] = float("nan")  # absmax drops NaN; the element itself clamps to fp8_min
    for ue in (False, True):
        bits, scales = ref_ops.per_token_group_quant_fp8(x, group_size, use_ue8m0=ue)
        out, sc = _ops().per_token_group_quant_fp8(x.to(DEV), group_size, use_ue8m0=ue)
        got_s, got = sc.cpu().numpy(), out.view(torch.uint8).cpu().numpy()
        if ue:  # ceil(log2f(.)) one ulp from a power of two may differ between libms: exclude those groups
            _, plain = ref_ops.per_token_group_quant_fp8(x, group_size)
            ok = ~ref_ops.ue8m0_near_tie(plain)
            assert ok.mean() > 0.9
            np.testing.assert_array_equal(got_s[ok], scales[ok])
            okx = np.repeat(ok, group_size, axis=1)
            np.testing.assert_array_equal(got[okx], bits[okx])
        else:
            np.testing.assert_array_equal(got_s, scales)
            np.testing.assert_array_equal(got, bits)
    assert got[1, 5] == 0xFE  # -448


def test_per_token_group_quant_errors_and_empty():
    ops = _ops()
    x = torch.zeros(4, 100, dtype=torch.bfloat16, device=DEV)
    with pytest.raises(AssertionError, match="divisible"):
        ops.per_token_group_quant_fp8(x, 64)
    q, s = ops.per_token_group_quant_fp8(torch.zeros(0, 128, dtype=torch.bfloat16, device=DEV), 128)
    assert q.shape == (0, 128) and s.shape == (0, 1)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.per_token_group_quant_fp8(torch.zeros(4, 128, dtype=torch.bfloat16), 128)
