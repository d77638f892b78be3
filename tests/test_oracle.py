"""CPU tests: pin oracle/ref_ops.py against the reference (golden fixtures, torch's e4m3 cast,
the reference's own python oracles)."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import ref_ops
from helpers import bits_to_bf16, make_kv_cache

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_e4m3_matches_torch_cast_all_codes():
    # every e4m3 code decodes to the value torch gives, and re-encodes to itself
    codes = np.arange(256, dtype=np.uint8)
    ours = ref_ops.e4m3_bits_to_f32(codes)
    theirs = torch.from_numpy(codes.copy()).view(torch.float8_e4m3fn).float().numpy()
    finite = ~np.isnan(theirs)
    assert np.array_equal(np.isnan(ours), np.isnan(theirs))
    assert np.array_equal(ours[finite], theirs[finite])
    re = ref_ops.f32_to_e4m3_bits(ours[finite])
    # -0.0 / +0.0 keep their sign bit
    assert np.array_equal(re, codes[finite])


def test_e4m3_rne_matches_torch_cast_random():
    g = np.random.default_rng(0)
    x = np.concatenate([
        g.uniform(-448, 448, 200000), g.normal(0, 1, 200000), g.normal(0, 1e-2, 100000),
        g.uniform(-2 ** -6, 2 ** -6, 100000),
    ]).astype(np.float32)
    # exact ties between neighbouring codes
    codes = np.arange(0, 0x7E, dtype=np.uint8)
    v = ref_ops.e4m3_bits_to_f32(codes).astype(np.float64)
    v2 = ref_ops.e4m3_bits_to_f32(codes + 1).astype(np.float64)
    ties = ((v + v2) / 2).astype(np.float32)
    x = np.concatenate([x, ties, -ties])
    ours = ref_ops.f32_to_e4m3_bits(x)
    theirs = torch.from_numpy(x).to(torch.float8_e4m3fn).view(torch.uint8).numpy()
    assert np.array_equal(ours, theirs)


def test_e4m3_saturates_to_finite():
    x = np.array([449.0, 1e9, np.inf, -1e9, -np.inf, 464.0, 480.0], dtype=np.float32)
    assert ref_ops.f32_to_e4m3_bits(x).tolist() == [0x7E, 0x7E, 0x7E, 0xFE, 0xFE, 0x7E, 0x7E]


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "cpu_attn_*.npz"))))
def test_attention_oracle_matches_reference_cpu_kernel(path):
    """oracle reshape_and_cache_flash + paged_attention_decode vs the outputs of the reference's own
    CPU kernels (fixtures made by tests/golden/make_golden.py)."""
    z = np.load(path)
    Hq, Hkv, D, bs, nb, win, has_alibi = [int(v) for v in z["meta"]]
    q = bits_to_bf16(z["query"])
    key, value = bits_to_bf16(z["key"]), bits_to_bf16(z["value"])
    _, kc, vc = make_kv_cache(nb, bs, Hkv, D, torch.bfloat16, "NHD")
    one = torch.ones(1)
    ref_ops.reshape_and_cache_flash(key, value, kc, vc, torch.from_numpy(z["slot_mapping"]), "auto",
                                    one, one)
    out = ref_ops.paged_attention_decode(
        q, kc, vc, torch.from_numpy(z["block_table"]), torch.from_numpy(z["seq_lens"]),
        float(z["scale"][0]), alibi_slopes=torch.from_numpy(z["alibi"]) if has_alibi else None,
        sliding_window=win or None, softcap=float(z["softcap"][0]) or None)
    golden = bits_to_bf16(z["out"]).float().numpy()
    # reference CPU kernel: bf16 output, fp32 accumulate (3e-5 vs naive fp32 before
    # the bf16 output rounding). bf16 rounding of |out| <= 1 => <= 2^-9 relative.
    np.testing.assert_allclose(out, golden, atol=4e-3, rtol=8e-3)
    assert np.abs(out - golden).mean() < 5e-4


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "cpu_varlen_*.npz"))))
def test_varlen_attention_oracle_matches_reference_cpu_kernel(path):
    """oracle paged_attention_varlen (query_len > 1: prefill / chunked prefill / mixed batches) vs the
    outputs of the reference's own CPU kernel cpu_attention_with_kv_cache on the same packed queries
    (fixtures made by tests/golden/make_golden.py, VARLEN_CASES)."""
    z = np.load(path)
    Hq, Hkv, D, bs, nb, win = [int(v) for v in z["meta"]]
    q = bits_to_bf16(z["query"])
    key, value = bits_to_bf16(z["key"]), bits_to_bf16(z["value"])
    _, kc, vc = make_kv_cache(nb, bs, Hkv, D, torch.bfloat16, "NHD")
    one = torch.ones(1)
    ref_ops.reshape_and_cache_flash(key, value, kc, vc, torch.from_numpy(z["slot_mapping"]), "auto", one, one)
    out = ref_ops.paged_attention_varlen(
        q, kc, vc, z["query_lens"].tolist(), z["seq_lens"].tolist(), torch.from_numpy(z["block_table"]),
        float(z["scale"][0]), sliding_window=win or None, softcap=float(z["softcap"][0]) or None)
    golden = bits_to_bf16(z["out"]).float().numpy()
    np.testing.assert_allclose(out, golden, atol=4e-3, rtol=8e-3)
    assert np.abs(out - golden).mean() < 5e-4


def test_varlen_oracle_reduces_to_decode_oracle():
    """query_len == 1 rows of the varlen oracle are the decode oracle (incl. LSE, ALiBi, FP8 scales)."""
    g = torch.Generator().manual_seed(3)
    lens = [77, 300, 16, 1]
    Hq, Hkv, D, bs = 8, 2, 64, 16
    nb = 4 * 19 + 3
    bt = torch.randperm(nb, generator=g)[: 4 * 19].to(torch.int32).view(4, 19)
    kc = torch.randint(0, 0x78, (nb, bs, Hkv, D), generator=g, dtype=torch.uint8)
    vc = torch.randint(0, 0x78, (nb, bs, Hkv, D), generator=g, dtype=torch.uint8)
    q = (torch.rand(4, Hq, D, generator=g) * 2 - 1).to(torch.bfloat16)
    ks, vs = torch.tensor([0.01]), torch.tensor([0.02])
    slopes = torch.tensor([0.5 ** (i + 1) for i in range(Hq)])
    a, la = ref_ops.paged_attention_varlen(q, kc, vc, [1] * 4, lens, bt, 0.125, k_scale=ks, v_scale=vs,
                                           alibi_slopes=slopes, sliding_window=100, return_lse=True)
    b, lb = ref_ops.paged_attention_decode(q, kc, vc, bt, torch.tensor(lens, dtype=torch.int32), 0.125,
                                           k_scale=ks, v_scale=vs, alibi_slopes=slopes, sliding_window=100,
                                           return_lse=True)
    np.testing.assert_allclose(a, b, atol=1e-6, rtol=1e-6)
    np.testing.assert_allclose(la, lb, atol=1e-5, rtol=1e-6)


@pytest.mark.parametrize("hidden", [17, 1024, 1025, 5137])
@pytest.mark.parametrize("tokens", [1, 7, 33])
def test_per_token_quant_restatement_vs_reference_python_oracle(hidden, tokens):
    torch.manual_seed(1)
    x = (torch.rand(tokens, hidden) * 2 - 1).mul(8).to(torch.bfloat16)
    ub = torch.tensor([5.0])
    for scale_ub in (None, ub):
        bits, scales = ref_ops.dynamic_per_token_scaled_fp8_quant(x, scale_ub)
        ref_q, ref_s = ref_ops.ref_dynamic_per_token_quant(x, scale_ub)
        np.testing.assert_array_equal(scales, ref_s.numpy())
        np.testing.assert_array_equal(bits, ref_q.view(torch.uint8).numpy())


def test_per_tensor_quant_restatement():
    torch.manual_seed(2)
    x = torch.randn(9, 130).to(torch.float16)
    bits, scale = ref_ops.dynamic_scaled_fp8_quant(x)
    assert scale[0] == np.float32(x.float().abs().max().item()) / np.float32(448.0)
    # same thing through the static path
    bits2 = ref_ops.static_scaled_fp8_quant(x, torch.from_numpy(scale))
    np.testing.assert_array_equal(bits, bits2)
    # dequantised error bounded by half an fp8 ulp (2^-4 relative) + subnormal step
    deq = ref_ops.e4m3_bits_to_f32(bits) * scale[0]
    err = np.abs(deq - x.float().numpy())
    assert (err <= np.abs(x.float().numpy()) * 2 ** -4 + scale[0] * 2 ** -10 + 1e-7).all()


def test_static_quant_group_shapes():
    torch.manual_seed(3)
    x = torch.randn(8, 256).to(torch.bfloat16)
    s2d = torch.rand(8, 2) + 0.5
    bits = ref_ops.static_scaled_fp8_quant(x, s2d)
    for i in range(8):
        for j in range(2):
            blk = ref_ops.static_scaled_fp8_quant(x[i:i + 1, j * 128:(j + 1) * 128], s2d[i, j].reshape(1))
            np.testing.assert_array_equal(bits[i:i + 1, j * 128:(j + 1) * 128], blk)
    per_ch = ref_ops.static_scaled_fp8_quant(x, torch.rand(256) + 0.5, group_shape=(-1, 1))
    assert per_ch.shape == (8, 256)


def test_slot_mapping_restatement():
    bt = np.arange(40, dtype=np.int32).reshape(4, 10)[:, ::-1].copy()
    req = np.array([0, 0, 1, 3, 3, 3])
    pos = np.array([0, 17, 159, 5, 16, 31])
    sm = ref_ops.compute_slot_mapping(bt, req, pos, 16)
    assert sm.dtype == np.int64
    assert sm.tolist() == [9 * 16, 8 * 16 + 1, 10 * 16 + 15, 39 * 16 + 5, 38 * 16, 38 * 16 + 15]


def test_reshape_and_cache_flash_oracle_fp8_and_skip():
    torch.manual_seed(4)
    T, H, D, bs, nb = 9, 3, 32, 16, 4
    key = torch.randn(T, H, D).to(torch.bfloat16)
    value = torch.randn(T, H, D).to(torch.bfloat16)
    slots = torch.tensor([5, -1, 17, 0, 63, 33, -1, 16, 2])
    for layout in ("NHD", "HND"):
        raw, kc, vc = make_kv_cache(nb, bs, H, D, torch.uint8, layout)
        ks, vs = torch.tensor([0.05, 0.1, 0.2]), torch.tensor([0.3, 0.02, 0.07])
        ref_ops.reshape_and_cache_flash(key, value, kc, vc, slots, "fp8", ks, vs)
        for t, s in enumerate(slots.tolist()):
            if s < 0:
                continue
            exp = (key[t].float() / ks[:, None]).clamp(-448, 448).to(torch.float8_e4m3fn).view(torch.uint8)
            assert torch.equal(kc[s // bs, s % bs], exp)
        assert int((raw != 0).any(dim=-1).sum()) <= 7 * H  # only 7 valid tokens were written


@pytest.mark.parametrize("group_size", [64, 128, 512])
@pytest.mark.parametrize("tokens,hidden", [(7, 512), (33, 4096), (5, 5120)])
def test_per_token_group_quant_restatement_vs_reference_python_oracle(tokens, hidden, group_size):
    """tests/kernels/quantization/test_block_fp8.py:57-79 grid; the reference's torch oracle
    (quant_utils.py:157-180) must give the same bytes and scales as the kernel restatement."""
    torch.manual_seed(0)
    x = torch.rand(tokens, hidden, dtype=torch.bfloat16) * 2 - 1
    x[0, :group_size] = 0  # eps path
    bits, scales = ref_ops.per_token_group_quant_fp8(x, group_size)
    q_ref, s_ref = ref_ops.ref_native_per_token_group_quant_fp8(x.float(), group_size)
    np.testing.assert_array_equal(scales, s_ref.numpy())
    np.testing.assert_array_equal(bits, q_ref.view(torch.uint8).numpy())
    assert scales[0, 0] == np.float32(1e-10) / np.float32(448.0)
    b8, s8 = ref_ops.per_token_group_quant_fp8(x, group_size, use_ue8m0=True)
    m, _ = np.frexp(s8)
    lo = np.maximum(scales, np.float32(1e-10))  # the kernel clamps before log2f
    assert np.all(m == 0.5) and np.all(s8 >= lo) and np.all(s8 < 2 * lo * (1 + 1e-6))


def test_block_scaled_mm_restatement_vs_reference_python_oracle():
    """native_w8a8_block_matmul (tests/kernels/quant_utils.py:91-155), restated with torch fp32 exactly as
    the reference writes it, against the float64 oracle (test_block_fp8.py:161-201 shape, N % 128 != 0)."""
    torch.manual_seed(0)
    M, N, K, bn, bk = 32, 576, 1024, 128, 128
    a = ((torch.rand(M, K) - 0.5) * 2 * 448).to(torch.float8_e4m3fn)
    b = ((torch.rand(N, K) - 0.5) * 2 * 448).to(torch.float8_e4m3fn)
    a_s = torch.rand(M, K // bk) * 1e-2
    b_s = torch.rand((N + bn - 1) // bn, K // bk) * 1e-2
    c = torch.zeros(M, N)
    for i in range(K // bk):
        for j in range((N + bn - 1) // bn):
            ns = slice(j * bn, min(N, (j + 1) * bn))
            c[:, ns] += (a[:, i * bk:(i + 1) * bk].float() @ b[ns, i * bk:(i + 1) * bk].float().t()) \
                * (a_s[:, i:i + 1] * b_s[j, i])
    got = ref_ops.w8a8_block_scaled_mm(a.view(torch.uint8).numpy(), b.view(torch.uint8).numpy(),
                                       a_s.numpy(), b_s.numpy())
    np.testing.assert_allclose(got, c.numpy(), rtol=2e-5, atol=1e-3)


@pytest.mark.parametrize("mode", ["k", "p", "kp"])
def test_top_k_top_p_restatement_vs_reference_golden(mode):
    """tests/golden/topk_topp_ref.npz = outputs of the reference's apply_top_k_top_p_pytorch
    (topk_topp_sampler.py:367-408), generated by tests/golden/make_sampler_golden.py."""
    import os
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "topk_topp_ref.npz"))
    x, k, p = torch.from_numpy(z["x"]), torch.from_numpy(z["k"]), torch.from_numpy(z["p"])
    got, _ = ref_ops.apply_top_k_top_p(x, k if "k" in mode else None, p if "p" in mode else None)
    np.testing.assert_array_equal(got.numpy(), z["out_" + mode])


def test_repetition_penalties_restatement_vs_reference_python():
    """apply_repetition_penalties_torch (vllm/_custom_ops.py:309-322), restated literally, vs the kernel restatement
    (fp32 logits: the two definitions agree bit for bit)."""
    torch.manual_seed(0)
    x = torch.randn(6, 500) * 4
    pm, om = torch.rand(6, 500) < 0.2, torch.rand(6, 500) < 0.2
    pen = torch.rand(6) + 0.7
    rp = pen.unsqueeze(1).repeat(1, 500)
    penalties = torch.where(pm | om, rp, 1.0)
    want = x * torch.where(x > 0, 1.0 / penalties, penalties)
    got = ref_ops.apply_repetition_penalties(x, pm, om, pen)
    np.testing.assert_allclose(got.numpy(), want.numpy(), rtol=2e-7, atol=0)
