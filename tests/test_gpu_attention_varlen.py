"""GPU parity: variable-length (prefill / chunked prefill / mixed) paged attention vs the oracle,
through the C ABI.

Mirrors tests/kernels/attention/test_triton_unified_attention.py:91-226 of the reference (same
seq_lens sets, head configurations, sliding windows, soft cap, FP8 KV with non-unit scales).
Tolerance (stated; north_star: bf16 attention output <= 2e-2 abs): atol 5e-3 + rtol 1e-2 against the
float64 oracle for bf16 / f16 outputs; FP8 outputs within one e4m3 step of the quantised oracle."""
import numpy as np
import pytest
import torch

from helpers import make_block_table, make_kv_cache, slots_for
from oracle import ref_ops

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
ATOL, RTOL = 5e-3, 1e-2


def _ops():
    from b200ops import ops
    return ops


@pytest.fixture(autouse=True, params=["auto", "mma"])
def prefill_impl(request, monkeypatch):
    """Every case runs twice: through the shape dispatch (tcgen05 kernel for 16-bit caches with head_size 128,
    mma.sync kernel otherwise) and with the mma.sync kernel forced (B200_PREFILL_IMPL, read per call)."""
    if request.param == "mma":
        monkeypatch.setenv("B200_PREFILL_IMPL", "mma")
    else:
        monkeypatch.delenv("B200_PREFILL_IMPL", raising=False)
    return request.param


def build(seq_lens, Hq, Hkv, D, bs, dtype, kv_fp8, seed=0, layout="NHD", k_scale=0.5, v_scale=0.25):
    """seq_lens: list of (query_len, kv_len) as in the reference test."""
    g = torch.Generator().manual_seed(seed)
    qlens = [a for a, _ in seq_lens]
    klens = [b for _, b in seq_lens]
    S = len(seq_lens)
    max_blocks = max(1, (max(klens) + bs - 1) // bs)
    nb = S * max_blocks + 5
    bt = make_block_table(klens, bs, nb, g)
    Tk = sum(klens)
    key = (torch.rand(Tk, Hkv, D, generator=g) * 2 - 1).to(dtype)
    value = (torch.rand(Tk, Hkv, D, generator=g) * 2 - 1).to(dtype)
    q = (torch.rand(sum(qlens), Hq, D, generator=g) * 2 - 1).to(dtype)
    cdt = torch.uint8 if kv_fp8 else dtype
    raw, kc, vc = make_kv_cache(nb, bs, Hkv, D, cdt, layout, True)
    if kv_fp8:
        raw.fill_(0x7F)       # NaN pattern in every unwritten slot: must never reach the output
    else:
        raw.fill_(float("nan"))
    ks = torch.tensor([k_scale / 64]) if kv_fp8 else None
    vs = torch.tensor([v_scale / 64]) if kv_fp8 else None
    one = torch.ones(1)
    ref_ops.reshape_and_cache_flash(key, value, kc, vc, slots_for(klens, bt, bs), "fp8" if kv_fp8 else "auto",
                                    ks if kv_fp8 else one, vs if kv_fp8 else one)
    cu = torch.zeros(S + 1, dtype=torch.int32)
    cu[1:] = torch.tensor(qlens).cumsum(0)
    return dict(q=q, raw=raw, kc=kc, vc=vc, bt=bt, qlens=qlens, klens=klens, cu=cu, ks=ks, vs=vs, D=D,
                layout=layout)


def run(c, scale, window=None, softcap=None, alibi=None, out_scale=None, lse=False):
    ops = _ops()
    raw = c["raw"].to(DEV)
    kv = raw if c["layout"] == "NHD" else raw.transpose(1, 2)
    D = c["D"]
    kc, vc = kv[..., :D], kv[..., D:]
    q = c["q"].to(DEV)
    out = torch.empty(q.shape, dtype=torch.float8_e4m3fn if out_scale is not None else q.dtype, device=DEV)
    lse_t = torch.empty(q.shape[0], q.shape[1], dtype=torch.float32, device=DEV) if lse else None
    ops.unified_attention(
        q, kc, vc, out, c["cu"].to(DEV), max(c["qlens"]), torch.tensor(c["klens"], dtype=torch.int32, device=DEV),
        max(c["klens"]), scale, True, (window - 1, 0) if window else (-1, -1), c["bt"].to(DEV),
        softcap or 0.0, None, None if c["ks"] is None else c["ks"].to(DEV),
        None if c["vs"] is None else c["vs"].to(DEV),
        alibi_slopes=None if alibi is None else alibi.to(DEV),
        output_scale=None if out_scale is None else out_scale.to(DEV), lse=lse_t)
    torch.cuda.synchronize()
    return (out, lse_t) if lse else out


def oracle(c, scale, window=None, softcap=None, alibi=None, lse=False):
    return ref_ops.paged_attention_varlen(c["q"], c["kc"], c["vc"], c["qlens"], c["klens"], c["bt"], scale,
                                          k_scale=c["ks"], v_scale=c["vs"], alibi_slopes=alibi,
                                          sliding_window=window, softcap=softcap, return_lse=lse)


def check(out, ref):
    got = out.float().cpu().numpy()
    assert np.isfinite(got).all(), "non-finite output"
    err = np.abs(got - ref)
    tol = ATOL + RTOL * np.abs(ref)
    assert (err <= tol).all(), f"max abs err {err.max():.3e} (worst excess {(err - tol).max():.3e})"


# reference: test_triton_unified_attention.py:91-106
SEQ_LENS = [[(1, 1328), (5, 18), (129, 463)], [(1, 523), (1, 37), (1, 2011)]]
NUM_HEADS = [(4, 4), (8, 2), (5, 1)]


@pytest.mark.parametrize("seq_lens", SEQ_LENS)
@pytest.mark.parametrize("heads", NUM_HEADS)
@pytest.mark.parametrize("D", [128, 64])
@pytest.mark.parametrize("bs", [16, 64])
@pytest.mark.parametrize("window", [None, 64, 256])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("softcap", [None, 50.0])
def test_reference_grid(seq_lens, heads, D, bs, window, dtype, softcap):
    if dtype is torch.float16 and (window == 64 or bs == 64):
        pytest.skip("thinned grid")
    c = build(seq_lens, heads[0], heads[1], D, bs, dtype, kv_fp8=False)
    scale = D ** -0.5
    check(run(c, scale, window, softcap), oracle(c, scale, window, softcap))


@pytest.mark.parametrize("seq_lens", SEQ_LENS + [[(64, 64), (200, 200), (1, 1), (17, 4000)]])
@pytest.mark.parametrize("heads", [(32, 8), (8, 2), (4, 4)])
@pytest.mark.parametrize("D", [128, 64])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("window", [None, 128])
def test_fp8_kv(seq_lens, heads, D, dtype, window):
    """FP8-E4M3 KV with non-unit per-tensor scales (test_triton_unified_attention.py:157-166)."""
    c = build(seq_lens, heads[0], heads[1], D, 16, dtype, kv_fp8=True)
    scale = D ** -0.5
    check(run(c, scale, window), oracle(c, scale, window))


@pytest.mark.parametrize("layout", ["NHD", "HND"])
@pytest.mark.parametrize("kv_fp8", [False, True])
def test_prefill_only_llama_shape(layout, kv_fp8):
    """Pure prefill (no context), Llama-3.1-8B head geometry, ragged prompt lengths incl. tile edges."""
    seq_lens = [(n, n) for n in (1, 15, 16, 17, 63, 64, 65, 127, 128, 129, 700, 1024)]
    c = build(seq_lens, 32, 8, 128, 16, torch.bfloat16, kv_fp8, layout=layout)
    scale = 128 ** -0.5
    out, lse = run(c, scale, lse=True)
    ref, ref_lse = oracle(c, scale, lse=True)
    check(out, ref)
    assert np.abs(lse.cpu().numpy() - ref_lse).max() <= 2e-3


def test_alibi_and_group_of_seven():
    """ALiBi slopes; Qwen2.5-0.5B geometry (14 query heads on 2 kv heads: 9 tokens x 7 heads per q block)."""
    c = build([(33, 180), (1, 97), (120, 120)], 14, 2, 64, 16, torch.bfloat16, kv_fp8=False)
    alibi = torch.tensor([2.0 ** (-(i + 1) / 2) for i in range(14)], dtype=torch.float32)
    scale = 64 ** -0.5
    check(run(c, scale, alibi=alibi), oracle(c, scale, alibi=alibi))


def test_decode_rows_match_decode_kernel():
    """query_len == 1 everywhere: same answer as the decode kernel on the same cache (within tolerance)."""
    ops = _ops()
    c = build([(1, 77), (1, 1500), (1, 16)], 32, 8, 128, 16, torch.bfloat16, kv_fp8=True)
    scale = 128 ** -0.5
    a = run(c, scale)
    raw = c["raw"].to(DEV)
    out = torch.empty_like(a)
    ops.paged_attention_decode(out, c["q"].to(DEV), raw[..., :128], raw[..., 128:], c["bt"].to(DEV),
                               torch.tensor(c["klens"], dtype=torch.int32, device=DEV), scale,
                               k_scale=c["ks"].to(DEV), v_scale=c["vs"].to(DEV))
    torch.cuda.synchronize()
    assert (a.float() - out.float()).abs().max().item() <= 1e-2


def test_fp8_output():
    """Fused static FP8 output (output_scale): e4m3(attention / output_scale). Compared with quantising the
    oracle's output: the attention tolerance (ATOL + RTOL |ref|, in units of output_scale) plus one e4m3 step."""
    c = build([(40, 300), (128, 128), (1, 50)], 8, 2, 128, 16, torch.bfloat16, kv_fp8=True)
    scale = 128 ** -0.5
    osc = torch.tensor([0.01])
    got = run(c, scale, out_scale=osc).view(torch.uint8).cpu().numpy()
    ref = oracle(c, scale)
    want = ref_ops.f32_to_e4m3_bits(np.clip(ref / np.float32(0.01), -448, 448).astype(np.float32))
    gv, wv = ref_ops.e4m3_bits_to_f32(got), ref_ops.e4m3_bits_to_f32(want)
    assert np.isfinite(gv).all()
    step = np.maximum(np.abs(wv) * 0.125, 2.0 ** -9)
    tol = step + (ATOL + RTOL * np.abs(ref)) / 0.01
    assert (np.abs(gv - wv) <= tol).all(), f"worst excess {(np.abs(gv - wv) - tol).max()}"
    assert (got == want).mean() > 0.8
    # and bit-exact against quantising the kernel's own bf16 output is not expected (the fused path
    # quantises the fp32 value, like the reference's out_scale_inv epilogue)


def test_rejects_bad_inputs():
    ops = _ops()
    c = build([(4, 20)], 4, 4, 128, 16, torch.bfloat16, kv_fp8=False)
    with pytest.raises(RuntimeError, match="causal"):
        q = c["q"].to(DEV)
        ops.unified_attention(q, c["kc"].to(DEV), c["vc"].to(DEV), torch.empty_like(q), c["cu"].to(DEV), 4,
                              torch.tensor([20], dtype=torch.int32, device=DEV), 20, 1.0, False, (-1, -1),
                              c["bt"].to(DEV), 0.0, None, None, None)
    with pytest.raises(RuntimeError, match="CUDA"):
        q = c["q"]
        ops.unified_attention(q, c["kc"], c["vc"], torch.empty_like(q), c["cu"], 4,
                              torch.tensor([20], dtype=torch.int32), 20, 1.0, True, (-1, -1), c["bt"], 0.0,
                              None, None, None)


import glob
import os

from helpers import bits_to_bf16

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "cpu_varlen_*.npz"))))
def test_vs_reference_cpu_kernel_golden(path):
    """CUDA kernel vs the committed outputs of the reference's own CPU attention kernel
    (csrc/cpu/cpu_attn.cpp cpu_attention_with_kv_cache, fixtures from tests/golden/make_golden.py)."""
    ops = _ops()
    z = np.load(path)
    Hq, Hkv, D, bs, nb, win = [int(v) for v in z["meta"]]
    q = bits_to_bf16(z["query"]).to(DEV)
    key, value = bits_to_bf16(z["key"]).to(DEV), bits_to_bf16(z["value"]).to(DEV)
    raw = torch.zeros(nb, bs, Hkv, 2 * D, dtype=torch.bfloat16, device=DEV)
    kc, vc = raw[..., :D], raw[..., D:]
    one = torch.ones(1, device=DEV)
    ops.reshape_and_cache_flash(key, value, kc, vc, torch.from_numpy(z["slot_mapping"]).to(DEV), "auto", one, one)
    qlens = z["query_lens"].tolist()
    cu = torch.zeros(len(qlens) + 1, dtype=torch.int32)
    cu[1:] = torch.tensor(qlens).cumsum(0)
    out = torch.empty_like(q)
    cap = float(z["softcap"][0])
    ops.unified_attention(q, kc, vc, out, cu.to(DEV), max(qlens), torch.from_numpy(z["seq_lens"]).to(DEV),
                          int(z["seq_lens"].max()), float(z["scale"][0]), True, (win - 1, 0) if win else (-1, -1),
                          torch.from_numpy(z["block_table"]).to(DEV), cap, None, None, None)
    torch.cuda.synchronize()
    golden = bits_to_bf16(z["out"]).float().numpy()
    got = out.float().cpu().numpy()
    # both sides round to bf16: two roundings of |out| <= 1.5
    np.testing.assert_allclose(got, golden, atol=8e-3, rtol=1.6e-2)
    assert np.abs(got - golden).mean() < 1e-3
