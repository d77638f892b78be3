"""GPU parity: paged-attention decode vs the oracle, through the C ABI.

Mirrors tests/kernels/attention/test_triton_unified_attention.py:91-226 (decode rows) and
tests/v1/attention/test_attention_backends.py of the reference. Tolerance (stated, north_star:
bf16 attention output <= 2e-2 abs): atol 5e-3 + rtol 1e-2 against the fp32 oracle."""
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


def build_case(lens, Hq, Hkv, D, bs, dtype, kv_fp8, layout="NHD", packed=True, seed=0,
               extra_blocks=7, k_scale=0.5, v_scale=0.25):
    g = torch.Generator().manual_seed(seed)
    S = len(lens)
    max_blocks = max(1, (max(lens) + bs - 1) // bs)
    nb = S * max_blocks + extra_blocks
    bt = make_block_table(lens, bs, nb, g)
    T = sum(lens)
    key = (torch.rand(T, Hkv, D, generator=g) * 2 - 1).to(dtype)
    value = (torch.rand(T, Hkv, D, generator=g) * 2 - 1).to(dtype)
    q = (torch.rand(S, Hq, D, generator=g) * 2 - 1).to(dtype)
    cdt = torch.uint8 if kv_fp8 else dtype
    raw, kc, vc = make_kv_cache(nb, bs, Hkv, D, cdt, layout, packed)
    # poison unwritten cache rows: NaN patterns must never leak into the output
    if kv_fp8:
        raw.fill_(0x7F)
    else:
        raw.fill_(float("nan"))
    ks = torch.tensor([k_scale / 64]) if kv_fp8 else None
    vs = torch.tensor([v_scale / 64]) if kv_fp8 else None
    one = torch.ones(1)
    ref_ops.reshape_and_cache_flash(key, value, kc, vc, slots_for(lens, bt, bs),
                                    "fp8" if kv_fp8 else "auto", ks if kv_fp8 else one,
                                    vs if kv_fp8 else one)
    return dict(q=q, raw=raw, kc=kc, vc=vc, bt=bt, lens=torch.tensor(lens, dtype=torch.int32),
                ks=ks, vs=vs, D=D, layout=layout, packed=packed)


def to_dev_views(c):
    raw = c["raw"].to(DEV)
    D = c["D"]
    if not c["packed"]:
        return raw, raw[0], raw[1]
    kv = raw if c["layout"] == "NHD" else raw.transpose(1, 2)
    return raw, kv[..., :D], kv[..., D:]


def run_gpu(c, scale, **kw):
    raw, kc, vc = to_dev_views(c)
    q = c["q"].to(DEV)
    out = torch.full_like(q, float("nan"))
    dev = lambda t: None if t is None else t.to(DEV)
    _ops().paged_attention_decode(out, q, kc, vc, c["bt"].to(DEV), c["lens"].to(DEV), scale,
                                  k_scale=dev(c["ks"]), v_scale=dev(c["vs"]), **kw)
    torch.cuda.synchronize()
    return out


def check(c, scale, **kw):
    lse_t = None
    if kw.pop("want_lse", False):
        lse_t = torch.empty(c["q"].shape[:2], dtype=torch.float32, device=DEV)
    okw = {k: v for k, v in kw.items()}
    out = run_gpu(c, scale, lse=lse_t, **{k: (v.to(DEV) if isinstance(v, torch.Tensor) else v)
                                          for k, v in kw.items()})
    ref = ref_ops.paged_attention_decode(c["q"], c["kc"], c["vc"], c["bt"], c["lens"], scale,
                                         k_scale=c["ks"], v_scale=c["vs"], return_lse=True, **okw)
    got = out.float().cpu().numpy()
    assert not np.isnan(got).any(), "NaN in attention output"
    np.testing.assert_allclose(got, ref[0], atol=ATOL, rtol=RTOL)
    if lse_t is not None:
        np.testing.assert_allclose(lse_t.cpu().numpy(), ref[1], atol=2e-3, rtol=1e-3)


@pytest.mark.parametrize("lens", [[523, 37, 2011], [1], [16, 15, 17, 1, 33, 257]])
@pytest.mark.parametrize("heads", [(4, 4), (8, 2), (5, 1), (32, 8), (14, 2), (64, 8), (32, 2)])
@pytest.mark.parametrize("D", [64, 128, 256])
@pytest.mark.parametrize("kv_fp8", [False, True])
def test_decode_basic(lens, heads, D, kv_fp8):
    if D == 256 and heads[0] // heads[1] > 8:
        pytest.skip("group > 8 at head_size 256 not built")
    c = build_case(lens, heads[0], heads[1], D, 16, torch.bfloat16, kv_fp8)
    check(c, D ** -0.5, want_lse=True)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("bs", [16, 32, 64])
@pytest.mark.parametrize("layout,packed", [("NHD", True), ("HND", True), ("NHD", False)])
@pytest.mark.parametrize("kv_fp8", [False, True])
def test_decode_layouts_block_sizes_dtypes(dtype, bs, layout, packed, kv_fp8):
    c = build_case([700, 45, 129, 64], 16, 4, 128, bs, dtype, kv_fp8, layout, packed, seed=1)
    check(c, 0.09)


@pytest.mark.parametrize("kv_fp8", [False, True])
def test_decode_long_context_split_kv(kv_fp8):
    """One long sequence is cut across many CTAs (split-KV) and merged by the reduce kernel;
    batch of short ones shares CTAs."""
    c = build_case([20000, 3, 9000, 77, 5000], 32, 8, 128, 16, torch.bfloat16, kv_fp8, seed=2)
    check(c, 128 ** -0.5, want_lse=True)
    c = build_case([32768], 8, 1, 128, 16, torch.bfloat16, kv_fp8, seed=3)
    check(c, 128 ** -0.5, want_lse=True)


@pytest.mark.parametrize("window", [64, 128, 1000])
@pytest.mark.parametrize("softcap", [None, 50.0])
def test_decode_sliding_window_softcap(window, softcap):
    c = build_case([523, 37, 2011, 128, 129], 8, 2, 128, 16, torch.bfloat16, False, seed=4)
    check(c, 128 ** -0.5, sliding_window=window, softcap=softcap)


def test_decode_alibi():
    Hq = 12
    slopes = torch.tensor([2 ** (-8.0 * (i + 1) / Hq) for i in range(Hq)], dtype=torch.float32)
    c = build_case([300, 41, 1025], Hq, 4, 128, 16, torch.bfloat16, False, seed=5)
    check(c, 128 ** -0.5, alibi_slopes=slopes)


def test_decode_empty_and_zero_length():
    c = build_case([0, 50, 0, 17], 8, 2, 128, 16, torch.bfloat16, False, seed=6)
    out = run_gpu(c, 0.1)
    assert torch.equal(out[0].cpu(), torch.zeros_like(out[0].cpu()))
    assert torch.equal(out[2].cpu(), torch.zeros_like(out[2].cpu()))
    ref = ref_ops.paged_attention_decode(c["q"], c["kc"], c["vc"], c["bt"], c["lens"], 0.1)
    np.testing.assert_allclose(out.float().cpu().numpy(), ref, atol=ATOL, rtol=RTOL)
    # zero sequences: no-op
    q = torch.empty(0, 8, 128, dtype=torch.bfloat16, device=DEV)
    raw, kc, vc = to_dev_views(c)
    _ops().paged_attention_decode(torch.empty_like(q), q, kc, vc,
                                  torch.empty(0, 4, dtype=torch.int32, device=DEV),
                                  torch.empty(0, dtype=torch.int32, device=DEV), 0.1)


def test_decode_strided_query_and_big_block_ids():
    """q is a view of a fused QKV tensor; physical block ids near num_blocks=32768 need 64-bit
    address arithmetic (test_triton_unified_attention.py:24-26)."""
    torch.manual_seed(7)
    S, Hq, Hkv, D, bs = 5, 8, 2, 128, 16
    lens = [100, 200, 31, 64, 513]
    nb = 32768
    c = build_case(lens, Hq, Hkv, D, bs, torch.bfloat16, True, seed=7)
    # relocate the blocks to the top of a large cache
    small_nb = c["raw"].shape[0]
    raw_big = torch.zeros(nb, bs, Hkv, 2 * D, dtype=torch.uint8, device=DEV)
    raw_big[nb - small_nb:] = c["raw"].to(DEV)
    bt = (c["bt"] + (nb - small_nb)).to(DEV)
    qkv = torch.randn(S, (Hq + 2 * Hkv) * D).to(torch.bfloat16)
    q = qkv[:, : Hq * D].view(S, Hq, D)
    c["q"] = q.contiguous()
    qg = qkv.to(DEV)[:, : Hq * D].view(S, Hq, D)
    out = torch.empty(S, Hq, D, dtype=torch.bfloat16, device=DEV)
    _ops().paged_attention_decode(out, qg, raw_big[..., :D], raw_big[..., D:], bt,
                                  c["lens"].to(DEV), 0.08, k_scale=c["ks"].to(DEV),
                                  v_scale=c["vs"].to(DEV))
    ref = ref_ops.paged_attention_decode(c["q"], c["kc"], c["vc"], c["bt"], c["lens"], 0.08,
                                         k_scale=c["ks"], v_scale=c["vs"])
    np.testing.assert_allclose(out.float().cpu().numpy(), ref, atol=ATOL, rtol=RTOL)


def test_decode_cuda_graph_replay_and_determinism():
    c = build_case([900, 5, 333, 2048], 32, 8, 128, 16, torch.bfloat16, True, seed=8)
    raw, kc, vc = to_dev_views(c)
    q = c["q"].to(DEV)
    bt, lens = c["bt"].to(DEV), c["lens"].to(DEV)
    ks, vs = c["ks"].to(DEV), c["vs"].to(DEV)
    out = torch.empty_like(q)
    ops = _ops()
    ops.paged_attention_decode(out, q, kc, vc, bt, lens, 0.09, k_scale=ks, v_scale=vs)
    torch.cuda.synchronize()
    eager = out.clone()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s):
            ops.paged_attention_decode(out, q, kc, vc, bt, lens, 0.09, k_scale=ks, v_scale=vs)
    out.zero_()
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, eager)
    # the graph reads seq_lens from device memory: shrink a sequence and replay
    lens[3] = 1000
    g.replay()
    torch.cuda.synchronize()
    c["lens"][3] = 1000
    ref = ref_ops.paged_attention_decode(c["q"], c["kc"], c["vc"], c["bt"], c["lens"], 0.09,
                                         k_scale=c["ks"], v_scale=c["vs"])
    np.testing.assert_allclose(out.float().cpu().numpy(), ref, atol=ATOL, rtol=RTOL)


def test_decode_full_size_properties():
    """BASELINE config 2 shape (256 seqs x 4096 ctx, 32/8 heads, D 128, FP8 KV): size-independent
    properties + oracle on a sample of sequences."""
    S, L, Hq, Hkv, D, bs = 256, 4096, 32, 8, 128, 16
    gen = torch.Generator(device=DEV).manual_seed(0)
    nblk = L // bs
    nb = S * nblk + 16
    raw = torch.randint(0, 0x78, (nb, bs, Hkv, 2 * D), dtype=torch.uint8, device=DEV, generator=gen)
    raw |= (torch.randint(0, 2, raw.shape, dtype=torch.uint8, device=DEV, generator=gen) << 7)
    # every V row of sequence 0 is the same vector -> its output must be that vector (x v_scale)
    bt = torch.randperm(nb, device=DEV, generator=gen)[: S * nblk].to(torch.int32).view(S, nblk)
    vrow = torch.randint(0x20, 0x50, (Hkv, D), dtype=torch.uint8, device=DEV, generator=gen)
    raw[bt[0].long(), :, :, D:] = vrow
    q = (torch.rand(S, Hq, D, device=DEV, generator=gen) * 2 - 1).mul(0.2).to(torch.bfloat16)
    lens = torch.full((S,), L, dtype=torch.int32, device=DEV)
    ks = torch.tensor([1.0 / 256], device=DEV)
    vs = torch.tensor([1.0 / 16], device=DEV)
    ops = _ops()
    out = torch.empty_like(q)
    ops.paged_attention_decode(out, q, raw[..., :D], raw[..., D:], bt, lens, D ** -0.5, k_scale=ks,
                               v_scale=vs)
    # (1) constant-V sequence
    exp0 = ref_ops.e4m3_bits_to_f32(vrow.cpu().numpy()) / 16.0
    exp0 = np.repeat(exp0, Hq // Hkv, axis=0)
    np.testing.assert_allclose(out[0].float().cpu().numpy(), exp0, atol=2e-2, rtol=1e-2)
    # (2) linearity in v_scale
    out2 = torch.empty_like(q)
    ops.paged_attention_decode(out2, q, raw[..., :D], raw[..., D:], bt, lens, D ** -0.5,
                               k_scale=ks, v_scale=vs * 2)
    np.testing.assert_allclose(out2.float().cpu().numpy(), 2 * out.float().cpu().numpy(),
                               atol=1e-3, rtol=8e-3)
    # (3) permuting the batch permutes the output (scheduling independence), bit-exact
    perm = torch.randperm(S, device=DEV, generator=gen)
    out3 = torch.empty_like(q)
    ops.paged_attention_decode(out3, q[perm].contiguous(), raw[..., :D], raw[..., D:],
                               bt[perm].contiguous(), lens, D ** -0.5, k_scale=ks, v_scale=vs)
    assert torch.allclose(out3.float(), out[perm].float(), atol=2e-3, rtol=1e-2)
    # (4) oracle on a sample of sequences (gather only their blocks)
    for s in (1, 100, 255):
        blocks = bt[s].long()
        sub = raw[blocks].cpu()
        sbt = torch.arange(nblk, dtype=torch.int32).view(1, -1)
        ref = ref_ops.paged_attention_decode(q[s:s + 1].cpu(), sub[..., :D], sub[..., D:], sbt,
                                             torch.tensor([L], dtype=torch.int32), D ** -0.5,
                                             k_scale=ks.cpu(), v_scale=vs.cpu())
        np.testing.assert_allclose(out[s:s + 1].float().cpu().numpy(), ref, atol=ATOL, rtol=RTOL)


@pytest.mark.parametrize("heads,D", [((32, 8), 128), ((14, 2), 64), ((16, 1), 128), ((8, 8), 256)])
@pytest.mark.parametrize("kv_fp8", [False, True])
def test_fused_fp8_output(heads, D, kv_fp8):
    """output_scale: the kernel writes e4m3(attention / scale) itself (AttentionImpl.forward's fused FP8
    output, backend.py:895-934; triton_unified_attention.py multiplies the fp32 accumulator by 1/scale
    and clamps). Both write paths are exercised: sequences owned by one CTA and split sequences (reduce
    kernel), plus an empty sequence. Bound: the decoded value is within the attention tolerance plus
    one e4m3 rounding step of the oracle."""
    Hq, Hkv = heads
    lens = [4096, 1, 700, 0, 33, 2049, 512, 17]
    c = build_case(lens, Hq, Hkv, D, 16, torch.bfloat16, kv_fp8, seed=11)
    scale = D ** -0.5
    ref = ref_ops.paged_attention_decode(c["q"], c["kc"], c["vc"], c["bt"], c["lens"], scale,
                                         k_scale=c["ks"], v_scale=c["vs"])
    oscale = torch.tensor([float(np.abs(ref).max()) / 200.0])
    raw, kc, vc = to_dev_views(c)
    out = torch.full((len(lens), Hq, D), 0x7F, dtype=torch.uint8, device=DEV).view(torch.float8_e4m3fn)
    dev = lambda t: None if t is None else t.to(DEV)
    _ops().paged_attention_decode(out, c["q"].to(DEV), kc, vc, c["bt"].to(DEV), c["lens"].to(DEV), scale,
                                  k_scale=dev(c["ks"]), v_scale=dev(c["vs"]), output_scale=oscale.to(DEV))
    torch.cuda.synchronize()
    got_bits = out.view(torch.uint8).cpu().numpy()
    assert not (got_bits == 0x7F).any() and not (got_bits == 0xFF).any()      # every byte written, no NaN
    got = ref_ops.e4m3_bits_to_f32(got_bits) * oscale.item()
    # e4m3 has 3 mantissa bits: half a step = 2^-4 relative; plus the attention tolerance
    tol = ATOL + RTOL * np.abs(ref) + 2.0 ** -4 * np.abs(ref) + oscale.item() * 2.0 ** -10
    assert (np.abs(got - ref) <= tol).all(), np.abs(got - ref).max()
    assert (got_bits[3] == 0).all()                                           # empty sequence -> zeros
    with pytest.raises(RuntimeError, match="fp8 exactly when output_scale"):
        _ops().paged_attention_decode(torch.empty(len(lens), Hq, D, dtype=torch.bfloat16, device=DEV),
                                      c["q"].to(DEV), kc, vc, c["bt"].to(DEV), c["lens"].to(DEV), scale,
                                      k_scale=dev(c["ks"]), v_scale=dev(c["vs"]), output_scale=oscale.to(DEV))
