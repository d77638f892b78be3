"""GPU parity: the NVFP4 fused-MoE operators (packed expert-sorted rows) vs the oracle.

get_cutlass_moe_mm_data / shuffle_rows / scaled_fp4_experts_quant / silu_and_mul_scaled_fp4_experts_quant /
cutlass_fp4_group_mm and the whole run_cutlass_moe_fp4 sequence. Mirrors tests/kernels/moe/test_nvfp4_moe.py of the
reference (there: the whole block vs dequantise + torch_moe at atol = rtol = 1e-1); the integer tables and byte
outputs are compared exactly here.
"""
import numpy as np
import pytest
import torch

from oracle import ref_ops

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _ops():
    from b200ops import ops
    return ops


def _moe():
    from b200ops import fused_moe
    return fused_moe


def _route(m, topk, E, seed, drop=False, skew=False):
    rng = np.random.default_rng(seed)
    if skew:      # most pairs on two experts, several experts empty
        p = np.full(E, 0.02 / max(E - 2, 1))
        p[0], p[E // 2] = 0.6, 0.38
        p /= p.sum()
    else:
        p = np.full(E, 1.0 / E)
    ids = np.stack([rng.choice(E, size=topk, replace=False, p=p) for _ in range(m)]).astype(np.int32)
    if drop:
        ids[rng.random(ids.shape) < 0.15] = -1
    return ids


def _mm_data(ids, E, n, k, blockscale=True, is_gated=True):
    ops = _ops()
    numel = ids.size
    t = torch.from_numpy(ids).to(DEV)
    eo = torch.full((E + 1,), -7, dtype=torch.int32, device=DEV)
    bso = torch.full((E + 1,), -7, dtype=torch.int32, device=DEV) if blockscale else None
    ps1 = torch.full((E, 3), -7, dtype=torch.int32, device=DEV)
    ps2 = torch.full((E, 3), -7, dtype=torch.int32, device=DEV)
    a_map = torch.full((numel,), -7, dtype=torch.int32, device=DEV)
    c_map = torch.full((numel,), -7, dtype=torch.int32, device=DEV)
    ops.get_cutlass_moe_mm_data(t, eo, ps1, ps2, a_map, c_map, E, n, k, bso, is_gated)
    return eo, ps1, ps2, a_map, c_map, bso


@pytest.mark.parametrize("m,topk,E", [(1, 1, 8), (2, 6, 40), (64, 8, 64), (224, 8, 256), (256, 2, 8), (1000, 4, 33),
                                      (4096, 8, 288)])
@pytest.mark.parametrize("drop", [False, True])
def test_get_cutlass_moe_mm_data(m, topk, E, drop):
    ids = _route(m, topk, E, seed=m + topk, drop=drop, skew=(E == 64))
    for blockscale in (True, False):
        for gated in (True, False):
            eo, ps1, ps2, a_map, c_map, bso = _mm_data(ids, E, 1024, 1536, blockscale, gated)
            swap = (not blockscale) and ids.size <= 64           # moe_data.cu:254-255
            ref = ref_ops.cutlass_moe_mm_data(ids, E, 1024, 1536, gated, swap, blockscale)
            np.testing.assert_array_equal(eo.cpu().numpy(), ref["expert_offsets"])
            np.testing.assert_array_equal(ps1.cpu().numpy(), ref["problem_sizes1"])
            np.testing.assert_array_equal(ps2.cpu().numpy(), ref["problem_sizes2"])
            # deterministic (ascending flat index inside an expert): the tables are equal, not only set-equal
            np.testing.assert_array_equal(a_map.cpu().numpy(), ref["a_map"])
            np.testing.assert_array_equal(c_map.cpu().numpy(), ref["c_map"])
            if blockscale:
                np.testing.assert_array_equal(bso.cpu().numpy(), ref["blockscale_offsets"])


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32, torch.uint8])
def test_shuffle_rows(dtype):
    ops = _ops()
    g = torch.Generator().manual_seed(3)
    src = (torch.randn(37, 256, generator=g) * 40).to(dtype)
    m = torch.randint(0, 37, (150,), generator=g).to(torch.int32)
    m[5], m[77] = -1, 37                       # no source row -> zeros
    out = ops.shuffle_rows(src.to(DEV), m.to(DEV)).cpu()
    exp = src[m.clamp(0, 36).long()].clone()
    exp[5] = 0
    exp[77] = 0
    assert torch.equal(out.view(torch.uint8), exp.view(torch.uint8))
    # strided source (a column slice of a wider buffer)
    wide = (torch.randn(37, 512, generator=g) * 40).to(dtype).to(DEV)
    out2 = ops.shuffle_rows(wide[:, 128:384], m.to(DEV)).cpu()
    exp2 = wide[:, 128:384].cpu()[m.clamp(0, 36).long()].clone()
    exp2[5] = 0
    exp2[77] = 0
    assert torch.equal(out2.view(torch.uint8), exp2.view(torch.uint8))


def _experts_case(m, topk, E, k, seed, skew=False):
    ids = _route(m, topk, E, seed, skew=skew)
    d = ref_ops.cutlass_moe_mm_data(ids, E, 64, k)
    rows = int(d["expert_offsets"][E])
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn(rows, k, generator=g) * torch.rand(rows, 1, generator=g) * 3).to(torch.bfloat16)
    gs = (448.0 * 6.0 / (torch.rand(E, generator=g) * 8 + 1)).float()
    return ids, d, x, gs


@pytest.mark.parametrize("m,topk,E,k", [(2, 6, 40, 1024), (64, 8, 64, 1536), (224, 8, 256, 1024), (256, 2, 8, 4096),
                                        (300, 1, 3, 160), (200, 2, 8, 14336)])
@pytest.mark.parametrize("gated", [False, True])
def test_scaled_fp4_experts_quant(m, topk, E, k, gated):
    ops = _ops()
    ids, d, x, gs = _experts_case(m, topk, E, k, seed=E + k, skew=(E == 8 and k == 14336))
    if gated:
        g = torch.Generator().manual_seed(9)
        x = torch.cat([x, torch.randn(x.shape, generator=g).to(x.dtype)], dim=1)
    eo = torch.from_numpy(d["expert_offsets"]).to(DEV)
    bso = torch.from_numpy(d["blockscale_offsets"]).to(DEV)
    fn = ops.silu_and_mul_scaled_fp4_experts_quant if gated else ops.scaled_fp4_experts_quant
    out, out_sf = fn(x.to(DEV), gs.to(DEV), eo, bso, topk)
    rows = x.shape[0]
    assert out.shape == (rows, k // 2) and out.dtype == torch.uint8 and out_sf.dtype == torch.float8_e4m3fn
    codes, sf_flat, sf_rows, y = ref_ops.nvfp4_experts_quant(x, gs.numpy(), d["expert_offsets"],
                                                             d["blockscale_offsets"], gated=gated, return_scaled=True)
    got_codes = ref_ops.nvfp4_unpack(out.cpu().numpy())
    got_flat = out_sf.view(torch.uint8).cpu().numpy().reshape(-1)
    rc = (k // 16 + 3) // 4 * 4
    # per expert: un-swizzle its own matrix at blockscale_offsets[e]; scales of valid rows are exact
    got_rows = np.zeros_like(sf_rows)
    for e in range(E):
        lo, hi = int(d["expert_offsets"][e]), int(d["expert_offsets"][e + 1])
        if hi == lo:
            continue
        base = int(d["blockscale_offsets"][e]) * rc
        n_pad = (hi - lo + 127) // 128 * 128
        got_rows[lo:hi] = ref_ops.unswizzle_sf(got_flat[base:base + n_pad * rc], hi - lo, k // 16)
    sf_mismatch = (got_rows != sf_rows).mean()
    assert sf_mismatch <= 1e-3, f"scale mismatch fraction {sf_mismatch}"
    same = np.repeat(got_rows == sf_rows, 16, axis=1)
    bad = (got_codes != codes) & same & ~ref_ops.nvfp4_near_tie(y)
    # the gated variant computes silu with __expf / __fdividef (the reference's own, nvfp4_utils.cuh:304-307)
    limit = 2e-3 * bad.size if gated else 0
    assert bad.sum() <= limit, f"{bad.sum()} fp4 codes differ away from a rounding tie"
    # bit-identity with the per-tensor quantiser (itself bit-identical to the wheel's kernel, test_gpu_vs_wheel.py):
    # expert e's rows with global scale gs[e] give the same codes and the same swizzled scale matrix
    single = ops.silu_and_mul_nvfp4_quant if gated else None
    for e in range(E):
        lo, hi = int(d["expert_offsets"][e]), int(d["expert_offsets"][e + 1])
        if hi == lo:
            continue
        xe = x[lo:hi].to(DEV)
        if gated:
            o1 = torch.empty(hi - lo, k // 2, dtype=torch.uint8, device=DEV)
            s1 = torch.zeros((hi - lo + 127) // 128 * 128, rc, dtype=torch.uint8, device=DEV)
            single(o1, s1.view(torch.float8_e4m3fn), xe, gs[e:e + 1].to(DEV))
        else:
            o1, s1 = ops.scaled_fp4_quant(xe, gs[e:e + 1].to(DEV))
        assert torch.equal(o1, out[lo:hi]), f"expert {e}: codes differ from the per-tensor quantiser"
        exp_sf = s1.view(torch.uint8).cpu().numpy().reshape(-1)
        base = int(d["blockscale_offsets"][e]) * rc
        mask = ref_ops.swizzle_sf(np.ones((hi - lo, k // 16), dtype=np.uint8)) != 0   # valid rows and columns
        np.testing.assert_array_equal(got_flat[base:base + exp_sf.size][mask], exp_sf[mask])
        if e > 8:
            break


def _rand_weights(E, n, k, seed):
    """NVFP4 expert weights through our own quantiser (generated on the device, one expert at a time): codes
    uint8 [E, n, k/2], swizzled scales per expert, global scales [E]."""
    ops = _ops()
    g = torch.Generator(device=DEV).manual_seed(seed)
    rn, rc = (n + 127) // 128 * 128, (k // 16 + 3) // 4 * 4
    q = torch.empty(E, n, k // 2, dtype=torch.uint8, device=DEV)
    sf = torch.empty(E, rn, rc, dtype=torch.uint8, device=DEV)
    gs = torch.empty(E, dtype=torch.float32, device=DEV)
    for e in range(E):
        w = (torch.randn(n, k, generator=g, device=DEV) / 8).to(torch.bfloat16)
        gs[e] = 448.0 * 6.0 / w.float().abs().max()
        qe, se = ops.scaled_fp4_quant(w, gs[e:e + 1])
        q[e] = qe
        sf[e] = se.view(torch.uint8).view(rn, rc)
    return None, q, sf.view(torch.float8_e4m3fn), gs.cpu()


def _unswizzle_w(sf, E, n, k):
    flat = sf.view(torch.uint8).cpu().numpy().reshape(E, -1)
    return np.stack([ref_ops.unswizzle_sf(flat[e], n, k // 16) for e in range(E)])


@pytest.mark.parametrize("m,topk,E,n,k", [(2, 6, 40, 256, 1024), (64, 8, 64, 512, 1024), (224, 8, 32, 384, 1536),
                                          (256, 2, 8, 1024, 512), (300, 1, 3, 200, 320), (77, 2, 4, 136, 256)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_cutlass_fp4_group_mm(m, topk, E, n, k, dtype):
    ops = _ops()
    ids = _route(m, topk, E, seed=n + k, skew=(E in (8, 32)))
    eo, ps1, ps2, a_map, c_map, bso = _mm_data(ids, E, n, k, True, False)
    rows = m * topk
    g = torch.Generator().manual_seed(5)
    x = torch.randn(rows, k, generator=g).to(dtype)
    a_gs = (448.0 * 6.0 / (torch.rand(E, generator=g) * 4 + 2)).float()
    a4, a_sf = ops.scaled_fp4_experts_quant(x.to(DEV), a_gs.to(DEV), eo, bso, topk)
    w, w4, w_sf, w_gs = _rand_weights(E, n, k, seed=11)
    alphas = (1.0 / (a_gs * w_gs)).to(DEV)
    out = torch.full((rows, n), float("nan"), dtype=dtype, device=DEV)
    ops.cutlass_fp4_moe_mm(out, a4, w4, a_sf, w_sf, alphas, ps1, eo[:-1], bso[:-1])
    # oracle on the SAME quantised bits
    off = eo.cpu().numpy()
    rc = (k // 16 + 3) // 4 * 4
    a_codes = ref_ops.nvfp4_unpack(a4.cpu().numpy())
    flat = a_sf.view(torch.uint8).cpu().numpy().reshape(-1)
    a_rows = np.zeros((rows, k // 16), dtype=np.uint8)
    bs = bso.cpu().numpy()
    for e in range(E):
        lo, hi = int(off[e]), int(off[e + 1])
        if hi > lo:
            n_pad = (hi - lo + 127) // 128 * 128
            a_rows[lo:hi] = ref_ops.unswizzle_sf(flat[int(bs[e]) * rc:int(bs[e]) * rc + n_pad * rc], hi - lo, k // 16)
    w_codes = np.stack([ref_ops.nvfp4_unpack(w4[e].cpu().numpy()) for e in range(E)])
    ref = ref_ops.nvfp4_group_mm(a_codes, a_rows, w_codes, _unswizzle_w(w_sf, E, n, k), alphas.cpu().numpy(), off)
    got = out.float().cpu().numpy()
    assert not np.isnan(got).any(), "a packed row was not written"
    tol = 2.0 ** -7 * np.abs(ref) + 1e-3 * np.abs(ref).max()
    bad = np.abs(got - ref) > tol
    assert not bad.any(), f"{bad.sum()} / {bad.size} outside tolerance, max err {np.abs(got - ref).max():.4g}"


def test_cutlass_fp4_group_mm_exact_small_integers_and_untouched_rows():
    """All block scales 1, codes in {0, +-1, +-2}, alphas powers of two: bit-exact whatever the accumulation
    order; experts with 0 / 1 / 129 rows; rows past the last expert (dropped pairs) stay untouched."""
    ops = _ops()
    E, n, k = 6, 384, 512
    counts = np.array([129, 0, 1, 64, 0, 70])
    off = np.concatenate([[0], np.cumsum(counts)]).astype(np.int32)
    bso = np.concatenate([[0], np.cumsum((counts + 127) // 128 * 128)]).astype(np.int32)
    rows = int(off[E]) + 5                       # 5 packed rows belong to nobody
    rng = np.random.default_rng(1)
    vals = np.array([0, 2, 4, 10, 12], dtype=np.uint8)
    ac = vals[rng.integers(0, 5, (rows, k))]
    wc = vals[rng.integers(0, 5, (E, n, k))]
    rc = k // 16
    a_sf = np.zeros(int(bso[E]) * rc, dtype=np.uint8)
    for e in range(E):
        if counts[e]:
            sw = ref_ops.swizzle_sf(np.full((counts[e], rc), 0x38, dtype=np.uint8))
            a_sf[int(bso[e]) * rc:int(bso[e]) * rc + sw.size] = sw
    # rows of the scale matrix past an expert's last row are undefined in the reference: poison them
    a_sf[a_sf == 0] = 0x7F                                             # NaN in e4m3
    w_sf = np.stack([ref_ops.swizzle_sf(np.full((n, rc), 0x38, dtype=np.uint8)) for _ in range(E)])
    alphas = np.array([1, 2, 0.5, 4, 1, 0.25], dtype=np.float32)
    ps = np.stack([counts, np.full(E, n), np.full(E, k)], axis=1).astype(np.int32)
    out = torch.full((rows, n), 777.0, dtype=torch.float16, device=DEV)
    ops.cutlass_fp4_group_mm(out, torch.from_numpy(ref_ops.nvfp4_pack(ac)).to(DEV),
                             torch.from_numpy(np.stack([ref_ops.nvfp4_pack(wc[e]) for e in range(E)])).to(DEV),
                             torch.from_numpy(a_sf).to(DEV).view(torch.float8_e4m3fn),
                             torch.from_numpy(w_sf).to(DEV).view(torch.float8_e4m3fn),
                             torch.from_numpy(alphas).to(DEV), torch.from_numpy(ps).to(DEV),
                             torch.from_numpy(off[:-1].copy()).to(DEV), torch.from_numpy(bso[:-1].copy()).to(DEV))
    one = np.full((rows, rc), 0x38, dtype=np.uint8)
    ref = ref_ops.nvfp4_group_mm(ac, one, wc, np.full((E, n, rc), 0x38, dtype=np.uint8), alphas, off)
    got = out.float().cpu().numpy()
    assert np.abs(ref).max() < 2048
    np.testing.assert_array_equal(got[:off[E]], ref[:off[E]])
    assert (got[off[E]:] == 777.0).all()


@pytest.mark.parametrize("m,n,k,e,topk", [(2, 1024, 1024, 40, 6), (64, 1024, 1024, 64, 8), (224, 1024, 1536, 40, 6),
                                          (64, 256, 1024, 256, 8), (256, 1408, 1024, 8, 2), (33, 512, 512, 16, 1)])
def test_cutlass_moe_fp4_block(m, n, k, e, topk):
    """The whole run_cutlass_moe_fp4 sequence vs (a) the operator-by-operator oracle and (b) the reference test's
    own criterion: dequantised weights / activations through a dense MoE at atol = rtol = 1e-1
    (tests/kernels/moe/test_nvfp4_moe.py:121-160)."""
    moe = _moe()
    g = torch.Generator().manual_seed(7)
    a = (torch.randn(m, k, generator=g) / 10).to(torch.bfloat16)
    w1, w1q, w1sf, w1gs = _rand_weights(e, 2 * n, k, seed=21)
    w2, w2q, w2sf, w2gs = _rand_weights(e, k, n, seed=22)
    score = torch.randn(m, e, generator=g)
    tw, ti = torch.topk(torch.softmax(score, dim=-1), topk, dim=-1)
    ti = ti.to(torch.int32)
    a1_gs = torch.ones(e)
    a2_gs = torch.ones(e)
    out = moe.cutlass_moe_fp4(a.to(DEV), a1_gs.to(DEV), w1q, w1sf, (1 / w1gs).to(DEV), a2_gs.to(DEV), w2q, w2sf,
                              (1 / w2gs).to(DEV), tw.to(DEV), ti.to(DEV)).cpu()
    w1c = np.stack([ref_ops.nvfp4_unpack(w1q[i].cpu().numpy()) for i in range(e)])
    w2c = np.stack([ref_ops.nvfp4_unpack(w2q[i].cpu().numpy()) for i in range(e)])
    w1s, w2s = _unswizzle_w(w1sf, e, 2 * n, k), _unswizzle_w(w2sf, e, k, n)
    ref = ref_ops.cutlass_moe_fp4(a, a1_gs.numpy(), w1c, w1s, (1 / w1gs).numpy(), a2_gs.numpy(), w2c, w2s,
                                  (1 / w2gs).numpy(), tw, ti.numpy())
    err = (out.float() - ref.float()).abs()
    scale = ref.float().abs().max().item()
    # W4A4 drift: a 1-ulp difference of an intermediate bf16 value can move a 4-bit code by a whole step
    assert err.max().item() <= 0.08 * scale + 1e-3, f"max err {err.max().item():.4g} (scale {scale:.4g})"
    assert err.mean().item() <= 0.01 * scale + 1e-4
    # (b) dense model on dequantised operands
    w1d = torch.from_numpy(np.stack([ref_ops.nvfp4_dequant(w1c[i], w1s[i], float(w1gs[i])) for i in range(e)]))
    w2d = torch.from_numpy(np.stack([ref_ops.nvfp4_dequant(w2c[i], w2s[i], float(w2gs[i])) for i in range(e)]))
    dense = torch.zeros(m, k)
    ac, asf = ref_ops.nvfp4_quant(a, 1.0)            # the reference test feeds the dequantised activations
    af = torch.from_numpy(ref_ops.nvfp4_dequant(ac, asf, 1.0)).to(torch.bfloat16).float()
    for t in range(m):
        for j in range(topk):
            ex = int(ti[t, j])
            h = w1d[ex] @ af[t]
            h = torch.nn.functional.silu(h[:n]) * h[n:]
            dense[t] += tw[t, j] * (w2d[ex] @ h)
    # the reference asserts every element at its own sizes; the dense model does not quantise the intermediate, so
    # at 344k outputs a handful of W4A4 outliers sit just outside: bound their number and their size
    outside = (out.float() - dense).abs() > 1e-1 + 1e-1 * dense.abs()
    assert outside.float().mean().item() <= 2e-5, f"{int(outside.sum())} elements outside atol = rtol = 1e-1"
    assert (out.float() - dense).abs().max().item() <= 0.2


def test_cutlass_moe_fp4_dropped_pairs_and_router_weight_on_input():
    moe = _moe()
    m, n, k, e = 48, 256, 512, 8
    g = torch.Generator().manual_seed(3)
    a = (torch.randn(m, k, generator=g) / 10).to(torch.bfloat16)
    _, w1q, w1sf, w1gs = _rand_weights(e, 2 * n, k, seed=31)
    _, w2q, w2sf, w2gs = _rand_weights(e, k, n, seed=32)
    tw = torch.rand(m, 2, generator=g)
    ti = torch.stack([torch.randperm(e, generator=g)[:2] for _ in range(m)]).to(torch.int32)
    one = torch.ones(e, device=DEV)
    args = (one, w1q, w1sf, (1 / w1gs).to(DEV), one, w2q, w2sf, (1 / w2gs).to(DEV))
    full = moe.cutlass_moe_fp4(a.to(DEV), *args, tw.to(DEV), ti.to(DEV)).float().cpu()
    # dropping the second choice of every token == running with topk = 1 on the first choice
    ti_drop = ti.clone()
    ti_drop[:, 1] = -1
    dropped = moe.cutlass_moe_fp4(a.to(DEV), *args, tw.to(DEV), ti_drop.to(DEV)).float().cpu()
    first = moe.cutlass_moe_fp4(a.to(DEV), *args, tw[:, :1].contiguous().to(DEV), ti[:, :1].contiguous().to(DEV)).float().cpu()
    assert torch.equal(dropped, first)
    assert not torch.equal(full, first)
    # router weight on the input (topk = 1): a * w quantised, no weighting at the end
    rw = moe.cutlass_moe_fp4(a.to(DEV), *args, tw[:, :1].contiguous().to(DEV), ti[:, :1].contiguous().to(DEV),
                             apply_router_weight_on_input=True).float().cpu()
    pre = (a * tw[:, :1].to(torch.bfloat16))
    exp = moe.cutlass_moe_fp4(pre.to(DEV), *args, torch.ones(m, 1, device=DEV), ti[:, :1].contiguous().to(DEV)).float().cpu()
    assert torch.equal(rw, exp)


def test_nvfp4_moe_validation():
    ops = _ops()
    z8 = torch.zeros(4, 32, dtype=torch.uint8, device=DEV)
    w8 = torch.zeros(2, 64, 32, dtype=torch.uint8, device=DEV)
    sf = torch.zeros(128, 4, dtype=torch.uint8, device=DEV).view(torch.float8_e4m3fn)
    wsf = torch.zeros(2, 128, 4, dtype=torch.uint8, device=DEV).view(torch.float8_e4m3fn)
    i32 = torch.zeros(2, dtype=torch.int32, device=DEV)
    ps = torch.zeros(2, 3, dtype=torch.int32, device=DEV)
    out = torch.zeros(4, 64, dtype=torch.bfloat16, device=DEV)
    with pytest.raises(RuntimeError, match="padded and swizzled per expert"):
        ops.cutlass_fp4_group_mm(out, z8, w8, sf, wsf[:1], torch.ones(2, device=DEV), ps, i32, i32)
    with pytest.raises(RuntimeError, match="int32"):
        ops.cutlass_fp4_group_mm(out, z8, w8, sf, wsf, torch.ones(2, device=DEV), ps.long(), i32, i32)
    with pytest.raises(RuntimeError, match="alphas must be float32"):
        ops.cutlass_fp4_group_mm(out, z8, w8, sf, wsf, torch.ones(2, device=DEV).half(), ps, i32, i32)
    x = torch.zeros(4, 40, dtype=torch.bfloat16, device=DEV)
    with pytest.raises(RuntimeError, match="multiple of 16"):
        ops.scaled_fp4_experts_quant(x, torch.ones(2, device=DEV), torch.zeros(3, dtype=torch.int32, device=DEV),
                                     torch.zeros(3, dtype=torch.int32, device=DEV), 1)
    with pytest.raises(RuntimeError, match="CUDA"):
        ops.shuffle_rows(torch.zeros(4, 64), torch.zeros(4, dtype=torch.int32))
