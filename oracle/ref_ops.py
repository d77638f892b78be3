"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's algorithms for the decode hot path.

Nothing under ``oracle/`` is imported by the product (``vllm-fork_b200/``); only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` use it,
and only as the checker / the timed CPU baseline.

Each function cites the reference file:line (relative to /root/reference) it restates.

Parity pinning (SURVEY.md section 8c -- the reference holds NO golden-vector files for this path):
  * paged attention / KV-cache write: pinned against outputs of the reference's OWN CPU kernels
    (csrc/cpu/cpu_attn.cpp, compiled unmodified into oracle/_ref/refcpu.so by
    oracle/build_ref_cpu_attn.sh) -> fixtures tests/golden/cpu_attn_*.npz made by
    tests/golden/make_golden.py; checked in tests/test_oracle.py.
  * e4m3 conversion: pinned bit-for-bit against torch's float8_e4m3fn cast for all finite inputs
    in range (tests/test_oracle.py) and against the E4M3 code table.
  * moe_align_block_size: pinned against the worked example in
    vllm/model_executor/layers/fused_moe/moe_align_block_size.py:59-72.
  * quantisers: restated from the CUDA sources; the reference's own oracle
    (tests/kernels/quant_utils.py:22-88) is restated beside them and cross-checked.
"""
from __future__ import annotations

import numpy as np
import torch

FP8_MAX = 448.0


# --------------------------------------------------------------------------------------------
# e4m3 (fn) conversion, RNE + saturate-to-finite.
# Reference: __nv_cvt_float_to_fp8(x, __NV_SATFINITE, __NV_E4M3)
#            csrc/quantization/w8a8/fp8/nvidia/quant_utils.cuh:492-503
# --------------------------------------------------------------------------------------------
def f32_to_e4m3_bits(x) -> np.ndarray:
    x = np.asarray(x, dtype=np.float32)
    sign = (np.signbit(x)).astype(np.uint8) << 7
    a = np.abs(x).astype(np.float64)
    nan = np.isnan(a)
    a = np.where(nan, 0.0, a)
    a = np.minimum(a, FP8_MAX)  # satfinite (also maps +-inf to +-448)
    out = np.zeros(a.shape, dtype=np.int64)
    sub = a < 2.0 ** -6
    # subnormals: multiples of 2^-9; rint = round-half-even; q == 8 rolls into the first normal
    out_sub = np.rint(a * 2.0 ** 9).astype(np.int64)
    m, e = np.frexp(np.where(sub, 1.0, a))  # a = m * 2^e, m in [0.5, 1)
    e = e - 1
    frac = m * 2.0 - 1.0  # [0, 1)
    q = np.rint(frac * 8.0).astype(np.int64)
    carry = q == 8
    e = np.where(carry, e + 1, e)
    q = np.where(carry, 0, q)
    out_norm = ((e + 7) << 3) | q
    out_norm = np.minimum(out_norm, 0x7E)
    out = np.where(sub, out_sub, out_norm)
    out = np.where(nan, 0x7F, out)
    return (out.astype(np.uint8) | sign).astype(np.uint8)


def e4m3_bits_to_f32(b) -> np.ndarray:
    b = np.asarray(b, dtype=np.uint8).astype(np.int64)
    s = np.where(b & 0x80, -1.0, 1.0)
    e = (b >> 3) & 0xF
    m = b & 7
    v = np.where(e == 0, m / 8.0 * 2.0 ** -6, (1.0 + m / 8.0) * np.exp2((e - 7).astype(np.float64)))
    v = np.where((e == 15) & (m == 7), np.nan, v)
    return (s * v).astype(np.float32)


def _f32(t: torch.Tensor) -> np.ndarray:
    return t.detach().to(torch.float32).cpu().numpy()


def _fminf(a, b):
    """CUDA fminf: returns the non-NaN operand."""
    return np.where(np.isnan(a), b, np.where(np.isnan(b), a, np.minimum(a, b)))


def _fmaxf(a, b):
    return np.where(np.isnan(a), b, np.where(np.isnan(b), a, np.maximum(a, b)))


def _scaled_fp8_conversion(val: np.ndarray, scale, inverted: bool) -> np.ndarray:
    """csrc/quantization/w8a8/fp8/common.cuh:58-79 (fp32 arithmetic, IEEE division)."""
    val = val.astype(np.float32)
    scale = np.asarray(scale, dtype=np.float32)
    with np.errstate(all="ignore"):
        x = (val * scale).astype(np.float32) if inverted else (val / scale).astype(np.float32)
    r = _fmaxf(np.float32(-FP8_MAX), _fminf(x, np.float32(FP8_MAX)))
    return f32_to_e4m3_bits(r)


# --------------------------------------------------------------------------------------------
# KV-cache write.  csrc/libtorch_stable/cache_kernels.cu:314-400 (kernel), :795-864 (host)
# --------------------------------------------------------------------------------------------
def reshape_and_cache_flash(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype,
                            k_scale, v_scale) -> None:
    """In-place on CPU tensors. key/value [T,H,D]; caches [nb, bs, H, D] views; slot < 0 skipped;
    num_tokens = slot_mapping.size(0). fp8: byte = e4m3(float(x) / scale) (true division)."""
    T = slot_mapping.shape[0]
    bs = key_cache.shape[1]
    H = key.shape[1]
    fp8 = kv_cache_dtype in ("fp8", "fp8_e4m3")
    ks = _f32(k_scale).reshape(-1)
    vs = _f32(v_scale).reshape(-1)
    sm = slot_mapping.cpu().tolist()
    for t in range(T):
        slot = sm[t]
        if slot < 0:
            continue
        blk, off = slot // bs, slot % bs
        if not fp8:
            key_cache[blk, off] = key[t]
            value_cache[blk, off] = value[t]
        else:
            kf, vf = _f32(key[t]), _f32(value[t])
            ksc = ks if ks.size == 1 else ks.reshape(H, 1)
            vsc = vs if vs.size == 1 else vs.reshape(H, 1)
            with np.errstate(all="ignore"):
                kb = f32_to_e4m3_bits((kf / ksc).astype(np.float32))
                vb = f32_to_e4m3_bits((vf / vsc).astype(np.float32))
            key_cache[blk, off].view(torch.uint8).copy_(torch.from_numpy(kb))
            value_cache[blk, off].view(torch.uint8).copy_(torch.from_numpy(vb))


def reshape_and_cache(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype, k_scale,
                      v_scale) -> None:
    """Legacy layout. csrc/libtorch_stable/cache_kernels.cu:254-312.
    key_cache [nb, H, D/x, bs, x]; value_cache [nb, H, D, bs]."""
    T = slot_mapping.shape[0]
    nb, H, Dx, bs, x = key_cache.shape
    fp8 = kv_cache_dtype in ("fp8", "fp8_e4m3")
    sm = slot_mapping.cpu().tolist()
    ks = float(_f32(k_scale).reshape(-1)[0])
    vs = float(_f32(v_scale).reshape(-1)[0])
    for t in range(T):
        slot = sm[t]
        if slot < 0:
            continue
        blk, off = slot // bs, slot % bs
        if not fp8:
            key_cache[blk, :, :, off, :] = key[t].reshape(H, Dx, x)
            value_cache[blk, :, :, off] = value[t]
        else:
            with np.errstate(all="ignore"):
                kb = f32_to_e4m3_bits((_f32(key[t]) / np.float32(ks)).astype(np.float32))
                vb = f32_to_e4m3_bits((_f32(value[t]) / np.float32(vs)).astype(np.float32))
            key_cache.view(torch.uint8)[blk, :, :, off, :] = torch.from_numpy(kb).reshape(H, Dx, x)
            value_cache.view(torch.uint8)[blk, :, :, off] = torch.from_numpy(vb)


# --------------------------------------------------------------------------------------------
# FP8 quantisers.  csrc/libtorch_stable/quantization/w8a8/fp8/common.cu:13-179
# Return (uint8 bits [M, K], float32 scales).
# --------------------------------------------------------------------------------------------
def static_scaled_fp8_quant(x: torch.Tensor, scale: torch.Tensor, group_shape=None):
    """common.cu:13-76 + host :183-335. out = e4m3(clamp(x * (1.0f / scale)))."""
    xf = _f32(x).reshape(-1, x.shape[-1])
    M, K = xf.shape
    sc = _f32(scale)
    if sc.size == 1:
        full = np.full((M, K), sc.reshape(-1)[0], dtype=np.float32)
    elif sc.ndim == 1:
        assert group_shape is not None
        gm = M if group_shape[0] == -1 else group_shape[0]
        gn = K if group_shape[1] == -1 else group_shape[1]
        grid = sc.reshape(M // gm, K // gn)
        full = np.repeat(np.repeat(grid, gm, axis=0), gn, axis=1)
    else:
        gm, gn = M // sc.shape[0], K // sc.shape[1]
        full = np.repeat(np.repeat(sc, gm, axis=0), gn, axis=1)
    with np.errstate(all="ignore"):
        inv = (np.float32(1.0) / full).astype(np.float32)
    return _scaled_fp8_conversion(xf, inv, True).reshape(x.shape)


def dynamic_scaled_fp8_quant(x: torch.Tensor):
    """common.cu:78-135 + host :337-381. scale = max_rows(rowmax / 448); out via 1/scale multiply."""
    xf = _f32(x).reshape(-1, x.shape[-1])
    rowmax = np.max(np.abs(xf), axis=1) if xf.size else np.zeros(0, np.float32)
    # fmaxf-based reduction starting from 0 drops NaNs
    rowmax = np.where(np.isnan(rowmax), np.float32(0), rowmax).astype(np.float32)
    scale = np.float32(0.0)
    for rm in rowmax:
        scale = max(scale, np.float32(rm / np.float32(FP8_MAX)))
    scale = np.float32(scale)
    with np.errstate(all="ignore"):
        inv = np.float32(1.0) / scale
    return _scaled_fp8_conversion(xf, inv, True).reshape(x.shape), np.array([scale], np.float32)


def dynamic_per_token_scaled_fp8_quant(x: torch.Tensor, scale_ub: torch.Tensor | None = None):
    """common.cu:137-179. scale_t = max(min(absmax_t, ub) / 448, 1/(448*512)); out = x / scale_t."""
    xf = _f32(x).reshape(-1, x.shape[-1])
    absmax = np.max(np.abs(xf), axis=1, keepdims=True).astype(np.float32)
    if scale_ub is not None:
        absmax = np.minimum(absmax, _f32(scale_ub).reshape(-1)[0]).astype(np.float32)
    min_sf = np.float32(1.0) / (np.float32(FP8_MAX) * np.float32(512.0))
    scales = np.maximum((absmax / np.float32(FP8_MAX)).astype(np.float32), min_sf).astype(np.float32)
    return _scaled_fp8_conversion(xf, scales, False).reshape(x.shape), scales


def ref_dynamic_per_token_quant(x: torch.Tensor, scale_ub: torch.Tensor | None = None):
    """The reference's own python oracle, tests/kernels/quant_utils.py:22-60 (restated): returns
    (torch fp8 tensor, scales) -- used to cross-check the restatement above."""
    qmax = torch.finfo(torch.float8_e4m3fn).max
    x32 = x.to(torch.float32)
    x_token_max = x32.abs().max(dim=-1, keepdim=True)[0]
    if scale_ub is not None:
        x_token_max = x_token_max.clamp(max=scale_ub.to(torch.float32))
    scales = (x_token_max / qmax).clamp(min=1.0 / (qmax * 512.0))
    out = (x32 / scales).clamp(-qmax, qmax).to(torch.float8_e4m3fn)
    return out, scales


def per_token_group_quant_fp8(x: torch.Tensor, group_size: int, eps: float = 1e-10,
                              use_ue8m0: bool = False):
    """csrc/libtorch_stable/quantization/w8a8/fp8/per_token_group_quant.cu:43-161.
    absmax (from eps, fmaxf drops NaN) / 448 (fp32 division) [-> 2^ceil(log2(.)) in fp32];
    q = fminf(fmaxf(x / y_s, -448), 448) -> e4m3.  Returns (uint8 bits like x, float32 [M, K/g])."""
    xf = _f32(x).reshape(-1, x.shape[-1])
    M, K = xf.shape
    g = xf.reshape(M, K // group_size, group_size)
    a = np.abs(g)
    a = np.where(np.isnan(a), np.float32(0), a)
    absmax = np.maximum(a.max(axis=2) if a.size else np.zeros((M, K // group_size), np.float32),
                        np.float32(eps)).astype(np.float32)
    ys = (absmax / np.float32(FP8_MAX)).astype(np.float32)
    if use_ue8m0:
        with np.errstate(all="ignore"):
            lg = np.log2(np.maximum(np.abs(ys), np.float32(1e-10)).astype(np.float32)).astype(np.float32)
        ys = np.exp2(np.ceil(lg).astype(np.float64)).astype(np.float32)
    with np.errstate(all="ignore"):
        q = (g / ys[:, :, None]).astype(np.float32)
    r = _fminf(_fmaxf(q, np.float32(-FP8_MAX)), np.float32(FP8_MAX))
    return f32_to_e4m3_bits(r).reshape(x.shape), ys


def ue8m0_near_tie(absmax_over_448: np.ndarray, rel: float = 1e-6) -> np.ndarray:
    """Groups whose scale sits within `rel` of a power of two: ceil(log2f(.)) may legitimately differ
    between libm implementations there (tests exclude them from the bit-exact check)."""
    m, _ = np.frexp(absmax_over_448.astype(np.float64))
    return (np.abs(m - 0.5) < rel) | (np.abs(m - 1.0) < rel)


def ref_native_per_token_group_quant_fp8(x: torch.Tensor, group_size: int, eps: float = 1e-10):
    """The reference's own torch oracle, tests/kernels/quant_utils.py:157-180 (restated) -- used to
    cross-check the restatement above."""
    fi = torch.finfo(torch.float8_e4m3fn)
    x_ = x.reshape(x.numel() // group_size, group_size)
    amax = x_.abs().max(dim=-1, keepdim=True)[0].clamp(min=eps).to(torch.float32)
    x_s = amax / fi.max
    x_q = (x_ / x_s).clamp(min=fi.min, max=fi.max).to(torch.float8_e4m3fn)
    return x_q.reshape(x.shape), x_s.reshape(x.shape[:-1] + (x.shape[-1] // group_size,))


def w8a8_block_scaled_mm(a_bits, b_bits_nk, a_s, b_s, block=(128, 128)) -> np.ndarray:
    """tests/kernels/quant_utils.py:91-155 (native_w8a8_block_matmul) restated in float64:
    C[:, nblk] = sum_kblk (A_kblk . B[nblk, kblk]^T) * a_s[:, kblk] * b_s[nblk, kblk].
    a_bits uint8 [M,K], b_bits_nk uint8 [N,K], a_s float32 [M, ceil(K/bk)], b_s float32
    [ceil(N/bn), ceil(K/bk)]; returns float32 [M,N] (no output rounding)."""
    a = e4m3_bits_to_f32(np.asarray(a_bits)).astype(np.float64)
    w = e4m3_bits_to_f32(np.asarray(b_bits_nk)).astype(np.float64)
    a_s = np.asarray(a_s, dtype=np.float64)
    b_s = np.asarray(b_s, dtype=np.float64)
    bn, bk = block
    M, K = a.shape
    N = w.shape[0]
    out = np.zeros((M, N), np.float64)
    for kb in range((K + bk - 1) // bk):
        ks = slice(kb * bk, min(K, (kb + 1) * bk))
        part = a[:, ks] @ w[:, ks].T
        scale_n = np.repeat(b_s[:, kb], bn)[:N]
        out += part * a_s[:, kb:kb + 1] * scale_n[None, :]
    return out.astype(np.float32)


# --------------------------------------------------------------------------------------------
# Slot mapping.  vllm/v1/worker/block_table.py:182-211, 379-441 (no-CP form and the DCP interleave branch)
# --------------------------------------------------------------------------------------------
PAD_SLOT_ID = -1


def compute_slot_mapping(block_table: np.ndarray, req_indices: np.ndarray, positions: np.ndarray,
                         block_size: int) -> np.ndarray:
    """slot = block_table[req, pos // bs] * bs + pos % bs  (int64)."""
    blk = block_table[req_indices, positions // block_size].astype(np.int64)
    return blk * block_size + positions % block_size


def compute_slot_mapping_dcp(block_table: np.ndarray, req_indices: np.ndarray, positions: np.ndarray,
                             block_size: int, cp_world_size: int, cp_rank: int,
                             cp_kv_cache_interleave_size: int = 1, kv_cache_block_size: int | None = None,
                             blocks_per_kv_block: int = 1) -> np.ndarray:
    """Slot mapping of ONE decode-context-parallel rank (vllm/v1/worker/block_table.py:413-441, the body of
    _compute_slot_mapping_kernel). The KV of a request is dealt to the ranks in runs of `interleave` tokens: a
    "virtual block" of kv_cache_block_size * world consecutive positions fills one local block on every rank.
    Positions that live on another rank get PAD_SLOT_ID (the cache write skips negative slots).
    Parity: no golden vector for this branch exists in the reference tests (tests/v1/worker/test_gpu_block_table.py
    only constructs BlockTable with interleave 1) - pinned by the hand-worked case and the partition properties in
    tests/test_oracle_dcp.py, and by get_dcp_local_seq_lens below (a second, independent reference function)."""
    kvbs = block_size if kv_cache_block_size is None else kv_cache_block_size
    W, I = int(cp_world_size), int(cp_kv_cache_interleave_size)
    pos = np.asarray(positions).astype(np.int64)
    virtual_block_size = kvbs * W                                              # :413
    vbi = pos // virtual_block_size                                            # :419
    vbo = pos - vbi * virtual_block_size                                       # :420
    is_local = (vbo // I) % W == cp_rank                                       # :421-423
    local_off = (vbo // (W * I)) * I + vbo % I                                 # :424-428
    block_idx = vbi * blocks_per_kv_block + local_off // block_size            # :430-433
    safe_idx = np.where(is_local, block_idx, 0)                                # masked load, other=0 (:434-438)
    blk = block_table[np.asarray(req_indices), safe_idx].astype(np.int64)
    slots = blk * block_size + local_off % block_size                          # :439-440
    return np.where(is_local, slots, PAD_SLOT_ID).astype(np.int64)             # :441


def dcp_local_seq_lens(seq_lens: np.ndarray, dcp_size: int, dcp_rank: int,
                       cp_kv_cache_interleave_size: int = 1) -> np.ndarray:
    """KV tokens of each sequence held by one rank (get_dcp_local_seq_lens,
    vllm/v1/attention/backends/utils.py:958-995, dcp_rank given)."""
    L = np.asarray(seq_lens).astype(np.int32)
    I = int(cp_kv_cache_interleave_size)
    base = L // I // dcp_size * I
    rem = np.clip(L - base * dcp_size - dcp_rank * I, 0, I)
    return (base + rem).astype(np.int32)


# --------------------------------------------------------------------------------------------
# Paged attention decode.
# Semantics: ref_paged_attn, tests/kernels/attention/test_triton_unified_attention.py:33-88 with
# query_len == 1 (flash layout [nb, bs, Hkv, D]); FP8 KV dequant = float(byte) * scale
# (triton_unified_attention.py k_descale / v_descale); ALiBi bias as in
# tests/kernels/attention/test_attention.py:96-103.
# Everything in fp32 (probabilities are NOT rounded to the V dtype -> the GPU tolerance covers it).
# --------------------------------------------------------------------------------------------
def _cache_f32(cache: torch.Tensor) -> np.ndarray:
    if cache.element_size() == 1:
        return e4m3_bits_to_f32(cache.view(torch.uint8).cpu().numpy())
    return _f32(cache)


def paged_attention_decode(query, key_cache, value_cache, block_table, seq_lens, scale,
                           k_scale=None, v_scale=None, alibi_slopes=None, sliding_window=None,
                           softcap=None, return_lse=False):
    """query [S,Hq,D]; caches [nb,bs,Hkv,D] (torch CPU, bf16/f16 or fp8/uint8 bytes).
    Returns float32 [S,Hq,D] (and natural-log LSE [S,Hq])."""
    q = _f32(query)
    S, Hq, D = q.shape
    kc, vc = _cache_f32(key_cache), _cache_f32(value_cache)
    nb, bs, Hkv, _ = kc.shape
    g = Hq // Hkv
    ks = 1.0 if k_scale is None else float(_f32(k_scale).reshape(-1)[0])
    vs = 1.0 if v_scale is None else float(_f32(v_scale).reshape(-1)[0])
    bt = block_table.cpu().numpy()
    sl = seq_lens.cpu().numpy()
    out = np.zeros((S, Hq, D), dtype=np.float32)
    lse = np.full((S, Hq), -np.inf, dtype=np.float32)
    for s in range(S):
        L = int(sl[s])
        if L <= 0:
            continue
        nblk = (L + bs - 1) // bs
        idx = bt[s, :nblk]
        k = kc[idx].reshape(-1, Hkv, D)[:L].astype(np.float64) * ks  # [L,Hkv,D]
        v = vc[idx].reshape(-1, Hkv, D)[:L].astype(np.float64) * vs
        k = np.repeat(k, g, axis=1)
        v = np.repeat(v, g, axis=1)
        att = np.einsum("hd,khd->hk", q[s].astype(np.float64) * scale, k)
        if softcap is not None and softcap > 0:
            att = softcap * np.tanh(att / softcap)
        if alibi_slopes is not None:
            pos = np.arange(L) - L + 1
            att = att + _f32(alibi_slopes).astype(np.float64)[:, None] * pos[None, :]
        if sliding_window is not None and sliding_window > 0 and L > sliding_window:
            att[:, : L - sliding_window] = -np.inf
        m = att.max(axis=1, keepdims=True)
        p = np.exp(att - m)
        den = p.sum(axis=1, keepdims=True)
        out[s] = np.einsum("hk,khd->hd", p / den, v).astype(np.float32)
        lse[s] = (m[:, 0] + np.log(den[:, 0])).astype(np.float32)
    if return_lse:
        return out, lse
    return out


# --------------------------------------------------------------------------------------------
# FP8 GEMM.  Baseline of the reference's own test: tests/kernels/utils.py:1130-1173
# (baseline_scaled_mm): torch.mm(scale_a * a.float(), scale_b * b.float()).to(out_dtype) + bias
# --------------------------------------------------------------------------------------------
def paged_attention_varlen(query, key_cache, value_cache, query_lens, kv_lens, block_table, scale,
                           k_scale=None, v_scale=None, alibi_slopes=None, sliding_window=None,
                           softcap=None, return_lse=False):
    """Variable-length causal paged attention (prefill / chunked prefill / mixed batches).
    Restates ref_paged_attn (tests/kernels/attention/test_triton_unified_attention.py:33-88): query
    [T,Hq,D] packed, query_lens / kv_lens python lists, caches [nb,bs,Hkv,D]; query token i of a
    sequence with context c = kv_len - query_len sees kv positions j <= c + i (triu diagonal
    kv_len - query_len + 1), sliding window keeps (c + i - j) < sliding_window (:71-79), soft-cap
    before masking (:80-81). FP8 dequant = fp8.float() * scale; ALiBi adds slope * (j - c)
    (apply_alibi_to_score, triton_attention_helpers.py:396-397). Returns float32 [T,Hq,D]
    (and natural-log LSE [T,Hq])."""
    q = _f32(query)
    T, Hq, D = q.shape
    kc, vc = _cache_f32(key_cache), _cache_f32(value_cache)
    nb, bs, Hkv, _ = kc.shape
    g = Hq // Hkv
    ks = 1.0 if k_scale is None else float(_f32(k_scale).reshape(-1)[0])
    vs = 1.0 if v_scale is None else float(_f32(v_scale).reshape(-1)[0])
    bt = block_table.cpu().numpy()
    out = np.zeros((T, Hq, D), dtype=np.float32)
    lse = np.full((T, Hq), -np.inf, dtype=np.float32)
    start = 0
    for s, (ql, L) in enumerate(zip(query_lens, kv_lens)):
        ql, L = int(ql), int(L)
        if ql == 0:
            continue
        ctx = L - ql
        nblk = (L + bs - 1) // bs
        idx = bt[s, :nblk]
        k = kc[idx].reshape(-1, Hkv, D)[:L].astype(np.float64) * ks
        v = vc[idx].reshape(-1, Hkv, D)[:L].astype(np.float64) * vs
        k = np.repeat(k, g, axis=1)
        v = np.repeat(v, g, axis=1)
        att = np.einsum("qhd,khd->hqk", q[start:start + ql].astype(np.float64) * scale, k)
        if softcap is not None and softcap > 0:
            att = softcap * np.tanh(att / softcap)
        qi = np.arange(ql)[:, None]
        kj = np.arange(L)[None, :]
        if alibi_slopes is not None:
            att = att + _f32(alibi_slopes).astype(np.float64)[:, None, None] * (kj - ctx)[None, :, :]
        mask = kj > ctx + qi
        if sliding_window is not None and sliding_window > 0:
            mask = mask | ((ctx + qi - kj) >= sliding_window)
        att = np.where(mask[None, :, :], -np.inf, att)
        m = att.max(axis=2, keepdims=True)
        p = np.exp(att - m)
        den = p.sum(axis=2, keepdims=True)
        out[start:start + ql] = np.einsum("hqk,khd->qhd", p / den, v).astype(np.float32)
        lse[start:start + ql] = (m[:, :, 0] + np.log(den[:, :, 0])).T.astype(np.float32)
        start += ql
    if return_lse:
        return out, lse
    return out


def scaled_mm(a_bits, b_bits_nk, scale_a, scale_b, bias=None) -> np.ndarray:
    """a_bits uint8 [M,K] (e4m3), b_bits_nk uint8 [N,K] (the weight; b = weight.T), scales float32
    ([1] / [M] / [N]); returns float64-accumulated float32 [M,N] (no output rounding)."""
    a = e4m3_bits_to_f32(np.asarray(a_bits)).astype(np.float64)
    w = e4m3_bits_to_f32(np.asarray(b_bits_nk)).astype(np.float64)
    sa = np.asarray(scale_a, dtype=np.float64).reshape(-1, 1)
    sb = np.asarray(scale_b, dtype=np.float64).reshape(1, -1)
    out = (a @ w.T) * sa * sb
    if bias is not None:
        out = out + np.asarray(bias, dtype=np.float64).reshape(1, -1)
    return out.astype(np.float32)


# --------------------------------------------------------------------------------------------
# silu_and_mul / silu_and_mul_quant.
# csrc/libtorch_stable/activation_kernels.cu:102-160 and quantization/activation_kernels.cu:33-104:
# silu is rounded to the activation dtype BEFORE the multiply, the product is rounded again.
# --------------------------------------------------------------------------------------------
def silu_and_mul(x: torch.Tensor) -> torch.Tensor:
    d = x.shape[-1] // 2
    xf = x[..., :d].to(torch.float32)
    s = (xf / (1.0 + torch.exp(-xf))).to(x.dtype)
    return (s.to(torch.float32) * x[..., d:].to(torch.float32)).to(x.dtype)


def silu_and_mul_quant(x: torch.Tensor, scale: torch.Tensor) -> np.ndarray:
    return static_scaled_fp8_quant(silu_and_mul(x), scale)


# --------------------------------------------------------------------------------------------
# RMSNorm family.  csrc/libtorch_stable/layernorm_kernels.cu, layernorm_quant_kernels.cu:21-156
# --------------------------------------------------------------------------------------------
def rms_norm(x: torch.Tensor, weight: torch.Tensor, eps: float, residual: torch.Tensor | None = None):
    """Returns (out, new_residual). z = T(x + r); out = T(z * rsqrt(mean(z^2)+eps) * w): fp32 product,
    ONE rounding (layernorm_kernels.cu:93 / :163 `static_cast<scalar_t>(x * s_variance * w)`)."""
    if residual is not None:
        x = (x.float() + residual.float()).to(x.dtype)
    xf = x.float()
    s = torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)
    out = (xf * s * weight.float()).to(x.dtype)
    return out, x


def rms_norm_static_fp8_quant(x, weight, scale, eps, residual=None):
    """out8 = e4m3(clamp(float(T(z * s * w)) * (1/scale))) (layernorm_quant_kernels.cu:66-71)."""
    if residual is not None:
        x = (x.float() + residual.float()).to(x.dtype)
    xf = x.float()
    s = torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)
    normed = (xf * s * weight.float()).to(x.dtype)
    return static_scaled_fp8_quant(normed, scale), x


# --------------------------------------------------------------------------------------------
# RoPE.  csrc/libtorch_stable/pos_encoding_kernels.cu:8-99
# --------------------------------------------------------------------------------------------
def rotary_embedding(positions, query, key, head_size, cos_sin_cache, is_neox):
    """Returns rotated (query, key) copies; query [T, H, D]."""
    rot = cos_sin_cache.shape[1]
    emb = rot // 2
    cs = cos_sin_cache[positions].float()
    cos, sin = cs[:, None, :emb], cs[:, None, emb:]

    def rot_one(t):
        t = t.clone()
        tf = t.float()
        if is_neox:
            x, y = tf[..., :emb], tf[..., emb:rot]
            t[..., :emb] = (x * cos - y * sin).to(t.dtype)
            t[..., emb:rot] = (y * cos + x * sin).to(t.dtype)
        else:
            x, y = tf[..., 0:rot:2], tf[..., 1:rot:2]
            t[..., 0:rot:2] = (x * cos - y * sin).to(t.dtype)
            t[..., 1:rot:2] = (y * cos + x * sin).to(t.dtype)
        return t

    return rot_one(query), (None if key is None else rot_one(key))


# --------------------------------------------------------------------------------------------
# MoE routing.
# moe_align_block_size: csrc/libtorch_stable/moe/moe_align_sum_kernels.cu:85-350 and the golden
# torch_moe_align_block_size of tests/kernels/moe/test_moe_align_block_size.py:92-181 (stable sort).
# --------------------------------------------------------------------------------------------
def moe_align_block_size(topk_ids: np.ndarray, block_size: int, num_experts: int,
                         expert_map: np.ndarray | None = None, pad_sorted_ids: bool = False,
                         ignore_invalid_experts: bool = False):
    """Returns (sorted_token_ids int32 [max_padded], expert_ids int32 [max_blocks],
    num_tokens_post_pad int). Pad value = numel, expert_ids tail = -1, stable order."""
    flat = np.asarray(topk_ids).reshape(-1).astype(np.int64)
    numel = flat.size
    max_padded = numel + num_experts * (block_size - 1)
    if pad_sorted_ids:
        max_padded = (max_padded + block_size - 1) // block_size * block_size
    if numel < num_experts:
        max_padded = min(numel * block_size, max_padded)
    max_blocks = (max_padded + block_size - 1) // block_size
    local = flat.copy()
    bad = (local < 0) | (local >= num_experts)
    if ignore_invalid_experts and expert_map is not None:
        m = np.asarray(expert_map).astype(np.int64)
        local = np.where(bad, -1, m[np.clip(local, 0, num_experts - 1)])
    else:
        local = np.where(bad, -1, local)
    sorted_ids = np.full(max_padded, numel, dtype=np.int32)
    expert_ids = np.full(max_blocks, -1, dtype=np.int32)
    pos = 0
    for e in range(num_experts):
        idx = np.nonzero(local == e)[0]  # ascending = stable
        n = idx.size
        padded = (n + block_size - 1) // block_size * block_size
        sorted_ids[pos:pos + n] = idx
        expert_ids[pos // block_size:(pos + padded) // block_size] = e
        pos += padded
    if expert_map is not None and not ignore_invalid_experts:
        m = np.asarray(expert_map).astype(np.int64)
        expert_ids = np.where(expert_ids >= 0, m[np.clip(expert_ids, 0, num_experts - 1)], -1).astype(np.int32)
    return sorted_ids, expert_ids, pos


def topk_softmax(gating: torch.Tensor, topk: int, renormalize: bool, bias: torch.Tensor | None = None):
    """csrc/libtorch_stable/moe/topk_softmax_kernels.cu:76-262 (softmax fp32, iterative arg-max with
    ties to the lower index, weights = softmax value, optional renormalise);
    token_expert_indices[m, k] = k * M + m."""
    g = gating.to(torch.float32)
    M, E = g.shape
    p = torch.softmax(g, dim=-1)
    sel = p + bias.float() if bias is not None else p.clone()
    ids = torch.empty(M, topk, dtype=torch.int32)
    w = torch.empty(M, topk, dtype=torch.float32)
    for k in range(topk):
        mx = sel.max(dim=-1, keepdim=True).values
        first = (sel == mx).float().argmax(dim=-1)  # lowest index among ties
        ids[:, k] = first.to(torch.int32)
        w[:, k] = p[torch.arange(M), first]
        sel[torch.arange(M), first] = -float("inf")
    if renormalize:
        den = w.sum(-1, keepdim=True)
        w = w * (1.0 / torch.where(den > 0, den, torch.ones_like(den)))
    tei = (torch.arange(topk, dtype=torch.int32)[None, :] * M + torch.arange(M, dtype=torch.int32)[:, None])
    return w, ids, tei


def moe_sum(x: torch.Tensor) -> torch.Tensor:
    """moe_align_sum_kernels.cu:352-420: fp32 sum over topk, rounded once."""
    return x.float().sum(dim=1).to(x.dtype)


# --------------------------------------------------------------------------------------------
# fused_experts (FP8 W8A8, per-tensor activation scales).
# Where quantisation happens: fused_moe.py:1764-1851; per-expert dense loop as torch_experts,
# tests/kernels/utils.py:912-1078.
# --------------------------------------------------------------------------------------------
def fused_experts_fp8(x: torch.Tensor, w1_bits: torch.Tensor, w2_bits: torch.Tensor,
                      topk_weights: torch.Tensor, topk_ids: torch.Tensor, w1_scale, w2_scale, a1_scale,
                      a2_scale, expert_map: torch.Tensor | None = None) -> torch.Tensor:
    """x [M,H] bf16; w1_bits uint8 [E,2I,H]; w2_bits uint8 [E,H,I]; scales float tensors
    ([E] weights, [1] activations). Returns float32 [M,H] (fp32 accumulate, bf16 intermediates)."""
    M, H = x.shape
    E = w1_bits.shape[0]
    x8 = torch.from_numpy(e4m3_bits_to_f32(static_scaled_fp8_quant(x, a1_scale)))
    a1 = float(a1_scale.reshape(-1)[0]); a2 = float(a2_scale.reshape(-1)[0])
    out = torch.zeros(M, H, dtype=torch.float32)
    deq = {}
    for t in range(M):
        for k in range(topk_ids.shape[1]):
            e = int(topk_ids[t, k])
            if e < 0:
                continue
            if expert_map is not None:
                e = int(expert_map[e])
                if e < 0:
                    continue
            if e not in deq:   # dequantise each expert once (same values, test run time only)
                deq[e] = (torch.from_numpy(e4m3_bits_to_f32(w1_bits[e].numpy())),
                          torch.from_numpy(e4m3_bits_to_f32(w2_bits[e].numpy())))
            w1, w2 = deq[e]
            h = (x8[t] @ w1.t()) * (a1 * float(w1_scale[e]))
            h = h.to(x.dtype)
            act = silu_and_mul(h[None, :])
            act8 = torch.from_numpy(e4m3_bits_to_f32(static_scaled_fp8_quant(act, a2_scale)))[0]
            o = (act8 @ w2.t()) * (a2 * float(w2_scale[e]))
            out[t] += float(topk_weights[t, k]) * o.to(x.dtype).float()
    return out


# --------------------------------------------------------------------------------------------
# NVFP4 quantisation.  csrc/libtorch_stable/quantization/fp4/nvfp4_utils.cuh:158-302 and the
# reference's own test oracle tests/kernels/quantization/test_nvfp4_quant.py:53-131
# (ref_nvfp4_quant, cast_to_fp4, recover_swizzled_scales), restated.
# The CUDA kernel uses rcp.approx.ftz (<= 1 ulp); this restatement uses exact reciprocals, so a
# 1-code difference is possible exactly at rounding boundaries.
# --------------------------------------------------------------------------------------------
E2M1_VALUES = np.array([0.0, 0.5, 1.0, 1.5, 2.0, 3.0, 4.0, 6.0], dtype=np.float32)


def f32_to_e2m1_code(x: np.ndarray) -> np.ndarray:
    """RNE to {0,.5,1,1.5,2,3,4,6} with saturation (cvt.rn.satfinite.e2m1x2.f32); ties to even
    mantissa: 0.25->0, 0.75->1, 1.25->1, 1.75->2, 2.5->2, 3.5->4, 5->4 (test_nvfp4_quant.py:84-95)."""
    a = np.abs(x).astype(np.float32)
    code = np.zeros(a.shape, dtype=np.uint8)
    code[(a > 0.25) & (a < 0.75)] = 1
    code[(a >= 0.75) & (a <= 1.25)] = 2
    code[(a > 1.25) & (a < 1.75)] = 3
    code[(a >= 1.75) & (a <= 2.5)] = 4
    code[(a > 2.5) & (a < 3.5)] = 5
    code[(a >= 3.5) & (a <= 5.0)] = 6
    code[a > 5.0] = 7
    return code | (np.signbit(x).astype(np.uint8) << 3)


def e2m1_code_to_f32(code: np.ndarray) -> np.ndarray:
    v = E2M1_VALUES[code & 7]
    return np.where(code & 8, -v, v).astype(np.float32)


def nvfp4_quant(x: torch.Tensor, global_scale: float, return_scaled: bool = False):
    """Returns (codes uint8 [M,K] one per element, scale bytes uint8 [M,K/16] row-major)."""
    xf = _f32(x)
    M, K = xf.shape
    blk = xf.reshape(M, K // 16, 16)
    amax = np.abs(blk).max(axis=-1).astype(np.float32)
    gs = np.float32(global_scale)
    sf = (gs * (amax * (np.float32(1.0) / np.float32(6.0))).astype(np.float32)).astype(np.float32)
    sf_bits = f32_to_e4m3_bits(sf)
    sfv = e4m3_bits_to_f32(sf_bits)
    with np.errstate(all="ignore"):
        os = np.where(sfv != 0, np.float32(1.0) / (sfv * (np.float32(1.0) / gs)).astype(np.float32), 0).astype(np.float32)
    y = (blk * os[..., None]).astype(np.float32)
    if return_scaled:
        return f32_to_e2m1_code(y).reshape(M, K), sf_bits, y.reshape(M, K)
    return f32_to_e2m1_code(y).reshape(M, K), sf_bits


E2M1_BOUNDARIES = np.array([0.25, 0.75, 1.25, 1.75, 2.5, 3.5, 5.0], dtype=np.float32)


def nvfp4_near_tie(y: np.ndarray, rel: float = 6e-7) -> np.ndarray:
    """True where |y| sits within `rel` (a few fp32 ulp = the error of the kernel's two rcp.approx.ftz
    against this restatement's exact reciprocals) of an e2m1 rounding boundary. Only there may the
    CUDA kernel's code legitimately differ from the restatement (by one e2m1 step)."""
    a = np.abs(y).astype(np.float32)[..., None]
    return (np.abs(a - E2M1_BOUNDARIES) <= rel * E2M1_BOUNDARIES).any(-1)


def nvfp4_pack(codes: np.ndarray) -> np.ndarray:
    return (codes[:, 0::2] | (codes[:, 1::2] << 4)).astype(np.uint8)


def nvfp4_unpack(packed: np.ndarray) -> np.ndarray:
    M, K2 = packed.shape
    out = np.empty((M, K2 * 2), dtype=np.uint8)
    out[:, 0::2] = packed & 0xF
    out[:, 1::2] = packed >> 4
    return out


def swizzle_sf(sf: np.ndarray) -> np.ndarray:
    """row-major [M, K/16] -> tcgen05 128x4 layout, flat [round_up(M,128) * round_up(K/16,4)], pads 0
    (offset formula nvfp4_utils.cuh:186-197; inverse of recover_swizzled_scales)."""
    M, C = sf.shape
    rm, rc = (M + 127) // 128 * 128, (C + 3) // 4 * 4
    pad = np.zeros((rm, rc), dtype=sf.dtype)
    pad[:M, :C] = sf
    t = pad.reshape(rm // 128, 4, 32, rc // 4, 4)      # [mTile, innerM, outerM, kTile, innerK]
    t = t.transpose(0, 3, 2, 1, 4)                      # [mTile, kTile, outerM, innerM, innerK]
    return np.ascontiguousarray(t).reshape(-1)


def unswizzle_sf(flat: np.ndarray, M: int, C: int) -> np.ndarray:
    rm, rc = (M + 127) // 128 * 128, (C + 3) // 4 * 4
    t = np.asarray(flat).reshape(rm // 128, rc // 4, 32, 4, 4).transpose(0, 3, 2, 1, 4)
    return np.ascontiguousarray(t).reshape(rm, rc)[:M, :C]


def nvfp4_dequant(codes: np.ndarray, sf_bits: np.ndarray, global_scale: float) -> np.ndarray:
    """codes uint8 [M,K] (one per element), sf_bits uint8 [M,K/16] row-major -> fp32 [M,K]:
    value = e2m1(code) * (e4m3(sf) / global_scale)   (tests/kernels/quantization/nvfp4_utils.py
    dequantize_nvfp4_to_dtype / test_nvfp4_scaled_mm.py:27-49)."""
    M, K = codes.shape
    v = e2m1_code_to_f32(codes).reshape(M, K // 16, 16)
    sf = (e4m3_bits_to_f32(sf_bits) / np.float32(global_scale)).astype(np.float32)
    return (v * sf[..., None]).astype(np.float32).reshape(M, K)


def nvfp4_scaled_mm(a_codes, a_sf_bits, b_codes, b_sf_bits, alpha: float) -> np.ndarray:
    """out = alpha * (a (.) sfa)(b (.) sfb)^T in float64-accumulated fp32: the exact value the block-scaled
    MMA approximates (products of e2m1 x e2m1 x e4m3 x e4m3 are exact in fp32; only the accumulation
    order differs). Equals get_ref_results (test_nvfp4_scaled_mm.py:27-49) with alpha = 1/(gs_a*gs_b)
    up to that function's rounding of the dequantised operands to bf16/f16."""
    M, K = a_codes.shape
    av = e2m1_code_to_f32(a_codes).reshape(M, K // 16, 16).astype(np.float64) * e4m3_bits_to_f32(a_sf_bits).astype(np.float64)[..., None]
    N = b_codes.shape[0]
    bv = e2m1_code_to_f32(b_codes).reshape(N, K // 16, 16).astype(np.float64) * e4m3_bits_to_f32(b_sf_bits).astype(np.float64)[..., None]
    return (np.float64(alpha) * (av.reshape(M, K) @ bv.reshape(N, K).T)).astype(np.float32)


# --------------------------------------------------------------------------------------------
# Legacy x-packed cache layout (paged_attention_v1/v2): ref_single_query_cached_kv_attention,
# tests/kernels/attention/test_attention.py:57-109 gathers k = key_cache[block, :, :, offset, :]
# .reshape(Hkv, D) and v = value_cache[block, :, :, offset]. Same gather, done once for the whole
# cache: the result is the flash layout paged_attention_decode() consumes.
# --------------------------------------------------------------------------------------------
def legacy_cache_to_flash(key_cache: torch.Tensor, value_cache: torch.Tensor):
    """key_cache [nb,Hkv,D/x,bs,x], value_cache [nb,Hkv,D,bs] -> ([nb,bs,Hkv,D], [nb,bs,Hkv,D])."""
    nb, Hkv, Dx, bs, x = key_cache.shape
    k = key_cache.permute(0, 3, 1, 2, 4).reshape(nb, bs, Hkv, Dx * x)
    v = value_cache.permute(0, 3, 1, 2)
    return k.contiguous(), v.contiguous()


def flash_cache_to_legacy(k: torch.Tensor, v: torch.Tensor):
    """inverse of legacy_cache_to_flash (test-input construction)."""
    nb, bs, Hkv, D = k.shape
    x = 16 // k.element_size()
    kc = k.reshape(nb, bs, Hkv, D // x, x).permute(0, 2, 3, 1, 4).contiguous()
    vc = v.permute(0, 2, 3, 1).contiguous()
    return kc, vc


def paged_attention_legacy(query, key_cache, value_cache, block_tables, seq_lens, scale,
                           k_scale=None, v_scale=None, alibi_slopes=None):
    k, v = legacy_cache_to_flash(key_cache, value_cache)
    return paged_attention_decode(query, k, v, block_tables, seq_lens, scale, k_scale=k_scale,
                                  v_scale=v_scale, alibi_slopes=alibi_slopes)


# --------------------------------------------------------------------------------------------
# merge_attn_states: csrc/libtorch_stable/attention/merge_attn_states.cu:13-200 (same case split,
# same operation order), vllm/v1/attention/ops/merge_attn_states.py:9-60.
# --------------------------------------------------------------------------------------------
def merge_attn_states(prefix_out, prefix_lse, suffix_out, suffix_lse, prefill_tokens_with_context=None,
                      output_scale=None):
    """prefix_out / suffix_out torch [T,H,D]; LSEs torch f32 [H,T]. Returns (out f32 [T,H,D] - or e4m3
    bits uint8 when output_scale is given -, out_lse f32 [H,T])."""
    po, so = _f32(prefix_out), _f32(suffix_out)
    T, H, D = po.shape
    pl = _f32(prefix_lse).T.copy()   # [T,H]
    sl = _f32(suffix_lse).T.copy()
    npre = T if prefill_tokens_with_context is None else int(prefill_tokens_with_context)
    pl = np.where(np.isinf(pl), -np.inf, pl).astype(np.float32)
    sl = np.where(np.isinf(sl), -np.inf, sl).astype(np.float32)
    mx = np.maximum(pl, sl)
    with np.errstate(all="ignore"):
        pe = np.exp((pl - mx).astype(np.float32)).astype(np.float32)
        se = np.exp((sl - mx).astype(np.float32)).astype(np.float32)
        tot = (pe + se).astype(np.float32)
        wp, ws = (pe / tot).astype(np.float32), (se / tot).astype(np.float32)
        merged = (po * wp[..., None]).astype(np.float32) + (so * ws[..., None]).astype(np.float32)
        lse = (np.log(tot) + mx).astype(np.float32)
    inf_case = np.isinf(mx)
    out = np.where(inf_case[..., None], po, merged).astype(np.float32)
    lse = np.where(inf_case, mx, lse).astype(np.float32)
    copy_case = np.arange(T)[:, None] >= npre
    out = np.where(copy_case[..., None], so, out).astype(np.float32)
    lse = np.where(copy_case, _f32(suffix_lse).T, lse).astype(np.float32)
    if output_scale is not None:
        inv = np.float32(1.0) / np.float32(_f32(output_scale).reshape(-1)[0])
        out = f32_to_e4m3_bits(np.clip((out * inv).astype(np.float32), -448.0, 448.0))
    return out, lse.T.copy()


def silu_and_mul_nvfp4_quant(x: torch.Tensor, global_scale: float, return_scaled: bool = False):
    """nvfp4_quant( T( silu_f32(gate) * up ) ): compute_silu_mul, nvfp4_utils.cuh:304-329 (one rounding
    to the activation dtype, unlike silu_and_mul), then cvt_warp_fp16_to_fp4."""
    d = x.shape[-1] // 2
    g, u = x[..., :d].float(), x[..., d:].float()
    act = ((g / (1.0 + torch.exp(-g))) * u).to(x.dtype)
    return nvfp4_quant(act, global_scale, return_scaled=return_scaled)


# --------------------------------------------------------------------------------------------
# Sampler-side ops.  vllm/v1/sample/ops/topk_topp_sampler.py:367-408 (apply_top_k_top_p_pytorch, the
# sort-based definition every other implementation in the reference is tested against) and
# vllm/_custom_ops.py:309-322 (apply_repetition_penalties_torch), restated with the same torch ops.
# --------------------------------------------------------------------------------------------
def apply_top_k_top_p(logits: torch.Tensor, k: torch.Tensor | None, p: torch.Tensor | None):
    """Returns (masked logits, info) where info carries what a test needs to exclude legitimate boundary
    cases: `kth` = the k-th largest value per row (ties of it are kept), `cum` = the ascending cumulative
    softmax mass at every token's sorted position scattered back to vocabulary order."""
    logits = logits.clone().float()
    info = {}
    if k is None and p is None:
        return logits, info
    logits_sort, logits_idx = logits.sort(dim=-1, descending=False, stable=True)
    if k is not None:
        top_k_mask = logits_sort.size(1) - k.to(torch.long)
        top_k_mask = logits_sort.gather(1, top_k_mask.clamp(min=0).unsqueeze(dim=1))
        info["kth"] = top_k_mask.squeeze(1).clone()
        no_k = (k.to(torch.long) >= logits_sort.size(1)) | (k.to(torch.long) <= 0)
        top_k_mask = torch.where(no_k.unsqueeze(1), torch.full_like(top_k_mask, -float("inf")), top_k_mask)
        logits_sort.masked_fill_(logits_sort < top_k_mask, -float("inf"))
    if p is not None:
        probs_sort = logits_sort.softmax(dim=-1)
        probs_sum = torch.cumsum(probs_sort, dim=-1)
        info["cum"] = torch.empty_like(probs_sum).scatter_(dim=-1, index=logits_idx, src=probs_sum)
        top_p_mask = probs_sum <= 1 - p.unsqueeze(dim=1)
        top_p_mask[:, -1] = False
        logits_sort.masked_fill_(top_p_mask, -float("inf"))
    return logits.scatter_(dim=-1, index=logits_idx, src=logits_sort), info


def apply_repetition_penalties(logits: torch.Tensor, prompt_mask: torch.Tensor, output_mask: torch.Tensor,
                               repetition_penalties: torch.Tensor) -> torch.Tensor:
    """csrc/libtorch_stable/sampler.cu:14-45: logit > 0 ? logit / penalty : logit * penalty in the logits dtype
    (fp32 operation, one rounding) on tokens of the prompt or the output."""
    pen = repetition_penalties.to(logits.dtype).float().unsqueeze(1)
    x = logits.float()
    y = torch.where(x > 0, x / pen, x * pen).to(logits.dtype)
    return torch.where(prompt_mask | output_mask, y, logits)


# --------------------------------------------------------------------------------------------
# Decode-context-parallel merge across ranks: vllm/v1/attention/ops/common.py
#   _correct_attn_cp_out_kernel :13-93, correct_attn_out :110-183, cp_lse_ag_out_rs :185-232,
#   cp_lse_ag_out_ar :235-258. No golden vectors exist in the reference for this path (its only implementation is
#   a Triton kernel); the restatement is pinned by the domain property tests/test_oracle_dcp.py checks:
#   attention over a context split across N ranks, merged, equals attention over the whole context
#   (the attention oracle itself is pinned to the reference CPU kernels).
# --------------------------------------------------------------------------------------------
def dcp_correct_attn_out(out, lses, cp_rank, is_lse_base_on_e=True):
    """out f32 [B,H,D] (this rank's partial), lses f32 [N,B,H] (all ranks). Returns (corrected out f32, lse f32 [B,H])
    in the kernel's operation order (fp32)."""
    out = _f32(out) if torch.is_tensor(out) else np.asarray(out, dtype=np.float32)
    l = _f32(lses) if torch.is_tensor(lses) else np.asarray(lses, dtype=np.float32)
    exp, log = (np.exp, np.log) if is_lse_base_on_e else (np.exp2, np.log2)
    with np.errstate(all="ignore"):
        ls = np.where(np.isnan(l) | (l == np.inf), -np.inf, l).astype(np.float32)
        mx = ls.max(axis=0)
        mx = np.where(mx == -np.inf, np.float32(0), mx).astype(np.float32)
        acc = exp((ls - mx).astype(np.float32)).astype(np.float32).sum(axis=0, dtype=np.float32)
        lse = (log(acc).astype(np.float32) + mx).astype(np.float32)
        fin = (l[cp_rank] - lse).astype(np.float32)
        fin = np.where(np.isnan(fin) | (fin == np.inf), -np.inf, fin).astype(np.float32)
        factor = exp(fin).astype(np.float32)
        o = (out * factor[..., None]).astype(np.float32)
        o = np.where(factor[..., None] == 0, np.float32(0), o)
    return o.astype(np.float32), lse


def dcp_lse_merge(outs, lses, dtype=torch.bfloat16, is_lse_base_on_e=True):
    """cp_lse_ag_out_ar on one host: outs list of N torch [B,H,D] partials (dtype), lses list of N f32 [B,H].
    Every corrected partial is rounded to `dtype` (the reference's kernel stores it in the output dtype before the
    collective), then summed in rank order with fp32 accumulation and rounded once. Returns (out torch dtype
    [B,H,D], lse f32 numpy [B,H]); the reduce-scatter form keeps heads [r*H/N, (r+1)*H/N) on rank r."""
    N = len(outs)
    L = np.stack([_f32(x) for x in lses])
    acc, lse = None, None
    for r in range(N):
        o, lse = dcp_correct_attn_out(outs[r], L, r, is_lse_base_on_e)
        o = torch.from_numpy(o).to(dtype).float().numpy()
        acc = o if acc is None else (acc + o).astype(np.float32)
    return torch.from_numpy(acc).to(dtype), lse


# --------------------------------------------------------------------------------------------
# NVFP4 fused MoE on packed expert-sorted rows.
#   get_cutlass_moe_mm_data   csrc/libtorch_stable/quantization/w8a8/cutlass/moe/moe_data.cu:17-117,236-289
#   scaled_fp4_experts_quant  csrc/libtorch_stable/quantization/fp4/nvfp4_experts_quant.cu:36-250
#   cutlass_fp4_group_mm      csrc/libtorch_stable/quantization/fp4/nvfp4_blockwise_moe_kernel.cu:676-746
#   run_cutlass_moe_fp4       vllm/model_executor/layers/fused_moe/experts/cutlass_moe.py:498-688
# and the reference test's own model of the whole block: dequantise, then torch_moe
# (tests/kernels/moe/test_nvfp4_moe.py:121-160, tests/kernels/utils.py torch_moe).
# --------------------------------------------------------------------------------------------
def cutlass_moe_mm_data(topk_ids: np.ndarray, num_experts: int, n: int, k: int, is_gated: bool = True,
                        swap_ab: bool = False, with_blockscale: bool = True):
    """Returns dict(expert_offsets [E+1], problem_sizes1/2 [E,3], a_map [numel] (-1 past the packed rows),
    c_map [numel], blockscale_offsets [E+1]). Rows inside an expert are in ascending flat-index order (the
    reference's compute_arg_sorts orders them by atomicAdd: only the per-expert SET is defined there)."""
    ids = np.asarray(topk_ids).astype(np.int64)
    topk = ids.shape[-1] if ids.ndim > 1 else 1
    flat = ids.reshape(-1)
    E = num_experts
    counts = np.array([(flat == e).sum() for e in range(E)], dtype=np.int64)
    off = np.concatenate([[0], np.cumsum(counts)]).astype(np.int32)
    bso = np.concatenate([[0], np.cumsum((counts + 127) // 128 * 128)]).astype(np.int32)
    n1 = 2 * n if is_gated else n
    ps1 = np.zeros((E, 3), dtype=np.int32)
    ps2 = np.zeros((E, 3), dtype=np.int32)
    for e in range(E):
        ps1[e] = (n1, counts[e], k) if swap_ab else (counts[e], n1, k)
        ps2[e] = (k, counts[e], n) if swap_ab else (counts[e], k, n)
    a_map = np.full(flat.size, -1, dtype=np.int32)
    c_map = np.full(flat.size, off[E], dtype=np.int32)
    fill = off[:-1].astype(np.int64).copy()
    for i, e in enumerate(flat):
        if 0 <= e < E:
            a_map[fill[e]] = i // topk
            c_map[i] = fill[e]
            fill[e] += 1
    return dict(expert_offsets=off, problem_sizes1=ps1, problem_sizes2=ps2, a_map=a_map, c_map=c_map,
                blockscale_offsets=bso if with_blockscale else None)


def nvfp4_experts_quant(x: torch.Tensor, global_scales: np.ndarray, expert_offsets: np.ndarray,
                        blockscale_offsets: np.ndarray, gated: bool = False, return_scaled: bool = False):
    """Per-expert nvfp4_quant of expert-sorted rows. Returns (codes uint8 [m, k] one per element, sf: flat uint8
    buffer of blockscale_offsets[E] * round_up(k/16, 4) bytes holding one swizzled matrix per expert -- rows
    past an expert's last row are zero here and UNDEFINED in the reference --, sf_rows: row-major scale bytes
    [m, k/16] (, y))."""
    E = len(expert_offsets) - 1
    m = int(expert_offsets[E])
    k = x.shape[1] // (2 if gated else 1)
    C = k // 16
    rc = (C + 3) // 4 * 4
    codes = np.zeros((m, k), dtype=np.uint8)
    sf_rows = np.zeros((m, C), dtype=np.uint8)
    ys = np.zeros((m, k), dtype=np.float32)
    sf = np.zeros(int(blockscale_offsets[E]) * rc, dtype=np.uint8)
    for e in range(E):
        lo, hi = int(expert_offsets[e]), int(expert_offsets[e + 1])
        if hi == lo:
            continue
        gs = 1.0 if global_scales is None else float(global_scales[e])
        fn = silu_and_mul_nvfp4_quant if gated else nvfp4_quant
        c, s, y = fn(x[lo:hi], gs, return_scaled=True)
        codes[lo:hi], sf_rows[lo:hi], ys[lo:hi] = c, s, y
        sw = swizzle_sf(s)
        base = int(blockscale_offsets[e]) * rc
        sf[base:base + sw.size] = sw
    if return_scaled:
        return codes, sf, sf_rows, ys
    return codes, sf, sf_rows


def nvfp4_group_mm(a_codes: np.ndarray, a_sf_rows: np.ndarray, w_codes: np.ndarray, w_sf_rows: np.ndarray,
                   alphas: np.ndarray, expert_offsets: np.ndarray) -> np.ndarray:
    """out[rows of e] = alphas[e] * (a_e (.) sfa_e)(w[e] (.) sfw[e])^T (nvfp4_scaled_mm per expert).
    a_codes [m, K], a_sf_rows [m, K/16], w_codes [E, N, K], w_sf_rows [E, N, K/16] (row-major scale bytes)."""
    E = w_codes.shape[0]
    out = np.zeros((a_codes.shape[0], w_codes.shape[1]), dtype=np.float32)
    for e in range(E):
        lo, hi = int(expert_offsets[e]), int(expert_offsets[e + 1])
        if hi > lo:
            out[lo:hi] = nvfp4_scaled_mm(a_codes[lo:hi], a_sf_rows[lo:hi], w_codes[e], w_sf_rows[e], float(alphas[e]))
    return out


def cutlass_moe_fp4(a: torch.Tensor, a1_gscale: np.ndarray, w1_codes: np.ndarray, w1_sf_rows: np.ndarray,
                    w1_alphas: np.ndarray, a2_gscale: np.ndarray, w2_codes: np.ndarray, w2_sf_rows: np.ndarray,
                    w2_alphas: np.ndarray, topk_weights: torch.Tensor, topk_ids: np.ndarray,
                    return_stages: bool = False):
    """run_cutlass_moe_fp4 operator by operator (cutlass_moe.py:587-688) with the GEMM outputs rounded to the
    activation dtype where the kernels round them. w*_codes one e2m1 code per element, w*_sf_rows row-major."""
    dt = a.dtype
    E, n2, k = w1_codes.shape
    n = n2 // 2
    m, topk = topk_ids.shape
    d = cutlass_moe_mm_data(topk_ids, E, n, k)
    off, bso = d["expert_offsets"], d["blockscale_offsets"]
    rows = int(off[E])
    rep_a = a[torch.from_numpy(d["a_map"][:rows].astype(np.int64))]
    c_a, _, s_a = nvfp4_experts_quant(rep_a, a1_gscale, off, bso)
    c1 = torch.from_numpy(nvfp4_group_mm(c_a, s_a, w1_codes, w1_sf_rows, w1_alphas, off)).to(dt)
    c_h, _, s_h = nvfp4_experts_quant(c1, a2_gscale, off, bso, gated=True)
    c3 = torch.from_numpy(nvfp4_group_mm(c_h, s_h, w2_codes, w2_sf_rows, w2_alphas, off)).to(dt)
    out = torch.zeros(m, k, dtype=torch.float32)
    c3f = c3.float()
    for t in range(m):
        for j in range(topk):
            r = int(d["c_map"][t * topk + j])
            if r < rows:
                out[t] += float(topk_weights[t, j]) * c3f[r]
    out = out.to(dt)
    if return_stages:
        return out, dict(data=d, rep_a=rep_a, c1=c1, c3=c3, a_codes=c_a, h_codes=c_h)
    return out


# --------------------------------------------------------------------------------------------
# fused_experts, the other FP8 quantisation schemes of fused_experts_impl (fused_moe.py:1764-1770, 1813-1819
# through moe_kernel_quantize_input / _fp8_quantize, fused_moe/utils.py:138-161):
#   "channel": per_channel_quant=True -> dynamic PER-TOKEN activation scales, weight scales [E, N];
#   "block":   block_shape=[128,128]  -> per-token-group (1x128) activation scales, weight scales [E, N/128, K/128];
#   "tensor":  per-tensor (the function above, kept for its callers).
# apply_router_weight_on_input (fused_moe.py:1757-1762, topk = 1): x * w rounded to the activation dtype first.
# Vectorised per expert; GEMM results rounded to the activation dtype where the kernels round them.
# --------------------------------------------------------------------------------------------
def fused_experts_fp8_scheme(x: torch.Tensor, w1_bits: torch.Tensor, w2_bits: torch.Tensor,
                             topk_weights: torch.Tensor, topk_ids: torch.Tensor, w1_scale: torch.Tensor,
                             w2_scale: torch.Tensor, scheme: str, a1_scale=None, a2_scale=None,
                             apply_router_weight_on_input: bool = False) -> torch.Tensor:
    assert scheme in ("tensor", "channel", "block")
    dt = x.dtype
    M, H = x.shape
    E, N1, _ = w1_bits.shape
    I = N1 // 2
    topk = topk_ids.shape[1]
    if apply_router_weight_on_input:
        assert topk == 1
        x = x * topk_weights.to(dt)

    def quant(t, static):
        """-> (dequantisable values fp64 [rows, K], scale array broadcastable per (row, k-block))"""
        if scheme == "block":
            bits, s = per_token_group_quant_fp8(t, 128)
            return e4m3_bits_to_f32(bits).astype(np.float64), s.astype(np.float64)
        if static is not None:
            return e4m3_bits_to_f32(static_scaled_fp8_quant(t, static)).astype(np.float64), \
                np.full((t.shape[0], 1), float(static.reshape(-1)[0]))
        if scheme == "channel":
            bits, s = dynamic_per_token_scaled_fp8_quant(t)
            return e4m3_bits_to_f32(bits).astype(np.float64), s.astype(np.float64)
        bits, s = dynamic_scaled_fp8_quant(t)
        return e4m3_bits_to_f32(bits).astype(np.float64), np.full((t.shape[0], 1), float(s[0]))

    def mm(a, a_s, wb, w_s):
        """a fp64 [r, K] with scales a_s; wb uint8 [N, K]; w_s per scheme -> fp32 [r, N] rounded to dt"""
        w = e4m3_bits_to_f32(wb.numpy()).astype(np.float64)
        if scheme == "block":
            K = a.shape[1]
            out = np.zeros((a.shape[0], w.shape[0]))
            ws = np.asarray(w_s, dtype=np.float64)            # [N/128, K/128]
            for kb in range((K + 127) // 128):
                ks = slice(kb * 128, min(K, (kb + 1) * 128))
                out += (a[:, ks] @ w[:, ks].T) * a_s[:, kb:kb + 1] * np.repeat(ws[:, kb], 128)[None, :w.shape[0]]
        else:
            ws = np.asarray(w_s, dtype=np.float64).reshape(1, -1)   # [1, 1] or [1, N]
            out = (a @ w.T) * a_s * ws
        return torch.from_numpy(out.astype(np.float32)).to(dt)

    x8, x_s = quant(x, a1_scale)
    flat = topk_ids.reshape(-1).numpy()
    inter = torch.zeros(M * topk, N1, dtype=dt)
    for e in range(E):
        rows = np.nonzero(flat == e)[0]
        if rows.size:
            tok = rows // topk
            inter[torch.from_numpy(rows)] = mm(x8[tok], x_s[tok], w1_bits[e], w1_scale[e].numpy())
    valid = flat >= 0
    act = torch.zeros(M * topk, I, dtype=dt)
    act[torch.from_numpy(valid)] = silu_and_mul(inter[torch.from_numpy(valid)])
    # the dynamic per-tensor scale reduces over the rows that exist (dropped pairs never enter the buffer)
    h8 = np.zeros((M * topk, I))
    h_s = np.zeros((M * topk, max(1, I // 128 if scheme == "block" else 1)))
    vi = np.nonzero(valid)[0]
    q, s = quant(act[torch.from_numpy(vi)], a2_scale)
    h8[vi], h_s[vi] = q, s
    out2 = torch.zeros(M * topk, H, dtype=dt)
    for e in range(E):
        rows = np.nonzero(flat == e)[0]
        if rows.size:
            out2[torch.from_numpy(rows)] = mm(h8[rows], h_s[rows], w2_bits[e], w2_scale[e].numpy())
    o = out2.float().view(M, topk, H)
    if not apply_router_weight_on_input:
        o = o * topk_weights.float().view(M, topk, 1)
    return o.sum(dim=1)
