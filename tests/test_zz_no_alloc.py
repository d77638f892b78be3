"""b200_paged_attention_decode_static: decode attention for callers that cannot allocate device memory.

Who needs it: a worker that gave all of its device memory away at start-up (weights first, every
remaining byte as KV-cache blocks) and therefore cannot cudaMalloc a split-KV workspace later, and
bindings from runtimes that have no device allocator at all (cgo, JNI, bare ctypes) and only hold the
buffers they were handed. For them the library carries an 8 MiB scratch array in its own device image;
the entry point takes the usual argument struct, ignores its workspace fields and allocates nothing.

What is checked here, through the raw C ABI (ctypes, no b200ops.ops wrapper in between):
  * the output equals the regular entry point's (same kernels, caller-owned workspace) on shapes that
    really use the scratch - long sequences cut across CTAs and merged by the reduce kernel;
  * two calls in a row on one stream, with DIFFERENT ragged batches and no synchronisation between them:
    the second call runs on the scratch the first one left behind (it is never cleared) and must not care;
  * a shape whose scratch exceeds the built-in array is refused with a clear message, nothing launched
    (needs no GPU: the size gate comes before any device access).

This file is named test_zz_* so that pytest collects it after every other test: the driver runs with -x,
and a failure here must not hide the parity tests. It passes alone in a fresh process (then the first call
below is the one that performs the per-device address look-up) and after the rest of the suite.
Tolerance: the suite's decode-attention tolerance (tests/test_gpu_attention.py: atol 5e-3 + rtol 1e-2)."""
import ctypes
import os

import numpy as np
import pytest
import torch

ATOL, RTOL = 5e-3, 1e-2
DEV = "cuda:0"
STATIC_WS_BYTES = 8 << 20   # B200_PAGED_ATTN_STATIC_WS_BYTES in include/b200ops.h


def _lib():
    from b200ops import _lib as L
    if not os.path.exists(L.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    return L, L.load()


def _args(L, out, q, kc, vc, bt, lens, scale, ks, vs, workspace):
    """Fill struct b200_paged_attn_args exactly as a foreign-language binding would: pointers and sizes."""
    p = lambda t: None if t is None else t.data_ptr()
    a = L.PagedAttnArgs()
    a.out, a.lse, a.q = p(out), None, p(q)
    a.k_cache, a.v_cache = p(kc), p(vc)
    a.block_table, a.seq_lens = p(bt), p(lens)
    a.k_scale, a.v_scale = p(ks), p(vs)
    a.alibi_slopes = None
    if workspace is None:
        a.workspace, a.workspace_bytes = None, 0
    else:
        a.workspace, a.workspace_bytes = workspace.data_ptr(), workspace.numel()
    S, Hq, D = q.shape
    nb, bs, Hkv, _ = kc.shape
    a.num_blocks = nb
    a.q_stride_s, a.q_stride_h = q.stride(0), q.stride(1)
    a.out_stride_s, a.out_stride_h = out.stride(0), out.stride(1)
    a.kv_block_stride, a.kv_token_stride, a.kv_head_stride = kc.stride(0), kc.stride(1), kc.stride(2)
    a.bt_stride = bt.stride(0)
    a.num_seqs, a.num_heads, a.num_kv_heads, a.head_size, a.block_size = S, Hq, Hkv, D, bs
    a.max_seq_len = 0
    a.q_dtype = L.BF16
    a.kv_dtype = L.FP8_E4M3 if kc.element_size() == 1 else L.BF16
    a.sliding_window = 0
    a.scale = float(scale)
    a.softcap = 0.0
    a.output_scale = None
    return a


def _device_case(lens, Hq, Hkv, D, kv_fp8, seed):
    """A ragged batch on the GPU, KV written by the CPU restatement of reshape_and_cache_flash."""
    from test_gpu_attention import build_case, to_dev_views
    c = build_case(lens, Hq, Hkv, D, 16, torch.bfloat16, kv_fp8, seed=seed)
    raw, kc, vc = to_dev_views(c)
    dev = lambda t: None if t is None else t.to(DEV)
    return dict(raw=raw, kc=kc, vc=vc, q=c["q"].to(DEV), bt=c["bt"].to(DEV), lens=c["lens"].to(DEV),
                ks=dev(c["ks"]), vs=dev(c["vs"]), scale=D ** -0.5)


@pytest.mark.gpu
@pytest.mark.parametrize("kv_fp8", [False, True])
def test_static_scratch_matches_regular_entry_twice_in_a_row(kv_fp8):
    L, lib = _lib()
    # both batches have sequences far longer than one CTA's share of tiles, so partials go through the scratch
    cases = [_device_case([20000, 3, 9000, 77, 5000], 32, 8, 128, kv_fp8, seed=11),
             _device_case([4097, 12000, 1, 640], 32, 8, 128, kv_fp8, seed=12)]
    stream = torch.cuda.current_stream().cuda_stream

    # expected: the regular entry point with a workspace the caller allocated (torch here)
    want = []
    for c in cases:
        S, Hq, D = c["q"].shape
        n = lib.b200_paged_attention_workspace_bytes(S, Hq, c["kc"].shape[2], D)
        assert 0 < n <= STATIC_WS_BYTES
        ws = torch.zeros(n, dtype=torch.uint8, device=DEV)
        out = torch.full_like(c["q"], float("nan"))
        a = _args(L, out, c["q"], c["kc"], c["vc"], c["bt"], c["lens"], c["scale"], c["ks"], c["vs"], ws)
        L.check(lib.b200_paged_attention_decode(ctypes.byref(a), stream), "paged_attention_decode")
        torch.cuda.synchronize()
        want.append(out)

    # the entry point under test: no workspace passed, two calls back to back, one synchronise at the end
    got = [torch.full_like(c["q"], float("nan")) for c in cases]
    for c, out in zip(cases, got):
        a = _args(L, out, c["q"], c["kc"], c["vc"], c["bt"], c["lens"], c["scale"], c["ks"], c["vs"], None)
        L.check(lib.b200_paged_attention_decode_static(ctypes.byref(a), stream), "paged_attention_decode_static")
    torch.cuda.synchronize()

    for i, (g, w) in enumerate(zip(got, want)):
        g, w = g.float().cpu().numpy(), w.float().cpu().numpy()
        assert not np.isnan(w).any(), "regular entry point produced NaN"
        assert not np.isnan(g).any(), f"call {i}: NaN in the output of the static-scratch entry point"
        np.testing.assert_allclose(g, w, atol=ATOL, rtol=RTOL, err_msg=f"call {i} differs from the regular entry")


def test_static_scratch_refuses_shapes_beyond_its_array():
    """64 heads of size 128 need 12.5 MiB of partials: more than the array. The refusal names both sizes
    and points at the regular entry point; it happens before any device access, so it is testable on CPU."""
    L, lib = _lib()
    a = L.PagedAttnArgs()
    a.num_seqs, a.num_heads, a.num_kv_heads, a.head_size, a.block_size = 8, 64, 8, 128, 16
    need = lib.b200_paged_attention_workspace_bytes(8, 64, 8, 128)
    assert need > STATIC_WS_BYTES
    rc = lib.b200_paged_attention_decode_static(ctypes.byref(a), None)
    assert rc == L.ERR_WORKSPACE
    msg = L.last_error()
    assert str(need) in msg and str(STATIC_WS_BYTES) in msg and "b200_paged_attention_decode" in msg
    # the largest documented shape is inside the limit
    assert lib.b200_paged_attention_workspace_bytes(16384, 40, 8, 128) <= STATIC_WS_BYTES
    assert lib.b200_paged_attention_decode_static(None, None) == L.ERR_INVALID
