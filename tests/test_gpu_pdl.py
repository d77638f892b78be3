"""Programmatic dependent launch (b200_set_pdl).
PDL changes WHEN a kernel's prologue runs, never its result: the same chain of hot-path kernels must produce
identical bytes at levels 0 / 1 / 2, eagerly and from a replayed CUDA graph. The split-KV attention (decode kernel +
reduce kernel, both launched programmatically) must stay correct over many back-to-back launches on one workspace
with different cut patterns."""
import numpy as np
import pytest
import torch

from b200ops import ops
from oracle import ref_ops
from helpers import make_block_table, make_kv_cache, slots_for

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _fp8(*shape, gen):
    return torch.randint(0, 0x78, shape, dtype=torch.uint8, device=DEV, generator=gen).view(torch.float8_e4m3fn)


class _Layer:
    """norm+quant -> qkv GEMM -> rope+cache -> attention (FP8 out) -> o GEMM -> add+norm+quant -> gate_up GEMM ->
    silu+quant -> down GEMM, the kernel chain of one decoder layer (b200ops/decoder.py)."""

    def __init__(self, M, seed):
        g = torch.Generator(device=DEV).manual_seed(seed)
        self.M, self.H, self.Hq, self.Hkv, self.D, self.I = M, 1024, 8, 2, 128, 2048
        H, Hq, Hkv, D, I = self.H, self.Hq, self.Hkv, self.D, self.I
        self.x = (torch.rand(M, H, device=DEV, generator=g) * 2 - 1).to(torch.bfloat16)
        self.w_norm = (torch.rand(H, device=DEV, generator=g) + 0.5).to(torch.bfloat16)
        self.wqkv, self.wo = _fp8((Hq + 2 * Hkv) * D, H, gen=g), _fp8(H, Hq * D, gen=g)
        self.wgu, self.wd = _fp8(2 * I, H, gen=g), _fp8(H, I, gen=g)
        self.s = {k: torch.tensor([v], device=DEV) for k, v in
                  dict(a_qkv=0.02, w=2e-3, a_o=0.01, a_gu=0.05, a_d=0.02, kv=0.05).items()}
        self.bs, self.ctx = 16, 700
        nblk = (self.ctx + self.bs - 1) // self.bs
        nb = M * nblk + 1
        cg = torch.Generator().manual_seed(seed)
        self.bt = make_block_table([self.ctx] * M, self.bs, nb, cg).to(DEV)
        self.kv = torch.randint(0, 0x60, (nb, self.bs, Hkv, 2 * D), dtype=torch.uint8, device=DEV, generator=g)
        self.lens = torch.full((M,), self.ctx, dtype=torch.int32, device=DEV)
        self.pos = torch.full((M,), self.ctx - 1, dtype=torch.int64, device=DEV)
        self.slots = (self.bt[:, (self.ctx - 1) // self.bs].long() * self.bs + (self.ctx - 1) % self.bs)
        inv = 1.0 / (10000.0 ** (torch.arange(0, D, 2, device=DEV).float() / D))
        f = torch.outer(torch.arange(1024, device=DEV).float(), inv)
        self.cos_sin = torch.cat([f.cos(), f.sin()], -1).to(torch.bfloat16)
        self.ws = ops.paged_attention_workspace(M, Hq, Hkv, D, torch.device(DEV))
        self.x8 = torch.empty(M, H, dtype=torch.float8_e4m3fn, device=DEV)
        self.qkv = torch.empty(M, (Hq + 2 * Hkv) * D, dtype=torch.bfloat16, device=DEV)
        self.a8 = torch.empty(M, Hq * D, dtype=torch.float8_e4m3fn, device=DEV)
        self.o = torch.empty(M, H, dtype=torch.bfloat16, device=DEV)
        self.gu = torch.empty(M, 2 * I, dtype=torch.bfloat16, device=DEV)
        self.h8 = torch.empty(M, I, dtype=torch.float8_e4m3fn, device=DEV)
        self.out = torch.empty(M, H, dtype=torch.bfloat16, device=DEV)
        self.res = torch.empty_like(self.x)
        self.kv0 = self.kv.clone()

    def reset(self):
        self.kv.copy_(self.kv0)
        self.res.copy_(self.x)
        for t in (self.x8, self.a8, self.h8):
            t.view(torch.uint8).zero_()
        for t in (self.qkv, self.o, self.gu, self.out):
            t.zero_()

    def run(self):
        M, Hq, Hkv, D, s = self.M, self.Hq, self.Hkv, self.D, self.s
        ops.rms_norm_static_fp8_quant(self.x8, self.x, self.w_norm, s["a_qkv"], 1e-5)
        ops.cutlass_scaled_mm(self.x8, self.wqkv.t(), s["a_qkv"], s["w"], torch.bfloat16, out=self.qkv)
        q = self.qkv[:, : Hq * D].view(M, Hq, D)
        k = self.qkv[:, Hq * D:(Hq + Hkv) * D].view(M, Hkv, D)
        v = self.qkv[:, (Hq + Hkv) * D:].view(M, Hkv, D)
        kc, vc = self.kv.view(torch.float8_e4m3fn)[..., :D], self.kv.view(torch.float8_e4m3fn)[..., D:]
        ops.rope_and_reshape_and_cache_flash(self.pos, q, k, v, D, self.cos_sin, True, kc, vc, self.slots, "fp8",
                                             s["kv"], s["kv"])
        ops.paged_attention_decode(self.a8.view(M, Hq, D), q, kc, vc, self.bt, self.lens, D ** -0.5,
                                   k_scale=s["kv"], v_scale=s["kv"], max_seq_len=self.ctx, workspace=self.ws,
                                   output_scale=s["a_o"])
        ops.cutlass_scaled_mm(self.a8, self.wo.t(), s["a_o"], s["w"], torch.bfloat16, out=self.o)
        ops.fused_add_rms_norm_static_fp8_quant(self.x8, self.o, self.res, self.w_norm, s["a_gu"], 1e-5)
        ops.cutlass_scaled_mm(self.x8, self.wgu.t(), s["a_gu"], s["w"], torch.bfloat16, out=self.gu)
        ops.silu_and_mul_quant(self.h8, self.gu, s["a_d"])
        ops.cutlass_scaled_mm(self.h8, self.wd.t(), s["a_d"], s["w"], torch.bfloat16, out=self.out)

    def snapshot(self):
        torch.cuda.synchronize()
        return [t.view(torch.uint8).clone() for t in (self.x8, self.qkv, self.a8, self.o, self.res, self.gu, self.h8,
                                                      self.out, self.kv)]


@pytest.mark.parametrize("M", [256, 64, 7])
def test_pdl_levels_give_identical_bytes(M):
    L = _Layer(M, seed=M)
    prev = ops.set_pdl(0)
    try:
        L.reset()
        L.run()
        ref = L.snapshot()
        assert float(ref[7].view(torch.bfloat16).float().abs().max()) > 0
        for level in (1, 2):
            ops.set_pdl(level)
            L.reset()
            for _ in range(3):          # back-to-back launches: prologues overlap the previous kernel's tail
                L.res.copy_(L.x)
                L.kv.copy_(L.kv0)
                L.run()
            got = L.snapshot()
            for i, (a, b) in enumerate(zip(ref, got)):
                assert torch.equal(a, b), f"PDL level {level}: tensor {i} differs from level 0"
            # the same chain captured in a CUDA graph (programmatic edges) and replayed
            L.reset()
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                L.run()
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            L.reset()
            with torch.cuda.graph(g):
                L.run()
            for _ in range(3):
                L.reset()
                g.replay()
            got = L.snapshot()
            for i, (a, b) in enumerate(zip(ref, got)):
                assert torch.equal(a, b), f"PDL level {level}, graph replay: tensor {i} differs from level 0"
    finally:
        ops.set_pdl(prev)


def test_split_kv_back_to_back_launches_on_one_workspace():
    """Ragged batches whose sequences are cut by CTA boundaries in different places, launched back to back on ONE
    workspace (eager and without host syncs in between): every launch must match the oracle - with programmatic
    launches the reduce kernel of launch i and the decode kernel of launch i+1 are resident together."""
    gen = torch.Generator().manual_seed(5)
    Hq, Hkv, D, bs = 32, 8, 128, 16
    cases = []
    for S, lo, hi in ((3, 3000, 9000), (40, 1, 900), (256, 1, 64), (1, 20000, 20001), (17, 0, 3000)):
        lens = torch.randint(lo, hi, (S,), generator=gen).tolist()
        nb = S * ((max(lens) + bs - 1) // bs) + 4
        bt = make_block_table(lens, bs, nb, gen)
        _, kc, vc = make_kv_cache(nb, bs, Hkv, D, torch.bfloat16)
        T = sum(lens)
        key = (torch.rand(T, Hkv, D, generator=gen) * 2 - 1).to(torch.bfloat16)
        val = (torch.rand(T, Hkv, D, generator=gen) * 2 - 1).to(torch.bfloat16)
        one = torch.ones(1)
        ref_ops.reshape_and_cache_flash(key, val, kc, vc, slots_for(lens, bt, bs), "auto", one, one)
        q = (torch.rand(S, Hq, D, generator=gen) * 2 - 1).to(torch.bfloat16)
        ref = ref_ops.paged_attention_decode(q, kc, vc, bt, torch.tensor(lens, dtype=torch.int32), D ** -0.5)
        cases.append((q.to(DEV), kc.to(DEV), vc.to(DEV), bt.to(DEV),
                      torch.tensor(lens, dtype=torch.int32, device=DEV), ref, lens))
    ws = ops.paged_attention_workspace(256, Hq, Hkv, D, torch.device(DEV))
    outs = []
    for rep in range(3):
        for q, kc, vc, bt, lens_d, ref, lens in cases:
            out = torch.full((q.shape[0], Hq, D), float("nan"), dtype=torch.bfloat16, device=DEV)
            ops.paged_attention_decode(out, q, kc, vc, bt, lens_d, D ** -0.5, max_seq_len=max(lens), workspace=ws)
            outs.append((out, ref, lens))
    torch.cuda.synchronize()
    for out, ref, lens in outs:
        o = out.float().cpu().numpy()
        assert np.isfinite(o).all()
        np.testing.assert_allclose(o, ref, atol=5e-3, rtol=1e-2)
        for s, L in enumerate(lens):
            if L == 0:
                assert not o[s].any()
