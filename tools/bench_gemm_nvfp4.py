"""Micro-benchmark of the NVFP4 W4A4 GEMM and the fused-MoE FP8 path (development tool, CUDA-graph timed,
weights rotated beyond L2). Usage: python tools/bench_gemm_nvfp4.py"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vllm-fork_b200"))
from b200ops import fused_moe, ops  # noqa: E402

DEV = torch.device("cuda:0")


def timed_graph(fn_list, reps=5):
    for f in fn_list[:2]:
        f()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for f in fn_list:
            f()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (reps * len(fn_list)) * 1e3


def nvfp4(M, N, K):
    ncopy = max(2, int(300e6 // (N * K // 2)) + 1)
    gs = torch.tensor([100.0], device=DEV)
    x = torch.randn(M, K, device=DEV).to(torch.bfloat16)
    a4, sfa = ops.scaled_fp4_quant(x, gs)
    ws = []
    for _ in range(ncopy):
        w = torch.randn(N, K, device=DEV).to(torch.bfloat16)
        ws.append(ops.scaled_fp4_quant(w, gs))
        del w
    alpha = torch.tensor([1e-4], device=DEV)
    out = torch.empty(M, N, dtype=torch.bfloat16, device=DEV)
    us = timed_graph([lambda i=i: ops.cutlass_scaled_fp4_mm(a4, ws[i][0], sfa, ws[i][1], alpha, torch.bfloat16, out=out)
                      for i in range(ncopy)])
    byts = (M * K + N * K) * 0.5625 + 2 * M * N
    print(json.dumps({"op": "nvfp4_gemm", "M": M, "N": N, "K": K, "us": round(us, 2),
                      "TFLOPs": round(2.0 * M * N * K / us / 1e6, 1), "GBps": round(byts / us / 1e3, 1)}), flush=True)


def moe(M, E, k, H, I):
    """Mixtral-shaped fused_experts (FP8 per-tensor), all experts local. C4: M=256, E=8, k=2."""
    ncopy = 2
    w1 = [torch.randint(0, 0x70, (E, 2 * I, H), dtype=torch.uint8, device=DEV).view(torch.float8_e4m3fn) for _ in range(ncopy)]
    w2 = [torch.randint(0, 0x70, (E, H, I), dtype=torch.uint8, device=DEV).view(torch.float8_e4m3fn) for _ in range(ncopy)]
    s1 = torch.full((E,), 1e-3, device=DEV)
    s2 = torch.full((E,), 1e-3, device=DEV)
    a1 = torch.tensor([0.01], device=DEV)
    a2 = torch.tensor([0.01], device=DEV)
    x = torch.randn(M, H, device=DEV).to(torch.bfloat16)
    gate = torch.randn(M, E, device=DEV)
    tw = torch.empty(M, k, device=DEV)
    ti = torch.empty(M, k, dtype=torch.int32, device=DEV)
    tt = torch.empty(M, k, dtype=torch.int32, device=DEV)
    fused_moe.topk_softmax(tw, ti, tt, gate, True)
    us = timed_graph([lambda i=i: fused_moe.fused_experts(x, w1[i], w2[i], tw, ti, use_fp8_w8a8=True, w1_scale=s1,
                                                          w2_scale=s2, a1_scale=a1, a2_scale=a2)
                      for i in range(ncopy)] * 3)
    flops = 6.0 * M * k * I * H
    byts = E * 3 * I * H
    print(json.dumps({"op": "fused_experts_fp8", "M": M, "E": E, "topk": k, "H": H, "I": I, "us": round(us, 1),
                      "TFLOPs": round(flops / us / 1e6, 1), "weight_GBps": round(byts / us / 1e3, 1)}), flush=True)


if __name__ == "__main__":
    for shp in [(256, 6144, 4096), (256, 4096, 4096), (256, 28672, 4096), (256, 4096, 14336), (32, 28672, 4096),
                (4096, 4096, 4096), (8192, 8192, 8192)]:
        nvfp4(*shp)
    moe(256, 8, 2, 4096, 14336)
    moe(256, 1, 1, 4096, 14336)   # one EP rank of C4 (every token routed to the local expert: upper bound of its load)
