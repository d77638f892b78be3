"""Micro-benchmark of the FP8 GEMM (development tool). Rotates over several weight copies so the
working set exceeds L2. Usage: python tools/bench_gemm.py [--shapes M,N,K;M,N,K ...]"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vllm-fork_b200"))
from b200ops import ops  # noqa: E402

DEFAULT = "256,6144,4096;256,4096,4096;256,28672,4096;256,4096,14336;128,8192,8192;64,4096,4096;16,4096,4096;4096,4096,4096;8192,8192,8192"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shapes", default=DEFAULT)
    ap.add_argument("--iters", type=int, default=30)
    ap.add_argument("--impl", default=None, help="v2 | v3 (default: ops.GEMM_IMPL)")
    ap.add_argument("--block", action="store_true", help="block-scaled GEMM (1x128 / 128x128 scales, csrc/gemm_fp8_block.cu)")
    ap.add_argument("--eager", action="store_true", help="no CUDA graph, 3 launches per shape (for ncu)")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    for shp in a.shapes.split(";"):
        M, N, K = [int(x) for x in shp.split(",")]
        ncopy = max(2, int(300e6 // (N * K)) + 1)
        ws = [torch.randint(0, 0x70, (N, K), dtype=torch.uint8, device=dev).view(torch.float8_e4m3fn)
              for _ in range(ncopy)]
        x = torch.randint(0, 0x70, (M, K), dtype=torch.uint8, device=dev).view(torch.float8_e4m3fn)
        sa = torch.ones(1, device=dev)
        sb = torch.ones(1, device=dev)
        if a.block:
            kb, nb = (K + 127) // 128, (N + 127) // 128
            sa = (torch.rand(kb, M, device=dev) + 0.5).t()   # column-major, as the reference passes it
            sb = torch.rand(kb, nb, device=dev) + 0.5
        out = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
        for i in range(3):
            ops.cutlass_scaled_mm(x, ws[i % ncopy].t(), sa, sb, torch.bfloat16, out=out, impl=a.impl)
        torch.cuda.synchronize()
        if a.eager:
            continue
        # capture one pass over all weight copies in a CUDA graph: python launch cost (~20 us per
        # call: ctypes + 3 tensor-map encodes) must not be part of a kernel number
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for i in range(ncopy):
                ops.cutlass_scaled_mm(x, ws[i].t(), sa, sb, torch.bfloat16, out=out, impl=a.impl)
        g.replay()
        torch.cuda.synchronize()
        reps = max(1, a.iters // ncopy)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(reps):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        a_iters = reps * ncopy
        us = e0.elapsed_time(e1) / a_iters * 1e3
        flops = 2.0 * M * N * K
        byts = M * K + N * K + 2 * M * N
        print(json.dumps({"impl": "block" if a.block else a.impl or ops.GEMM_IMPL, "M": M, "N": N, "K": K, "us": round(us, 2), "TFLOPs": round(flops / us / 1e6, 1),
                          "GBps": round(byts / us / 1e3, 1)}))


if __name__ == "__main__":
    main()
