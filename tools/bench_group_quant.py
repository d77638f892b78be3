"""Micro-benchmark of per_token_group_fp8_quant (development tool): CUDA-graph timed, rotating over inputs larger
than L2. Algorithmic bytes = M*K*(b_in + 1) + 4*M*K/group."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vllm-fork_b200"))
from b200ops import ops  # noqa: E402


def main():
    dev = "cuda:0"
    for M, K in ((256, 7168), (4096, 7168), (16384, 7168)):
        n = max(2, int(400e6 // (M * K * 2)) + 1)
        xs = [torch.randn(M, K, device=dev, dtype=torch.bfloat16) for _ in range(n)]
        q = torch.empty(M, K, device=dev, dtype=torch.float8_e4m3fn)
        for col in (False, True):
            for x in xs[:2]:
                ops.per_token_group_quant_fp8(x, 128, column_major_scales=col, out_q=q)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                for x in xs:
                    ops.per_token_group_quant_fp8(x, 128, column_major_scales=col, out_q=q)
            g.replay()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5):
                g.replay()
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) / (5 * n) * 1e3
            byts = M * K * 3 + 4 * M * K // 128
            print(json.dumps({"M": M, "K": K, "group": 128, "column_major_scales": col, "us": round(us, 2),
                              "GBps": round(byts / us / 1e3, 1)}))


if __name__ == "__main__":
    main()
