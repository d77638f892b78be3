"""Micro-benchmark of the variable-length (prefill) attention kernel (development tool).
usage: python tools/bench_prefill.py [--seqs 8 --qlen 4096 --ctx 0 --fp8]"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vllm-fork_b200"))
from b200ops import ops  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seqs", type=int, default=8)
    ap.add_argument("--qlen", type=int, default=4096)
    ap.add_argument("--ctx", type=int, default=0)
    ap.add_argument("--heads", type=int, default=32)
    ap.add_argument("--kv-heads", type=int, default=8)
    ap.add_argument("--head-size", type=int, default=128)
    ap.add_argument("--iters", type=int, default=10)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    S, Q, C, Hq, Hkv, D, bs = a.seqs, a.qlen, a.ctx, a.heads, a.kv_heads, a.head_size, 16
    L = Q + C
    mb = (L + bs - 1) // bs
    nb = S * mb + 1
    bt = torch.randperm(nb, device=dev)[: S * mb].to(torch.int32).view(S, mb)
    cu = (torch.arange(S + 1, device=dev) * Q).to(torch.int32)
    sl = torch.full((S,), L, dtype=torch.int32, device=dev)
    q = (torch.rand(S * Q, Hq, D, device=dev) * 2 - 1).to(torch.bfloat16)
    out = torch.empty_like(q)
    pairs = S * sum(C + i + 1 for i in range(Q))
    flops = 4.0 * D * Hq * pairs
    for fp8 in (False, True):
        if fp8:
            raw = torch.randint(0, 0x70, (nb, bs, Hkv, 2 * D), dtype=torch.uint8, device=dev)
            ks = torch.tensor([0.01], device=dev)
            vs = torch.tensor([0.01], device=dev)
        else:
            raw = (torch.rand(nb, bs, Hkv, 2 * D, device=dev) * 2 - 1).to(torch.bfloat16)
            ks = vs = None
        kc, vc = raw[..., :D], raw[..., D:]

        def run():
            ops.unified_attention(q, kc, vc, out, cu, Q, sl, L, D ** -0.5, True, (-1, -1), bt, 0.0, None, ks, vs)
        for _ in range(3):
            run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.iters):
            run()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / a.iters
        print(json.dumps({"kv": "fp8" if fp8 else "bf16", "seqs": S, "qlen": Q, "ctx": C, "ms": round(ms, 3),
                          "TFLOPs": round(flops / ms / 1e9, 1)}))


if __name__ == "__main__":
    main()
