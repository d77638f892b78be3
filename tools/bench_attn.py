"""Micro-benchmark of the paged-attention decode kernel alone (development tool; bench.py is the
judged benchmark). Usage: python tools/bench_attn.py [--seqs 256 --ctx 4096 --hq 32 --hkv 8 --d 128
--kv fp8|bf16 --bs 16 --iters 20 --ragged]"""
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
    ap.add_argument("--seqs", type=int, default=256)
    ap.add_argument("--ctx", type=int, default=4096)
    ap.add_argument("--hq", type=int, default=32)
    ap.add_argument("--hkv", type=int, default=8)
    ap.add_argument("--d", type=int, default=128)
    ap.add_argument("--bs", type=int, default=16)
    ap.add_argument("--kv", default="fp8")
    ap.add_argument("--layout", default="NHD")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--ncache", type=int, default=2)
    ap.add_argument("--ragged", action="store_true")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    S, L, Hq, Hkv, D, bs = a.seqs, a.ctx, a.hq, a.hkv, a.d, a.bs
    g = torch.Generator(device=dev).manual_seed(0)
    nblk = (L + bs - 1) // bs
    nb = int(S * nblk * 1.1) + 1
    fp8 = a.kv == "fp8"
    caches = []
    for _ in range(a.ncache):
        if a.layout == "NHD":
            shape = (nb, bs, Hkv, 2 * D)
        else:
            shape = (nb, Hkv, bs, 2 * D)
        if fp8:
            raw = torch.randint(0, 0x78, shape, dtype=torch.uint8, device=dev, generator=g)
        else:
            raw = (torch.rand(shape, device=dev, generator=g) * 2 - 1).mul(D ** -0.5).to(torch.bfloat16)
        kv = raw if a.layout == "NHD" else raw.transpose(1, 2)
        caches.append((kv[..., :D], kv[..., D:]))
    bt = torch.randperm(nb, device=dev, generator=g)[: S * nblk].to(torch.int32).view(S, nblk)
    if a.ragged:
        lens = torch.randint(L // 2, L + 1, (S,), device=dev, generator=g).to(torch.int32)
    else:
        lens = torch.full((S,), L, dtype=torch.int32, device=dev)
    q = (torch.rand(S, Hq, D, device=dev, generator=g) * 2 - 1).mul(D ** -0.5).to(torch.bfloat16)
    out = torch.empty_like(q)
    ks = torch.tensor([1.0 / 64], device=dev) if fp8 else None
    vs = torch.tensor([1.0 / 64], device=dev) if fp8 else None
    es = 1 if fp8 else 2
    kv_bytes = int(lens.sum().item()) * 2 * Hkv * D * es
    alg_bytes = kv_bytes + S * Hq * D * 4 + int(((lens + bs - 1) // bs).sum().item()) * 4 + 4 * S

    def run(i):
        kc, vc = caches[i % a.ncache]
        ops.paged_attention_decode(out, q, kc, vc, bt, lens, D ** -0.5, k_scale=ks, v_scale=vs)

    for i in range(5):
        run(i)
    torch.cuda.synchronize()
    times = []
    for i in range(a.iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        run(i)
        e1.record()
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1))
    times.sort()
    med, mn = times[len(times) // 2], times[0]
    print(json.dumps({"shape": vars(a), "alg_bytes": alg_bytes, "ms_median": med, "ms_min": mn,
                      "GBps_median": alg_bytes / med / 1e6, "GBps_best": alg_bytes / mn / 1e6}))


if __name__ == "__main__":
    main()
