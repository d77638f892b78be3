"""Micro-benchmark of the sampler-side kernels (development tool): greedy argmax and top-k / top-p masking at
batch 256 x vocab 128256, beside the torch ops the reference uses for the same step (logits.argmax; the sort-based
apply_top_k_top_p_pytorch restated in oracle/ is NOT timed here - torch.sort + softmax + cumsum + scatter are)."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vllm-fork_b200"))
from b200ops import ops  # noqa: E402


def timeit(fn, iters=20):
    """CUDA-graph replay of `fn` (10 calls per graph): python / ctypes launch cost (~30 us per call) must not be
    part of a kernel number."""
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(10):
            fn()
    g.replay()
    torch.cuda.synchronize()
    return _time(g.replay, max(1, iters // 10)) / 10


def _time(fn, iters):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


def main():
    dev = "cuda:0"
    B, V = 256, 128256
    g = torch.Generator(device=dev).manual_seed(0)
    lb = (torch.randn(B, V, device=dev, generator=g) * 3).to(torch.bfloat16)
    lf = lb.float()
    k = torch.randint(1, 100, (B,), device=dev, generator=g, dtype=torch.int32)
    p = torch.rand(B, device=dev, generator=g) * 0.9 + 0.05
    out = torch.empty(B, dtype=torch.int64, device=dev)
    res = {"batch": B, "vocab": V}
    res["argmax_bf16_us"] = round(timeit(lambda: ops.greedy_sample(lb, out=out)), 1)
    res["torch_argmax_bf16_us"] = round(timeit(lambda: lb.argmax(-1)), 1)
    res["argmax_bf16_GBps"] = round(B * V * 2 / res["argmax_bf16_us"] / 1e3, 1)
    work = lf.clone()

    def ours():
        work.copy_(lf)
        ops.apply_top_k_top_p(work, k, p)

    def torch_sort():
        x = lf.clone()
        s, idx = x.sort(dim=-1, descending=False)
        kth = s.gather(1, (V - k.long()).unsqueeze(1))
        s.masked_fill_(s < kth, -float("inf"))
        pr = s.softmax(-1)
        cs = torch.cumsum(pr, -1)
        m = cs <= 1 - p.unsqueeze(1)
        m[:, -1] = False
        s.masked_fill_(m, -float("inf"))
        x.scatter_(-1, idx, s)

    copy_us = timeit(lambda: work.copy_(lf))
    res["top_k_top_p_us"] = round(timeit(ours) - copy_us, 1)
    res["torch_sort_based_us"] = round(timeit(torch_sort, 5) - copy_us, 1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
