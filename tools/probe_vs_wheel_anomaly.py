"""Why does tools/bench_vs_wheel.py time the v5 GEMM at 34 us on 256 x 4096 x 14336 when tools/bench_gemm.py times the
same call at 25 us? Times the call (a) in a clean process state, (b) after loading the wheel's libraries, (c) after
running one of the wheel's GEMMs, (d) with more / fewer rotated weight copies."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vllm-fork_b200"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
from b200ops import ops  # noqa: E402
import bench_vs_wheel as bvw  # noqa: E402

DEV = torch.device("cuda:0")


def t_ours(M, N, K, ncopy, impl=None):
    ws = [bvw.rand_fp8(N, K) for _ in range(ncopy)]
    x = bvw.rand_fp8(M, K)
    sa, sb = torch.ones(1, device=DEV), torch.ones(1, device=DEV)
    out = torch.empty(M, N, dtype=torch.bfloat16, device=DEV)
    return round(bvw.graph_time(lambda i: ops.cutlass_scaled_mm(x, ws[i].t(), sa, sb, torch.bfloat16, out=out, impl=impl), ncopy), 2)


def main():
    shapes = [(256, 4096, 14336), (128, 1280, 8192), (256, 4096, 4096)]
    for tag in ("clean", "wheel loaded", "after a wheel GEMM"):
        if tag == "wheel loaded":
            W = bvw.load_wheel()
        if tag == "after a wheel GEMM":
            x = bvw.rand_fp8(256, 4096); w = bvw.rand_fp8(4096, 4096)
            o = torch.empty(256, 4096, dtype=torch.bfloat16, device=DEV)
            one = torch.ones(1, device=DEV)
            W._C.cutlass_scaled_mm(o, x, w.t(), one, one, None)
            torch.cuda.synchronize()
        for (M, N, K) in shapes:
            nc = max(2, int(300e6 // (N * K)) + 1)
            print(tag, (M, N, K), "ncopy", nc, "auto", t_ours(M, N, K, nc), "v5", t_ours(M, N, K, nc, "v5"),
                  "v2", t_ours(M, N, K, nc, "v2"), "ncopy 8:", t_ours(M, N, K, 8), flush=True)


if __name__ == "__main__":
    main()
