"""Development tool: per-pair timeline of the v3 FP8 GEMM (needs a library built with
`make EXTRA=-DB200_V3_TRACE`). Prints, per CTA pair and piece, cycles spent by the MMA thread
(wait for a free TMEM buffer, issue) and by the epilogue (wait for the accumulator, drain, rest)."""
import ctypes
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vllm-fork_b200"))
from b200ops import _lib, ops  # noqa: E402


def main():
    shapes = sys.argv[1] if len(sys.argv) > 1 else "256,28672,4096;4096,4096,4096"
    lib = _lib.load()
    raw = ctypes.CDLL(os.path.join(ROOT, "vllm-fork_b200", "b200ops", "libb200ops.so"))
    dev = torch.device("cuda:0")
    trace = torch.zeros(74 * 2 * 8 * 4, dtype=torch.int64, device=dev)
    for shp in shapes.split(";"):
        M, N, K = [int(x) for x in shp.split(",")]
        w = torch.randint(0, 0x70, (N, K), dtype=torch.uint8, device=dev).view(torch.float8_e4m3fn)
        x = torch.randint(0, 0x70, (M, K), dtype=torch.uint8, device=dev).view(torch.float8_e4m3fn)
        sa = torch.ones(1, device=dev)
        sb = torch.ones(1, device=dev)
        out = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
        raw.b200_fp8_gemm_v3_set_trace(ctypes.c_void_p(0))
        for _ in range(3):
            ops.cutlass_scaled_mm(x, w.t(), sa, sb, torch.bfloat16, out=out, impl="v3")
        torch.cuda.synchronize()
        trace.zero_()
        raw.b200_fp8_gemm_v3_set_trace(ctypes.c_void_p(trace.data_ptr()))
        ops.cutlass_scaled_mm(x, w.t(), sa, sb, torch.bfloat16, out=out, impl="v3")
        torch.cuda.synchronize()
        raw.b200_fp8_gemm_v3_set_trace(ctypes.c_void_p(0))
        t = trace.cpu().view(74, 2, 8, 4)
        print(f"== {M}x{N}x{K}")
        for pair in (0, 1, 2, 36, 37, 72, 73):
            t0 = int(t[pair, 0, 0, 0])
            rows = []
            for n in range(8):
                m = [int(v) for v in t[pair, 0, n]]
                e = [int(v) for v in t[pair, 1, n]]
                if m[0] == 0 and e[0] == 0:
                    break
                rows.append(f"  piece {n}: MMA wait_tempty@{m[0]-t0} got@{m[1]-t0} issued@{m[2]-t0} | "
                            f"EPI wait_tfull@{e[0]-t0} got@{e[1]-t0} released@{e[2]-t0} done@{e[3]-t0}")
            print(f" pair {pair}:")
            print("\n".join(rows))


if __name__ == "__main__":
    main()
