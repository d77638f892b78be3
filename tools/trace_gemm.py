"""Development tool: per-CTA timeline of the v2 FP8 GEMM (decode shapes) from %globaltimer probes
(b200_debug_set_trace). For each probe the distribution over CTAs of (stamp - first CTA start), in us.
usage: python tools/trace_gemm.py "256,4096,4096;256,6144,4096" [BN ...]   (BN forced through B200_GEMM_BN needs
a fresh process: the tool re-executes itself per BN)"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vllm-fork_b200"))
NAMES = ["entry", "setup done", "first TMA issued", "first k-block landed", "last MMA committed",
         "accumulator ready (epi)", "tile staged in smem (epi)", "exit"]


def run(shapes):
    import torch
    from b200ops import _lib, ops
    lib = _lib.load()
    dev = torch.device("cuda:0")
    for shp in shapes.split(";"):
        M, N, K = [int(x) for x in shp.split(",")]
        ws = [torch.randint(0, 0x70, (N, K), dtype=torch.uint8, device=dev).view(torch.float8_e4m3fn) for _ in range(8)]
        x = torch.randint(0, 0x70, (M, K), dtype=torch.uint8, device=dev).view(torch.float8_e4m3fn)
        one = torch.ones(1, device=dev)
        out = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
        for w in ws[:4]:
            ops.cutlass_scaled_mm(x, w.t(), one, one, torch.bfloat16, out=out, impl="v2")
        torch.cuda.synchronize()
        trace = torch.zeros(4096 * 8 + 4 * 64, dtype=torch.int64, device=dev)
        lib.b200_debug_set_trace(trace.data_ptr())
        ops.cutlass_scaled_mm(x, ws[5].t(), one, one, torch.bfloat16, out=out, impl="v2")
        torch.cuda.synchronize()
        lib.b200_debug_set_trace(None)
        fine = trace.cpu()[32768:].view(4, 64)
        t = trace.cpu()[:32768].view(-1, 8)
        t = t[t[:, 0] > 0].double()
        t0 = t[:, 0].min()
        print(f"== M={M} N={N} K={K}  BN={os.environ.get('B200_GEMM_BN', 'auto')} stages={os.environ.get('B200_GEMM_STAGES', 'auto')}"
              f"  CTAs={t.shape[0]}  span={(t[:, 7].max() - t0) / 1e3:.2f} us")
        for i, n in enumerate(NAMES):
            c = (t[:, i] - t0) / 1e3
            print(f"   {n:28s} min {c.min():6.2f}  median {c.median():6.2f}  max {c.max():6.2f} us")
        if os.environ.get("TRACE_FINE"):
            nkb = min(64, (K + 127) // 128)
            c0 = int(fine[0, 0])
            print("   CTA(0,0) clock64 per k-block: producer got-empty / issued | MMA got-full / issued+committed (cycles since first)")
            for kb in range(nkb):
                print(f"     kb {kb:2d}: P {int(fine[0, kb]) - c0:7d} {int(fine[1, kb]) - c0:7d} | M {int(fine[2, kb]) - c0:7d} {int(fine[3, kb]) - c0:7d}")
        # graph-timed average over rotated weights
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for w in ws:
                ops.cutlass_scaled_mm(x, w.t(), one, one, torch.bfloat16, out=out, impl="v2")
        g.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            g.replay()
        e1.record(); torch.cuda.synchronize()
        print(f"   graph-timed: {e0.elapsed_time(e1) / 40 * 1e3:.2f} us per call")


if __name__ == "__main__":
    shapes = sys.argv[1] if len(sys.argv) > 1 else "256,4096,4096;256,6144,4096;256,4096,14336"
    if os.environ.get("_TRACE_CHILD"):
        run(shapes)
    else:
        for bn in (sys.argv[2:] or ["auto"]):
            env = dict(os.environ, _TRACE_CHILD="1")
            if bn != "auto":
                if ":" in bn:
                    bn, st = bn.split(":")
                    env["B200_GEMM_STAGES"] = st
                env["B200_GEMM_BN"] = bn
            subprocess.run([sys.executable, __file__, shapes], env=env)
