"""Latency table of the NVLink all-reduce vs NCCL (CUDA-graph timed, device time).
torchrun --nproc-per-node N --master-addr 127.0.0.1 tools/bench_allreduce.py"""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "vllm-fork_b200")):
    sys.path.insert(0, p)
from b200ops.all_reduce import B200AllReduce  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(rank)
    dev = torch.device(f"cuda:{rank}")
    dist.init_process_group("nccl", device_id=dev)
    ar = B200AllReduce(None, dev, max_size=16 << 20, heap_size=96 << 20)
    if rank == 0:
        print(json.dumps({"multicast": ar.multicast_note}), flush=True)

    def timeit(fn, iters=20, reps=5):
        s2 = torch.cuda.Stream()
        s2.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s2):
            for _ in range(3):
                fn()
        torch.cuda.current_stream().wait_stream(s2)
        torch.cuda.synchronize(); dist.barrier()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            for _ in range(iters):
                fn()
        gr.replay(); torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            gr.replay()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / (iters * reps) * 1e3

    for nbytes in (16, 4096, 65536, 262144, 1 << 20, 2 << 20, 8 << 20):
        n = nbytes // 2
        start = ar._heap_used
        x = ar.alloc(n, torch.bfloat16); x.fill_(1.0)
        o = torch.empty(n, dtype=torch.bfloat16, device=dev)
        xn = torch.ones(n, dtype=torch.bfloat16, device=dev)
        t1 = timeit(lambda: ar.all_reduce(x, out=o, algo=1))
        t2 = timeit(lambda: ar.all_reduce(x, out=o, algo=2))
        ts = timeit(lambda: ar.all_reduce(xn, out=o, algo=1))
        tn = timeit(lambda: dist.all_reduce(xn))
        row = {"bytes": nbytes, "world": world, "one_shot_us": round(t1, 2), "two_shot_us": round(t2, 2),
               "one_shot_staged_us": round(ts, 2), "nccl_us": round(tn, 2)}
        if ar.multicast:   # NVSwitch in-network reduction: private output / output broadcast by the switch
            oh = ar.alloc(n, torch.bfloat16)
            row["multimem_us"] = round(timeit(lambda: ar.all_reduce(x, out=o, algo=3)), 2)
            row["multimem_bcast_us"] = round(timeit(lambda: ar.all_reduce(x, out=oh, algo=3)), 2)
        if nbytes >= 65536 and n % 4096 == 0:
            # the fused reduce-scatter + residual add + RMSNorm + FP8 quant + all-gather kernel on [M, 4096]
            M, H = n // 4096, 4096
            w = torch.ones(H, dtype=torch.bfloat16, device=dev)
            shard = torch.zeros((M + world - 1) // world, H, dtype=torch.bfloat16, device=dev)
            sc = torch.tensor([0.05], device=dev)
            o8 = ar.alloc((M, H), torch.float8_e4m3fn)
            x2 = x.view(M, H)
            row["fused_norm_quant_us"] = round(timeit(lambda: ar.fused_add_rms_norm(o8, x2, shard, w, 1e-5, sc)), 2)
        if rank == 0:
            print(json.dumps(row), flush=True)
        torch.cuda.synchronize(); dist.barrier()
        ar._heap_used = start
    ar.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
