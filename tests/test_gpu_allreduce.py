"""GPU parity (needs >= 2 GPUs; skipped on a 1-GPU box): the NVLink all-reduce and the fused
all-reduce + residual + RMSNorm + FP8-quant kernel vs their definitions.

Mirrors tests/distributed/test_custom_all_reduce.py of the reference (eager + CUDA-graph all-reduce of
random tensors of several sizes / dtypes against NCCL) - here the expected value is computed exactly
(fp32 sum in rank order, rounded once), so the comparison is bit-exact, and NCCL is only the
timing baseline. One process per GPU (torch.multiprocessing spawn, rendezvous on 127.0.0.1).
"""
import os
import socket
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank_input(r: int, numel: int, dtype, seed: int) -> torch.Tensor:
    g = torch.Generator().manual_seed(seed * 100 + r)
    return (torch.rand(numel, generator=g) * 2 - 1).to(dtype)


def _expected_sum(world: int, numel: int, dtype, seed: int) -> torch.Tensor:
    acc = torch.zeros(numel, dtype=torch.float32)
    for r in range(world):
        acc = acc + _rank_input(r, numel, dtype, seed).float()  # rank order, fp32: the kernel's order
    return acc.to(dtype)


def _worker(rank: int, world: int, port: int, q, multicast=False):
    try:
        for p in (ROOT, os.path.join(ROOT, "vllm-fork_b200")):
            if p not in sys.path:
                sys.path.insert(0, p)
        import torch.distributed as dist
        from b200ops import ops
        from b200ops.all_reduce import B200AllReduce
        torch.cuda.set_device(rank)
        dev = torch.device(f"cuda:{rank}")
        dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world,
                                device_id=dev)
        # multicast=False: the peer-pull data plane, whose results are bit-exact by construction (fp32 sum in rank
        # order). multicast=None: the NVSwitch multimem plane when the box offers a multicast mapping - the switch's
        # accumulation order and rounding are its own, so the fused kernel is checked within one ulp / one FP8 step
        ar = B200AllReduce(None, dev, max_size=8 << 20, heap_size=48 << 20, multicast=multicast)
        log = ["multicast plane" if ar.multicast else "peer-pull plane"]
        # ---- 1. plain all-reduce: staged and zero-copy inputs, every algorithm, bit-exact
        seed = 0
        for dtype in (torch.bfloat16, torch.float16, torch.float32):
            for numel in (8, 4096, 256 * 4096, 1024 * 1024 + 8):
                for algo in (1, 2, 0):
                    for zero_copy in (False, True):
                        seed += 1
                        x = _rank_input(rank, numel, dtype, seed).to(dev)
                        if zero_copy:
                            start = ar._heap_used
                            buf = ar.alloc(numel, dtype)
                            buf.copy_(x)
                            x = buf
                        out = ar.all_reduce(x, algo=algo)
                        exp = _expected_sum(world, numel, dtype, seed)
                        assert torch.equal(out.cpu(), exp), f"all_reduce mismatch {dtype} n={numel} algo={algo} zc={zero_copy}"
                        if zero_copy:
                            torch.cuda.synchronize()
                            ar._heap_used = start  # bump allocator: release (same on every rank)
        log.append("all_reduce bit-exact")
        # ---- 1b. STAGED inputs while a GEMM on another stream keeps most SMs busy: the CTAs of the collective
        # start at different times, so a CTA must only read packs its peer counterpart staged (the barriers
        # order same-index CTAs across ranks, nothing else); part = 65537 packs is not a multiple of the stride
        busy = torch.randn(4096, 4096, device=dev, dtype=torch.bfloat16)
        side0 = torch.cuda.Stream()
        for it in range(6):
            for algo, numel in ((2, 1024 * 1024 + 8), (1, 1024 * 1024 + 8), (2, 3 * 4096 + 8)):
                seed += 1
                x = _rank_input(rank, numel, torch.bfloat16, seed).to(dev)
                exp = _expected_sum(world, numel, torch.bfloat16, seed)
                torch.cuda.synchronize(); dist.barrier()
                with torch.cuda.stream(side0):
                    for _ in range(2 + (rank + it) % 3):   # ranks are busy for different lengths
                        torch.matmul(busy, busy)
                out = ar.all_reduce(x, algo=algo)
                # a second staged collective right behind the first, asking to skip the end barrier: the
                # library must keep it (the staging area is reused)
                out2 = ar.all_reduce(x, algo=1, skip_end_barrier=True)
                out3 = ar.all_reduce(out2 * 0 + x, algo=1)
                assert torch.equal(out.cpu(), exp), f"staged all_reduce under load mismatch algo={algo} n={numel}"
                assert torch.equal(out2.cpu(), exp) and torch.equal(out3.cpu(), exp), "staged skip_end mismatch"
        log.append("staged all_reduce under SM contention bit-exact")
        # ---- 2. CUDA graph: capture several calls, replay many times with changing inputs
        numel = 256 * 4096
        xg = ar.alloc(numel, torch.bfloat16)
        og = torch.empty(numel, dtype=torch.bfloat16, device=dev)
        og2 = torch.empty(numel, dtype=torch.bfloat16, device=dev)
        xs = torch.empty(numel, dtype=torch.bfloat16, device=dev)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            ar.all_reduce(xg, out=og, algo=2)
            ar.all_reduce(xs, out=og2, algo=1)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            ar.all_reduce(xg, out=og, algo=2)
            ar.all_reduce(xs, out=og2, algo=1)
            ar.all_reduce(xg, out=og2, algo=0)
        for it in range(5):
            xg.copy_(_rank_input(rank, numel, torch.bfloat16, 1000 + it).to(dev))
            xs.copy_(_rank_input(rank, numel, torch.bfloat16, 2000 + it).to(dev))
            graph.replay()
            torch.cuda.synchronize()
            assert torch.equal(og.cpu(), _expected_sum(world, numel, torch.bfloat16, 1000 + it))
            assert torch.equal(og2.cpu(), _expected_sum(world, numel, torch.bfloat16, 1000 + it))
        log.append("graph replay bit-exact")
        # ---- 3. fused reduce-scatter + add + RMSNorm + FP8 quant + all-gather == unfused pair, bit-exact
        for dtype in (torch.bfloat16, torch.float16):
            for (M, H) in ((256, 4096), (128, 8192), (7, 896), (1, 4096), (33, 2048)):
                for quant in (True, False):
                    seed += 1
                    x = _rank_input(rank, M * H, dtype, seed).view(M, H).to(dev)
                    g = torch.Generator().manual_seed(seed)
                    res_full = (torch.rand(M, H, generator=g) * 2 - 1).to(dtype).to(dev)
                    w = (torch.rand(H, generator=g) + 0.5).to(dtype).to(dev)
                    scale = torch.tensor([0.02], device=dev) if quant else None
                    # unfused reference pair on this rank (our own single-GPU ops, already oracle-checked)
                    summed = ar.all_reduce(x.view(-1)).view(M, H)
                    res_ref = res_full.clone()
                    if quant:
                        out_ref = torch.empty(M, H, dtype=torch.float8_e4m3fn, device=dev)
                        ops.fused_add_rms_norm_static_fp8_quant(out_ref, summed, res_ref, w, scale, 1e-5)
                    else:
                        ops.fused_add_rms_norm(summed, res_ref, w, 1e-5)
                        out_ref = summed
                    lo, hi = ar.rows_of_rank(M)
                    mp = (M + world - 1) // world
                    shard = torch.zeros(mp, H, dtype=dtype, device=dev)
                    shard[: hi - lo] = res_full[lo:hi]
                    start = ar._heap_used
                    out = ar.alloc((M, H), torch.float8_e4m3fn if quant else dtype)
                    out.view(torch.uint8).fill_(0x55)
                    torch.cuda.synchronize()
                    dist.barrier()
                    ar.fused_add_rms_norm(out, x, shard, w, 1e-5, scale)
                    torch.cuda.synchronize()
                    dist.barrier()
                    if ar.multicast:
                        # in-switch sum: one ulp of the SUM (|sum| < 2 * world) survives the residual add, is scaled by
                        # rsqrt(mean z^2) * w (<= ~3 here) and by 1 / scale (50) when quantising (bench.py tp_check)
                        o1 = out.view(torch.float8_e4m3fn).float() if quant else out.float()
                        o2 = out_ref.float()
                        kw = max(1.0, world / 2.0)
                        bad = (o1 - o2).abs() > (0.13 * o2.abs() + 2.5 * kw if quant else 2 ** -6 * o2.abs() + 0.05 * kw)
                        assert not bool(bad.any()), f"fused norm (multimem) out of tolerance {dtype} {(M, H)} quant={quant}"
                        torch.testing.assert_close(shard[: hi - lo].float(), res_ref[lo:hi].float(), rtol=2 ** -7,
                                                   atol=2 ** -7 * kw)
                    else:
                        assert torch.equal(out.view(torch.uint8).cpu(), out_ref.view(torch.uint8).cpu()), \
                            f"fused norm out mismatch {dtype} {(M, H)} quant={quant}"
                        assert torch.equal(shard[: hi - lo].cpu(), res_ref[lo:hi].cpu()), "fused norm residual mismatch"
                    ar._heap_used = start
        log.append("fused add+rmsnorm(+fp8) == unfused, " + ("within the switch's rounding" if ar.multicast else "bit-exact"))
        # ---- 3b. decode-context-parallel merge (cp_lse_ag_out_rs / _ar) in one NVLink kernel vs the oracle
        from oracle import ref_ops
        for dtype, tol in ((torch.bfloat16, 2 ** -7), (torch.float16, 2 ** -10), (torch.float32, 1e-5)):
            for (B, H, D) in ((256, 32, 128), (5, 8, 64), (1, 16, 128), (33, 8, 256)):
                for base_e in (True, False):
                    seed += 1
                    outs = [(torch.rand(B, H, D, generator=torch.Generator().manual_seed(seed * 16 + r)) * 2 - 1).to(dtype)
                            for r in range(world)]
                    lses = [torch.randn(B, H, generator=torch.Generator().manual_seed(seed * 16 + 8 + r)) * 4
                            for r in range(world)]
                    lses[0][0, 0] = float("-inf")
                    lses[world - 1][B - 1, H - 1] = float("nan")
                    if B > 1:
                        for r in range(world):
                            lses[r][1, 2] = float("-inf")      # nobody holds a token of this (row, head)
                    exp_out, exp_lse = ref_ops.dcp_lse_merge(outs, lses, dtype=dtype, is_lse_base_on_e=base_e)
                    for in_heap in (False, True):
                        start = ar._heap_used
                        o, l = outs[rank].to(dev), lses[rank].to(dev)
                        if in_heap:
                            ob, lb = ar.alloc((B, H, D), dtype), ar.alloc((B, H), torch.float32)
                            ob.copy_(o); lb.copy_(l)
                            o, l = ob, lb
                        torch.cuda.synchronize(); dist.barrier()
                        full, lse_full = ar.dcp_lse_merge(o, l, reduce_scatter=False, return_lse=True,
                                                          is_lse_base_on_e=base_e)
                        part, lse_part = ar.dcp_lse_merge(o, l, reduce_scatter=True, return_lse=True,
                                                          is_lse_base_on_e=base_e)
                        torch.cuda.synchronize(); dist.barrier()
                        hs = H // world
                        sl = slice(rank * hs, (rank + 1) * hs)
                        assert torch.equal(part, full[:, sl]) and torch.equal(lse_part, lse_full[:, sl])
                        torch.testing.assert_close(full.float().cpu(), exp_out.float(), rtol=2 * tol, atol=2 * tol)
                        el = torch.from_numpy(exp_lse)
                        fin = torch.isfinite(el)
                        assert torch.equal(torch.isfinite(lse_full.cpu()), fin)
                        torch.testing.assert_close(lse_full.cpu()[fin], el[fin], rtol=1e-5, atol=1e-5)
                        # all-reduce form: bit-identical on every rank
                        mine = full.contiguous().view(torch.uint8).view(-1).to(torch.int32).sum().reshape(1)
                        both = [torch.empty_like(mine) for _ in range(world)]
                        dist.all_gather(both, mine)
                        assert all(int(b) == int(mine) for b in both)
                        if in_heap:
                            torch.cuda.synchronize()
                            ar._heap_used = start
        log.append("dcp_lse_merge (rs / ar) == oracle within output rounding, identical on all ranks")
        # ---- 4. latency vs NCCL at the decode message size (2 MiB bf16)
        numel = 256 * 4096
        xz = ar.alloc(numel, torch.bfloat16)
        xz.copy_(_rank_input(rank, numel, torch.bfloat16, 7).to(dev))
        o = torch.empty_like(xz)
        xn = xz.clone()

        def timeit(fn, iters=20, reps=10):
            # a CUDA graph of `iters` back-to-back calls: device time without host launch overhead
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
            gr.replay()
            torch.cuda.synchronize(); dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                gr.replay()
            e1.record(); torch.cuda.synchronize()
            return e0.elapsed_time(e1) / (iters * reps) * 1e3

        t1 = timeit(lambda: ar.all_reduce(xz, out=o, algo=1))
        t2 = timeit(lambda: ar.all_reduce(xz, out=o, algo=2))
        tn = timeit(lambda: dist.all_reduce(xn))
        x8 = ar.alloc((256, 4096), torch.float8_e4m3fn)
        wn = torch.ones(4096, dtype=torch.bfloat16, device=dev)
        shard = torch.zeros((256 + world - 1) // world, 4096, dtype=torch.bfloat16, device=dev)
        sc = torch.tensor([0.02], device=dev)
        tf = timeit(lambda: ar.fused_add_rms_norm(x8, xz.view(256, 4096), shard, wn, 1e-5, sc))
        resf = torch.zeros(256, 4096, dtype=torch.bfloat16, device=dev)
        x8b = torch.empty(256, 4096, dtype=torch.float8_e4m3fn, device=dev)

        def unfused():
            dist.all_reduce(xn)
            ops.fused_add_rms_norm_static_fp8_quant(x8b, xn.view(256, 4096), resf, wn, sc, 1e-5)

        tu = timeit(unfused)
        log.append(f"2 MiB bf16 all-reduce, world {world}: one-shot {t1:.1f} us, two-shot {t2:.1f} us, NCCL {tn:.1f} us; "
                   f"fused AR+add+rmsnorm+fp8 {tf:.1f} us vs NCCL + norm kernel {tu:.1f} us")
        ar.close()
        dist.destroy_process_group()
        q.put((rank, "ok", log))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, "fail", [repr(e), traceback.format_exc()]))


@pytest.mark.parametrize("multicast", [False, None], ids=["peer_pull", "nvls_if_available"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_nvlink_allreduce(world, multicast):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q, multicast)) for r in range(world)]
    for p in procs:
        p.start()
    results = []
    try:
        for _ in range(world):
            results.append(q.get(timeout=600))
    finally:
        for p in procs:
            p.join(timeout=30)
            if p.is_alive():
                p.kill()
    for rank, status, log in sorted(results):
        print(f"[rank {rank}] {status}: " + " | ".join(log))
    assert all(s == "ok" for _, s, _ in results), results
