"""CPU, world_size 2 over gloo: the second expert-parallel form (all-gather tokens -> local experts -> reduce-scatter,
prepare_finalize/naive_dp_ep.py:112-210) equals the un-partitioned MoE block for every rank's own tokens, with equal and
with uneven token counts per rank, with gathered per-token scales, and with router weights applied on the input."""
import os
import socket
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _dense_moe(x, w1, w2, tw, ti, emap=None, weighted=True):
    """float64 MoE on the experts this rank holds (emap: global id -> local id or -1)."""
    out = torch.zeros(x.shape[0], w2.shape[1], dtype=torch.float64)
    I = w1.shape[1] // 2
    for t in range(x.shape[0]):
        for j in range(ti.shape[1]):
            e = int(ti[t, j])
            if emap is not None:
                e = int(emap[e])
            if e < 0:
                continue
            h = w1[e] @ x[t]
            h = torch.nn.functional.silu(h[:I]) * h[I:]
            out[t] += (tw[t, j] if weighted else 1.0) * (w2[e] @ h)
    return out


def _worker(rank, world, port, q):
    try:
        for p in (ROOT, os.path.join(ROOT, "vllm-fork_b200")):
            if p not in sys.path:
                sys.path.insert(0, p)
        import torch.distributed as dist
        from b200ops import ep, tp
        dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
        grp = dist.group.WORLD
        g = torch.Generator().manual_seed(0)
        E, H, I, topk = 8, 32, 24, 2
        w1 = torch.randn(E, 2 * I, H, generator=g, dtype=torch.float64)
        w2 = torch.randn(E, H, I, generator=g, dtype=torch.float64)
        lo, hi = tp.expert_shard(E, rank, world)
        emap = tp.expert_map(E, rank, world)
        for sizes in ([6, 6], [3, 8]):
            M = sum(sizes)
            x = torch.randn(M, H, generator=g, dtype=torch.float64)
            tw = torch.rand(M, topk, generator=g, dtype=torch.float64)
            ti = torch.stack([torch.randperm(E, generator=g)[:topk] for _ in range(M)]).to(torch.int32)
            full = _dense_moe(x, w1, w2, tw, ti)
            r0 = sum(sizes[:rank])
            sl = slice(r0, r0 + sizes[rank])
            seen = {}

            def experts_fn(a, s, w, i):
                seen["rows"] = a.shape[0]
                seen["scale"] = s
                if s is not None and s.numel() > 1:
                    a = a * s                      # "dequantise" with the gathered per-token scales
                return _dense_moe(a, w1[lo:hi], w2[lo:hi], w, i, emap)

            # (a) no quantisation
            out = ep.dp_ep_fused_experts(x[sl], tw[sl], ti[sl], grp, experts_fn, sizes=sizes)
            assert seen["rows"] == M and seen["scale"] is None
            np.testing.assert_allclose(out.numpy(), full[sl].numpy(), rtol=1e-11, atol=1e-9)
            # (b) per-token "quantisation": the scales travel with the tokens
            def quant(a):
                s = a.abs().amax(dim=1, keepdim=True)
                return a / s, s
            out = ep.dp_ep_fused_experts(x[sl], tw[sl], ti[sl], grp, experts_fn, sizes=sizes, quantize=quant)
            assert tuple(seen["scale"].shape) == (M, 1)
            np.testing.assert_allclose(out.numpy(), full[sl].numpy(), rtol=1e-11, atol=1e-9)
            # (c) a static (1-element) scale is replicated, not gathered
            out = ep.dp_ep_fused_experts(x[sl], tw[sl], ti[sl], grp, experts_fn, sizes=sizes,
                                         quantize=lambda a: (a, torch.ones(1, dtype=torch.float64)))
            assert seen["scale"].numel() == 1
            np.testing.assert_allclose(out.numpy(), full[sl].numpy(), rtol=1e-11, atol=1e-9)
        # (d) router weight on the input, topk = 1
        M = 8
        x = torch.randn(M, H, generator=g, dtype=torch.float64)
        tw = torch.rand(M, 1, generator=g, dtype=torch.float64)
        ti = torch.randint(0, E, (M, 1), generator=g).to(torch.int32)
        full = _dense_moe(x * tw, w1, w2, tw, ti, weighted=False)
        sl = slice(rank * 4, rank * 4 + 4)
        out = ep.dp_ep_fused_experts(x[sl], tw[sl], ti[sl], grp,
                                     lambda a, s, w, i: _dense_moe(a, w1[lo:hi], w2[lo:hi], w, i, emap, weighted=False),
                                     apply_router_weight_on_input=True)
        np.testing.assert_allclose(out.numpy(), full[sl].numpy(), rtol=1e-11, atol=1e-9)
        # (e) validation
        try:
            ep.all_gatherv(x[sl], grp, sizes=[1, 2])
            raise AssertionError("sizes mismatch accepted")
        except ValueError:
            pass
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, "".join(traceback.format_exception(e))))


@pytest.mark.timeout(300)
def test_dp_ep_dispatch_combine_world2():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in res:
        assert msg == "ok", f"rank {rank}: {msg}"
