"""CPU, world_size 2 over gloo: the host-side tensor-parallel logic of the decode path.
(1) column / row / qkv sharding (b200ops/tp.py) + one all-reduce reproduces the unsharded layer
    (oracle arithmetic in float64 so the comparison is about the partitioning, not rounding);
(2) the sequence-parallel formulation of the fused epilogue - reduce-scatter rows, add residual shard,
    RMSNorm + FP8 quant on the owner, all-gather - equals all-reduce followed by
    fused_add_rms_norm_static_fp8_quant on every rank, BIT-exactly (oracle/ref_ops.py arithmetic),
    including a token count that does not divide by the world size;
(3) B200AllReduce refuses to come up without CUDA (no CPU fallback);
(4) expert parallelism: per-rank fused_experts through expert_map + one all-reduce equals the full MoE block."""
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


def _worker(rank, world, port, q):
    try:
        for p in (ROOT, os.path.join(ROOT, "vllm-fork_b200")):
            if p not in sys.path:
                sys.path.insert(0, p)
        import torch.distributed as dist
        from b200ops import tp
        from oracle import ref_ops
        dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
        g = torch.Generator().manual_seed(0)  # same full tensors on every rank
        M, H, Hq, Hkv, D, I = 5, 64, 8, 4, 16, 96
        x = torch.randn(M, H, generator=g, dtype=torch.float64)
        wq, wk, wv = (torch.randn(n * D, H, generator=g, dtype=torch.float64) for n in (Hq, Hkv, Hkv))
        wo = torch.randn(H, Hq * D, generator=g, dtype=torch.float64)
        wg, wu = torch.randn(I, H, generator=g, dtype=torch.float64), torch.randn(I, H, generator=g, dtype=torch.float64)
        wd = torch.randn(H, I, generator=g, dtype=torch.float64)
        # ---- (1) attention block: qkv column-parallel by head, "attention" = per-head op, o_proj row-parallel
        sq, sk, sv = tp.qkv_shard(Hq, Hkv, D, rank, world)
        ql, kl, vl = x @ wq[sq].T, x @ wk[sk].T, x @ wv[sv].T
        grp = Hq // Hkv
        hq_l = Hq // world
        # stand-in for attention that needs q head h and kv head h // grp together: elementwise q * (k + v)
        kv_l = (kl + vl).view(M, Hkv // world, D).repeat_interleave(grp, dim=1)
        a_l = (ql.view(M, hq_l, D) * kv_l).reshape(M, hq_l * D)
        part = a_l @ wo[:, tp.row_shard(Hq * D, rank, world)].T
        dist.all_reduce(part)
        kv_f = ((x @ wk.T) + (x @ wv.T)).view(M, Hkv, D).repeat_interleave(grp, dim=1)
        full = ((x @ wq.T).view(M, Hq, D) * kv_f).reshape(M, Hq * D) @ wo.T
        np.testing.assert_allclose(part.numpy(), full.numpy(), rtol=1e-12, atol=1e-9)
        # MLP: gate_up column-parallel, down row-parallel
        cs = tp.column_shard(I, rank, world)
        h_l = torch.nn.functional.silu(x @ wg[cs].T) * (x @ wu[cs].T)
        part = h_l @ wd[:, tp.row_shard(I, rank, world)].T
        dist.all_reduce(part)
        full = (torch.nn.functional.silu(x @ wg.T) * (x @ wu.T)) @ wd.T
        np.testing.assert_allclose(part.numpy(), full.numpy(), rtol=1e-12, atol=1e-9)
        # ---- (2) sequence-parallel fused epilogue == all-reduce + fused_add_rms_norm_static_fp8_quant
        for M2 in (6, 5, 1):
            gi = torch.Generator().manual_seed(100 + rank)
            mine = torch.randn(M2, H, generator=gi).to(torch.bfloat16)      # this rank's GEMM partial
            g2 = torch.Generator().manual_seed(7)
            residual = torch.randn(M2, H, generator=g2).to(torch.bfloat16)
            w = (torch.rand(H, generator=g2) + 0.5).to(torch.bfloat16)
            scale = torch.tensor([0.03])
            parts = [torch.empty_like(mine) for _ in range(world)]
            dist.all_gather(parts, mine)
            acc = torch.zeros(M2, H)
            for p_ in parts:                                               # fp32, rank order, one rounding
                acc = acc + p_.float()
            summed = acc.to(torch.bfloat16)
            ref_bits, ref_res = ref_ops.rms_norm_static_fp8_quant(summed, w, scale, 1e-5, residual=residual.clone())
            lo, hi = tp.rows_of_rank(M2, rank, world)
            own_bits, own_res = (ref_ops.rms_norm_static_fp8_quant(summed[lo:hi], w, scale, 1e-5,
                                                                   residual=residual[lo:hi].clone())
                                 if hi > lo else (np.zeros((0, H), np.uint8), residual[lo:hi]))
            per = (M2 + world - 1) // world
            pad = torch.zeros(per, H, dtype=torch.uint8)
            pad[: hi - lo] = torch.from_numpy(np.asarray(own_bits).reshape(hi - lo, H))
            gathered = [torch.empty_like(pad) for _ in range(world)]
            dist.all_gather(gathered, pad)
            out = torch.cat(gathered)[:M2]
            np.testing.assert_array_equal(out.numpy(), np.asarray(ref_bits).reshape(M2, H))
            np.testing.assert_array_equal(torch.as_tensor(own_res).float().numpy(),
                                          torch.as_tensor(ref_res).float().numpy()[lo:hi])
        los = [tp.rows_of_rank(5, r, 2) for r in range(2)]
        assert los == [(0, 3), (3, 5)] and tp.rows_of_rank(1, 1, 2) == (1, 1)
        # ---- (4) expert parallel (BASELINE configs[3]): every rank holds all tokens, runs fused_experts on its
        # own experts through expert_map (non-local slots contribute zero) and one all-reduce restores the full
        # MoE output (fused_moe.py:1821-1822) - oracle arithmetic, bit-identical FP8 inputs on both ranks
        E, topk, Hm, Im, Mm = 4, 2, 64, 32, 7
        xm = (torch.randn(Mm, Hm, generator=g) * 0.5).to(torch.bfloat16)
        w1 = torch.randint(0, 0x60, (E, 2 * Im, Hm), generator=g, dtype=torch.uint8)
        w2 = torch.randint(0, 0x60, (E, Hm, Im), generator=g, dtype=torch.uint8)
        w1s, w2s = torch.rand(E, generator=g) * 0.01 + 0.005, torch.rand(E, generator=g) * 0.01 + 0.005
        a1, a2 = torch.tensor([0.02]), torch.tensor([0.05])
        tw, ids, _ = ref_ops.topk_softmax(torch.randn(Mm, E, generator=g), topk, True)
        lo, hi = tp.expert_shard(E, rank, world)
        em = tp.expert_map(E, rank, world)
        assert em.tolist() == [i - lo if lo <= i < hi else -1 for i in range(E)]
        part = ref_ops.fused_experts_fp8(xm, w1[lo:hi], w2[lo:hi], tw, ids, w1s[lo:hi], w2s[lo:hi], a1, a2,
                                         expert_map=em).double()
        dist.all_reduce(part)
        full = ref_ops.fused_experts_fp8(xm, w1, w2, tw, ids, w1s, w2s, a1, a2).double()
        np.testing.assert_allclose(part.numpy(), full.numpy(), rtol=1e-6, atol=1e-6)
        assert tp.expert_map(E, 0, 1) is None
        # ---- (3) no CPU fallback for the NVLink collective
        from b200ops.all_reduce import B200AllReduce
        try:
            B200AllReduce(None, "cuda:0")
            raise AssertionError("B200AllReduce came up without CUDA")
        except RuntimeError as e:
            assert "no CPU fallback" in str(e) or "CUDA" in str(e), e
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, "ok", ""))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, "fail", repr(e) + "\n" + traceback.format_exc()))


def test_tp_sharding_and_fused_epilogue_world2():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in range(2)]
    for p in procs:
        p.join(timeout=30)
    assert all(s == "ok" for _, s, _ in res), res


def test_shard_validation():
    sys.path.insert(0, os.path.join(ROOT, "vllm-fork_b200"))
    from b200ops import tp
    with pytest.raises(ValueError):
        tp.column_shard(10, 0, 4)
    with pytest.raises(ValueError):
        tp.qkv_shard(32, 2, 128, 0, 4)
    assert tp.qkv_shard(32, 8, 128, 3, 8) == (slice(1536, 2048), slice(384, 512), slice(384, 512))
