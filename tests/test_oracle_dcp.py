"""CPU: the decode-context-parallel merge (cp_lse_ag_out_rs / _ar, vllm/v1/attention/ops/common.py:13-258).

(1) Pins the restatement by the domain property: paged attention over a context whose tokens are dealt out to N
    "ranks" (interleaved by token, the DCP layout, or in contiguous shards), each shard attended separately with
    the reference-pinned attention oracle, then merged by LSE == attention over the whole context; including
    ranks that hold NO token of a sequence (LSE = -inf) and log2-based LSEs.
(2) world_size-2 gloo: all_gather(lse) -> correct_attn_out -> all_reduce / reduce-scatter across processes equals
    the single-host restatement (the same arithmetic the NVLink kernel implements), bit-exactly."""
import os
import socket
import sys

import numpy as np
import pytest
import torch

from oracle import ref_ops

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _shard_attention(q, k, v, scale, N, interleave):
    """full attention and per-rank partial (out, lse) with keys dealt to N ranks. q [B,H,D], k/v [B,L,H,D]."""
    B, H, D = q.shape
    L = k.shape[1]
    s = np.einsum("bhd,blhd->bhl", q, k) * scale
    m = s.max(-1, keepdims=True)
    p = np.exp(s - m)
    full = np.einsum("bhl,blhd->bhd", p / p.sum(-1, keepdims=True), v)
    outs, lses = [], []
    for r in range(N):
        idx = np.arange(L)[(np.arange(L) % N == r) if interleave else (np.arange(L) * N // L == r)]
        if len(idx) == 0:
            outs.append(np.zeros((B, H, D))), lses.append(np.full((B, H), -np.inf))
            continue
        sr = s[:, :, idx]
        mr = sr.max(-1, keepdims=True)
        pr = np.exp(sr - mr)
        outs.append(np.einsum("bhl,blhd->bhd", pr / pr.sum(-1, keepdims=True), v[:, idx]))
        lses.append((np.log(pr.sum(-1)) + mr[..., 0]))
    return full, outs, lses


@pytest.mark.parametrize("N", [2, 4, 8])
@pytest.mark.parametrize("interleave", [True, False])
@pytest.mark.parametrize("base_e", [True, False])
def test_merge_of_context_shards_equals_full_attention(N, interleave, base_e):
    rng = np.random.default_rng(N)
    B, H, D, L = 3, 8, 16, 5 if N == 8 else 37       # L < N: some ranks hold nothing
    q, k, v = rng.standard_normal((B, H, D)), rng.standard_normal((B, L, H, D)), rng.standard_normal((B, L, H, D))
    full, outs, lses = _shard_attention(q, k, v, D ** -0.5, N, interleave)
    if not base_e:
        lses = [l / np.log(2.0) for l in lses]
    merged, lse = ref_ops.dcp_lse_merge([torch.from_numpy(o).float() for o in outs],
                                        [torch.from_numpy(l).float() for l in lses], dtype=torch.float32,
                                        is_lse_base_on_e=base_e)
    np.testing.assert_allclose(merged.numpy(), full, rtol=2e-5, atol=2e-6)
    s = np.einsum("bhd,blhd->bhl", q, k) * D ** -0.5
    full_lse = np.log(np.exp(s - s.max(-1, keepdims=True)).sum(-1)) + s.max(-1)
    np.testing.assert_allclose(lse, full_lse if base_e else full_lse / np.log(2.0), rtol=2e-5, atol=2e-6)


def test_sanitised_lses_and_empty_everywhere():
    """NaN / +inf partial LSEs count as -inf (kernel :54, :83-87); a row with no tokens anywhere gives out 0."""
    B, H, D = 2, 2, 8
    outs = [torch.ones(B, H, D), 2 * torch.ones(B, H, D)]
    l0 = torch.tensor([[0.0, float("nan")], [float("-inf"), float("inf")]])
    l1 = torch.tensor([[0.0, 1.0], [float("-inf"), 2.0]])
    merged, lse = ref_ops.dcp_lse_merge(outs, [l0, l1], dtype=torch.float32)
    np.testing.assert_allclose(merged[0, 0].numpy(), 1.5)           # equal weights
    np.testing.assert_allclose(merged[0, 1].numpy(), 2.0)           # NaN partial dropped
    np.testing.assert_allclose(merged[1, 0].numpy(), 0.0)           # nobody has tokens
    np.testing.assert_allclose(merged[1, 1].numpy(), 2.0)           # +inf partial dropped
    assert lse[1, 0] == -np.inf and lse[0, 1] == np.float32(1.0)


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
        from oracle import ref_ops as R
        dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
        B, H, D = 5, 8, 32
        outs = [(torch.rand(B, H, D, generator=torch.Generator().manual_seed(10 + r)) * 2 - 1).to(torch.bfloat16)
                for r in range(world)]
        lses = [torch.randn(B, H, generator=torch.Generator().manual_seed(20 + r)) * 3 for r in range(world)]
        lses[1][0, 0] = float("-inf")
        # the reference's data flow on this rank: all_gather(lse) -> correct_attn_out -> collective on out
        gathered = [torch.empty(B, H) for _ in range(world)]
        dist.all_gather(gathered, lses[rank].contiguous())
        corr, lse = R.dcp_correct_attn_out(outs[rank], torch.stack(gathered), rank)
        corr = torch.from_numpy(corr).to(torch.bfloat16)                    # stored in the output dtype
        parts = [torch.empty_like(corr) for _ in range(world)]
        dist.all_gather(parts, corr)
        acc = torch.zeros(B, H, D)
        for p in parts:                                                      # rank order, fp32, one rounding
            acc = acc + p.float()
        ar = acc.to(torch.bfloat16)
        hs = H // world
        rs = ar[:, rank * hs:(rank + 1) * hs]                                # reduce_scatter(dim=1) keeps my heads
        exp_out, exp_lse = R.dcp_lse_merge(outs, lses, dtype=torch.bfloat16)
        assert torch.equal(ar, exp_out) and np.array_equal(lse, exp_lse)
        assert torch.equal(rs, exp_out[:, rank * hs:(rank + 1) * hs])
        dist.destroy_process_group()
        q.put((rank, "ok", ""))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, "fail", repr(e) + traceback.format_exc()))


def test_dcp_merge_world2_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in range(2)]
    for p in procs:
        p.join(timeout=30)
    assert all(s == "ok" for _, s, _ in res), res


# ---- DCP KV interleave: which rank stores which token (block_table.py:413-441) -------------------------------
def test_dcp_slot_mapping_hand_worked_case():
    """world 2, interleave 1, block_size 4: a virtual block is 8 positions; rank 0 keeps the even ones, packed."""
    bt = np.array([[7, 3]], dtype=np.int32)
    pos = np.arange(12)
    req = np.zeros(12, dtype=np.int64)
    r0 = ref_ops.compute_slot_mapping_dcp(bt, req, pos, 4, 2, 0, 1)
    r1 = ref_ops.compute_slot_mapping_dcp(bt, req, pos, 4, 2, 1, 1)
    assert r0.tolist() == [28, -1, 29, -1, 30, -1, 31, -1, 12, -1, 13, -1]
    assert r1.tolist() == [-1, 28, -1, 29, -1, 30, -1, 31, -1, 12, -1, 13]
    # interleave 2: runs of two tokens alternate between the ranks
    r0 = ref_ops.compute_slot_mapping_dcp(bt, req, pos, 4, 2, 0, 2)
    assert r0.tolist() == [28, 29, -1, -1, 30, 31, -1, -1, 12, 13, -1, -1]
    # a single rank is the plain mapping
    np.testing.assert_array_equal(ref_ops.compute_slot_mapping_dcp(bt, req[:8], pos[:8], 4, 1, 0, 1),
                                  ref_ops.compute_slot_mapping(bt, req[:8], pos[:8], 4))


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("interleave", [1, 4, 16])
@pytest.mark.parametrize("bs", [16, 32])
def test_dcp_slot_mapping_partitions_every_sequence(world, interleave, bs):
    """Every position is stored by exactly one rank; a rank's slots are distinct, fill its blocks from the
    front, and their count is what get_dcp_local_seq_lens says that rank must attend over."""
    if bs % interleave:
        pytest.skip("the reference requires block_size % interleave == 0")
    rng = np.random.default_rng(world * 100 + interleave + bs)
    lens = np.array([1, interleave * world - 1, 777, 2048, 5], dtype=np.int32)
    nblk = int((lens.max() + bs * world - 1) // (bs * world))
    bt = rng.permutation(len(lens) * nblk).astype(np.int32).reshape(len(lens), nblk)
    for s, L in enumerate(lens):
        pos = np.arange(L)
        req = np.full(L, s)
        per_rank = [ref_ops.compute_slot_mapping_dcp(bt, req, pos, bs, world, r, interleave) for r in range(world)]
        owners = np.stack([m >= 0 for m in per_rank]).sum(0)
        assert (owners == 1).all()
        for r, m in enumerate(per_rank):
            loc = m[m >= 0]
            assert len(np.unique(loc)) == len(loc)
            assert len(loc) == int(ref_ops.dcp_local_seq_lens(lens[s:s + 1], world, r, interleave)[0])
            # local token j (in position order) sits at local offset j: block_table[s, j // bs] * bs + j % bs
            j = np.arange(len(loc))
            np.testing.assert_array_equal(loc, bt[s, j // bs].astype(np.int64) * bs + j % bs)


def test_dcp_sharded_cache_attention_merges_to_the_unsharded_result():
    """Oracle-level end to end (flash_attn.py:859-872,1177): KV written through each rank's DCP slot mapping,
    decode attention over each rank's local cache with its local lengths, LSE merge == attention over one cache."""
    g = torch.Generator().manual_seed(5)
    world, interleave, bs, Hq, Hkv, D = 4, 2, 16, 8, 2, 64
    lens = [3, 129, 700]
    S, nblk = len(lens), 44   # the unsharded cache needs ceil(700 / 16) blocks per row
    key = [(torch.rand(L, Hkv, D, generator=g) * 2 - 1).to(torch.bfloat16) for L in lens]
    val = [(torch.rand(L, Hkv, D, generator=g) * 2 - 1).to(torch.bfloat16) for L in lens]
    q = (torch.rand(S, Hq, D, generator=g) * 2 - 1).to(torch.bfloat16)
    bt = torch.randperm(S * nblk, generator=g).to(torch.int32).view(S, nblk)
    one = torch.ones(1)
    req = np.concatenate([np.full(L, s) for s, L in enumerate(lens)])
    pos = np.concatenate([np.arange(L) for L in lens])

    def attend(slots, local_lens):
        kc = torch.zeros(S * nblk, bs, Hkv, D, dtype=torch.bfloat16)
        vc = torch.zeros_like(kc)
        ref_ops.reshape_and_cache_flash(torch.cat(key), torch.cat(val), kc, vc, torch.from_numpy(slots), "auto", one, one)
        return ref_ops.paged_attention_decode(q, kc, vc, bt, torch.tensor(local_lens, dtype=torch.int32), D ** -0.5,
                                              return_lse=True)

    full, _ = attend(ref_ops.compute_slot_mapping(bt.numpy(), req, pos, bs), lens)
    outs, lses = [], []
    for r in range(world):
        slots = ref_ops.compute_slot_mapping_dcp(bt.numpy(), req, pos, bs, world, r, interleave)
        o, l = attend(slots, ref_ops.dcp_local_seq_lens(np.array(lens), world, r, interleave).tolist())
        outs.append(torch.from_numpy(o).float()), lses.append(torch.from_numpy(l.astype(np.float32)))
    merged, _ = ref_ops.dcp_lse_merge(outs, lses, dtype=torch.float32)
    np.testing.assert_allclose(merged.numpy(), full, atol=2e-5, rtol=1e-5)
