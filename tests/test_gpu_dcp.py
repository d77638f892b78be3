"""GPU parity: decode context parallelism end to end on ONE device (the ranks are simulated in turn).

The reference shards the KV of every sequence across the DCP ranks by token (block_table.py:413-441), lets each
rank attend over its own cache with LSE output and merges the partials by LSE (flash_attn.py:859-872, 1177;
ops/common.py:13-258). Here every "rank" writes its cache through reshape_and_cache_flash with the oracle's DCP
slot mapping (other ranks' tokens are negative slots, which the kernel must skip), runs paged_attention_decode
with that rank's local lengths (some are 0: output zeros, LSE -inf), and the partials are merged pairwise with
merge_attn_states. The result must equal the CUDA attention over the unsharded cache and the fp32 oracle.
Tolerance: north_star's bf16 attention bound, 2e-2 abs (two extra bf16 roundings of the partials)."""
import numpy as np
import pytest
import torch

from oracle import ref_ops

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.parametrize("world,interleave", [(2, 1), (2, 16), (4, 1), (4, 4)])
@pytest.mark.parametrize("kv_fp8", [False, True])
def test_dcp_sharded_decode_and_lse_merge_equals_unsharded(world, interleave, kv_fp8):
    from b200ops import ops
    g = torch.Generator().manual_seed(world * 10 + interleave)
    bs, Hq, Hkv, D = 16, 16, 4, 128
    lens = [1, 3, 129, 700, 2049]
    S = len(lens)
    nblk = (max(lens) + bs - 1) // bs
    key = torch.cat([(torch.rand(L, Hkv, D, generator=g) * 2 - 1) for L in lens]).to(torch.bfloat16)
    val = torch.cat([(torch.rand(L, Hkv, D, generator=g) * 2 - 1) for L in lens]).to(torch.bfloat16)
    q = (torch.rand(S, Hq, D, generator=g) * 2 - 1).to(torch.bfloat16)
    bt = torch.randperm(S * nblk, generator=g).to(torch.int32).view(S, nblk)
    req = np.concatenate([np.full(L, s) for s, L in enumerate(lens)])
    pos = np.concatenate([np.arange(L) for L in lens])
    ks = torch.tensor([0.5 / 64]) if kv_fp8 else torch.ones(1)
    vs = torch.tensor([0.25 / 64]) if kv_fp8 else torch.ones(1)
    cdt = torch.uint8 if kv_fp8 else torch.bfloat16
    scale = D ** -0.5

    def attend(slots, local_lens):
        raw = torch.zeros(S * nblk, bs, Hkv, 2 * D, dtype=cdt, device=DEV)
        kc, vc = raw[..., :D], raw[..., D:]
        ops.reshape_and_cache_flash(key.to(DEV), val.to(DEV), kc, vc, torch.from_numpy(slots).to(DEV),
                                    "fp8" if kv_fp8 else "auto", ks.to(DEV), vs.to(DEV))
        out = torch.full((S, Hq, D), float("nan"), dtype=torch.bfloat16, device=DEV)
        lse = torch.full((S, Hq), float("nan"), dtype=torch.float32, device=DEV)
        ops.paged_attention_decode(out, q.to(DEV), kc, vc, bt.to(DEV),
                                   torch.tensor(local_lens, dtype=torch.int32, device=DEV), scale,
                                   k_scale=ks.to(DEV) if kv_fp8 else None, v_scale=vs.to(DEV) if kv_fp8 else None,
                                   lse=lse)
        torch.cuda.synchronize()
        return out, lse, (kc.cpu(), vc.cpu())

    full, _, (kc_cpu, vc_cpu) = attend(ref_ops.compute_slot_mapping(bt.numpy(), req, pos, bs), lens)
    want = ref_ops.paged_attention_decode(q, kc_cpu, vc_cpu, bt, torch.tensor(lens, dtype=torch.int32), scale,
                                          k_scale=ks if kv_fp8 else None, v_scale=vs if kv_fp8 else None)

    acc_o, acc_l = None, None
    held = np.zeros(S, dtype=np.int64)
    for r in range(world):
        slots = ref_ops.compute_slot_mapping_dcp(bt.numpy(), req, pos, bs, world, r, interleave)
        local = ref_ops.dcp_local_seq_lens(np.array(lens), world, r, interleave)
        held += local
        o, l, _ = attend(slots, local.tolist())
        if acc_o is None:
            acc_o, acc_l = o, l
            continue
        merged = torch.empty_like(o)
        merged_lse = torch.empty(Hq, S, dtype=torch.float32, device=DEV)
        ops.merge_attn_states(merged, acc_o, acc_l.t().contiguous(), o, l.t().contiguous(), output_lse=merged_lse)
        torch.cuda.synchronize()
        acc_o, acc_l = merged, merged_lse.t().contiguous()
    assert held.tolist() == lens
    got = acc_o.float().cpu().numpy()
    assert not np.isnan(got).any()
    np.testing.assert_allclose(got, full.float().cpu().numpy(), atol=2e-2, rtol=0)
    np.testing.assert_allclose(got, want, atol=2e-2, rtol=0)
