"""Eager launches of one kernel at one shape, for `ncu -k regex:<kernel> -s 2 -c 1 python tools/ncu_targets.py ...`.
    python tools/ncu_targets.py fp8 M N K [impl]      # cutlass_scaled_mm (impl v2 | v3 | v4 | default dispatch)
    python tools/ncu_targets.py nvfp4 M N K           # cutlass_scaled_fp4_mm
    python tools/ncu_targets.py moe M E topk H I      # fused_experts (FP8 per-tensor)
    python tools/ncu_targets.py attn S L Hq Hkv       # paged_attention_decode, FP8 KV, D 128
    python tools/ncu_targets.py groupquant M K        # per_token_group_fp8_quant g=128
No timing is printed: numbers taken under a profiler are never bench values."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vllm-fork_b200"))
from b200ops import fused_moe, ops  # noqa: E402

DEV = torch.device("cuda:0")


def rand_fp8(*shape):
    return torch.randint(0, 0x70, shape, dtype=torch.uint8, device=DEV).view(torch.float8_e4m3fn)


def main():
    what, a = sys.argv[1], sys.argv[2:]
    reps = 4
    if what == "fp8":
        M, N, K = (int(v) for v in a[:3])
        impl = a[3] if len(a) > 3 and a[3] != "default" else None
        ws = [rand_fp8(N, K) for _ in range(reps)]
        x, one = rand_fp8(M, K), torch.ones(1, device=DEV)
        out = torch.empty(M, N, dtype=torch.bfloat16, device=DEV)
        for w in ws:
            ops.cutlass_scaled_mm(x, w.t(), one, one, torch.bfloat16, out=out, impl=impl)
    elif what == "nvfp4":
        M, N, K = (int(v) for v in a[:3])
        gs = torch.ones(1, device=DEV)
        x4, xsf = ops.scaled_fp4_quant(torch.randn(M, K, device=DEV, dtype=torch.bfloat16), gs)
        ws = [ops.scaled_fp4_quant(torch.randn(N, K, device=DEV, dtype=torch.bfloat16), gs) for _ in range(reps)]
        out = torch.empty(M, N, dtype=torch.bfloat16, device=DEV)
        for w4, wsf in ws:
            ops.cutlass_scaled_fp4_mm(x4, w4, xsf, wsf, gs, torch.bfloat16, out=out)
    elif what == "moe":
        M, E, k, H, I = (int(v) for v in a[:5])
        w1, w2 = rand_fp8(E, 2 * I, H), rand_fp8(E, H, I)
        s = torch.full((E,), 0.01, device=DEV)
        x = torch.randn(M, H, device=DEV, dtype=torch.bfloat16)
        tw, ti = torch.topk(torch.softmax(torch.randn(M, E, device=DEV), -1), k)
        a1 = torch.tensor([0.02], device=DEV)
        for _ in range(reps):
            fused_moe.fused_experts(x, w1, w2, tw, ti.to(torch.int32), use_fp8_w8a8=True, w1_scale=s, w2_scale=s,
                                    a1_scale=a1, a2_scale=a1)
    elif what == "attn":
        S, L, Hq, Hkv = (int(v) for v in a[:4])
        D, bs = 128, 16
        nblk = (L + bs - 1) // bs
        nb = S * nblk
        one = torch.tensor([0.05], device=DEV)
        bt = torch.randperm(nb, device=DEV).to(torch.int32).view(S, nblk)
        lens = torch.full((S,), L, dtype=torch.int32, device=DEV)
        q = torch.randn(S, Hq, D, device=DEV, dtype=torch.bfloat16)
        out = torch.empty(S, Hq, D, device=DEV, dtype=torch.bfloat16)
        ws = ops.paged_attention_workspace(S, Hq, Hkv, D, DEV)
        kvs = [torch.randint(0, 0x70, (nb, bs, Hkv, 2 * D), dtype=torch.uint8, device=DEV) for _ in range(2)]
        for i in range(reps):
            kv = kvs[i % 2]
            ops.paged_attention_decode(out, q, kv[..., :D], kv[..., D:], bt, lens, D ** -0.5, k_scale=one, v_scale=one,
                                       max_seq_len=L, workspace=ws)
    elif what == "groupquant":
        M, K = (int(v) for v in a[:2])
        fi = torch.finfo(torch.float8_e4m3fn)
        o8 = torch.empty(M, K, dtype=torch.float8_e4m3fn, device=DEV)
        s = torch.empty(M, K // 128, device=DEV)
        for _ in range(reps):
            x = torch.randn(M, K, device=DEV, dtype=torch.bfloat16)
            ops.per_token_group_fp8_quant(x, o8, s, 128, 1e-10, fi.min, fi.max, False)
    else:
        raise SystemExit(__doc__)
    torch.cuda.synchronize()


if __name__ == "__main__":
    main()
