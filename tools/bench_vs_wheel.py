"""Head-to-head timing against the reference's OWN sm_100 kernels (the vllm wheel on the box: a build of upstream
vLLM whose torch.ops._C / _moe_C ops are what SURVEY.md 2c rows K1/K4/K5/K6/K8' say "must beat").

Same tensors for both sides, every call captured in a CUDA graph (no host launch cost in a kernel number), the big
operand rotated over enough copies to exceed the 126 MB L2, CUDA-event timed. The wheel is LIBRARY code used as the
race opponent only; nothing under b200ops imports it.

    python tools/bench_vs_wheel.py [--out profiles/<round>_vs_wheel.md] [--only gemm,fp4,quant,attn,moe]

Prints one JSON line per case and writes a markdown table (ratio = wheel_us / ours_us: > 1 means ours is faster)."""
import argparse
import importlib.util
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vllm-fork_b200"))
from b200ops import fused_moe, ops  # noqa: E402

DEV = torch.device("cuda:0")


def load_wheel():
    spec = importlib.util.find_spec("vllm")
    d = os.path.dirname(spec.origin)
    for lib in ("_C.abi3.so", "_C_stable_libtorch.abi3.so", "_moe_C.abi3.so"):
        p = os.path.join(d, lib)
        if os.path.exists(p):
            torch.ops.load_library(p)
    torch.ops._C.cutlass_scaled_mm  # noqa: B018
    return torch.ops


def graph_time(fn_of_i, ncopy, iters=30):
    """us per call of fn_of_i(i), i cycling over `ncopy` operand copies, inside one CUDA graph."""
    for i in range(min(ncopy, 3)):
        fn_of_i(i)
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        fn_of_i(0)
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(ncopy):
            fn_of_i(i)
    g.replay()
    torch.cuda.synchronize()
    reps = max(2, iters // ncopy)
    best = 1e30
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / (reps * ncopy) * 1e3)
    return best


ROWS = []


def report(group, case, ours_us, wheel_us, note=""):
    r = {"group": group, "case": case, "ours_us": round(ours_us, 2),
         "wheel_us": None if wheel_us is None else round(wheel_us, 2),
         "ratio_wheel_over_ours": None if wheel_us is None else round(wheel_us / ours_us, 3), "note": note}
    ROWS.append(r)
    print(json.dumps(r), flush=True)


def try_time(fn, ncopy):
    try:
        return graph_time(fn, ncopy), ""
    except Exception as e:  # noqa: BLE001
        torch.cuda.synchronize()
        return None, repr(e)[:120]


def rand_fp8(*shape):
    return torch.randint(0, 0x70, shape, dtype=torch.uint8, device=DEV).view(torch.float8_e4m3fn)


# --------------------------------------------------------------------------------------------------- FP8 GEMM
def bench_gemm(W):
    # (N, K): C2 (Llama-8B TP1) qkv / o / gate_up / down; C3 (Llama-70B TP8 and TP2) shards
    shapes = [("8B qkv", 6144, 4096), ("8B o", 4096, 4096), ("8B gate_up", 28672, 4096), ("8B down", 4096, 14336),
              ("70B/tp8 qkv", 1280, 8192), ("70B/tp8 o", 8192, 1024), ("70B/tp8 gate_up", 7168, 8192),
              ("70B/tp8 down", 8192, 3584), ("70B/tp2 gate_up", 28672, 8192), ("70B/tp2 down", 8192, 14336)]
    for name, N, K in shapes:
        for M in (16, 64, 128, 256, 4096):
            if M == 4096 and not name.startswith("8B"):
                continue
            ncopy = max(2, int(300e6 // (N * K)) + 1)
            ws = [rand_fp8(N, K) for _ in range(ncopy)]
            x = rand_fp8(M, K)
            sa, sb = torch.ones(1, device=DEV), torch.ones(1, device=DEV)
            out = torch.empty(M, N, dtype=torch.bfloat16, device=DEV)
            ref = torch.empty(M, N, dtype=torch.bfloat16, device=DEV)
            t_o = graph_time(lambda i: ops.cutlass_scaled_mm(x, ws[i].t(), sa, sb, torch.bfloat16, out=out), ncopy)
            t_w, why = try_time(lambda i: W._C.cutlass_scaled_mm(ref, x, ws[i].t(), sa, sb, None), ncopy)
            byts = M * K + N * K + 2 * M * N
            report("cutlass_scaled_mm", f"{name} M={M} N={N} K={K}", t_o, t_w,
                   why or f"ours ({ops.LAST_GEMM_IMPL}) {byts / t_o / 1e3:.0f} GB/s, {2.0 * M * N * K / t_o / 1e6:.0f} TFLOP/s")
            del ws


# --------------------------------------------------------------------------------------------------- NVFP4
def bench_fp4(W):
    for name, N, K in [("8B qkv", 6144, 4096), ("8B o", 4096, 4096), ("8B gate_up", 28672, 4096),
                       ("8B down", 4096, 14336)]:
        for M in (16, 64, 256, 4096):
            ncopy = max(2, int(300e6 // (N * K // 2)) + 1)
            gs = torch.ones(1, device=DEV)
            wsrc = torch.randn(N, K, device=DEV, dtype=torch.bfloat16)
            w4, wsf = ops.scaled_fp4_quant(wsrc, gs)
            w4s = [w4.clone() for _ in range(ncopy)]
            x4, xsf = ops.scaled_fp4_quant(torch.randn(M, K, device=DEV, dtype=torch.bfloat16), gs)
            alpha = torch.ones(1, device=DEV)
            out = torch.empty(M, N, dtype=torch.bfloat16, device=DEV)
            ref = torch.empty(M, N, dtype=torch.bfloat16, device=DEV)
            t_o = graph_time(lambda i: ops.cutlass_scaled_fp4_mm(x4, w4s[i], xsf, wsf, alpha, torch.bfloat16, out=out),
                             ncopy)
            t_w, why = try_time(lambda i: W._C.cutlass_scaled_fp4_mm(ref, x4, w4s[i], xsf, wsf, alpha), ncopy)
            report("cutlass_scaled_fp4_mm", f"{name} M={M} N={N} K={K}", t_o, t_w,
                   why or f"ours {2.0 * M * N * K / t_o / 1e6:.0f} TFLOP/s")
            del w4s, wsrc


# --------------------------------------------------------------------------------------------------- quantisers
def bench_quant(W):
    for M, K in ((256, 4096), (256, 14336), (4096, 4096), (16384, 7168)):
        ncopy = max(2, int(300e6 // (M * K * 2)) + 1)
        xs = [torch.randn(M, K, device=DEV, dtype=torch.bfloat16) for _ in range(ncopy)]
        gs = torch.ones(1, device=DEV)
        o4 = torch.empty(M, K // 2, dtype=torch.uint8, device=DEV)
        osf = ops.create_fp4_scale_tensor(M, K, DEV, True)
        t_o = graph_time(lambda i: ops.scaled_fp4_quant(xs[i], gs, True, output=o4, output_scale=osf), ncopy)
        try:
            W._C.scaled_fp4_quant.out  # noqa: B018
            t_w, why = try_time(lambda i: W._C.scaled_fp4_quant.out(xs[i], gs, True, output=o4, output_scale=osf), ncopy)
        except Exception:  # older wheels: functional form only (allocates inside the graph pool)
            t_w, why = try_time(lambda i: W._C.scaled_fp4_quant(xs[i], gs, True), ncopy)
        report("scaled_fp4_quant", f"M={M} K={K} bf16", t_o, t_w, why or f"ours {M * K * 2.5625 / t_o / 1e3:.0f} GB/s")
        o8 = torch.empty(M, K, dtype=torch.float8_e4m3fn, device=DEV)
        sc = torch.tensor([0.02], device=DEV)
        t_o = graph_time(lambda i: ops.static_scaled_fp8_quant(o8, xs[i], sc), ncopy)
        t_w, why = try_time(lambda i: W._C.static_scaled_fp8_quant(o8, xs[i], sc), ncopy)
        report("static_scaled_fp8_quant", f"M={M} K={K} bf16", t_o, t_w, why or f"ours {M * K * 3 / t_o / 1e3:.0f} GB/s")
        ts = torch.empty(M, 1, device=DEV)
        t_o = graph_time(lambda i: ops.dynamic_per_token_scaled_fp8_quant(o8, xs[i], ts, None), ncopy)
        t_w, why = try_time(lambda i: W._C.dynamic_per_token_scaled_fp8_quant(o8, xs[i], ts, None), ncopy)
        report("dynamic_per_token_scaled_fp8_quant", f"M={M} K={K} bf16", t_o, t_w, why)
        if K % 128 == 0:
            s = torch.empty(M, K // 128, device=DEV)
            fi = torch.finfo(torch.float8_e4m3fn)
            t_o = graph_time(lambda i: ops.per_token_group_fp8_quant(xs[i], o8, s, 128, 1e-10, fi.min, fi.max, False), ncopy)
            t_w, why = try_time(lambda i: W._C.per_token_group_fp8_quant(xs[i], o8, s, 128, 1e-10, fi.min, fi.max, False,
                                                                         False, False), ncopy)
            report("per_token_group_fp8_quant", f"M={M} K={K} g=128 bf16", t_o, t_w,
                   why or f"ours {M * K * 3 / t_o / 1e3:.0f} GB/s")
        del xs


# --------------------------------------------------------------------------------------------------- attention
def bench_attn(W):
    """paged_attention_v2 (legacy x-packed layout) ours vs wheel on identical caches; and the flash-layout decode
    kernel on the same KV content (what the reference's backend would run through Triton / FlashAttention)."""
    for name, S, L, Hq, Hkv, fp8 in (("C2 8B", 256, 4096, 32, 8, True), ("C2 8B bf16 KV", 128, 4096, 32, 8, False),
                                     ("TP8 shard", 256, 4096, 4, 1, True), ("32k split", 32, 32768, 32, 8, True)):
        D, bs, x = 128, 16, 16 if fp8 else 8
        nblk = L // bs
        nb = S * nblk
        cdt = torch.uint8 if fp8 else torch.bfloat16
        kc = (torch.randint(0, 0x70, (nb, Hkv, D // x, bs, x), dtype=torch.uint8, device=DEV) if fp8 else
              torch.randn(nb, Hkv, D // x, bs, x, device=DEV, dtype=cdt))
        vc = (torch.randint(0, 0x70, (nb, Hkv, D, bs), dtype=torch.uint8, device=DEV) if fp8 else
              torch.randn(nb, Hkv, D, bs, device=DEV, dtype=cdt))
        bt = torch.randperm(nb, device=DEV).to(torch.int32).view(S, nblk)
        lens = torch.full((S,), L, dtype=torch.int32, device=DEV)
        q = torch.randn(S, Hq, D, device=DEV, dtype=torch.bfloat16)
        one = torch.tensor([0.05], device=DEV)
        P = (L + 511) // 512
        es, ml = torch.empty(S, Hq, P, device=DEV), torch.empty(S, Hq, P, device=DEV)
        tmp = torch.empty(S, Hq, P, D, device=DEV, dtype=torch.bfloat16)
        out = torch.empty(S, Hq, D, device=DEV, dtype=torch.bfloat16)
        tail = (Hkv, D ** -0.5, bt, lens, bs, L, None, "fp8" if fp8 else "auto", one, one, 0, 0, 0, 64, 0)
        byts = S * 2 * L * Hkv * D * (1 if fp8 else 2)
        t_o = graph_time(lambda i: ops.paged_attention_v2(out, es, ml, tmp, q, kc, vc, *tail), 1, iters=5)
        t_w, why = try_time(lambda i: W._C.paged_attention_v2(out, es, ml, tmp, q, kc, vc, *tail), 1)
        report("paged_attention_v2 (legacy layout)", f"{name} S={S} L={L} Hq={Hq} Hkv={Hkv}", t_o, t_w,
               why or f"ours {byts / t_o / 1e3:.0f} GB/s, wheel {byts / t_w / 1e3:.0f} GB/s")
        del kc, vc
        kv = (torch.randint(0, 0x70, (nb, bs, Hkv, 2 * D), dtype=torch.uint8, device=DEV) if fp8 else
              torch.randn(nb, bs, Hkv, 2 * D, device=DEV, dtype=cdt))
        ws = ops.paged_attention_workspace(S, Hq, Hkv, D, DEV)
        t_f = graph_time(lambda i: ops.paged_attention_decode(out, q, kv[..., :D], kv[..., D:], bt, lens, D ** -0.5,
                                                              k_scale=one if fp8 else None, v_scale=one if fp8 else None,
                                                              max_seq_len=L, workspace=ws), 1, iters=5)
        report("paged_attention_decode (flash layout) vs wheel paged_attention_v2", f"{name} S={S} L={L} Hq={Hq} Hkv={Hkv}",
               t_f, t_w, f"ours {byts / t_f / 1e3:.0f} GB/s")
        del kv


# --------------------------------------------------------------------------------------------------- MoE
def bench_moe(W):
    """C4 (Mixtral): 256 tokens, top-2 of 8, H 4096, I 14336, FP8 per-tensor. Ours: fused_experts whole;
    wheel: cutlass_moe_mm for the two grouped GEMMs (the heavy part) with its own problem-size tables."""
    M, E, k, H, I = 256, 8, 2, 4096, 14336
    w1 = rand_fp8(E, 2 * I, H)
    w2 = rand_fp8(E, H, I)
    w1s = torch.full((E,), 0.01, device=DEV)
    w2s = torch.full((E,), 0.01, device=DEV)
    x = torch.randn(M, H, device=DEV, dtype=torch.bfloat16)
    gate = torch.randn(M, E, device=DEV)
    tw, ti = torch.topk(torch.softmax(gate, -1), k)
    ti = ti.to(torch.int32)
    a1, a2 = torch.tensor([0.02], device=DEV), torch.tensor([0.05], device=DEV)
    t_o = graph_time(lambda i: fused_moe.fused_experts(x, w1, w2, tw, ti, use_fp8_w8a8=True, w1_scale=w1s, w2_scale=w2s,
                                                       a1_scale=a1, a2_scale=a2), 1, iters=6)
    report("fused_experts (whole block)", f"C4 M={M} E={E} top{k} H={H} I={I}", t_o, None,
           f"ours {(w1.numel() + w2.numel()) / t_o / 1e3:.0f} GB/s of expert weights")
    try:
        eo = torch.empty(E + 1, dtype=torch.int32, device=DEV)
        ps1 = torch.empty(E, 3, dtype=torch.int32, device=DEV)
        ps2 = torch.empty(E, 3, dtype=torch.int32, device=DEV)
        a_map = torch.empty(M * k, dtype=torch.int32, device=DEV)
        c_map = torch.empty(M * k, dtype=torch.int32, device=DEV)
        W._C.get_cutlass_moe_mm_data(ti, eo, ps1, ps2, a_map, c_map, E, 2 * I, H, None, True)
        xq = rand_fp8(M * k, H)
        hq = rand_fp8(M * k, I)
        c1 = torch.empty(M * k, 2 * I, dtype=torch.bfloat16, device=DEV)
        c2 = torch.empty(M * k, H, dtype=torch.bfloat16, device=DEV)
        sa = torch.ones(1, device=DEV)
        ab1 = torch.full((E,), H, dtype=torch.int64, device=DEV)
        c1s = torch.full((E,), 2 * I, dtype=torch.int64, device=DEV)
        ab2 = torch.full((E,), I, dtype=torch.int64, device=DEV)
        c2s = torch.full((E,), H, dtype=torch.int64, device=DEV)
        w1t, w2t = w1.transpose(1, 2), w2.transpose(1, 2)   # [E, K, N] column-major views, as cutlass_moe.py passes

        def wheel_pair(i):
            W._C.cutlass_moe_mm(c1, xq, w1t, sa, w1s.view(E, 1, 1), eo[:-1], ps1, ab1, ab1, c1s, False, False)
            W._C.cutlass_moe_mm(c2, hq, w2t, sa, w2s.view(E, 1, 1), eo[:-1], ps2, ab2, ab2, c2s, False, False)

        t_w, why = try_time(wheel_pair, 1)
        # ours: the two grouped GEMMs alone, for the like-for-like line
        bs = 64
        sorted_ids, expert_ids, npp, inv = fused_moe.moe_align_block_size(ti, bs, E, None, pad_sorted_ids=True,
                                                                            ignore_invalid_experts=True,
                                                                            return_inv_perm=True)
        P = sorted_ids.numel()
        a8 = rand_fp8(P, H)
        h8 = rand_fp8(P, I)
        o1 = torch.empty(P, 2 * I, dtype=torch.bfloat16, device=DEV)
        o2 = torch.empty(P, H, dtype=torch.bfloat16, device=DEV)
        from b200ops import _lib
        from b200ops.ops import _dt, _p, _stream
        lib = _lib.load()

        def ours_pair(i):
            lib.b200_moe_grouped_gemm_fp8(_p(o1), _p(a8), _p(w1), _p(expert_ids), _p(npp), _p(a1), _p(w1s), None, P, E,
                                          2 * I, H, bs, 0, 0, _dt(o1), _stream())
            lib.b200_moe_grouped_gemm_fp8(_p(o2), _p(h8), _p(w2), _p(expert_ids), _p(npp), _p(a2), _p(w2s), None, P, E,
                                          H, I, bs, 0, 0, _dt(o2), _stream())

        t_p = graph_time(ours_pair, 1, iters=6)
        report("grouped GEMM pair vs cutlass_moe_mm x2", f"C4 M={M} E={E} top{k}", t_p, t_w,
               why or f"ours {(w1.numel() + w2.numel()) / t_p / 1e3:.0f} GB/s")
    except Exception as e:  # noqa: BLE001
        report("grouped GEMM pair vs cutlass_moe_mm x2", "C4", t_o, None, "wheel op unavailable: " + repr(e)[:150])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    ap.add_argument("--only", default="gemm,fp4,quant,attn,moe")
    a = ap.parse_args()
    W = load_wheel()
    for name, fn in (("gemm", bench_gemm), ("fp4", bench_fp4), ("quant", bench_quant), ("attn", bench_attn),
                     ("moe", bench_moe)):
        if name in a.only.split(","):
            try:
                fn(W)
            except Exception as e:  # noqa: BLE001
                print(json.dumps({"group": name, "failed": repr(e)[:300]}), flush=True)
                torch.cuda.synchronize()
    if a.out:
        with open(a.out, "w") as f:
            f.write("# Head-to-head vs the reference's sm_100 kernels (vllm wheel on the box), 1 x B200\n\n"
                    "`python tools/bench_vs_wheel.py` - same tensors, CUDA-graph timed, large operand rotated beyond L2, "
                    "best of 3. ratio = wheel_us / ours_us (> 1: ours faster).\n\n"
                    "| op | case | ours us | wheel us | ratio | note |\n|---|---|---|---|---|---|\n")
            for r in ROWS:
                f.write(f"| {r['group']} | {r['case']} | {r['ours_us']} | {r['wheel_us']} | "
                        f"{r['ratio_wheel_over_ours']} | {r['note']} |\n")


if __name__ == "__main__":
    main()
