"""Generate tests/golden/cpu_attn_*.npz from the reference's OWN CPU paged attention.

Run in the build container (needs oracle/_ref/refcpu.so, built from /root/reference by
oracle/build_ref_cpu_attn.sh):   python tests/golden/make_golden.py

Each fixture holds seeded inputs (query, per-token key/value, slot_mapping, block_table, seq_lens)
and the output of cpu_attn_reshape_and_cache + cpu_attention_with_kv_cache
(/root/reference/csrc/cpu/cpu_attn.cpp:88-254) on them. bf16 tensors are stored as uint16 bit
patterns. tests/test_oracle.py replays the inputs through oracle/ref_ops.py.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import refcpu  # noqa: E402

CASES = {
    # name: (seq_lens, Hq, Hkv, D, block_size, sliding_window, softcap, alibi, kv_split)
    "c1_qwen_shape": ([512, 37, 300, 129, 1, 16, 511, 200], 14, 2, 64, 16, None, None, False, True),
    "llama_shape": ([523, 37, 201, 64], 8, 2, 128, 16, None, None, False, False),
    "window_softcap": ([300, 77, 130], 4, 4, 128, 16, 128, 50.0, False, False),
    "alibi_mqa": ([90, 257, 33], 5, 1, 128, 32, None, None, True, True),
}


# variable-length (prefill / chunked prefill / mixed) cases for the query_len > 1 path:
# name: (seq_lens [(query_len, kv_len)], Hq, Hkv, D, block_size, sliding_window, softcap)
VARLEN_CASES = {
    "mixed_llama": ([(1, 523), (5, 18), (129, 463), (64, 64)], 8, 2, 128, 16, None, None),
    "prefill_qwen": ([(33, 33), (200, 200), (1, 1), (17, 300)], 14, 2, 64, 16, None, None),
    "window_softcap": ([(70, 300), (130, 130), (1, 77)], 4, 4, 128, 16, 128, 50.0),
}


def bits(t: torch.Tensor) -> np.ndarray:
    return t.contiguous().view(torch.int16).numpy().view(np.uint16)


def alibi_slopes(n):
    base = 2 ** (-8.0 / n)
    return torch.tensor([base ** (i + 1) for i in range(n)], dtype=torch.float32)


def main():
    out_dir = os.path.dirname(os.path.abspath(__file__))
    for name, (lens, Hq, Hkv, D, bs, win, cap, alibi, split) in CASES.items():
        torch.manual_seed(0)
        S = len(lens)
        dtype = torch.bfloat16
        max_blocks = (max(lens) + bs - 1) // bs
        num_blocks = S * max_blocks + 3
        perm = torch.randperm(num_blocks)[: S * max_blocks].to(torch.int32).view(S, max_blocks)
        scale = D ** -0.5
        q = (torch.rand(S, Hq, D) * 2 - 1).mul(1.5).to(dtype)
        T = sum(lens)
        key = (torch.rand(T, Hkv, D) * 2 - 1).to(dtype)
        value = (torch.rand(T, Hkv, D) * 2 - 1).to(dtype)
        slots = []
        for s, L in enumerate(lens):
            pos = torch.arange(L)
            slots.append(perm[s, pos // bs].to(torch.int64) * bs + pos % bs)
        slot_mapping = torch.cat(slots)
        isa = refcpu.pick_isa(bs, D)
        kc, vc = refcpu.make_packed_cache(num_blocks, Hkv, bs, D, dtype)
        kc.zero_(); vc.zero_()
        refcpu.reshape_and_cache(key, value, kc, vc, slot_mapping, isa)
        sl = torch.tensor(lens, dtype=torch.int32)
        qsl = torch.arange(S + 1, dtype=torch.int32)
        md = refcpu.scheduler_metadata(S, Hq, Hkv, D, sl, dtype, qsl, isa, enable_kv_split=split,
                                       sliding_window=win if win else -1)
        out = torch.empty_like(q)
        slopes = alibi_slopes(Hq) if alibi else None
        refcpu.attention(q, kc, vc, out, qsl, sl, scale, perm, md, alibi_slopes=slopes,
                         sliding_window=win if win else -1, softcap=cap if cap else 0.0)
        np.savez_compressed(
            os.path.join(out_dir, f"cpu_attn_{name}.npz"),
            query=bits(q), key=bits(key), value=bits(value), slot_mapping=slot_mapping.numpy(),
            block_table=perm.numpy(), seq_lens=sl.numpy(), out=bits(out),
            meta=np.array([Hq, Hkv, D, bs, num_blocks, win or 0, int(alibi)], dtype=np.int64),
            softcap=np.array([cap or 0.0], dtype=np.float32),
            scale=np.array([scale], dtype=np.float64),
            alibi=(slopes.numpy() if alibi else np.zeros(0, np.float32)))
        print(name, "ok", tuple(out.shape), float(out.float().abs().mean()))

    for name, (seq_lens, Hq, Hkv, D, bs, win, cap) in VARLEN_CASES.items():
        torch.manual_seed(0)
        qlens = [a for a, _ in seq_lens]
        lens = [b for _, b in seq_lens]
        S = len(lens)
        dtype = torch.bfloat16
        max_blocks = (max(lens) + bs - 1) // bs
        num_blocks = S * max_blocks + 3
        perm = torch.randperm(num_blocks)[: S * max_blocks].to(torch.int32).view(S, max_blocks)
        scale = D ** -0.5
        q = (torch.rand(sum(qlens), Hq, D) * 2 - 1).mul(1.5).to(dtype)
        T = sum(lens)
        key = (torch.rand(T, Hkv, D) * 2 - 1).to(dtype)
        value = (torch.rand(T, Hkv, D) * 2 - 1).to(dtype)
        slots = []
        for s, L in enumerate(lens):
            pos = torch.arange(L)
            slots.append(perm[s, pos // bs].to(torch.int64) * bs + pos % bs)
        slot_mapping = torch.cat(slots)
        isa = refcpu.pick_isa(bs, D)
        kc, vc = refcpu.make_packed_cache(num_blocks, Hkv, bs, D, dtype)
        kc.zero_(); vc.zero_()
        refcpu.reshape_and_cache(key, value, kc, vc, slot_mapping, isa)
        sl = torch.tensor(lens, dtype=torch.int32)
        qsl = torch.zeros(S + 1, dtype=torch.int32)
        qsl[1:] = torch.tensor(qlens).cumsum(0)
        md = refcpu.scheduler_metadata(S, Hq, Hkv, D, sl, dtype, qsl, isa, enable_kv_split=False,
                                       sliding_window=win if win else -1)
        out = torch.empty_like(q)
        refcpu.attention(q, kc, vc, out, qsl, sl, scale, perm, md, sliding_window=win if win else -1,
                         softcap=cap if cap else 0.0)
        np.savez_compressed(
            os.path.join(out_dir, f"cpu_varlen_{name}.npz"),
            query=bits(q), key=bits(key), value=bits(value), slot_mapping=slot_mapping.numpy(),
            block_table=perm.numpy(), seq_lens=sl.numpy(), query_lens=np.array(qlens, dtype=np.int32),
            out=bits(out), meta=np.array([Hq, Hkv, D, bs, num_blocks, win or 0], dtype=np.int64),
            softcap=np.array([cap or 0.0], dtype=np.float32), scale=np.array([scale], dtype=np.float64))
        print("varlen", name, "ok", tuple(out.shape), float(out.float().abs().mean()))


if __name__ == "__main__":
    main()
