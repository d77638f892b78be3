"""Same-box A/B of the paged-attention decode path: the current library against the round-2 starting point
(tools/_ab_old/libattn_old.so = commit 3a004ea's attention.cu, built by hand: `git show 3a004ea:vllm-fork_b200/csrc/...`
+ `nvcc -shared`; decode kernel + reduce kernel, plain launches). Both are timed as a CUDA graph of 32 launches over 2 alternating KV caches (device time, us per
attention call incl. the old version's second launch). Development tool; needs tools/_ab_old built."""
import ctypes
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vllm-fork_b200"))
from b200ops import _lib, ops  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    lib = _lib.load()
    old = ctypes.CDLL(os.path.join(ROOT, "tools", "_ab_old", "libattn_old.so"))
    old.b200_paged_attention_decode.restype = ctypes.c_int
    old.b200_paged_attention_decode.argtypes = [ctypes.POINTER(_lib.PagedAttnArgs), ctypes.c_void_p]
    new_fn = lib.b200_paged_attention_decode
    for (S, L, Hq, Hkv) in ((256, 4096, 32, 8), (256, 4096, 4, 1), (32, 32768, 32, 8), (128, 8192, 8, 1)):
        D, bs = 128, 16
        nblk = (L + bs - 1) // bs
        nb = S * nblk
        one = torch.tensor([0.05], device=dev)
        bt = torch.randperm(nb, device=dev).to(torch.int32).view(S, nblk)
        lens = torch.full((S,), L, dtype=torch.int32, device=dev)
        q = torch.randn(S, Hq, D, device=dev, dtype=torch.bfloat16)
        kvs = [torch.randint(0, 0x70, (nb, bs, Hkv, 2 * D), dtype=torch.uint8, device=dev) for _ in range(2)]
        ws = torch.zeros(ops.paged_attention_workspace(S, Hq, Hkv, D, dev).numel() + 8192, dtype=torch.uint8, device=dev)
        outs = {}
        res = {}
        for rnd in range(2):               # A B A B: drift between the two shows up as a difference between rounds
            for name, fn in (("old", old.b200_paged_attention_decode), ("new", new_fn)):
                lib.b200_paged_attention_decode = fn
                ws.zero_()
                out = torch.empty(S, Hq, D, device=dev, dtype=torch.bfloat16)

                def call(i):
                    kv = kvs[i % 2]
                    ops.paged_attention_decode(out, q, kv[..., :D], kv[..., D:], bt, lens, D ** -0.5, k_scale=one,
                                               v_scale=one, max_seq_len=L, workspace=ws)
                for i in range(3):
                    call(i)
                torch.cuda.synchronize()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    for i in range(32):
                        call(i)
                g.replay()
                torch.cuda.synchronize()
                best = 1e9
                for _ in range(5):
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    g.replay()
                    e1.record()
                    torch.cuda.synchronize()
                    best = min(best, e0.elapsed_time(e1) / 32 * 1e3)
                res.setdefault(name, []).append(round(best, 2))
                outs[name] = out.clone()
        lib.b200_paged_attention_decode = new_fn
        same = bool(torch.equal(outs["old"], outs["new"]))
        print(json.dumps({"S": S, "L": L, "Hq": Hq, "Hkv": Hkv, "us_old_decode_plus_reduce": res["old"],
                          "us_current": res["new"], "outputs_bit_identical": same}), flush=True)


if __name__ == "__main__":
    main()
