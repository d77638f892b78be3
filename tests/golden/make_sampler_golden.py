"""Generates tests/golden/topk_topp_ref.npz by EXECUTING the reference's own apply_top_k_top_p_pytorch
(vllm/v1/sample/ops/topk_topp_sampler.py:367-408; pure torch, its source text is exec'd from /root/reference so no
vllm import is needed). Run once in the build container: python tests/golden/make_sampler_golden.py"""
import os

import numpy as np
import torch

REF = "/root/reference/vllm/v1/sample/ops/topk_topp_sampler.py"
src = open(REF).read()
ns = {"torch": torch}
exec(src[src.index("def apply_top_k_top_p_pytorch("):src.index("def apply_top_k_only(")], ns)
f = ns["apply_top_k_top_p_pytorch"]
g = torch.Generator().manual_seed(1234)
x = torch.randn(12, 2000, generator=g) * 3
k = torch.randint(1, 64, (12,), generator=g)
k[0] = 2000
p = torch.rand(12, generator=g) * 0.9 + 0.05
out = {"x": x.numpy(), "k": k.numpy(), "p": p.numpy()}
for name, (kk, pp) in {"k": (k, None), "p": (None, p), "kp": (k, p)}.items():
    out["out_" + name] = f(x.clone(), kk, pp).numpy()
np.savez_compressed(os.path.join(os.path.dirname(__file__), "topk_topp_ref.npz"), **out)
print("wrote topk_topp_ref.npz")
