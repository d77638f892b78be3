"""Shared input builders for the parity tests (recipes of SURVEY.md section 8d)."""
import numpy as np
import torch


def bits_to_bf16(a: np.ndarray) -> torch.Tensor:
    return torch.from_numpy(a.view(np.int16).copy()).view(torch.bfloat16)


def make_block_table(lens, block_size, num_blocks, gen):
    """Random permutation of physical blocks WITHOUT repetition."""
    S = len(lens)
    max_blocks = max(1, (max(lens) + block_size - 1) // block_size)
    assert num_blocks >= S * max_blocks
    perm = torch.randperm(num_blocks, generator=gen)[: S * max_blocks]
    return perm.to(torch.int32).view(S, max_blocks)


def slots_for(lens, block_table, block_size):
    out = []
    for s, L in enumerate(lens):
        pos = torch.arange(L)
        out.append(block_table[s, pos // block_size].to(torch.int64) * block_size + pos % block_size)
    return torch.cat(out) if out else torch.zeros(0, dtype=torch.int64)


def make_kv_cache(num_blocks, block_size, Hkv, D, dtype, layout="NHD", packed=True):
    """Cache in the reference fork's layout: logical (nb, Hkv, bs, 2D), K|V packed in the last dim,
    physical order NHD or HND (flash_attn.py:139-173). Returns (raw, key_cache, value_cache) with
    the caches as [nb, bs, Hkv, D] views. packed=False gives the (2, nb, bs, Hkv, D) split layout."""
    if not packed:
        raw = torch.zeros(2, num_blocks, block_size, Hkv, D, dtype=dtype)
        return raw, raw[0], raw[1]
    if layout == "NHD":
        raw = torch.zeros(num_blocks, block_size, Hkv, 2 * D, dtype=dtype)
        kv = raw
    else:
        raw = torch.zeros(num_blocks, Hkv, block_size, 2 * D, dtype=dtype)
        kv = raw.transpose(1, 2)
    return raw, kv[..., :D], kv[..., D:]
