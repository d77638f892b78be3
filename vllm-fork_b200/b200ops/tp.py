"""Tensor-parallel partitioning rules of the decode path (pure host logic, importable without CUDA).

Mirrors vllm/model_executor/layers/linear.py: ColumnParallelLinear (:419) / QKVParallelLinear (:1022)
shard the OUTPUT features (weight rows), RowParallelLinear (:1613) shards the INPUT features (weight
columns) and is followed by one all-reduce (:1766-1767). Attention shards by KV head with no
communication. The sequence-parallel split of the fused epilogue (csrc/allreduce_fused.cuh)
assigns contiguous row blocks of ceil(M / world) tokens to the ranks.
"""
from __future__ import annotations


def column_shard(n_out: int, rank: int, world: int) -> slice:
    """rows of a [n_out, k] weight owned by `rank` (ColumnParallelLinear)."""
    if n_out % world:
        raise ValueError(f"output size {n_out} is not divisible by tp size {world}")
    per = n_out // world
    return slice(rank * per, (rank + 1) * per)


def row_shard(k_in: int, rank: int, world: int) -> slice:
    """columns of a [n, k_in] weight owned by `rank` (RowParallelLinear)."""
    if k_in % world:
        raise ValueError(f"input size {k_in} is not divisible by tp size {world}")
    per = k_in // world
    return slice(rank * per, (rank + 1) * per)


def qkv_shard(num_heads: int, num_kv_heads: int, head_size: int, rank: int, world: int):
    """(q rows, k rows, v rows) of the separate q/k/v weights owned by `rank` (QKVParallelLinear):
    heads are dealt out contiguously, so query head h and its kv head h // group land on one rank."""
    if num_heads % world or num_kv_heads % world:
        raise ValueError("num_heads and num_kv_heads must be divisible by the tp size (no KV-head replication)")
    hq, hkv = num_heads // world, num_kv_heads // world
    return (slice(rank * hq * head_size, (rank + 1) * hq * head_size),
            slice(rank * hkv * head_size, (rank + 1) * hkv * head_size),
            slice(rank * hkv * head_size, (rank + 1) * hkv * head_size))


def rows_of_rank(num_tokens: int, rank: int, world: int) -> tuple[int, int]:
    """[lo, hi) token rows that `rank` normalises in the fused reduce-scatter epilogue."""
    per = (num_tokens + world - 1) // world
    return min(num_tokens, rank * per), min(num_tokens, (rank + 1) * per)


def expert_shard(num_experts: int, rank: int, world: int) -> tuple[int, int]:
    """Expert parallelism: rank r owns the contiguous experts [lo, hi) whole (linear placement of
    fused_moe/layer.py determine_expert_map)."""
    if num_experts % world != 0:
        raise ValueError(f"expert parallel: {num_experts} experts do not divide over {world} ranks")
    per = num_experts // world
    return rank * per, (rank + 1) * per


def expert_map(num_experts: int, rank: int, world: int):
    """int32 [num_experts]: global expert id -> local id on this rank, or -1 (the `expert_map` argument of
    fused_experts / moe_align_block_size). None when there is one rank."""
    import torch
    if world == 1:
        return None
    lo, hi = expert_shard(num_experts, rank, world)
    em = torch.full((num_experts,), -1, dtype=torch.int32)
    em[lo:hi] = torch.arange(hi - lo, dtype=torch.int32)
    return em
