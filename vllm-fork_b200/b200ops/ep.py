"""Expert parallelism over data-parallel attention: the "naive" dispatch / combine of the reference.

Reference: vllm/model_executor/layers/fused_moe/prepare_finalize/naive_dp_ep.py:71-211
(MoEPrepareAndFinalizeNaiveDPEPModular: prepare = quantise + dispatch, finalize = weight-and-reduce + combine) on top
of the EP group's dispatch / combine (vllm/distributed/device_communicators/base_device_communicator.py: all-gather of
the tokens, their scales, top-k weights and ids along dim 0; reduce-scatter of the expert outputs along dim 0).

Every rank holds ITS tokens (data-parallel attention, or the sequence-parallel row shard of a TP group) and ITS
experts. dispatch all-gathers the (already quantised) tokens and routing tables so every rank sees all tokens; each
rank runs `fused_experts` with its `expert_map` (pairs routed elsewhere are dropped: their contribution is zero);
combine reduce-scatters the partial outputs so every rank gets the full sum for its own tokens. Against the first EP
form (replicated tokens, all-reduce of [M, H]) the combine moves (n-1)/n of one activation instead of 2(n-1)/n.

The collectives here are plain torch.distributed calls (NCCL on GPUs; gloo in the CPU tests, where reduce_scatter is
emulated by all_reduce + slice exactly like the reference's "AR" fallback). Inside the decoder driver the same form
runs on the fused NVLink kernel instead (decoder.py, comm="fused": reduce-scatter + residual + RMSNorm + all-gather in
one launch). Pure host logic: importable without CUDA.
"""
from __future__ import annotations

from typing import Callable, Sequence

import torch
import torch.distributed as dist


def _sizes(group, n_local: int, sizes: Sequence[int] | None) -> list[int]:
    world = dist.get_world_size(group)
    if sizes is None:
        return [n_local] * world
    if len(sizes) != world or sizes[dist.get_rank(group)] != n_local:
        raise ValueError(f"ep: sizes {list(sizes)} do not describe this rank's {n_local} tokens")
    return list(sizes)


def all_gatherv(t: torch.Tensor, group, sizes: Sequence[int] | None = None) -> torch.Tensor:
    """concatenate every rank's rows along dim 0 (rows per rank = sizes[r]; equal when sizes is None)."""
    sz = _sizes(group, t.shape[0], sizes)
    t = t.contiguous()
    if len(set(sz)) == 1:
        out = torch.empty((sum(sz),) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        dist.all_gather_into_tensor(out, t, group=group)
        return out
    # uneven: pad every rank's block to the largest, gather, drop the pads (works on every backend)
    mx = max(sz)
    padded = torch.zeros((mx,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    padded[: t.shape[0]] = t
    out = torch.empty((mx * len(sz),) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    dist.all_gather_into_tensor(out, padded, group=group)
    return torch.cat([out[r * mx: r * mx + n] for r, n in enumerate(sz)], dim=0)


def reduce_scatterv(t: torch.Tensor, group, sizes: Sequence[int] | None = None) -> torch.Tensor:
    """sum over the ranks, keep this rank's rows (dim 0 split by sizes)."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    if sizes is None:
        if t.shape[0] % world:
            raise ValueError(f"ep: {t.shape[0]} rows do not divide over {world} ranks (pass sizes)")
        sz = [t.shape[0] // world] * world
    else:
        sz = list(sizes)
        if sum(sz) != t.shape[0]:
            raise ValueError(f"ep: sizes {sz} do not sum to {t.shape[0]} rows")
    t = t.contiguous()
    lo = sum(sz[:rank])
    if len(set(sz)) == 1 and dist.get_backend(group) == "nccl":
        out = torch.empty((sz[rank],) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        dist.reduce_scatter_tensor(out, t, group=group)
        return out
    # uneven shards, or a backend without reduce-scatter (gloo): all-reduce and slice (the reference's AR fallback)
    full = t.clone()
    dist.all_reduce(full, group=group)
    return full[lo:lo + sz[rank]].clone()


class NaiveDPEPPrepareFinalize:
    """MoEPrepareAndFinalizeNaiveDPEPModular (naive_dp_ep.py:71-211) with the quantiser and the experts passed in.

    quantize(a1) -> (a1q, a1q_scale or None): moe_kernel_quantize_input for the configured scheme (None = the input
    quantisation is deferred to the expert kernel, `defer_input_quant`). A scale that is a 0-d / 1-element tensor is
    static (replicated on every rank) and is not gathered (naive_dp_ep.py:46-50)."""

    def __init__(self, group, sizes: Sequence[int] | None = None,
                 quantize: Callable[[torch.Tensor], tuple[torch.Tensor, torch.Tensor | None]] | None = None):
        self.group = group
        self.sizes = None if sizes is None else list(sizes)
        self.quantize = quantize

    def prepare(self, a1: torch.Tensor, topk_weights: torch.Tensor, topk_ids: torch.Tensor,
                apply_router_weight_on_input: bool = False):
        """-> (a1q [M_total, H], a1q_scale, topk_ids [M_total, k], topk_weights [M_total, k])"""
        if apply_router_weight_on_input:
            if topk_ids.size(1) != 1:
                raise RuntimeError("apply_router_weight_on_input is only implemented for topk=1")
            a1 = a1 * topk_weights.to(a1.dtype)
        a1q, scale = (a1, None) if self.quantize is None else self.quantize(a1)
        gather_scale = scale is not None and scale.numel() > 1
        a1q_all = all_gatherv(a1q.view(torch.uint8) if a1q.element_size() == 1 else a1q, self.group, self.sizes)
        if a1q.element_size() == 1:
            a1q_all = a1q_all.view(a1q.dtype)
        tw_all = all_gatherv(topk_weights, self.group, self.sizes)
        ti_all = all_gatherv(topk_ids, self.group, self.sizes)
        if gather_scale:
            scale = all_gatherv(scale.reshape(a1.shape[0], -1), self.group, self.sizes)
        return a1q_all, scale, ti_all, tw_all

    def finalize(self, fused_expert_output: torch.Tensor) -> torch.Tensor:
        """fused_expert_output [M_total, H]: this rank's experts' weighted partial sums for ALL tokens (what
        fused_experts with an expert_map returns) -> the complete output rows of this rank's tokens."""
        return reduce_scatterv(fused_expert_output, self.group, self.sizes)


def dp_ep_fused_experts(hidden_local: torch.Tensor, topk_weights: torch.Tensor, topk_ids: torch.Tensor, group,
                        experts_fn: Callable[..., torch.Tensor], sizes: Sequence[int] | None = None,
                        quantize=None, apply_router_weight_on_input: bool = False) -> torch.Tensor:
    """The whole block: prepare (dispatch) -> experts_fn(a1q, a1q_scale, topk_weights, topk_ids) on the gathered
    tokens with this rank's experts -> finalize (combine). experts_fn is e.g.
    ``lambda a, s, w, i: fused_experts(a, w1_local, w2_local, w, i, use_fp8_w8a8=True, expert_map=emap, ...)``."""
    pf = NaiveDPEPPrepareFinalize(group, sizes, quantize)
    a1q, scale, ti, tw = pf.prepare(hidden_local, topk_weights, topk_ids, apply_router_weight_on_input)
    out = experts_fn(a1q, scale, tw, ti)
    return pf.finalize(out)
