"""Synthetic decoder driver: one decode step of a Llama-shaped FP8 W8A8 + FP8-KV model built ONLY from
the hot-path operators of this package (the role gpu_model_runner._model_forward plays in the
reference, vllm/v1/worker/gpu_model_runner.py:3927, for the layer loop of
vllm/model_executor/models/llama.py:222-328).

Per layer:
    RMSNorm -> static FP8 quant -> qkv GEMM -> RoPE -> KV-cache write -> paged attention ->
    static FP8 quant -> o_proj GEMM (-> all-reduce) -> add+RMSNorm -> quant -> gate_up GEMM ->
    silu_and_mul_quant -> down GEMM (-> all-reduce)
then final norm -> bf16 lm_head -> greedy argmax.

Embedding lookup / bf16 lm_head are outside the path named by north_star and use plain torch ops (a library
GEMM for lm_head); greedy sampling is this package's argmax kernel; everything else - including the SURVEY.md section 8f rank-1 rows
RMSNorm(+residual)+FP8-quant and RoPE - is libb200ops.so.
Tensor parallelism: attention heads and GEMM N/K are sharded over the ranks
of `tp_group`, one all-reduce after o_proj and down_proj (linear.py:1766-1767). `comm` picks the rung
of the SURVEY 8e ladder: "nccl" (torch.distributed, NVLS), "nvlink" (own one-shot all-reduce over peer
memory, b200ops/all_reduce.py, GEMM outputs in the symmetric heap = zero copy), "fused" (one kernel:
reduce-scatter + residual add + RMSNorm + FP8 quant + all-gather; the residual stream is then kept
row-sharded across the ranks).
"""
from __future__ import annotations

import dataclasses
import contextlib
import os

import torch

from . import fused_moe, ops, tp


@dataclasses.dataclass(frozen=True)
class ModelShape:
    name: str
    hidden: int
    layers: int
    heads: int
    kv_heads: int
    head_size: int
    inter: int
    vocab: int
    rope_theta: float = 500000.0
    eps: float = 1e-5
    experts: int = 0   # > 0: sparse MoE MLP (Mixtral: models/mixtral.py:76-160 -> FusedMoE), `inter` per expert
    topk: int = 0


LLAMA_3_1_8B = ModelShape("Llama-3.1-8B-shaped", 4096, 32, 32, 8, 128, 14336, 128256)
LLAMA_3_1_70B = ModelShape("Llama-3.1-70B-shaped", 8192, 80, 64, 8, 128, 28672, 128256)
QWEN2_5_0_5B = ModelShape("Qwen2.5-0.5B-shaped", 896, 24, 14, 2, 64, 4864, 151936, 1000000.0, 1e-6)
MIXTRAL_8X7B = ModelShape("Mixtral-8x7B-shaped", 4096, 32, 32, 8, 128, 14336, 32000, 1000000.0, 1e-5, 8, 2)
TINY = ModelShape("tiny-llama-shaped", 512, 2, 8, 2, 128, 1024, 1024)
# tiny shape whose heads / kv heads / inter divide by 8 ranks (bench.py's TP=N vs TP=1 self-check)
TINY_TP = ModelShape("tiny-llama-shaped-tp8", 1024, 2, 16, 8, 128, 2048, 1024)
TINY_MOE = ModelShape("tiny-mixtral-shaped", 512, 2, 8, 2, 128, 512, 1024, 1000000.0, 1e-5, 4, 2)
TINY_MOE_EP = ModelShape("tiny-mixtral-shaped-ep8", 1024, 2, 16, 8, 128, 512, 1024, 1000000.0, 1e-5, 8, 2)

# kernels launched per call of each op (our own kernels only), for the bench's gpu_launches claim
LAUNCHES = {"quant_static": 1, "gemm": 1, "cache_write": 1, "attention": 2, "silu_quant": 1,
            # moe_align + permute/quant + grouped GEMM + silu/quant + grouped GEMM + unpermute/reduce
            "topk_softmax": 1, "fused_experts": 6}


def _quant_weight(w: torch.Tensor):
    """per-tensor FP8 weight quantisation, scale = amax / 448."""
    scale = (w.abs().amax().float() / 448.0).clamp(min=1e-12).reshape(1)
    q = (w.float() / scale).clamp(-448, 448).to(torch.float8_e4m3fn)
    return q, scale


class DecoderModel:
    def __init__(self, shape: ModelShape, max_seqs: int, max_ctx: int, block_size: int = 16,
                 device: str | torch.device = "cuda:0", tp_rank: int = 0, tp_size: int = 1,
                 tp_group=None, seed: int = 1234, kv_fp8: bool = True, num_blocks: int | None = None,
                 comm: str = "nccl", quant: str = "fp8", multicast: bool | None = None,
                 pdl: int | str | None = "auto"):
        self.shape = s = shape
        # Programmatic dependent launch between the kernels of the step. "auto": level 2 on one GPU (the GEMMs request
        # their first weight tiles before waiting for the previous kernel - the weights are never written after init;
        # +13 % on a launch-bound step, profiles/r2s_small_pdl*.json), level 1 with tensor parallelism, where the NVLink
        # collectives themselves stay plain launches (B200_PDL_AR, csrc/common.cuh): measured at 2 ranks
        # (profiles/r3e_*.json) that is 1.8 % faster than no programmatic launch, while collectives that take part in
        # it make the step 1-2 % slower (and one such run ended in the all-reduce's bounded-spin trap). B200_PDL in
        # the environment wins (A/B runs); pdl=None leaves the library setting alone.
        # The level is a property of THIS model, applied around its own launches in forward_device() and restored
        # afterwards: level 2 is a promise about this model's weights and must not leak to other callers of the
        # process-global b200_set_pdl (a later cutlass_scaled_mm whose `b` was produced on-stream would inherit it).
        if pdl == "auto":
            pdl = 2 if tp_size == 1 else 1
        lib_level = ops.set_pdl(0)
        ops.set_pdl(lib_level)
        self.pdl = lib_level if (pdl is None or "B200_PDL" in os.environ) else max(0, min(2, int(pdl)))
        assert quant in ("fp8", "nvfp4")
        self.quant = quant   # "nvfp4": W4A4 linears (scaled_fp4_quant -> cutlass_scaled_fp4_mm), FP8 KV cache
        self.dev = torch.device(device)
        self.tp_rank, self.tp_size, self.tp_group = tp_rank, tp_size, tp_group
        assert s.heads % tp_size == 0 and s.kv_heads % tp_size == 0 and s.inter % tp_size == 0
        self.moe = s.experts > 0
        if self.moe:
            assert s.experts % tp_size == 0, "expert parallel: experts must divide over the ranks"
            assert quant == "fp8", "MoE layers: FP8 fused_experts only"
            self.e_lo, self.e_hi = tp.expert_shard(s.experts, tp_rank, tp_size)
        self.hq, self.hkv = s.heads // tp_size, s.kv_heads // tp_size
        self.inter = s.inter if s.experts > 0 else s.inter // tp_size
        self.D, self.bs = s.head_size, block_size
        self.max_seqs, self.max_ctx = max_seqs, max_ctx
        self.blocks_per_seq = (max_ctx + block_size - 1) // block_size
        self.num_blocks = num_blocks or max_seqs * self.blocks_per_seq + 8
        self.kv_fp8 = kv_fp8
        self.launches_per_step = 0
        assert comm in ("nccl", "nvlink", "nvls", "fused")
        self.comm = comm if tp_size > 1 else "nccl"
        self.multicast = multicast
        self.ar = None
        g = torch.Generator(device=self.dev).manual_seed(seed)

        def rnd(*shp, fan_in):
            return (torch.rand(*shp, device=self.dev, generator=g) * 2 - 1) * (1.7 / fan_in ** 0.5)

        H, D = s.hidden, s.head_size
        self.embed = rnd(s.vocab, H, fan_in=1.0).mul(0.5).to(torch.bfloat16)
        self.lm_head = rnd(s.vocab, H, fan_in=H).to(torch.bfloat16)
        self.final_norm = torch.ones(H, device=self.dev, dtype=torch.bfloat16)
        self.layers = []
        one = lambda v: torch.tensor([v], device=self.dev, dtype=torch.float32)
        for _ in range(s.layers):
            L = {}
            # every rank draws the full tensors with the same generator and keeps its shard, so a
            # TP run computes the same function as the single-GPU run
            wq = rnd(s.heads * D, H, fan_in=H)
            wk = rnd(s.kv_heads * D, H, fan_in=H)
            wv = rnd(s.kv_heads * D, H, fan_in=H)
            sq, sk, sv = tp.qkv_shard(s.heads, s.kv_heads, D, tp_rank, tp_size)
            wqkv = torch.cat([wq[sq], wk[sk], wv[sv]])
            wo = rnd(H, s.heads * D, fan_in=s.heads * D)[:, tp.row_shard(s.heads * D, tp_rank, tp_size)]
            # per-tensor scales must be identical on all ranks: derive them from the full tensors
            if quant == "nvfp4":
                self._init_nvfp4_layer(L, wqkv, wo, rnd, one)
                L["in_norm"] = torch.ones(H, device=self.dev, dtype=torch.bfloat16)
                L["post_norm"] = torch.ones(H, device=self.dev, dtype=torch.bfloat16)
                L["k_scale"], L["v_scale"] = one(4.0 / 448), one(4.0 / 448)
                cdt = torch.uint8 if kv_fp8 else torch.bfloat16
                L["kv"] = torch.zeros(self.num_blocks, block_size, self.hkv, 2 * D, dtype=cdt, device=self.dev)
                self.layers.append(L)
                del wq, wk, wv, wqkv, wo
                continue
            L["wqkv"], L["wqkv_s"] = _quant_weight(wqkv.contiguous())
            L["wo"], L["wo_s"] = _quant_weight(wo.contiguous())
            if self.moe:
                # expert parallel: rank r owns experts [r*E/ep, (r+1)*E/ep) whole (no TP inside
                # an expert); every rank draws all experts from the same generator and keeps its own
                L["gate_w_t"] = rnd(s.experts, H, fan_in=H).to(torch.bfloat16).t().contiguous()
                w1, w1s, w2, w2s = [], [], [], []
                for e in range(s.experts):
                    wg = rnd(2 * s.inter, H, fan_in=H)
                    wd = rnd(H, s.inter, fan_in=s.inter)
                    if self.e_lo <= e < self.e_hi:
                        q1, s1 = _quant_weight(wg)
                        q2, s2 = _quant_weight(wd)
                        w1.append(q1); w1s.append(s1); w2.append(q2); w2s.append(s2)
                    del wg, wd
                L["w1"], L["w1_s"] = torch.stack(w1), torch.cat(w1s)
                L["w2"], L["w2_s"] = torch.stack(w2), torch.cat(w2s)
            else:
                cs = tp.column_shard(s.inter, tp_rank, tp_size)
                wg = rnd(s.inter, H, fan_in=H)[cs]
                wu = rnd(s.inter, H, fan_in=H)[cs]
                wd = rnd(H, s.inter, fan_in=s.inter)[:, tp.row_shard(s.inter, tp_rank, tp_size)]
                L["wgu"], L["wgu_s"] = _quant_weight(torch.cat([wg, wu]).contiguous())
                L["wd"], L["wd_s"] = _quant_weight(wd.contiguous())
                del wg, wu, wd
            L["in_norm"] = torch.ones(H, device=self.dev, dtype=torch.bfloat16)
            L["post_norm"] = torch.ones(H, device=self.dev, dtype=torch.bfloat16)
            # static activation scales (FP8 "static" checkpoints carry input_scale per linear)
            L["a_qkv"], L["a_o"], L["a_gu"], L["a_d"] = one(4.0 / 448), one(2.0 / 448), one(4.0 / 448), one(8.0 / 448)
            L["k_scale"], L["v_scale"] = one(4.0 / 448), one(4.0 / 448)
            cdt = torch.uint8 if kv_fp8 else torch.bfloat16
            # reference layout: (num_blocks, block_size, kv_heads, 2*head_size), K|V packed (NHD)
            L["kv"] = torch.zeros(self.num_blocks, block_size, self.hkv, 2 * D, dtype=cdt, device=self.dev)
            self.layers.append(L)
            del wq, wk, wv, wqkv, wo
        # rotary cache [max_pos, D] (cos | sin), neox style (rotary_embedding/base.py)
        inv = 1.0 / (s.rope_theta ** (torch.arange(0, D, 2, device=self.dev).float() / D))
        t = torch.arange(max_ctx + 1, device=self.dev).float()
        fr = torch.outer(t, inv)
        self.cos_sin = torch.cat([fr.cos(), fr.sin()], dim=-1).to(torch.bfloat16).contiguous()
        self.residual = torch.empty(max_seqs, H, dtype=torch.bfloat16, device=self.dev)
        # persistent step buffers (CUDA-graph friendly: fixed addresses)
        S = max_seqs
        self.token_ids = torch.zeros(S, dtype=torch.int64, device=self.dev)
        self.positions = torch.zeros(S, dtype=torch.int64, device=self.dev)
        self.slot_mapping = torch.zeros(S, dtype=torch.int64, device=self.dev)
        self.seq_lens = torch.zeros(S, dtype=torch.int32, device=self.dev)
        self.block_table = torch.zeros(S, self.blocks_per_seq, dtype=torch.int32, device=self.dev)
        self.next_tokens = torch.zeros(S, dtype=torch.int64, device=self.dev)
        self.workspace = ops.paged_attention_workspace(S, self.hq, self.hkv, D, self.dev)
        # FP8 GEMM inputs: normalised hidden states [S, H] and attention output [S, hq*D] (both contiguous)
        self.x8 = torch.empty(S, H, dtype=torch.float8_e4m3fn, device=self.dev)
        self.a8 = torch.empty(S, self.hq * D, dtype=torch.float8_e4m3fn, device=self.dev)
        self.h8 = torch.empty(S, self.inter, dtype=torch.float8_e4m3fn, device=self.dev)
        self.qkv = torch.empty(S, (self.hq + 2 * self.hkv) * D, dtype=torch.bfloat16, device=self.dev)
        self.attn_out = torch.empty(S, self.hq, D, dtype=torch.bfloat16, device=self.dev)
        self.proj = torch.empty(S, H, dtype=torch.bfloat16, device=self.dev)
        self.proj_in = [self.proj, self.proj]   # GEMM outputs feeding the collective (o_proj, down_proj)
        if self.comm != "nccl":
            from .all_reduce import B200AllReduce
            self.ar = B200AllReduce(tp_group, self.dev, max_size=max(8 << 20, S * H * 2),
                                    heap_size=6 * S * H * 2 + (1 << 20), multicast=self.multicast)
            if self.comm == "nvls":
                if not self.ar.multicast:
                    raise RuntimeError(f"comm=nvls needs an NVSwitch multicast mapping: {self.ar.multicast_note}")
                # the reduced activations are broadcast by the switch into every rank's copy of this buffer
                self.proj = self.ar.alloc((S, H), torch.bfloat16)
            # alternating GEMM-output buffers in the symmetric heap: peers read them in place, and the
            # start barrier of the next collective doubles as the end barrier of this one
            self.proj_in = [self.ar.alloc((S, H), torch.bfloat16), self.ar.alloc((S, H), torch.bfloat16)]
            if self.comm == "fused":
                self.x8_sym = self.ar.alloc((S, H), torch.float8_e4m3fn)
                self.final_sym = self.ar.alloc((S, H), torch.bfloat16)
                # MoE layers, second EP form (all-gather tokens -> local experts -> reduce-scatter,
                # prepare_finalize/naive_dp_ep.py:112-210 with is_sequence_parallel): the gathered, normalised
                # bf16 tokens every rank routes and feeds to its local experts
                self.xn_sym = self.ar.alloc((S, H), torch.bfloat16) if self.moe else None
                lo, hi = self.ar.rows_of_rank(S)
                self.row_lo, self.row_hi = lo, hi
                self.res_shard = torch.zeros((S + tp_size - 1) // tp_size, H, dtype=torch.bfloat16, device=self.dev)
        self.gu = None if self.moe else torch.empty(S, 2 * self.inter, dtype=torch.bfloat16, device=self.dev)
        if self.moe:
            E, k = s.experts, s.topk
            self.router_logits = torch.empty(S, E, dtype=torch.bfloat16, device=self.dev)
            self.topk_w = torch.empty(S, k, dtype=torch.float32, device=self.dev)
            self.topk_ids = torch.empty(S, k, dtype=torch.int32, device=self.dev)
            self.tok_exp_idx = torch.empty(S, k, dtype=torch.int32, device=self.dev)
            # global expert id -> local id or -1 (fused_moe/layer.py determine_expert_map)
            em = tp.expert_map(E, tp_rank, tp_size)
            self.expert_map = None if em is None else em.to(self.dev)
        # gate_up GEMM + silu_and_mul + FP8 quant in one kernel. Measured in round 1: at
        # batch 256 the GEMM is tensor-bound and the fused epilogue (silu on 4 warps of 112 CTAs, not
        # overlapped) costs more than the separate full-GPU activation kernel: 16.55 vs 15.95 ms/step.
        # It pays only while the GEMM is a weight stream (small batches).
        if quant == "nvfp4":
            assert comm != "fused", "NVFP4 linears: comm must be nccl or nvlink"
            self.xn = torch.empty(S, H, dtype=torch.bfloat16, device=self.dev)
            self.h4 = torch.empty(S, self.inter // 2, dtype=torch.uint8, device=self.dev)
            self.h_sf = ops.create_fp4_scale_tensor(S, self.inter, self.dev, True)
        self.fuse_gate_up = max_seqs <= 64 and self.inter % 128 == 0 and not self.moe
        self.fuse_attn_quant = os.environ.get("B200_FUSE_ATTN_QUANT", "1") == "1"
        self.routing_log = None   # list -> forward_device() appends every MoE layer's sorted top-k ids (eager only)
        self.attn_hook = None  # bench instrumentation: called around the attention launch
        self.graph = None

    def _init_nvfp4_layer(self, L, wqkv, wo, rnd, one):
        """NVFP4 (W4A4) linears, ModelOpt convention (quantization/modelopt.py ModelOptNvFp4LinearMethod):
        weight_global_scale = 448*6/amax(w), input_global_scale = 448*6/amax(x) (static), block scales e4m3 per
        16 elements in the swizzled layout, alpha = 1/(input_gs * weight_gs). Weights are quantised with this
        package's own scaled_fp4_quant (bit-identical to the reference's kernel, tests/test_gpu_vs_wheel.py)."""
        s, H = self.shape, self.shape.hidden
        cs = tp.column_shard(s.inter, self.tp_rank, self.tp_size)
        wg = rnd(s.inter, H, fan_in=H)[cs]
        wu = rnd(s.inter, H, fan_in=H)[cs]
        wd = rnd(H, s.inter, fan_in=s.inter)[:, tp.row_shard(s.inter, self.tp_rank, self.tp_size)]
        amax_x = {"qkv": 4.0, "o": 2.0, "gu": 4.0, "d": 8.0}
        for name, w in (("qkv", wqkv), ("o", wo), ("gu", torch.cat([wg, wu])), ("d", wd)):
            w = w.contiguous()
            w_gs = (448.0 * 6.0 / w.abs().amax().float().clamp(min=1e-12)).reshape(1)
            q, sf = ops.scaled_fp4_quant(w.to(torch.bfloat16), w_gs)
            a_gs = one(448.0 * 6.0 / amax_x[name])
            L["w4_" + name], L["wsf_" + name], L["wgs_" + name] = q, sf, w_gs
            L["ags_" + name] = a_gs
            L["alpha_" + name] = (1.0 / (a_gs * w_gs)).to(torch.float32)
        del wg, wu, wd

    def _fp4_linear(self, x, L, name, out):
        x4, x_sf = ops.scaled_fp4_quant(x, L["ags_" + name])
        ops.cutlass_scaled_fp4_mm(x4, L["w4_" + name], x_sf, L["wsf_" + name], L["alpha_" + name],
                                  torch.bfloat16, out=out)
        return 2

    def _forward_nvfp4(self) -> None:
        """The same layer loop with NVFP4 W4A4 linears (BASELINE configs[4]): norms emit bf16, every linear is
        scaled_fp4_quant -> cutlass_scaled_fp4_mm, the MLP activation is silu_and_mul_nvfp4_quant."""
        s, D, S = self.shape, self.D, self.max_seqs
        n = 0
        q = self.qkv[:, : self.hq * D].view(S, self.hq, D)
        k = self.qkv[:, self.hq * D:(self.hq + self.hkv) * D].view(S, self.hkv, D)
        v = self.qkv[:, (self.hq + self.hkv) * D:].view(S, self.hkv, D)
        pa, pb = self.proj_in
        red = self.proj
        torch.index_select(self.embed, 0, self.token_ids, out=self.residual)
        for li, L in enumerate(self.layers):
            if li == 0:
                ops.rms_norm(self.xn, self.residual, L["in_norm"], s.eps); n += 1
                xin = self.xn
            else:
                ops.fused_add_rms_norm(red, self.residual, L["in_norm"], s.eps); n += 1
                xin = red
            n += self._fp4_linear(xin, L, "qkv", self.qkv)
            kc, vc = L["kv"][..., :D], L["kv"][..., D:]
            ops.rope_and_reshape_and_cache_flash(self.positions, q, k, v, D, self.cos_sin, True, kc, vc,
                                                 self.slot_mapping, "fp8" if self.kv_fp8 else "auto",
                                                 L["k_scale"], L["v_scale"])
            n += 1
            if self.attn_hook:
                self.attn_hook(0)
            n += self.attention_layer(L)
            if self.attn_hook:
                self.attn_hook(1)
            n += self._fp4_linear(self.attn_out.view(S, self.hq * D), L, "o", pa)
            red = self._all_reduce(pa)
            n += 1 if self.ar is not None else 0
            ops.fused_add_rms_norm(red, self.residual, L["post_norm"], s.eps); n += 1
            n += self._fp4_linear(red, L, "gu", self.gu)
            ops.silu_and_mul_nvfp4_quant(self.h4, self.h_sf, self.gu, L["ags_d"]); n += 1
            ops.cutlass_scaled_fp4_mm(self.h4, L["w4_d"], self.h_sf.view(torch.float8_e4m3fn), L["wsf_d"],
                                      L["alpha_d"], torch.bfloat16, out=pb); n += 1
            red = self._all_reduce(pb)
            n += 1 if self.ar is not None else 0
        ops.fused_add_rms_norm(red, self.residual, self.final_norm, s.eps); n += 1
        logits = red @ self.lm_head.t()
        self.last_logits = logits
        ops.greedy_sample(logits, out=self.next_tokens); n += 1   # Sampler.greedy_sample
        self.launches_per_step = n

    def _all_reduce(self, t):
        """sum over the TP group; returns the tensor holding the result."""
        if self.tp_size == 1:
            return t
        if self.comm == "nccl":
            torch.distributed.all_reduce(t, group=self.tp_group)
            return t
        if self.comm == "nvls":
            return self.ar.all_reduce(t, out=self.proj, algo=3)
        return self.ar.all_reduce(t, out=self.proj, algo=1, skip_end_barrier=True)

    # ---- one decode step over the persistent device buffers
    def forward_device(self) -> None:
        """Launch one decode step (eagerly or under capture) at this model's programmatic-launch level."""
        with self.pdl_scope():
            self._forward_device()

    @contextlib.contextmanager
    def pdl_scope(self):
        """Launches inside the block use this model's programmatic-launch level; the library's level is restored on
        exit (also when a launch raises)."""
        prev = ops.set_pdl(self.pdl)
        try:
            yield
        finally:
            ops.set_pdl(prev)

    def _forward_device(self) -> None:
        if self.quant == "nvfp4":
            return self._forward_nvfp4()
        s, D, S = self.shape, self.D, self.max_seqs
        H = s.hidden
        n = 0
        x8 = self.x8
        a8 = self.a8
        q = self.qkv[:, : self.hq * D].view(S, self.hq, D)
        k = self.qkv[:, self.hq * D:(self.hq + self.hkv) * D].view(S, self.hkv, D)
        v = self.qkv[:, (self.hq + self.hkv) * D:].view(S, self.hkv, D)
        fused = self.comm == "fused"
        if fused:
            x8 = self.x8_sym
        pa, pb = self.proj_in
        red = self.proj
        torch.index_select(self.embed, 0, self.token_ids, out=self.residual)
        if fused:  # the residual stream lives row-sharded across the ranks from here on
            self.res_shard[: self.row_hi - self.row_lo].copy_(self.residual[self.row_lo:self.row_hi])
        for li, L in enumerate(self.layers):
            if li == 0:
                ops.rms_norm_static_fp8_quant(x8, self.residual, L["in_norm"], L["a_qkv"], s.eps)
                n += 1
            elif fused:
                self.ar.fused_add_rms_norm(x8, pb, self.res_shard, L["in_norm"], s.eps, L["a_qkv"])
                n += 1
            else:
                ops.fused_add_rms_norm_static_fp8_quant(x8, red, self.residual, L["in_norm"],
                                                        L["a_qkv"], s.eps)
                n += 1
            ops.cutlass_scaled_mm(x8, L["wqkv"].t(), L["a_qkv"], L["wqkv_s"], torch.bfloat16, out=self.qkv)
            n += LAUNCHES["gemm"]
            kc, vc = L["kv"][..., :D], L["kv"][..., D:]
            # RoPE + KV-cache write in one launch (do_rope_and_kv_cache_update, backend.py:950-1006)
            ops.rope_and_reshape_and_cache_flash(self.positions, q, k, v, D, self.cos_sin, True, kc, vc,
                                                 self.slot_mapping, "fp8" if self.kv_fp8 else "auto",
                                                 L["k_scale"], L["v_scale"])
            n += 1
            if self.attn_hook:
                self.attn_hook(0)
            n += self.attention_layer(L)
            if self.attn_hook:
                self.attn_hook(1)
            if not self.fuse_attn_quant:
                ops.static_scaled_fp8_quant(a8, self.attn_out.view(S, self.hq * D), L["a_o"]); n += 1
            ops.cutlass_scaled_mm(a8, L["wo"].t(), L["a_o"], L["wo_s"], torch.bfloat16, out=pa); n += 1
            if self.moe:
                # MixtralMoE.forward (models/mixtral.py:145-160): router logits (bf16 replicated linear: library
                # GEMM) -> topk_softmax (renormalised) -> fused_experts on the local experts -> all-reduce
                if fused:
                    # second EP form: ONE kernel reduce-scatters the o_proj partials, adds the row-sharded residual,
                    # normalises and all-gathers the bf16 tokens (dispatch); the experts' partial sums are
                    # reduce-scattered by the next layer's fused kernel (combine): no all-reduce of [S, H]
                    self.ar.fused_add_rms_norm(self.xn_sym, pa, self.res_shard, L["post_norm"], s.eps, None); n += 1
                    red = self.xn_sym
                else:
                    red = self._all_reduce(pa)
                    n += 1 if self.ar is not None else 0
                    ops.fused_add_rms_norm(red, self.residual, L["post_norm"], s.eps); n += 1
                torch.matmul(red, L["gate_w_t"], out=self.router_logits)
                fused_moe.topk_softmax(self.topk_w, self.topk_ids, self.tok_exp_idx, self.router_logits, True)
                n += LAUNCHES["topk_softmax"]
                if self.routing_log is not None:   # bench.py tp_check: which rows were routed differently from TP=1
                    self.routing_log.append(self.topk_ids.sort(dim=1).values.cpu())
                mo = fused_moe.fused_experts(red, L["w1"], L["w2"], self.topk_w, self.topk_ids, use_fp8_w8a8=True,
                                             global_num_experts=s.experts, expert_map=self.expert_map,
                                             w1_scale=L["w1_s"], w2_scale=L["w2_s"], a1_scale=L["a_gu"],
                                             a2_scale=L["a_d"])
                n += LAUNCHES["fused_experts"]
                pb.copy_(mo)
                if not fused:
                    red = self._all_reduce(pb)
                    n += 1 if self.ar is not None else 0
                continue
            if fused:
                self.ar.fused_add_rms_norm(x8, pa, self.res_shard, L["post_norm"], s.eps, L["a_gu"]); n += 1
            else:
                red = self._all_reduce(pa)
                n += 1 if self.ar is not None else 0
                ops.fused_add_rms_norm_static_fp8_quant(x8, red, self.residual, L["post_norm"],
                                                        L["a_gu"], s.eps); n += 1
            if self.fuse_gate_up:
                ops.scaled_mm_silu_and_mul_quant(self.h8, x8, L["wgu"], L["a_gu"], L["wgu_s"], L["a_d"],
                                                 torch.bfloat16); n += 1
            else:
                ops.cutlass_scaled_mm(x8, L["wgu"].t(), L["a_gu"], L["wgu_s"], torch.bfloat16, out=self.gu); n += 1
                ops.silu_and_mul_quant(self.h8, self.gu, L["a_d"]); n += LAUNCHES["silu_quant"]
            ops.cutlass_scaled_mm(self.h8, L["wd"].t(), L["a_d"], L["wd_s"], torch.bfloat16, out=pb); n += 1
            if not fused:
                red = self._all_reduce(pb)
                n += 1 if self.ar is not None else 0
        if fused:
            self.ar.fused_add_rms_norm(self.final_sym, pb, self.res_shard, self.final_norm, s.eps, None); n += 1
            red = self.final_sym
        else:
            ops.fused_add_rms_norm(red, self.residual, self.final_norm, s.eps); n += 1
        logits = red @ self.lm_head.t()
        self.last_logits = logits
        ops.greedy_sample(logits, out=self.next_tokens); n += 1   # Sampler.greedy_sample
        self.launches_per_step = n

    def attention_layer(self, L) -> int:
        """The paged-attention decode call of one layer over the persistent buffers (q from self.qkv, the
        layer's KV cache); forward_device() and bench.py's roofline timing share it. Returns launches."""
        S, D = self.max_seqs, self.D
        q = self.qkv[:, : self.hq * D].view(S, self.hq, D)
        kc, vc = L["kv"][..., :D], L["kv"][..., D:]
        fp8_out = self.quant == "fp8" and self.fuse_attn_quant
        # fp8_out: the attention kernel writes the statically quantised o_proj input itself (fused FP8 output)
        ops.paged_attention_decode(self.a8.view(S, self.hq, D) if fp8_out else self.attn_out, q, kc, vc,
                                   self.block_table, self.seq_lens, D ** -0.5,
                                   k_scale=L["k_scale"] if self.kv_fp8 else None,
                                   v_scale=L["v_scale"] if self.kv_fp8 else None,
                                   max_seq_len=self.max_ctx, workspace=self.workspace,
                                   output_scale=L["a_o"] if fp8_out else None)
        return LAUNCHES["attention"]

    def capture(self) -> None:
        """Capture forward_device() in a CUDA graph (decode steps are launch-bound otherwise)."""
        side = torch.cuda.Stream(device=self.dev)
        side.wait_stream(torch.cuda.current_stream(self.dev))
        with torch.cuda.stream(side):
            for _ in range(2):
                self.forward_device()
        torch.cuda.current_stream(self.dev).wait_stream(side)
        torch.cuda.synchronize(self.dev)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.forward_device()

    def run_device(self) -> None:
        if self.graph is not None:
            self.graph.replay()
        else:
            self.forward_device()

    # ---- public step: host (pinned) inputs in, host tokens out
    def step(self, host_in: dict, host_out: torch.Tensor) -> torch.Tensor:
        """host_in: pinned CPU tensors token_ids, positions, slot_mapping [S] int64, seq_lens [S] int32,
        block_table [S, blocks_per_seq] int32. host_out: pinned int64 [S]. Returns host_out after the
        stream has been synchronised (the sampled token ids are needed by the caller's scheduler)."""
        self.token_ids.copy_(host_in["token_ids"], non_blocking=True)
        self.positions.copy_(host_in["positions"], non_blocking=True)
        self.slot_mapping.copy_(host_in["slot_mapping"], non_blocking=True)
        self.seq_lens.copy_(host_in["seq_lens"], non_blocking=True)
        self.block_table.copy_(host_in["block_table"], non_blocking=True)
        self.run_device()
        host_out.copy_(self.next_tokens, non_blocking=True)
        torch.cuda.current_stream(self.dev).synchronize()
        return host_out

    def h2d_bytes(self) -> int:
        S = self.max_seqs
        return S * 8 * 3 + S * 4 + S * self.blocks_per_seq * 4

    def d2h_bytes(self) -> int:
        return self.max_seqs * 8
