#!/usr/bin/env python
"""Headline benchmark: decode tokens/s of a Llama-3.1-8B-shaped FP8 W8A8 + FP8-KV model,
batch 256, 4k context (BASELINE.json configs[1]) on N B200s (tensor parallel), plus the roofline of
the dominant kernel (paged-attention decode, HBM-bound) and the reference's CPU attention timed on
the host cores.

    python bench.py --gpus N --steps K --warmup W            # our arm (one rank per GPU via torchrun for N>1)
    python bench.py --impl reference --gpus N --steps K ...  # the reference's own CPU implementation

One JSON line on stdout (rank 0). See DESIGN.md section 6 for every field.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "vllm-fork_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402

METRIC = "decode_tokens_per_s"
UNIT = "tokens/s"
BATCH, CTX, BLOCK = 256, 4096, 16


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=BATCH)
    ap.add_argument("--ctx", type=int, default=CTX)
    ap.add_argument("--model", default="llama-3.1-8b")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--profile-step", action="store_true",
                    help="bracket ONE eager decode step with cudaProfilerStart/Stop and exit "
                         "(for `ncu --profile-from-start off`; prints no bench line)")
    ap.add_argument("--comm", default="auto", choices=["auto", "nccl", "nvlink", "fused"],
                    help="TP collective (N>1): NCCL all-reduce, own NVLink all-reduce, or the fused "
                         "reduce-scatter+add+RMSNorm+FP8-quant+all-gather kernel")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--quant", default="fp8", choices=["fp8", "nvfp4"],
                    help="linear layers: FP8 W8A8 (configs[1]) or NVFP4 W4A4 (configs[4]); the KV cache stays FP8")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU every 100 ms during the timed region."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_evt = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.1)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


def model_shape(name):
    from b200ops import decoder
    return {"llama-3.1-8b": decoder.LLAMA_3_1_8B, "tiny": decoder.TINY,
            "qwen2.5-0.5b": decoder.QWEN2_5_0_5B, "llama-3.1-70b": decoder.LLAMA_3_1_70B,
            "mixtral-8x7b": decoder.MIXTRAL_8X7B, "tiny-moe": decoder.TINY_MOE}[name]


# ------------------------------------------------------------------------------------------------
# CPU baseline / reference arm: the reference's own CPU paged attention (oracle/_ref/refcpu.so, built
# unmodified from /root/reference/csrc/cpu) + torch CPU bf16 linears for the GEMMs (the reference's
# CPU GEMMs need oneDNN, not buildable offline -- BASELINE.md section 4.5).
# ------------------------------------------------------------------------------------------------
class CpuReferenceDecoder:
    def __init__(self, shape, sample_seqs: int, ctx: int, sample_layers: int):
        from oracle import refcpu
        self.refcpu = refcpu
        refcpu.load()
        self.shape, self.S, self.ctx, self.layers = shape, sample_seqs, ctx, sample_layers
        s, D, bs, S = shape, shape.head_size, BLOCK, sample_seqs
        torch.manual_seed(0)
        nblk = (ctx + bs - 1) // bs
        nb = S * nblk + 1
        self.isa = refcpu.pick_isa(bs, D)
        self.kv = []
        for _ in range(sample_layers):
            kc, vc = refcpu.make_packed_cache(nb, s.kv_heads, bs, D, torch.bfloat16)
            kc.uniform_(-0.1, 0.1)
            vc.uniform_(-0.1, 0.1)
            self.kv.append((kc, vc))
        self.bt = torch.randperm(nb)[: S * nblk].to(torch.int32).view(S, nblk)
        self.seq_lens = torch.full((S,), ctx, dtype=torch.int32)
        self.qsl = torch.arange(S + 1, dtype=torch.int32)
        self.slots = (self.bt[:, (ctx - 1) // bs].long() * bs + (ctx - 1) % bs)
        self.md = refcpu.scheduler_metadata(S, s.heads, s.kv_heads, D, self.seq_lens, torch.bfloat16,
                                            self.qsl, self.isa, enable_kv_split=True)
        H = s.hidden
        self.wqkv = torch.randn((s.heads + 2 * s.kv_heads) * D, H).mul(0.02).to(torch.bfloat16)
        self.wo = torch.randn(H, s.heads * D).mul(0.02).to(torch.bfloat16)
        self.wgu = torch.randn(2 * s.inter, H).mul(0.02).to(torch.bfloat16)
        self.wd = torch.randn(H, s.inter).mul(0.02).to(torch.bfloat16)
        self.x = torch.randn(S, H).to(torch.bfloat16)

    def step(self):
        """one decode step over `sample_layers` layers of the sample batch; returns seconds"""
        s, D, S = self.shape, self.shape.head_size, self.S
        F = torch.nn.functional
        t0 = time.perf_counter()
        x = self.x
        for kc, vc in self.kv:
            qkv = F.linear(x, self.wqkv)
            q = qkv[:, : s.heads * D].reshape(S, s.heads, D)
            k = qkv[:, s.heads * D:(s.heads + s.kv_heads) * D].reshape(S, s.kv_heads, D)
            v = qkv[:, (s.heads + s.kv_heads) * D:].reshape(S, s.kv_heads, D)
            self.refcpu.reshape_and_cache(k.contiguous(), v.contiguous(), kc, vc, self.slots, self.isa)
            out = torch.empty_like(q)
            self.refcpu.attention(q.contiguous(), kc, vc, out, self.qsl, self.seq_lens, D ** -0.5,
                                  self.bt, self.md)
            x = x + F.linear(out.reshape(S, -1), self.wo)
            gu = F.linear(x, self.wgu)
            h = F.silu(gu[:, : s.inter]) * gu[:, s.inter:]
            x = x + F.linear(h, self.wd)
        return time.perf_counter() - t0


def cpu_reference_run(shape, steps, warmup, ctx):
    """Times the reference CPU path on a bounded sample and extrapolates to the full layer count."""
    sample_seqs, sample_layers = 8, 2
    dec = CpuReferenceDecoder(shape, sample_seqs, ctx, sample_layers)
    for _ in range(warmup):
        dec.step()
    ts = [dec.step() for _ in range(steps)]
    t_layer = (sum(ts) / len(ts)) / sample_layers
    t_step = t_layer * shape.layers  # embedding / lm_head excluded (in favour of the CPU)
    cores = torch.get_num_threads()
    return {"value": sample_seqs / t_step, "unit": UNIT, "cores": cores, "kind": "reference",
            "sample": f"{sample_seqs} of {BATCH} sequences x {ctx}-token context, {sample_layers} of "
                      f"{shape.layers} layers timed (reference csrc/cpu attention ISA {dec.isa}, bf16 KV, "
                      f"+ torch CPU bf16 linears), extrapolated to {shape.layers} layers; "
                      f"{steps} steps after {warmup} warm-up",
            "ms_per_step": t_step * 1e3}


def reference_arm(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # rank 0 alone runs the CPU reference
    shape = model_shape(a.model)
    from oracle import refcpu
    if not refcpu.available():
        print(json.dumps({"impl": "reference", "unavailable":
                          "oracle/_ref/refcpu.so not built (needs /root/reference at build time)"}))
        return
    cb = cpu_reference_run(shape, max(1, a.steps), max(1, a.warmup), a.ctx)
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT,
            "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": cb["ms_per_step"],
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic",
            "config": {"workload": f"{shape.name} decode, batch {a.batch}, ctx {a.ctx}, block {BLOCK} "
                                   "(BASELINE.json configs[1]); CPU sample, see cpu_baseline.sample"},
            "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def main():
    a = parse()
    if a.impl == "reference":
        return reference_arm(a)
    from b200ops import decoder
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    group = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
        group = dist.group.WORLD
    if a.comm == "auto":
        # measured (gpurun_out/bench_tp{2,4,8}_*.json): the one-shot all-reduce wins at 2 ranks (8.8 us per
        # 2 MiB), the fused reduce-scatter + norm + quant + all-gather kernel from 4 ranks up (19 us at 8
        # ranks where a one-shot pulls 14 MiB per rank and takes 32 us)
        a.comm = "nvlink" if world <= 2 else "fused"
    shape = model_shape(a.model)
    if (shape.experts > 0 or a.quant == "nvfp4") and a.comm == "fused":
        a.comm = "nvlink"   # MoE layers (expert parallel): the all-reduce follows fused_experts, no norm fusion
    S, ctx = a.batch, a.ctx
    model = decoder.DecoderModel(shape, S, ctx, BLOCK, dev, tp_rank=rank, tp_size=world, tp_group=group,
                                 comm=a.comm, quant=a.quant)
    D = shape.head_size
    # ---- synthetic state: KV caches full of random (non-NaN) fp8 codes; block table = random
    # permutation without repetition
    g = torch.Generator(device=dev).manual_seed(1)
    for L in model.layers:
        kv = L["kv"]
        kv.view(-1).copy_(torch.randint(0, 0x70, (kv.numel(),), dtype=torch.uint8, device=dev, generator=g))
        kv.view(-1).bitwise_or_(torch.randint(0, 2, (kv.numel(),), dtype=torch.uint8, device=dev, generator=g) << 7)
    nblk = model.blocks_per_seq
    cg = torch.Generator().manual_seed(2)
    bt = torch.randperm(model.num_blocks, generator=cg)[: S * nblk].to(torch.int32).view(S, nblk)
    pin = lambda t: t.pin_memory()
    host = []
    for i in range(4):  # a few distinct host input sets, cycled
        host.append({
            "token_ids": pin(torch.randint(0, shape.vocab, (S,), generator=cg)),
            "positions": pin(torch.full((S,), ctx - 1, dtype=torch.int64)),
            "slot_mapping": pin(bt[:, (ctx - 1) // BLOCK].long() * BLOCK + (ctx - 1) % BLOCK),
            "seq_lens": pin(torch.full((S,), ctx, dtype=torch.int32)),
            "block_table": pin(bt.clone()),
        })
    host_out = pin(torch.zeros(S, dtype=torch.int64))
    model.step(host[0], host_out)  # eager once (sets launches_per_step, loads the inputs)
    launches = model.launches_per_step
    if a.profile_step:
        model.step(host[1], host_out)
        torch.cuda.synchronize(dev)
        torch.cuda.profiler.start()
        model.step(host[2], host_out)
        torch.cuda.synchronize(dev)
        torch.cuda.profiler.stop()
        return
    if not a.no_graph:
        model.capture()

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize(dev)

    def timed(fn, steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for i in range(steps):
            fn(i)
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    # ---- (1) device-resident throughput: inputs already in HBM, K graph replays back to back
    for i in range(a.warmup):
        model.run_device()
    sampler = ClockSampler(local)
    sampler.start()
    ms_dev = timed(lambda i: model.run_device(), a.steps)
    # ---- (2) end to end through the public API: pinned host inputs -> H2D -> step -> D2H tokens
    for i in range(a.warmup):
        model.step(host[i % len(host)], host_out)
    ms_e2e = timed(lambda i: model.step(host[i % len(host)], host_out), a.steps)
    clocks = sampler.stop()

    # ---- (3) roofline of the dominant kernel: CUDA events around every paged-attention launch of an
    # eager (non-graph) pass over the same steps, on the launching stream
    evs = []

    def hook(which):
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        evs.append(e)

    model.attn_hook = hook
    saved_graph, model.graph = model.graph, None
    barrier()
    for i in range(min(a.steps, 5)):
        model.run_device()
    torch.cuda.synchronize(dev)
    model.attn_hook, model.graph = None, saved_graph
    durs = [evs[i].elapsed_time(evs[i + 1]) for i in range(0, len(evs), 2)]
    attn_ms = sum(durs) / len(durs)
    es = 1  # fp8 KV
    alg_bytes = (S * 2 * ctx * model.hkv * D * es + S * model.hq * D * (2 + 2) + S * nblk * 4 + 4 * S)
    peak, peak_src = peaks()
    achieved = alg_bytes / (attn_ms * 1e-3) / 1e9
    # dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of this kernel on this exact workload, from
    # the ncu --set full capture summarised in profiles/r1b_ncu_full.md (a number taken under the profiler is
    # only used for the byte count, never for time)
    traffic, traffic_src = None, None
    if world == 1 and S == 256 and ctx == 4096 and shape.name.startswith("Llama-3.1-8B"):
        traffic, traffic_src = 2149912000 + 5132544, "profiles/r1b_ncu_full.md (ncu, per launch)"
    step_ms = ms_dev / a.steps
    value = S / (step_ms * 1e-3)
    e2e_value = S / (ms_e2e / a.steps * 1e-3)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": step_ms, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None,
        "dtype": ("nvfp4 (W4A4 block-scaled GEMMs, fp32 accumulate; FP8 KV; bf16 activations)" if a.quant == "nvfp4"
                  else "fp8_e4m3 (W8A8 GEMMs, fp32 accumulate; FP8 KV; bf16 activations)"),
        "data": "synthetic",
        "config": {"workload": f"{shape.name} decode step, " + ("NVFP4 W4A4" if a.quant == "nvfp4" else "FP8 W8A8") +
                               f" + FP8 KV cache, batch {S}, "
                               f"context {ctx}, block_size {BLOCK} " +
                               ("(BASELINE.json configs[3]: FP8 fused_moe, %d experts top-%d, expert parallel; "
                                "context length not given there)" % (shape.experts, shape.topk)
                                if shape.experts > 0 else
                                "(BASELINE.json configs[4])" if a.quant == "nvfp4" else "(BASELINE.json configs[1])"),
                   "parallelism": (f"tp{world} attention + ep{world} experts" if shape.experts > 0 else f"tp{world}"), "tp_collective": (a.comm if world > 1 else None), "weights": "random init (dummy), per-tensor FP8 scales",
                   "cuda_graph": not a.no_graph,
                   "l2": "inputs larger than L2 (KV cache %.1f GB + weights per GPU)" %
                         (sum(L["kv"].numel() for L in model.layers) / 1e9),
                   "context": "seq_len fixed at ctx every step (the new token is written at position ctx-1)"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": model.h2d_bytes(),
                "d2h_bytes_per_step": model.d2h_bytes(), "ms_per_step": ms_e2e / a.steps},
        "gpu_launches": launches * a.steps,
        "gpu_launches_per_step": launches,
        "clocks": clocks,
        "roofline": {"kernel": "paged_attn_decode_kernel<128,1,BF16_FP8>", "bound": "hbm",
                     "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "peak_source": peak_src, "traffic": traffic, "traffic_source": traffic_src,
                     "algorithmic_bytes_per_launch": alg_bytes, "ms_per_launch": attn_ms,
                     "launches_per_step": shape.layers,
                     "share_of_step": attn_ms * shape.layers / step_ms},
    }
    if rank == 0 and world == 1 and not a.no_cpu_baseline and shape.experts == 0:
        try:
            from oracle import refcpu
            if refcpu.available():
                cb = cpu_reference_run(shape, 3, 1, ctx)
                line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
            else:
                line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference",
                                        "sample": "oracle/_ref/refcpu.so not built"}
        except Exception as e:  # the baseline must never take the GPU number down with it
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference",
                                    "sample": f"failed: {e}"}
    if rank == 0:
        print(json.dumps(line))
    sys.stdout.flush()
    if world > 1:
        # tear down in a fixed order and leave without running NCCL's destructors: destroying a process
        # group whose collectives were captured in a live CUDA graph can hang at interpreter exit
        model.graph = None
        torch.cuda.synchronize(dev)
        torch.distributed.barrier()
        if model.ar is not None:
            model.ar.close()
        sys.stderr.flush()
        os._exit(0)


if __name__ == "__main__":
    main()
