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

if "reference" in sys.argv and any(a == "--impl" or a.startswith("--impl=") for a in sys.argv):
    # torch.distributed.run exports OMP_NUM_THREADS=1 to every rank; the reference arm is a CPU run that
    # must use all the host threads it can, and the OpenMP runtimes read the variable when they load
    for _v in ("OMP_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ.pop(_v, None)

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
    ap.add_argument("--comm", default="auto", choices=["auto", "nccl", "nvlink", "nvls", "fused"],
                    help="TP collective (N>1): NCCL all-reduce, own NVLink peer-pull all-reduce, own NVSwitch multimem "
                         "all-reduce, or the fused reduce-scatter+add+RMSNorm+FP8-quant+all-gather kernel (in-switch "
                         "reduction when a multicast mapping exists)")
    ap.add_argument("--no-multicast", action="store_true", help="never use the NVSwitch multicast mapping")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-tp-check", action="store_true",
                    help="skip the multi-GPU correctness self-check that runs before the timed region at N>1")
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


def measured_traffic(key):
    """(bytes per launch, source) of the dominant kernel from profiles/traffic.json, or (None, None)."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        e = json.load(open(p))[key]
        return int(e["dram_bytes_read"]) + int(e["dram_bytes_write"]), e["source"]
    except Exception:
        return None, None


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


def host_threads() -> int:
    """Threads the CPU arm uses: the physical cores this process may run on (what torch picks by itself when
    OMP_NUM_THREADS is unset; set explicitly so a launcher's OMP_NUM_THREADS=1 cannot shrink it)."""
    try:
        avail = len(os.sched_getaffinity(0))
    except AttributeError:
        avail = os.cpu_count() or 1
    try:
        import psutil
        phys = psutil.cpu_count(logical=False) or avail
    except Exception:
        phys = avail
    return max(1, min(avail, phys))


def cpu_reference_run(shape, steps, warmup, ctx, sample_seqs=32, sample_layers=4, batch=BATCH):
    """Times the reference CPU path on a bounded sample and extrapolates to the full layer count."""
    cores = host_threads()
    torch.set_num_threads(cores)   # ATen intra-op pool + omp_set_num_threads of the shared OpenMP runtime
    sample_seqs = min(sample_seqs, batch)
    sample_layers = min(sample_layers, shape.layers)
    if sample_seqs * ctx * shape.layers <= 8 * 512 * 64:
        sample_layers = shape.layers      # small workloads (BASELINE configs[0]) are timed whole: no extrapolation
    dec = CpuReferenceDecoder(shape, sample_seqs, ctx, sample_layers)
    for _ in range(warmup):
        dec.step()
    ts = [dec.step() for _ in range(steps)]
    t_sample = sum(ts) / len(ts)           # one timed sample step (what the wall clock saw)
    t_layer = t_sample / sample_layers
    t_step = t_layer * shape.layers  # embedding / lm_head excluded (in favour of the CPU)
    cores = torch.get_num_threads()
    return {"value": sample_seqs / t_step, "unit": UNIT, "cores": cores, "kind": "reference",
            "timed_ms_per_sample_step": t_sample * 1e3, "extrapolated_ms_per_step": t_step * 1e3,
            "sample": f"{sample_seqs} of {batch} sequences x {ctx}-token context, {sample_layers} of "
                      f"{shape.layers} layers timed (reference csrc/cpu attention ISA {dec.isa}, bf16 KV, "
                      f"+ torch CPU bf16 linears), extrapolated to {shape.layers} layers; "
                      f"{steps} steps after {warmup} warm-up",
            "ms_per_step": t_step * 1e3}


def baseline_config_tag(shape, quant):
    if shape.experts > 0:
        return "BASELINE.json configs[3]"
    if quant == "nvfp4":
        return "BASELINE.json configs[4]"
    return {"Llama-3.1-8B-shaped": "BASELINE.json configs[1]", "Llama-3.1-70B-shaped": "BASELINE.json configs[2]",
            "Qwen2.5-0.5B-shaped": "BASELINE.json configs[0]"}.get(shape.name, "not a BASELINE.json config")


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
    cb = cpu_reference_run(shape, max(1, a.steps), max(1, a.warmup), a.ctx, batch=a.batch)
    # ms_per_step = what was actually timed (one sample step); the full-workload figure the value is derived
    # from is reported separately, so steps x ms_per_step is the wall time of the timed region
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT,
            "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": cb["timed_ms_per_sample_step"],
            "extrapolated_ms_per_step": cb["extrapolated_ms_per_step"],
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic",
            "config": {"workload": f"{shape.name} decode, batch {a.batch}, ctx {a.ctx}, block {BLOCK} "
                                   f"({baseline_config_tag(shape, a.quant)}); CPU sample, see cpu_baseline.sample"},
            "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def tp_self_check(rank, world, dev, group):
    """Correctness of the multi-GPU data plane, run before the timed region at world > 1 (the driver's GPU test
    box has one GPU, so this is where the NVLink collectives are verified under the driver's eyes):
    (1) B200AllReduce one-shot / two-shot / auto, zero-copy (symmetric heap) and staged inputs, launched while a
        GEMM on a side stream keeps the SMs busy (CTAs of the collective start at different times), against the
        exact definition: fp32 sum in rank order of the NCCL-all-gathered inputs, rounded once -> bit-exact;
    (2) the fused reduce-scatter + add + RMSNorm + FP8-quant + all-gather kernel against
        torch.distributed.all_gather -> rank-order sum -> ops.fused_add_rms_norm_static_fp8_quant -> bit-exact;
    (3) two decode steps of a tiny Llama-shaped model at TP=world (every comm rung this world size supports)
        against the same model at TP=1 on rank 0 (every rank draws the full weights from one generator):
        greedy tokens must match, except where the TP=1 logits of the two candidates are within 4 bf16 ulps.
    Raises on mismatch; returns the dict printed as "tp_check"."""
    import torch.distributed as dist
    from b200ops import decoder, ops
    from b200ops.all_reduce import B200AllReduce
    res = {}
    ar = B200AllReduce(group, dev, max_size=8 << 20, heap_size=32 << 20)
    busy_a = torch.randn(4096, 4096, device=dev, dtype=torch.bfloat16)
    side = torch.cuda.Stream(device=dev)

    def keep_sms_busy():
        with torch.cuda.stream(side):
            for _ in range(2):
                torch.matmul(busy_a, busy_a)

    def rank_input(numel, dtype, seed):
        g = torch.Generator(device=dev).manual_seed(seed * 64 + rank)
        return (torch.rand(numel, device=dev, generator=g) * 2 - 1).to(dtype)

    def exact_sum(x):
        parts = [torch.empty_like(x) for _ in range(world)]
        dist.all_gather(parts, x, group=group)
        acc = torch.zeros(x.numel(), dtype=torch.float32, device=dev)
        for p in parts:                      # rank order 0..N-1, the kernels' order
            acc = acc + p.float().view(-1)
        return acc.to(x.dtype).view(x.shape)

    seed, cases = 0, 0
    for dtype in (torch.bfloat16, torch.float16, torch.float32):
        for numel in (4096, 256 * 4096, 1024 * 1024 + 8):
            for algo in (1, 2, 0):
                for zero_copy in (False, True):
                    seed += 1
                    x = rank_input(numel, dtype, seed)
                    used = ar._heap_used
                    if zero_copy:
                        buf = ar.alloc(numel, dtype)
                        buf.copy_(x)
                        x = buf
                    exp = exact_sum(x)
                    torch.cuda.synchronize(dev)
                    keep_sms_busy()
                    out = ar.all_reduce(x, algo=algo)
                    torch.cuda.synchronize(dev)
                    if not torch.equal(out, exp):
                        raise AssertionError(f"tp_check: all_reduce mismatch {dtype} n={numel} algo={algo} "
                                             f"zero_copy={zero_copy} rank={rank}")
                    ar._heap_used = used
                    cases += 1
    res["allreduce"] = "bit-exact"
    res["allreduce_cases"] = cases
    res["multicast"] = ar.multicast_note
    if ar.multicast:
        # (1b) NVSwitch in-network reduction (multimem.ld_reduce / multimem.st): the switch's accumulation order is
        # not the rank order, so equality is within one rounding of the output dtype; with the output in the
        # symmetric heap the switch broadcasts ONE result: bit-identical on every rank
        cases = 0
        for dtype, tol in ((torch.bfloat16, 2 ** -7), (torch.float16, 2 ** -10), (torch.float32, 2 ** -20)):
            for numel in (4096, 256 * 4096, 1024 * 1024 + 8):
                for zero_copy in (False, True):
                    for out_in_heap in (False, True):
                        seed += 1
                        used = ar._heap_used
                        x = rank_input(numel, dtype, seed)
                        if zero_copy:
                            buf = ar.alloc(numel, dtype)
                            buf.copy_(x)
                            x = buf
                        out = ar.alloc(numel, dtype) if out_in_heap else torch.empty(numel, dtype=dtype, device=dev)
                        exp = exact_sum(x)
                        torch.cuda.synchronize(dev)
                        dist.barrier(group=group)
                        keep_sms_busy()
                        ar.all_reduce(x, out=out, algo=3)
                        torch.cuda.synchronize(dev)
                        if not torch.allclose(out.float(), exp.float(), rtol=tol, atol=tol):
                            raise AssertionError(f"tp_check: multimem all_reduce mismatch {dtype} n={numel} "
                                                 f"zero_copy={zero_copy} out_in_heap={out_in_heap} rank={rank}")
                        if out_in_heap:
                            mine = out.view(torch.uint8).to(torch.int64).sum().reshape(1)
                            alls = [torch.empty_like(mine) for _ in range(world)]
                            dist.all_gather(alls, mine, group=group)
                            if any(int(a_) != int(mine) for a_ in alls):
                                raise AssertionError("tp_check: multimem all_reduce result differs between ranks")
                        dist.barrier(group=group)
                        ar._heap_used = used
                        cases += 1
        res["multimem_allreduce"] = "within one output rounding of the exact sum; identical on all ranks"
        res["multimem_cases"] = cases
    if world in (2, 4, 8):
        cases = 0
        for dtype in (torch.bfloat16, torch.float16):
            for (M, H) in ((256, 4096), (128, 8192), (33, 2048), (1, 4096)):
                for quant in (True, False):
                    for staged in (True, False):
                        seed += 1
                        used = ar._heap_used
                        x = rank_input(M * H, dtype, seed).view(M, H)
                        if not staged:
                            buf = ar.alloc((M, H), dtype)
                            buf.copy_(x)
                            x = buf
                        g = torch.Generator(device=dev).manual_seed(seed)    # same on every rank
                        res_full = (torch.rand(M, H, device=dev, generator=g) * 2 - 1).to(dtype)
                        w = (torch.rand(H, device=dev, generator=g) + 0.5).to(dtype)
                        scale = torch.tensor([0.02], device=dev) if quant else None
                        summed = exact_sum(x)
                        res_ref = res_full.clone()
                        if quant:
                            out_ref = torch.empty(M, H, dtype=torch.float8_e4m3fn, device=dev)
                            ops.fused_add_rms_norm_static_fp8_quant(out_ref, summed, res_ref, w, scale, 1e-5)
                        else:
                            ops.fused_add_rms_norm(summed, res_ref, w, 1e-5)
                            out_ref = summed
                        lo, hi = ar.rows_of_rank(M)
                        shard = torch.zeros((M + world - 1) // world, H, dtype=dtype, device=dev)
                        shard[: hi - lo] = res_full[lo:hi]
                        out = ar.alloc((M, H), torch.float8_e4m3fn if quant else dtype)
                        out.view(torch.uint8).fill_(0x55)
                        torch.cuda.synchronize(dev)
                        dist.barrier(group=group)
                        keep_sms_busy()
                        ar.fused_add_rms_norm(out, x, shard, w, 1e-5, scale)
                        torch.cuda.synchronize(dev)
                        dist.barrier(group=group)
                        if ar.multicast:
                            # in-switch sum (measured on this box even at 2 ranks: ~14 % of the bf16 sums differ from
                            # RN(fp32(a) + fp32(b)) by one ulp - the NVSwitch ALU does not round like the SM): the
                            # residual stream agrees within one ulp of T, every FP8 code within one e4m3 step
                            o1 = out.view(torch.float8_e4m3fn).float() if quant else out.float()
                            o2 = out_ref.float()
                            # one ulp of the SUM (|sum| < 4: 2^-6) survives the residual add as an absolute error, is
                            # scaled by rsqrt(mean z^2) * w (<= ~3 here) and by 1 / scale (50) when quantising
                            # (|sum| grows with the number of ranks and its ulp with it: the absolute terms, measured at
                            # 2 ranks, scale by world / 2)
                            kw = max(1.0, world / 2.0)
                            bad = ((o1 - o2).abs() > (0.13 * o2.abs() + 2.5 * kw if quant else 2 ** -6 * o2.abs() + 0.05 * kw))
                            ok = not bool(bad.any()) and torch.allclose(
                                shard[: hi - lo].float(), res_ref[lo:hi].float(), rtol=2 ** -7, atol=2 ** -7 * kw)
                        else:
                            ok = torch.equal(out.view(torch.uint8), out_ref.view(torch.uint8)) and \
                                torch.equal(shard[: hi - lo], res_ref[lo:hi])
                        if not ok:
                            o1 = out.view(torch.float8_e4m3fn).float() if quant else out.float()
                            nbad = int((out.view(torch.uint8) != out_ref.view(torch.uint8)).sum())
                            rbad = int((shard[: hi - lo] != res_ref[lo:hi]).sum())
                            rows_bad = (out.view(torch.uint8) != out_ref.view(torch.uint8)).any(dim=1).nonzero().flatten()[:8].tolist()
                            raise AssertionError(f"tp_check: fused all-reduce+norm mismatch {dtype} {(M, H)} "
                                                 f"quant={quant} staged={staged} rank={rank} multicast={ar.multicast}: "
                                                 f"{nbad} of {out.numel() * out.element_size()} output bytes differ (rows {rows_bad}), "
                                                 f"{rbad} residual elements differ, max |diff| "
                                                 f"{float((o1 - out_ref.float()).abs().max()):.4g}")
                        ar._heap_used = used
                        cases += 1
        res["fused_allreduce_rmsnorm"] = ("in-switch reduction: residual within one ulp, FP8 codes within one step of the unfused pair"
                                          if ar.multicast else "bit-exact")
        res["fused_cases"] = cases
    # ---- (2b) decode-context-parallel merge (one NVLink kernel) vs all_gather + the definition in torch fp32
    cases = 0
    for dtype, tol in ((torch.bfloat16, 2 ** -7), (torch.float16, 2 ** -10)):
        for (B, H, D) in ((256, 32, 128), (7, 8, 64)):
            seed += 1
            g = torch.Generator(device=dev).manual_seed(seed * 64 + rank)
            o = (torch.rand(B, H, D, device=dev, generator=g) * 2 - 1).to(dtype)
            l = torch.randn(B, H, device=dev, generator=g) * 4
            if rank == 0:
                l[0, 0] = float("-inf")
            outs = [torch.empty_like(o) for _ in range(world)]
            lses = [torch.empty_like(l) for _ in range(world)]
            dist.all_gather(outs, o, group=group)
            dist.all_gather(lses, l, group=group)
            L = torch.stack(lses)
            lse = torch.logsumexp(L, dim=0)
            acc = torch.zeros(B, H, D, device=dev)
            for r in range(world):
                f = torch.exp(L[r] - lse)
                acc = acc + (outs[r].float() * f[..., None]).to(dtype).float()
            exp = acc.to(dtype)
            torch.cuda.synchronize(dev)
            dist.barrier(group=group)
            keep_sms_busy()
            got, got_lse = ar.dcp_lse_merge(o, l, reduce_scatter=True, return_lse=True)
            torch.cuda.synchronize(dev)
            hs = H // world
            sl = slice(rank * hs, (rank + 1) * hs)
            if not torch.allclose(got.float(), exp[:, sl].float(), rtol=2 * tol, atol=2 * tol) or \
                    not torch.allclose(got_lse, lse[:, sl], rtol=1e-5, atol=1e-5):
                raise AssertionError(f"tp_check: dcp_lse_merge mismatch {dtype} {(B, H, D)} rank={rank}")
            cases += 1
    res["dcp_lse_merge"] = "within output rounding of the definition"
    res["dcp_cases"] = cases
    have_mc = ar.multicast
    torch.cuda.synchronize(dev)
    ar.close()
    if have_mc and world in (2, 4, 8):
        # the peer-pull data plane (no multicast) must still be bit-exact: the fused kernel against the unfused pair
        ar = B200AllReduce(group, dev, max_size=8 << 20, heap_size=32 << 20, multicast=False)
        for (M, H) in ((256, 4096), (33, 2048)):
            seed += 1
            x = ar.alloc((M, H), torch.bfloat16)
            x.copy_(rank_input(M * H, torch.bfloat16, seed).view(M, H))
            g = torch.Generator(device=dev).manual_seed(seed)
            res_full = (torch.rand(M, H, device=dev, generator=g) * 2 - 1).to(torch.bfloat16)
            w = (torch.rand(H, device=dev, generator=g) + 0.5).to(torch.bfloat16)
            scale = torch.tensor([0.02], device=dev)
            summed = exact_sum(x)
            res_ref = res_full.clone()
            out_ref = torch.empty(M, H, dtype=torch.float8_e4m3fn, device=dev)
            ops.fused_add_rms_norm_static_fp8_quant(out_ref, summed, res_ref, w, scale, 1e-5)
            lo, hi = ar.rows_of_rank(M)
            shard = torch.zeros((M + world - 1) // world, H, dtype=torch.bfloat16, device=dev)
            shard[: hi - lo] = res_full[lo:hi]
            out = ar.alloc((M, H), torch.float8_e4m3fn)
            torch.cuda.synchronize(dev)
            dist.barrier(group=group)
            ar.fused_add_rms_norm(out, x, shard, w, 1e-5, scale)
            torch.cuda.synchronize(dev)
            dist.barrier(group=group)
            if not torch.equal(out.view(torch.uint8), out_ref.view(torch.uint8)) or \
                    not torch.equal(shard[: hi - lo], res_ref[lo:hi]):
                raise AssertionError(f"tp_check: peer-pull fused all-reduce+norm not bit-exact {(M, H)} rank={rank}")
        res["fused_allreduce_rmsnorm_peer_pull"] = "bit-exact"
        torch.cuda.synchronize(dev)
        ar.close()
    del busy_a

    # ---- (3) tiny model, TP=world vs TP=1
    # dense (TP) and sparse-MoE (TP attention + expert parallel; comm=fused is the second EP form:
    # all-gather tokens -> local experts -> reduce-scatter) tiny models
    tiny_notes = []
    for shape in (decoder.TINY_TP, decoder.TINY_MOE_EP):
        S, ctx, bs = 48, 64, BLOCK
        cg = torch.Generator().manual_seed(5)
        nblk = (ctx + bs - 1) // bs
        nb = S * nblk + 8
        bt = torch.randperm(nb, generator=cg)[: S * nblk].to(torch.int32).view(S, nblk)
        tok0 = torch.randint(0, shape.vocab, (S,), generator=cg)
        lens0 = torch.randint(1, ctx - 2, (S,), generator=cg)
        # one global FP8 KV prefix per layer (random non-NaN e4m3 codes); a TP rank keeps its contiguous kv heads
        kv_gen = torch.Generator().manual_seed(9)
        kv_full = [(torch.randint(0, 0x70, (nb, bs, shape.kv_heads, 2 * shape.head_size), generator=kv_gen) |
                    (torch.randint(0, 2, (nb, bs, shape.kv_heads, 2 * shape.head_size), generator=kv_gen) << 7)
                    ).to(torch.uint8) for _ in range(shape.layers)]

        def run(model, tokens_forced=None):
            """two eager decode steps from the same synthetic state; returns (tokens[2, S], logits[2, S, V])"""
            for li, L in enumerate(model.layers):
                lo = model.tp_rank * model.hkv
                L["kv"].copy_(kv_full[li][:, :, lo:lo + model.hkv])
            toks, logs, routes = [], [], []
            tok = tok0.clone()
            for step in range(2):
                lens = lens0 + step
                pos = lens - 1
                model.token_ids.copy_(tok)
                model.positions.copy_(pos)
                model.seq_lens.copy_(lens.to(torch.int32))
                model.block_table.copy_(bt)
                model.slot_mapping.copy_(bt[torch.arange(S), pos // bs].long() * bs + pos % bs)
                model.routing_log = [] if shape.experts > 0 else None
                model.forward_device()
                torch.cuda.synchronize(dev)
                toks.append(model.next_tokens.cpu().clone())
                logs.append(model.last_logits.float().cpu().clone())
                routes.append(torch.stack(model.routing_log) if shape.experts > 0 else torch.zeros(1, S, 1))
                model.routing_log = None
                tok = toks[-1] if tokens_forced is None else tokens_forced[step]
            return torch.stack(toks), torch.stack(logs), torch.stack(routes)

        ref_t = ref_l = None
        if rank == 0:
            m1 = decoder.DecoderModel(shape, S, ctx, bs, dev, num_blocks=nb)
            ref_t, ref_l, ref_r = run(m1)
            del m1
        forced = [torch.empty(S, dtype=torch.int64, device=dev) for _ in range(2)]
        if rank == 0:
            for i in range(2):
                forced[i].copy_(ref_t[i])
        for i in range(2):
            dist.broadcast(forced[i], 0, group=group)
        forced = [f.cpu() for f in forced]
        # (comm, multicast): the peer-pull data plane (multicast False) and, where the box has it, the NVLS one
        cfgs = [("nccl", None), ("nvlink", False)] + ([("fused", False)] if world in (2, 4, 8) else [])
        if have_mc:
            cfgs += [("nvls", None)] + ([("fused", None)] if world in (2, 4, 8) else [])
        comms = [c_ + ("+multimem" if (m_ is None and c_ != "nccl") else "") for c_, m_ in cfgs]
        near_ties, identical, total, worst_rel, flips = 0, 0, 0, 0.0, 0
        logits_by_comm = {}
        for (comm, mcast), label in zip(cfgs, comms):
            mN = decoder.DecoderModel(shape, S, ctx, bs, dev, tp_rank=rank, tp_size=world, tp_group=group,
                                      comm=comm, num_blocks=nb, multicast=mcast)
            t, lg, rt = run(mN, forced)      # teacher-forced with the TP=1 tokens so step 2 sees the same context
            logits_by_comm[label] = lg
            comm = label
            if mN.ar is not None:
                torch.cuda.synchronize(dev)
                mN.ar.close()
            del mN
            if rank == 0:
                # TP changes where the bf16 / FP8 roundings fall (per-rank partial sums are rounded before the
                # all-reduce, and the next layer re-quantises to FP8), so logits agree to FP8 noise, not bit for bit:
                # require a small relative error, and that a differing greedy token is explained by it (the TP=1
                # margin between the two candidates is below twice that row's largest logit deviation)
                for step in range(2):
                    d = (lg[step] - ref_l[step]).abs()
                    # sparse MoE: a router near-tie resolved the other way sends a token to another expert and changes
                    # its whole row; such rows (top-k id sets different from TP=1 in some layer) are counted, bounded
                    # and left out of the numeric comparison
                    kept = (rt[step] == ref_r[step]).all(dim=2).all(dim=0)
                    flips += int((~kept).sum())
                    d = d[kept]
                    rel = float(d.norm() / ref_l[step][kept].norm())
                    worst_rel = max(worst_rel, rel)
                    if rel > 5e-2:
                        raise AssertionError(f"tp_check: tiny model TP={world} comm={comm} step {step}: logits differ "
                                             f"from TP=1 by {rel:.3e} (relative Frobenius)")
                    same = (t[step] == ref_t[step]) | ~kept
                    identical += int((same & kept).sum())
                    total += int(kept.sum())
                    dfull = (lg[step] - ref_l[step]).abs()
                    for sidx in (~same).nonzero().flatten().tolist():
                        a, b = ref_l[step, sidx, ref_t[step, sidx]], ref_l[step, sidx, t[step, sidx]]
                        if float(a - b) > 2.0 * float(dfull[sidx].max()):
                            raise AssertionError(f"tp_check: tiny model TP={world} comm={comm} step {step} seq {sidx}: "
                                                 f"token {int(t[step, sidx])} vs TP=1 {int(ref_t[step, sidx])} (TP=1 logits "
                                                 f"{float(a):.5f} vs {float(b):.5f}, max logit deviation {float(dfull[sidx].max()):.5f})")
                        near_ties += 1
        # the fused reduce-scatter+norm+quant+all-gather kernel is arithmetically identical to all-reduce -> norm+quant
        # with the rank-order NVLink all-reduce: the two decoders must produce the same logits bit for bit
        if "fused" in logits_by_comm and not torch.equal(logits_by_comm["fused"], logits_by_comm["nvlink"]):  # peer-pull
            raise AssertionError(f"tp_check: tiny model TP={world}: comm=fused and comm=nvlink logits are not bit-identical")
        if world == 2 and not torch.equal(logits_by_comm["nccl"], logits_by_comm["nvlink"]):
            raise AssertionError("tp_check: tiny model TP=2: NCCL and NVLink all-reduce decoders differ (a+b is order-free)")
        if rank == 0 and identical < 0.9 * total:
            raise AssertionError(f"tp_check: only {identical} of {total} greedy tokens equal TP=1")
        if rank == 0 and flips > 0.15 * 2 * S * len(cfgs):
            raise AssertionError(f"tp_check: {flips} rows routed differently from TP=1 (more than near-ties explain)")
        tiny_notes.append(f"{shape.name}, {S} seqs x 2 steps x comm {'/'.join(comms)}: {identical} of {total} greedy "
                          f"tokens identical to TP=1, {near_ties} differ inside the measured logit deviation "
                          f"(near-ties); logits vs TP=1 rel err <= {worst_rel:.2e}; peer-pull fused == nvlink logits "
                          f"bit-identical" + (f"; {flips} rows left out (router near-ties routed to another expert)"
                                              if shape.experts > 0 else ""))
    res["tiny_tokens"] = "match"
    res["tiny_model"] = tiny_notes[0]
    res["tiny_moe_model"] = tiny_notes[1]
    return res


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
    tp_check = None
    if world > 1 and not a.no_tp_check:
        try:
            tp_check = tp_self_check(rank, world, dev, group)
        except Exception as e:
            import traceback
            traceback.print_exc()
            print(json.dumps({"tp_check": {"failed": repr(e)}, "rank": rank}), flush=True)
            sys.stderr.flush()
            os._exit(3)
    if a.quant == "nvfp4" and a.comm == "fused":
        a.comm = "nvlink"   # NVFP4 linears take bf16 activations: no FP8-quantising norm fusion
    S, ctx = a.batch, a.ctx
    model = decoder.DecoderModel(shape, S, ctx, BLOCK, dev, tp_rank=rank, tp_size=world, tp_group=group,
                                 comm=a.comm, quant=a.quant, multicast=False if a.no_multicast else None)
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

    # ---- (3) roofline of the dominant kernel: a CUDA graph holding ONLY the paged-attention launches of the
    # step (one per layer, each over that layer's own KV cache: 2.1 GB per launch, nothing survives in L2),
    # replayed and timed with CUDA events on the launching stream. Device time per launch without host launch
    # overhead (eager events around a ~70 us kernel at TP=8 would mostly measure Python).
    side = torch.cuda.Stream(device=dev)
    side.wait_stream(torch.cuda.current_stream(dev))
    with torch.cuda.stream(side), model.pdl_scope():
        for L in model.layers[:2]:
            model.attention_layer(L)
    torch.cuda.current_stream(dev).wait_stream(side)
    torch.cuda.synchronize(dev)
    attn_graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(attn_graph), model.pdl_scope():   # same launch level as inside the step
        for L in model.layers:
            model.attention_layer(L)
    attn_graph.replay()
    torch.cuda.synchronize(dev)
    reps = max(1, min(a.steps, 5))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        attn_graph.replay()
    e1.record()
    torch.cuda.synchronize(dev)
    attn_ms = e0.elapsed_time(e1) / (reps * shape.layers)
    del attn_graph
    es = 1  # fp8 KV
    alg_bytes = (S * 2 * ctx * model.hkv * D * es + S * model.hq * D * (2 + 2) + S * nblk * 4 + 4 * S)
    peak, peak_src = peaks()
    achieved = alg_bytes / (attn_ms * 1e-3) / 1e9
    # dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of this kernel on this exact workload: parsed at
    # run time from the committed ncu --set full summary (profiles/traffic.json names the .md it was extracted
    # from); a number taken under the profiler is only used for the byte count, never for time. null when no
    # capture of this workload is committed.
    traffic, traffic_src = measured_traffic(f"{shape.name}|tp{world}|batch{S}|ctx{ctx}|{a.quant}")
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
                                f"({baseline_config_tag(shape, a.quant)})"),
                   "parallelism": (f"tp{world} attention + ep{world} experts" if shape.experts > 0 else f"tp{world}"), "tp_collective": (a.comm if world > 1 else None),
                   "nvswitch_multicast": (bool(model.ar.multicast) if model.ar is not None else None), "weights": "random init (dummy), per-tensor FP8 scales",
                   "cuda_graph": not a.no_graph, "pdl_level": model.pdl,
                   "l2": ("inputs larger than L2 (KV cache %.1f GB + weights per GPU)" if
                          sum(L["kv"].numel() for L in model.layers) > 256e6 else
                          "working set of %.2f GB KV + weights is L2-sized: a parity-case line, not a bandwidth claim") %
                         (sum(L["kv"].numel() for L in model.layers) / 1e9),
                   "context": "seq_len fixed at ctx every step (the new token is written at position ctx-1)"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": model.h2d_bytes(),
                "d2h_bytes_per_step": model.d2h_bytes(), "ms_per_step": ms_e2e / a.steps},
        "gpu_launches": launches * a.steps,
        "gpu_launches_per_step": launches,
        "clocks": clocks,
        "roofline": {"kernel": "paged_attn_decode_kernel<128,1,BF16_FP8>", "bound": "hbm",
                     "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "timing": "CUDA graph of the step's attention launches (one per layer, each on its own "
                               "KV cache), CUDA events, per launch",
                     "peak_source": peak_src, "traffic": traffic, "traffic_source": traffic_src,
                     "algorithmic_bytes_per_launch": alg_bytes, "ms_per_launch": attn_ms,
                     "launches_per_step": shape.layers,
                     "share_of_step": attn_ms * shape.layers / step_ms},
    }
    if tp_check is not None:
        line["tp_check"] = tp_check
    if rank == 0 and world == 1 and not a.no_cpu_baseline and shape.experts == 0:
        try:
            from oracle import refcpu
            if refcpu.available():
                cb = cpu_reference_run(shape, 3, 1, ctx, batch=S)
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
