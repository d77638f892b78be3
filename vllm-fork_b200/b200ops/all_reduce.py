"""NVLink all-reduce for tensor parallelism: host side of csrc/allreduce.cu.

Mirrors the reference's ``CustomAllreduce`` (vllm/distributed/device_communicators/
custom_all_reduce.py:56-560): one instance per TP group and process, IPC buffers exchanged through
the (CPU / NCCL) process group at construction, ``should_custom_ar`` / ``custom_all_reduce`` /
``all_reduce`` / ``close`` with the same meaning. Additions that are B200-first:

* ``alloc(shape, dtype)`` hands out tensors that live in the SYMMETRIC heap (same offset on every
  rank, SPMD allocation order), which peers then reduce in place - no staging copy and no graph-buffer
  registration step (the reference's ``register_graph_buffers``);
* ``fused_add_rms_norm(...)`` = reduce-scatter + residual add + RMSNorm (+ static FP8 quant) +
  all-gather in one kernel (the reference's fuse_allreduce_rms target, docs/design/fusions.md:23);
* NVSwitch in-network reduction (NVLS): when torch symmetric memory can give the buffer a multicast mapping
  (the plumbing the reference uses too: device_communicators/symm_mem.py:38-150), ``all_reduce(algo=3)`` and the
  fused kernel reduce with ``multimem.ld_reduce`` inside the switch and broadcast with ``multimem.st`` - our own
  kernels, torch only allocates and exchanges the handles. Without it the buffer is cudaMalloc + CUDA IPC and
  the kernels use peer loads / stores.

There is no fallback inside this class: the caller decides between this and NCCL with
``should_custom_ar`` exactly like ``CudaCommunicator.all_reduce`` (cuda_communicator.py:275-341).
"""
from __future__ import annotations

import ctypes as C
import os

import torch
import torch.distributed as dist

from . import _lib, tp
from .ops import _dt, _p, _stream

_SUPPORTED_WORLD_SIZES = [2, 4, 6, 8]


class _DevMem:
    """A raw device range exposed through __cuda_array_interface__ so torch can wrap it (no copy)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {
            "shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3, "strides": None}


class B200AllReduce:
    def __init__(self, group: dist.ProcessGroup | None, device: torch.device | int | str,
                 max_size: int = 8 * 1024 * 1024, heap_size: int = 64 * 1024 * 1024,
                 multicast: bool | None = None):
        """group: process group used ONLY to exchange the memory handles (any backend for the IPC path).
        max_size: largest message (bytes) a single all_reduce may carry (staging + scratch size).
        heap_size: bytes of symmetric heap for alloc().
        multicast: None = use an NVSwitch multicast mapping when torch symmetric memory provides one
        (B200_AR_SYMM=0 disables), False = never (cudaMalloc + CUDA IPC), True = require it."""
        self.disabled = True
        self.multicast = False
        self.multicast_note = "not requested"
        self._symm = None
        self._ctx = None
        self._own = None
        self._peers: list[int] = []
        self.group = group
        self.rank = dist.get_rank(group)
        self.world_size = dist.get_world_size(group)
        if isinstance(device, int):
            device = torch.device(f"cuda:{device}")
        self.device = torch.device(device)
        if not torch.cuda.is_available():
            raise RuntimeError("B200AllReduce needs CUDA devices with peer access (no CPU fallback)")
        if self.world_size == 1:
            return  # like the reference: nothing to do for a single rank
        if self.world_size not in _SUPPORTED_WORLD_SIZES:
            raise RuntimeError(f"custom allreduce only supports num gpus in {_SUPPORTED_WORLD_SIZES}, "
                               f"got {self.world_size}")
        lib = _lib.load()
        self.max_size = (max_size + 1023) // 1024 * 1024
        meta = lib.b200_allreduce_meta_bytes()
        total = meta + 2 * self.max_size + (heap_size + 1023) // 1024 * 1024
        with torch.cuda.device(self.device):
            ptrs = (C.c_void_p * self.world_size)()
            mc_ptr = 0
            want_symm = multicast is not False and os.environ.get("B200_AR_SYMM", "1") == "1"
            if want_symm:
                mc_ptr = self._alloc_symmetric(group, total, ptrs)
                # The symmetric path is taken by ALL ranks or by none: a rank that fell back alone would enter
                # the IPC handle exchange below while its peers go straight to b200_allreduce_init + barrier
                # (mismatched collectives = hang). Same for the multicast mapping.
                flags: list = [None] * self.world_size
                dist.all_gather_object(flags, (self._symm is not None, bool(mc_ptr)), group=group)
                if not all(f[0] for f in flags):
                    if self._symm is not None:
                        self.multicast_note = "torch symmetric memory failed on a peer rank: CUDA-IPC path on all"
                    self._symm = None
                    mc_ptr = 0
                elif not all(f[1] for f in flags):
                    mc_ptr = 0
            if multicast is True and not mc_ptr:
                raise RuntimeError(f"B200AllReduce: no NVSwitch multicast mapping ({self.multicast_note})")
            if self._symm is None:
                own = C.c_void_p()
                handle = (C.c_ubyte * 64)()
                _lib.check(lib.b200_ipc_alloc(total, C.byref(own), handle), "b200_ipc_alloc")
                self._own = own.value
                handles: list[bytes | None] = [None] * self.world_size
                dist.all_gather_object(handles, bytes(handle), group=group)
                for r, h in enumerate(handles):
                    if r == self.rank:
                        ptrs[r] = self._own
                    else:
                        p = C.c_void_p()
                        buf = (C.c_ubyte * 64).from_buffer_copy(h)
                        _lib.check(lib.b200_ipc_open(buf, C.byref(p)), "b200_ipc_open")
                        ptrs[r] = p.value
                        self._peers.append(p.value)
            ctx = C.c_void_p()
            _lib.check(lib.b200_allreduce_init(C.byref(ctx), self.rank, self.world_size, ptrs, total,
                                               self.max_size), "b200_allreduce_init")
            self._ctx = ctx.value
            if mc_ptr:
                _lib.check(lib.b200_allreduce_set_multicast(self._ctx, mc_ptr), "b200_allreduce_set_multicast")
                self.multicast = True
            off, nbytes = C.c_size_t(), C.c_size_t()
            _lib.check(lib.b200_allreduce_heap(self._ctx, C.byref(off), C.byref(nbytes)), "b200_allreduce_heap")
            self._heap = torch.as_tensor(_DevMem(self._own + off.value, nbytes.value), device=self.device)
            self._heap_used = 0
        dist.barrier(group=group)  # every rank has zeroed its signals and opened its peers
        self.disabled = False

    def _alloc_symmetric(self, group, total: int, ptrs) -> int:
        """Allocate the symmetric buffer through torch symmetric memory (cuMemCreate + handle exchange + multicast
        bind: plumbing) and fill `ptrs` with every rank's mapping. Returns the multicast address or 0; on any
        failure leaves self._symm None so that the caller takes the cudaMalloc + CUDA-IPC path."""
        try:
            import warnings
            import torch.distributed._symmetric_memory as symm
            pg = group if group is not None else dist.group.WORLD
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                if hasattr(symm, "enable_symm_mem_for_group"):
                    try:
                        symm.enable_symm_mem_for_group(pg.group_name)
                    except Exception:
                        pass
                t = symm.empty(total, dtype=torch.uint8, device=self.device)
                hdl = symm.rendezvous(t, pg)
            t.zero_()
            torch.cuda.synchronize(self.device)
            bufs = list(hdl.buffer_ptrs)
            if len(bufs) != self.world_size or any(int(b) % 1024 for b in bufs):
                raise RuntimeError("unexpected buffer_ptrs from symmetric memory")
            for r in range(self.world_size):
                ptrs[r] = int(bufs[r])
            self._own = int(bufs[self.rank])
            self._symm = (t, hdl)
            mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
            self.multicast_note = "torch symmetric memory, multicast_ptr %s" % ("mapped" if mc else "unavailable")
            return mc if mc % 1024 == 0 else 0
        except Exception as e:  # noqa: BLE001
            self._symm = None
            self.multicast_note = f"torch symmetric memory unavailable: {e!r}"[:200]
            return 0

    # ---- symmetric heap ---------------------------------------------------------------------
    def alloc(self, shape, dtype: torch.dtype) -> torch.Tensor:
        """Tensor inside the symmetric heap. All ranks must allocate the same sequence of sizes."""
        shape = tuple(shape) if not isinstance(shape, int) else (shape,)
        n = 1
        for s in shape:
            n *= s
        nbytes = n * torch.empty((), dtype=dtype).element_size()
        start = (self._heap_used + 1023) // 1024 * 1024
        if start + nbytes > self._heap.numel():
            raise RuntimeError(f"B200AllReduce.alloc: symmetric heap exhausted ({self._heap.numel()} bytes)")
        self._heap_used = start + nbytes
        return self._heap[start:start + nbytes].view(dtype).view(shape)

    # ---- reference-compatible surface -------------------------------------------------------
    def should_custom_ar(self, inp: torch.Tensor) -> bool:
        if self.disabled:
            return False
        nbytes = inp.numel() * inp.element_size()
        if nbytes % 16 != 0 or not inp.is_contiguous():
            return False
        if inp.dtype not in (torch.float32, torch.float16, torch.bfloat16):
            return False
        return nbytes <= self.max_size

    def all_reduce(self, inp: torch.Tensor, *, out: torch.Tensor | None = None, algo: int = 0,
                   skip_end_barrier: bool = False) -> torch.Tensor:
        """out-of-place sum over the group; algo 0 = auto, 1 = one-shot, 2 = two-shot, 3 = NVSwitch multimem
        (self.multicast must be True; `out` allocated with alloc() is broadcast by the switch and is then
        bit-identical on every rank).
        skip_end_barrier: the caller promises not to overwrite `inp` before the next collective of this
        object has started on every rank (alternating input buffers) - saves one NVLink rendezvous."""
        if skip_end_barrier:
            algo |= 4
        if self.disabled:
            raise RuntimeError("B200AllReduce is disabled (world size 1)")
        if not inp.is_cuda or not inp.is_contiguous():
            raise RuntimeError("all_reduce: input must be a contiguous CUDA tensor (no CPU fallback)")
        if out is None:
            out = torch.empty_like(inp)
        rc = _lib.load().b200_allreduce_run(self._ctx, _p(inp), _p(out), inp.numel(), _dt(inp), algo, _stream())
        _lib.check(rc, "all_reduce")
        return out

    def custom_all_reduce(self, input: torch.Tensor) -> torch.Tensor | None:
        if not self.should_custom_ar(input):
            return None
        return self.all_reduce(input)

    def fused_add_rms_norm(self, out: torch.Tensor, inp: torch.Tensor, residual_shard: torch.Tensor | None,
                           weight: torch.Tensor, epsilon: float, scale: torch.Tensor | None = None) -> None:
        """out[all rows] = quant(rmsnorm(allreduce(inp) + residual)); see b200_allreduce_fused_add_rms_norm.
        residual_shard: this rank's rows [ceil(M / world), hidden], updated in place."""
        if self.disabled:
            raise RuntimeError("B200AllReduce is disabled (world size 1)")
        M, H = inp.shape
        mp = (M + self.world_size - 1) // self.world_size
        if residual_shard is not None and (tuple(residual_shard.shape) != (mp, H) or not residual_shard.is_contiguous()):
            raise RuntimeError(f"fused_add_rms_norm: residual_shard must be contiguous [{mp}, {H}]")
        if not (inp.is_contiguous() and out.is_contiguous()) or tuple(out.shape) != (M, H):
            raise RuntimeError("fused_add_rms_norm: inp / out must be contiguous [M, hidden]")
        if scale is None and out.dtype != inp.dtype:
            raise RuntimeError("fused_add_rms_norm: without a scale, out must have the input dtype")
        if scale is not None and out.element_size() != 1:
            raise RuntimeError("fused_add_rms_norm: with a scale, out must be fp8")
        rc = _lib.load().b200_allreduce_fused_add_rms_norm(
            self._ctx, _p(out), _p(inp), _p(residual_shard), _p(weight), _p(scale), float(epsilon), M, H,
            _dt(inp), _stream())
        _lib.check(rc, "allreduce_fused_add_rms_norm")

    def dcp_lse_merge(self, cp_attn_out: torch.Tensor, cp_attn_lse: torch.Tensor, *, reduce_scatter: bool = True,
                      return_lse: bool = False, is_lse_base_on_e: bool = True):
        """cp_lse_ag_out_rs (reduce_scatter=True) / cp_lse_ag_out_ar (False) of vllm/v1/attention/ops/common.py:185-258:
        merge the partial attention outputs [B, H, D] that every rank of the (decode-)context-parallel group
        computed over ITS shard of the KV cache, weighted by their log-sum-exps [B, H], in one NVLink kernel.
        Inputs that already live in the symmetric heap (alloc()) are read in place by the peers; others are
        copied into heap scratch first (one device copy). Returns out [B, H / world, D] (this rank's heads) or
        [B, H, D], plus the merged LSE when return_lse."""
        if self.disabled:
            raise RuntimeError("B200AllReduce is disabled (world size 1)")
        if cp_attn_out.dim() == 4 and cp_attn_out.shape[1] == 1:
            cp_attn_out = cp_attn_out.squeeze(1)
        if cp_attn_out.dim() != 3 or tuple(cp_attn_lse.shape) != tuple(cp_attn_out.shape[:2]):
            raise RuntimeError(f"dcp_lse_merge: expected out [B,H,D] and lse [B,H], got {tuple(cp_attn_out.shape)} "
                               f"and {tuple(cp_attn_lse.shape)}")
        if not cp_attn_out.is_cuda:
            raise RuntimeError("dcp_lse_merge: inputs must be CUDA tensors (no CPU fallback)")
        B, H, D = cp_attn_out.shape
        if reduce_scatter and H % self.world_size:
            raise RuntimeError(f"dcp_lse_merge: {H} heads do not divide over {self.world_size} ranks")
        o_in, l_in = cp_attn_out, cp_attn_lse
        if not (self._in_heap(o_in) and o_in.is_contiguous()):
            o_in = self._scratch("dcp_out", (B, H, D), cp_attn_out.dtype)
            o_in.copy_(cp_attn_out)
        if not (self._in_heap(l_in) and l_in.is_contiguous() and l_in.dtype == torch.float32):
            l_in = self._scratch("dcp_lse", (B, H), torch.float32)
            l_in.copy_(cp_attn_lse)
        hs = H // self.world_size if reduce_scatter else H
        out = torch.empty(B, hs, D, dtype=cp_attn_out.dtype, device=self.device)
        lse = torch.empty(B, hs, dtype=torch.float32, device=self.device) if return_lse else None
        rc = _lib.load().b200_dcp_lse_merge(self._ctx, _p(out), _p(lse), _p(o_in), _p(l_in), B, H, D, _dt(o_in),
                                            0 if reduce_scatter else 1, 1 if is_lse_base_on_e else 0, _stream())
        _lib.check(rc, "dcp_lse_merge")
        return (out, lse) if return_lse else out

    def _in_heap(self, t: torch.Tensor) -> bool:
        lo = self._heap.data_ptr()
        return lo <= t.data_ptr() and t.data_ptr() + t.numel() * t.element_size() <= lo + self._heap.numel()

    def _scratch(self, name: str, shape, dtype: torch.dtype) -> torch.Tensor:
        """persistent heap scratch by (name, bytes): allocated on first use, in the same order on every rank"""
        n = 1
        for s in shape:
            n *= s
        key = (name, n * torch.empty((), dtype=dtype).element_size())
        if not hasattr(self, "_scratch_bufs"):
            self._scratch_bufs = {}
        if key not in self._scratch_bufs:
            self._scratch_bufs[key] = self.alloc(key[1], torch.uint8)
        return self._scratch_bufs[key].view(dtype).view(shape)

    def rows_of_rank(self, M: int) -> tuple[int, int]:
        return tp.rows_of_rank(M, self.rank, self.world_size)

    def close(self) -> None:
        if self._ctx is None:
            return
        lib = _lib.load()
        torch.cuda.synchronize(self.device)
        if self.group is not None:
            dist.barrier(group=self.group)  # nobody frees memory a peer may still be reading
        lib.b200_allreduce_destroy(self._ctx)
        for p in self._peers:
            lib.b200_ipc_close(p)
        self._heap = None
        if self._symm is None:
            lib.b200_ipc_free(self._own)
        self._symm = None   # torch frees the symmetric allocation with the last reference
        self._ctx, self._peers, self._own, self.disabled = None, [], None, True

    def __del__(self):
        try:
            if self._ctx is not None and not torch.cuda.is_initialized():
                return
        except Exception:
            return
