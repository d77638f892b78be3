"""NVLink all-reduce for tensor parallelism: host side of csrc/allreduce.cu.

Mirrors the reference's ``CustomAllreduce`` (vllm/distributed/device_communicators/
custom_all_reduce.py:56-560): one instance per TP group and process, IPC buffers exchanged through
the (CPU / NCCL) process group at construction, ``should_custom_ar`` / ``custom_all_reduce`` /
``all_reduce`` / ``close`` with the same meaning. Additions that are B200-first:

* ``alloc(shape, dtype)`` hands out tensors that live in the SYMMETRIC heap (same offset on every
  rank, SPMD allocation order), which peers then reduce in place - no staging copy and no graph-buffer
  registration step (the reference's ``register_graph_buffers``);
* ``fused_add_rms_norm(...)`` = reduce-scatter + residual add + RMSNorm (+ static FP8 quant) +
  all-gather in one kernel (the reference's fuse_allreduce_rms target, docs/design/fusions.md:23).

There is no fallback inside this class: the caller decides between this and NCCL with
``should_custom_ar`` exactly like ``CudaCommunicator.all_reduce`` (cuda_communicator.py:275-341).
"""
from __future__ import annotations

import ctypes as C

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
                 max_size: int = 8 * 1024 * 1024, heap_size: int = 64 * 1024 * 1024):
        """group: process group used ONLY to exchange the 64-byte IPC handles (any backend).
        max_size: largest message (bytes) a single all_reduce may carry (staging + scratch size).
        heap_size: bytes of symmetric heap for alloc()."""
        self.disabled = True
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
            own = C.c_void_p()
            handle = (C.c_ubyte * 64)()
            _lib.check(lib.b200_ipc_alloc(total, C.byref(own), handle), "b200_ipc_alloc")
            self._own = own.value
            handles: list[bytes | None] = [None] * self.world_size
            dist.all_gather_object(handles, bytes(handle), group=group)
            ptrs = (C.c_void_p * self.world_size)()
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
            off, nbytes = C.c_size_t(), C.c_size_t()
            _lib.check(lib.b200_allreduce_heap(self._ctx, C.byref(off), C.byref(nbytes)), "b200_allreduce_heap")
            self._heap = torch.as_tensor(_DevMem(self._own + off.value, nbytes.value), device=self.device)
            self._heap_used = 0
        dist.barrier(group=group)  # every rank has zeroed its signals and opened its peers
        self.disabled = False

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
        """out-of-place sum over the group; algo 0 = auto, 1 = one-shot, 2 = two-shot.
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
        lib.b200_ipc_free(self._own)
        self._ctx, self._peers, self._own, self.disabled = None, [], None, True

    def __del__(self):
        try:
            if self._ctx is not None and not torch.cuda.is_initialized():
                return
        except Exception:
            return
