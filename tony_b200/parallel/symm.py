"""Symmetric (peer-mapped) device memory for one-process-per-GPU jobs on an NVSwitch domain.

Every rank allocates the same buffers through the kernel library (``tony_symm_alloc``),
exports a CUDA-IPC handle, and maps every peer's allocation, so kernels can issue ordinary
loads/stores to any rank's memory over NVLink.  Handle exchange rides on the process group
the launcher's env contract already set up (``torch.distributed``; the reference exports
INIT_METHOD/RANK/WORLD for exactly this purpose —
tony-core/src/main/java/com/linkedin/tony/runtime/PyTorchRuntime.java:45-56).
The reference has no data plane; this replaces what NCCL would set up.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional

import torch

from ..ops import _lib
from ..ops._lib import MAX_RANKS, FusedComm, KernelError, check

_DTYPE_STR = {torch.bfloat16: None, torch.float32: "<f4", torch.uint8: "|u1", torch.int32: "<i4",
              torch.uint16: "<u2", torch.int16: "<i2"}


class _RawCuda:
    """Minimal ``__cuda_array_interface__`` carrier so torch can wrap a raw device pointer."""

    def __init__(self, ptr: int, nbytes: int, owner=None):
        self.__cuda_array_interface__ = {
            "shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3, "strides": None,
        }
        self._owner = owner


def tensor_from_ptr(ptr: int, nbytes: int, dtype: torch.dtype, device, owner=None) -> torch.Tensor:
    """Zero-copy 1-D tensor over ``nbytes`` of device memory at ``ptr``."""
    raw = torch.as_tensor(_RawCuda(ptr, nbytes, owner), device=device)
    return raw.view(dtype)


class SymmBuffer:
    """One symmetric allocation: ``local`` tensor + ``peer_ptrs[r]`` for every rank."""

    def __init__(self, nbytes: int, dtype: torch.dtype, device, group=None):
        import torch.distributed as dist

        self.device = torch.device(device)
        self.nbytes = int(nbytes)
        self.dtype = dtype
        self.world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        self.rank = dist.get_rank(group) if self.world > 1 else 0
        if self.world > MAX_RANKS:
            raise KernelError(f"symmetric memory supports up to {MAX_RANKS} ranks per NVSwitch domain")
        lib = _lib.lib()
        self._opened: List[int] = []
        with torch.cuda.device(self.device):
            p = ctypes.c_void_p()
            check(lib.tony_symm_alloc(self.nbytes, ctypes.byref(p)), "tony_symm_alloc")
            self.ptr = int(p.value)
            self.peer_ptrs: List[int] = [0] * self.world
            self.peer_ptrs[self.rank] = self.ptr
            if self.world > 1:
                handle = (ctypes.c_ubyte * 64)()
                check(lib.tony_ipc_get_handle(ctypes.c_void_p(self.ptr), handle), "tony_ipc_get_handle")
                mine = bytes(handle)
                gathered: List[Optional[bytes]] = [None] * self.world
                dist.all_gather_object(gathered, mine, group=group)
                for r, h in enumerate(gathered):
                    if r == self.rank:
                        continue
                    buf = (ctypes.c_ubyte * 64).from_buffer_copy(h)
                    q = ctypes.c_void_p()
                    check(lib.tony_ipc_open(buf, ctypes.byref(q)), f"tony_ipc_open(rank {r})")
                    self.peer_ptrs[r] = int(q.value)
                    self._opened.append(int(q.value))
        self.local = tensor_from_ptr(self.ptr, self.nbytes, dtype, self.device, owner=self)

    def peer_tensor(self, r: int) -> torch.Tensor:
        """Tensor view of rank ``r``'s buffer (debug / tests; kernels use the raw pointers)."""
        return tensor_from_ptr(self.peer_ptrs[r], self.nbytes, self.dtype, self.device, owner=self)

    def close(self):
        lib = _lib.lib()
        for q in self._opened:
            lib.tony_ipc_close(ctypes.c_void_p(q))
        self._opened = []
        if self.ptr:
            lib.tony_symm_free(ctypes.c_void_p(self.ptr))
            self.ptr = 0


class SymmComm:
    """Gradient + parameter + flag buffers of all ranks, packaged as ``tony_fused_comm``."""

    FLAG_BYTES = 1024  # 2*8 flag words, counter, status, padding

    def __init__(self, numel: int, device, group=None):
        import torch.distributed as dist

        self.device = torch.device(device)
        self.group = group
        self.world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        self.rank = dist.get_rank(group) if self.world > 1 else 0
        self.numel = int(numel)
        self.epoch = 0
        if self.world > 1:
            self.grad_buf = SymmBuffer(self.numel * 2, torch.bfloat16, device, group)
            self.param_buf = SymmBuffer(self.numel * 2, torch.bfloat16, device, group)
            self.flag_buf = SymmBuffer(self.FLAG_BYTES, torch.int32, device, group)
            self.grad = self.grad_buf.local
            self.param = self.param_buf.local
            flags_local = self.flag_buf.ptr
            grad_ptrs, param_ptrs = self.grad_buf.peer_ptrs, self.param_buf.peer_ptrs
            flag_ptrs = self.flag_buf.peer_ptrs
        else:
            self.grad = torch.zeros(self.numel, dtype=torch.bfloat16, device=device)
            self.param = torch.zeros(self.numel, dtype=torch.bfloat16, device=device)
            self._flags = torch.zeros(self.FLAG_BYTES // 4, dtype=torch.int32, device=device)
            flags_local = self._flags.data_ptr()
            grad_ptrs, param_ptrs, flag_ptrs = [self.grad.data_ptr()], [self.param.data_ptr()], [flags_local]
        c = FusedComm()
        c.rank, c.world = self.rank, self.world
        for r in range(self.world):
            c.grad[r] = grad_ptrs[r]
            c.param[r] = param_ptrs[r]
            c.flags[r] = flag_ptrs[r]
        c.counter = flags_local + 2 * MAX_RANKS * 4
        c.status = flags_local + 2 * MAX_RANKS * 4 + 64
        self.c = c
        self._status_view = tensor_from_ptr(int(c.status), 4, torch.int32, self.device, owner=self)
        if self.world > 1:
            dist.barrier(group=group)  # every rank has mapped every peer before first use

    def next_epoch(self) -> int:
        self.epoch += 1
        return self.epoch

    def check_status(self):
        """Raise if a fused collective gave up waiting for a peer (host sync point)."""
        code = int(self._status_view.item())
        if code != 0:
            raise KernelError(f"fused collective reported status {code} (5 = a peer never arrived)")

    def barrier(self, stream=None):
        if self.world == 1:
            return
        s = ctypes.c_void_p(stream.cuda_stream) if stream is not None else _lib.cur_stream()
        check(_lib.lib().tony_symm_barrier(ctypes.byref(self.c), self.next_epoch(), s), "tony_symm_barrier")

    def allreduce_(self, begin: int, count: int, scale: float = 1.0, num_ctas: int = 0, stream=None):
        """param[begin:begin+count] on every rank = scale * sum_r grad_r[begin:begin+count]."""
        s = ctypes.c_void_p(stream.cuda_stream) if stream is not None else _lib.cur_stream()
        check(_lib.lib().tony_symm_allreduce_bf16(ctypes.byref(self.c), begin, count, scale,
                                                  self.next_epoch(), num_ctas, s),
              "tony_symm_allreduce_bf16")

    def close(self):
        if self.world > 1:
            for b in (self.grad_buf, self.param_buf, self.flag_buf):
                b.close()
