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
import os
import socket
import tempfile
import threading
import uuid
from typing import List, Optional

import torch

from ..ops import _lib
from ..ops._lib import MAX_RANKS, FusedComm, KernelError, NvlsComm, check

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


def symm_mode() -> str:
    """``ipc`` (cudaMalloc + CUDA IPC) or ``vmm`` (cuMemCreate + fd exchange + NVLS multicast).
    TONY_B200_SYMM=ipc|vmm|auto; auto picks vmm when the device supports multicast."""
    mode = os.environ.get("TONY_B200_SYMM", "auto").lower()
    if mode == "auto":
        try:
            mode = "vmm" if _lib.lib().tony_nvls_supported() else "ipc"
        except Exception:  # noqa: BLE001
            mode = "ipc"
    return mode


def _exchange_fds(my_fd: Optional[int], rank: int, world: int, group, only_from: Optional[int] = None):
    """Pass POSIX fds between the ranks of one node over unix sockets (SCM_RIGHTS).
    Returns ``fds[r]`` = a duplicate of rank r's fd valid in this process (own slot = my_fd).
    With ``only_from=k`` only rank k serves its fd."""
    import torch.distributed as dist

    token = [uuid.uuid4().hex if rank == 0 else None]
    dist.broadcast_object_list(token, src=0, group=group)
    base = os.path.join(tempfile.gettempdir(), f"tony_symm_{token[0]}")
    serving = my_fd is not None and (only_from is None or only_from == rank)
    srv = None
    if serving:
        srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        srv.bind(f"{base}.{rank}")
        srv.listen(world)
    dist.barrier(group=group)

    def serve():
        for _ in range(world - 1):
            conn, _ = srv.accept()
            socket.send_fds(conn, [b"f"], [my_fd])
            conn.close()

    t = None
    if serving and world > 1:
        t = threading.Thread(target=serve, daemon=True)
        t.start()
    fds = [-1] * world
    fds[rank] = my_fd if my_fd is not None else -1
    sources = range(world) if only_from is None else [only_from]
    for r in sources:
        if r == rank:
            continue
        c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        c.connect(f"{base}.{r}")
        _, got, _, _ = socket.recv_fds(c, 16, 1)
        c.close()
        fds[r] = got[0]
    if t is not None:
        t.join()
    dist.barrier(group=group)
    if srv is not None:
        srv.close()
        try:
            os.unlink(f"{base}.{rank}")
        except OSError:
            pass
    return fds


class VmmBuffer:
    """Symmetric allocation on CUDA VMM: unicast mapping of every rank's copy + one NVLS
    multicast mapping (``mc_ptr``) of all copies.  Same attributes as :class:`SymmBuffer`."""

    def __init__(self, nbytes: int, dtype: torch.dtype, device, group=None):
        import torch.distributed as dist

        self.device = torch.device(device)
        self.dtype = dtype
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        lib = _lib.lib()
        with torch.cuda.device(self.device):
            padded = ctypes.c_size_t()
            check(lib.tony_vmm_granularity(self.world, int(nbytes), ctypes.byref(padded)), "tony_vmm_granularity")
            self.padded = int(padded.value)
            self.nbytes = int(nbytes)
            h, fd = ctypes.c_ulonglong(), ctypes.c_int()
            check(lib.tony_vmm_alloc(self.padded, ctypes.byref(h), ctypes.byref(fd)), "tony_vmm_alloc")
            self.handle = int(h.value)
            fds = _exchange_fds(int(fd.value), self.rank, self.world, group)
            self.peer_ptrs: List[int] = [0] * self.world
            self._handles = {}
            for r in range(self.world):
                hr = self.handle
                if r != self.rank:
                    hh = ctypes.c_ulonglong()
                    check(lib.tony_vmm_import(fds[r], ctypes.byref(hh)), f"tony_vmm_import(rank {r})")
                    hr = int(hh.value)
                    os.close(fds[r])
                    self._handles[r] = hr
                p = ctypes.c_void_p()
                check(lib.tony_vmm_map(hr, self.padded, ctypes.byref(p)), f"tony_vmm_map(rank {r})")
                self.peer_ptrs[r] = int(p.value)
            os.close(int(fd.value))
            self.ptr = self.peer_ptrs[self.rank]
            # multicast object: created by rank 0, imported by the others
            mch, mcfd = ctypes.c_ulonglong(), ctypes.c_int(-1)
            if self.rank == 0:
                check(lib.tony_mc_create(self.padded, self.world, ctypes.byref(mch), ctypes.byref(mcfd)), "tony_mc_create")
            mfds = _exchange_fds(int(mcfd.value) if self.rank == 0 else None, self.rank, self.world, group, only_from=0)
            if self.rank != 0:
                check(lib.tony_vmm_import(mfds[0], ctypes.byref(mch)), "tony_vmm_import(multicast)")
                os.close(mfds[0])
            else:
                os.close(int(mcfd.value))
            self.mc_handle = int(mch.value)
            check(lib.tony_mc_add_device(self.mc_handle), "tony_mc_add_device")
            dist.barrier(group=group)
            check(lib.tony_mc_bind(self.mc_handle, self.handle, self.padded), "tony_mc_bind")
            dist.barrier(group=group)
            q = ctypes.c_void_p()
            check(lib.tony_vmm_map(self.mc_handle, self.padded, ctypes.byref(q)), "tony_vmm_map(multicast)")
            self.mc_ptr = int(q.value)
        self.local = tensor_from_ptr(self.ptr, self.nbytes, dtype, self.device, owner=self)
        self.local.zero_()
        torch.cuda.synchronize(self.device)
        dist.barrier(group=group)

    def peer_tensor(self, r: int) -> torch.Tensor:
        return tensor_from_ptr(self.peer_ptrs[r], self.nbytes, self.dtype, self.device, owner=self)

    def close(self):
        lib = _lib.lib()
        if getattr(self, "mc_ptr", 0):
            lib.tony_vmm_unmap(ctypes.c_void_p(self.mc_ptr), self.padded)
            self.mc_ptr = 0
        for p in getattr(self, "peer_ptrs", []):
            if p:
                lib.tony_vmm_unmap(ctypes.c_void_p(p), self.padded)
        self.peer_ptrs = []
        for h in list(getattr(self, "_handles", {}).values()) + [getattr(self, "mc_handle", 0), getattr(self, "handle", 0)]:
            if h:
                lib.tony_vmm_release(h)
        self._handles, self.mc_handle, self.handle = {}, 0, 0


class SymmBuffer:
    """One symmetric allocation: ``local`` tensor + ``peer_ptrs[r]`` for every rank."""

    mc_ptr = 0  # the IPC path has no multicast mapping

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
        self.nvls = None
        self.nvls_calls = 0
        if self.world > 1:
            Buf = VmmBuffer if symm_mode() == "vmm" else SymmBuffer
            self.grad_buf = Buf(self.numel * 2, torch.bfloat16, device, group)
            self.param_buf = Buf(self.numel * 2, torch.bfloat16, device, group)
            self.flag_buf = Buf(self.FLAG_BYTES, torch.int32, device, group)
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
        if self.world > 1 and self.grad_buf.mc_ptr and os.environ.get("TONY_B200_NVLS", "1") != "0":
            n = NvlsComm()
            n.rank, n.world = self.rank, self.world
            n.grad_mc, n.param_mc, n.param_local = self.grad_buf.mc_ptr, self.param_buf.mc_ptr, self.param_buf.ptr
            n.flags_mc, n.flags_local = self.flag_buf.mc_ptr, self.flag_buf.ptr
            n.counter, n.status = c.counter, c.status
            self.nvls = n
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

    def next_nvls_call(self) -> int:
        self.nvls_calls += 1
        return self.nvls_calls

    def allreduce_(self, begin: int, count: int, scale: float = 1.0, num_ctas: int = 0, stream=None):
        """param[begin:begin+count] on every rank = scale * sum_r grad_r[begin:begin+count]."""
        s = ctypes.c_void_p(stream.cuda_stream) if stream is not None else _lib.cur_stream()
        if self.nvls is not None:
            check(_lib.lib().tony_fused_reduce_optim_nvls(
                ctypes.byref(self.nvls), begin, count, None, None, None, 0, -1, 0.0, 0.0, 0.0, 0.0, 0.0, 1.0, 1.0,
                scale, None, 1, self.next_nvls_call(), num_ctas, s), "tony_fused_reduce_optim_nvls(allreduce)")
            return
        check(_lib.lib().tony_symm_allreduce_bf16(ctypes.byref(self.c), begin, count, scale,
                                                  self.next_epoch(), num_ctas, s),
              "tony_symm_allreduce_bf16")

    COLL = {"allreduce": 0, "reduce_scatter": 1, "allgather": 2, "broadcast": 3}

    def collective_(self, op: str, begin_bytes: int, nbytes: int, *, fp32: bool = False, scale: float = 1.0,
                    root: int = 0, num_ctas: int = 0, stream=None):
        """Stand-alone collective over the byte range ``[begin_bytes, begin_bytes + nbytes)`` of every rank's
        ``grad`` (input) and ``param`` (output) buffers; ``nbytes % (world * 16) == 0``.  See
        ``tony_symm_collective`` in the C header for the four operations."""
        s = ctypes.c_void_p(stream.cuda_stream) if stream is not None else _lib.cur_stream()
        check(_lib.lib().tony_symm_collective(ctypes.byref(self.c), self.COLL[op], int(fp32), begin_bytes, nbytes,
                                              scale, root, self.next_epoch(), num_ctas, s), f"tony_symm_collective({op})")

    def broadcast_(self, begin: int, count: int, root: int = 0, num_ctas: int = 0, stream=None):
        """``param[begin:begin+count]`` on every rank = the root's ``param[begin:begin+count]`` (bf16-sized
        elements; any payload).  With NVLS the root publishes through the switch with multimem.st; otherwise
        it stages through its ``grad`` buffer and every rank pulls over NVLink."""
        s = ctypes.c_void_p(stream.cuda_stream) if stream is not None else _lib.cur_stream()
        if self.world == 1:
            return
        if self.nvls is not None:
            check(_lib.lib().tony_nvls_broadcast(ctypes.byref(self.nvls), begin, count, root, self.next_nvls_call(),
                                                 num_ctas, s), "tony_nvls_broadcast")
            return
        if self.rank == root:
            self.grad[begin:begin + count].copy_(self.param[begin:begin + count])
        self.collective_("broadcast", begin * 2, count * 2, root=root, num_ctas=num_ctas, stream=stream)

    def close(self):
        if self.world > 1:
            for b in (self.grad_buf, self.param_buf, self.flag_buf):
                b.close()
