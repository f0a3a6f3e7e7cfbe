"""Symmetric peer memory for one-process-per-GPU jobs.

Each rank allocates one buffer with the same layout, exports it with CUDA IPC and
maps every peer's copy, so kernels can load from and store to any rank's buffer
through NVLink.  ``torch.distributed`` is only used to exchange the 64-byte IPC
handles at set-up.  Regions of the local buffer, and of the mapped peer buffers,
are exposed as ordinary torch tensors (zero-copy views).
"""
import ctypes as C

import torch
import torch.distributed as dist

from .._native import knn_lib


class _RawCuda:
    """Minimal __cuda_array_interface__ holder so torch can alias raw device memory."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3}


def _lib():
    lib = knn_lib()
    lib.dmlp_last_error.restype = C.c_char_p
    lib.dmlp_symm_alloc.argtypes = [C.c_size_t, C.POINTER(C.c_void_p), C.c_char_p]
    lib.dmlp_symm_open.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
    lib.dmlp_symm_close.argtypes = [C.c_void_p]
    lib.dmlp_symm_free.argtypes = [C.c_void_p]
    lib.dmlp_symm_signal.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_uint32, C.c_void_p]
    lib.dmlp_symm_wait.argtypes = [C.c_void_p, C.c_int, C.c_uint32, C.c_double, C.c_void_p, C.c_void_p]
    return lib


def _check(rc):
    if rc != 0:
        raise RuntimeError("dmlp_symm: " + _lib().dmlp_last_error().decode())


class Layout:
    """Named, 256-byte aligned regions inside a symmetric buffer."""

    def __init__(self):
        self.off, self.size, self.total = {}, {}, 0

    def add(self, name, nbytes):
        self.off[name] = self.total
        self.size[name] = nbytes
        self.total += (nbytes + 255) // 256 * 256
        return self.off[name]


class SymmetricBuffer:
    def __init__(self, layout, device, group=None):
        self.layout, self.device, self.group = layout, torch.device(device), group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        lib = _lib()
        with torch.cuda.device(self.device):
            ptr, handle = C.c_void_p(), C.create_string_buffer(64)
            _check(lib.dmlp_symm_alloc(layout.total, C.byref(ptr), handle))
            self.local_ptr = ptr.value
            handles = [None] * self.world
            if self.world > 1:
                dist.all_gather_object(handles, bytes(handle.raw), group=group)
            self.peer_ptrs, self._opened = [], []
            for r in range(self.world):
                if r == self.rank:
                    self.peer_ptrs.append(self.local_ptr)
                else:
                    p = C.c_void_p()
                    _check(lib.dmlp_symm_open(handles[r], C.byref(p)))
                    self.peer_ptrs.append(p.value)
                    self._opened.append(p.value)
            self._keep = []
            self._bytes = [self._alias(p, layout.total) for p in self.peer_ptrs]
            if self.world > 1:
                dist.barrier(group=group)

    def _alias(self, ptr, nbytes):
        holder = _RawCuda(ptr, nbytes)
        self._keep.append(holder)
        return torch.as_tensor(holder, device=self.device)

    def view(self, name, dtype, shape=None, rank=None):
        """Tensor view of region `name` in this rank's buffer (or in rank `rank`'s, as mapped here)."""
        raw = self._bytes[self.rank if rank is None else rank]
        off, size = self.layout.off[name], self.layout.size[name]
        t = raw[off:off + size].view(dtype)
        return t.view(shape) if shape is not None else t

    def pointer_table(self, name):
        """Device tensor of `world` pointers: region `name` in every rank's buffer."""
        off = self.layout.off[name]
        return torch.tensor([p + off for p in self.peer_ptrs], dtype=torch.int64, device=self.device)

    def signal(self, table, value):
        _check(_lib().dmlp_symm_signal(C.c_void_p(table.data_ptr()), self.world, self.rank, value,
                                       C.c_void_p(torch.cuda.current_stream().cuda_stream)))

    def wait(self, name, value, status=None, timeout=20.0):
        _check(_lib().dmlp_symm_wait(C.c_void_p(self.local_ptr + self.layout.off[name]), self.world, value, timeout,
                                     C.c_void_p(status.data_ptr() if status is not None else 0),
                                     C.c_void_p(torch.cuda.current_stream().cuda_stream)))

    # -- in-kernel signalling: argument blocks for the kernels that carry the flag protocol themselves
    def signal_args(self, name, value):
        """`signal` block for a kernel launch: when the launch has finished, its last block stores
        `value` into this rank's word of region `name` (a flag array) in every rank's buffer."""
        from ..ops.cuda import SignalArgs
        if not hasattr(self, "_done_counters"):
            self._done_counters, self._next_counter = torch.zeros(64, dtype=torch.int32, device=self.device), {}
        slot = self._next_counter.setdefault(name, len(self._next_counter))   # one counter per flag array
        sg = SignalArgs()
        off = self.layout.off[name]
        for r, ptr in enumerate(self.peer_ptrs):
            sg.peer_flags[r] = ptr + off
        sg.world, sg.rank, sg.value = self.world, self.rank, value
        sg.done_counter = self._done_counters.data_ptr() + 4 * slot
        return sg

    def wait_args(self, name, value, status=None, timeout=20.0):
        """`wait` block for a kernel launch: every block first spins until all words of the local
        flag array `name` have reached `value`."""
        from ..ops.cuda import WaitArgs
        wa = WaitArgs()
        wa.flags = self.local_ptr + self.layout.off[name]
        wa.world, wa.value, wa.timeout_seconds = self.world, value, timeout
        wa.status = status.data_ptr() if status is not None else 0
        return wa

    def close(self):
        lib = _lib()
        torch.cuda.synchronize(self.device)
        if self.world > 1 and dist.is_initialized():
            dist.barrier(group=self.group)
        for p in self._opened:
            lib.dmlp_symm_close(C.c_void_p(p))
        self._opened = []
        if self.local_ptr:
            lib.dmlp_symm_free(C.c_void_p(self.local_ptr))
            self.local_ptr = 0
