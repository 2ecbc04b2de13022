"""Device buffers readable by every rank of one box (CUDA IPC over NVLink; multi-GPU only).

The reference is single-device; this supports ShardedCollider(mode="sharded_build"), where a
rank's kernels read pieces of ANOTHER rank's traversal data in place instead of receiving a
broadcast of all of it.  A SharedBuffer is a plain cudaMalloc allocation of this rank (wrapped
as a uint8 torch tensor) plus the addresses under which the same-sized allocations of the other
ranks are mapped into this process."""
import ctypes as C

import torch
import torch.distributed as dist

from . import _lib


class _RawCuda:
    """Minimal __cuda_array_interface__ carrier so that torch can view a raw device pointer."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False),
                                         "version": 3}


class SharedBuffer:
    def __init__(self, ctx, nbytes, group=None):
        self.device = _lib.device_of(ctx)
        self.nbytes = max(int(nbytes), 1)
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        L = _lib.lib()
        with torch.cuda.device(self.device):
            p = C.c_void_p()
            _lib.check(L.col_ipc_alloc(C.byref(p), self.nbytes), "col_ipc_alloc")
            self._ptr = p.value
            handle = C.create_string_buffer(64)
            _lib.check(L.col_ipc_export(self._ptr, handle), "col_ipc_export")
            handles = [None] * self.world
            dist.all_gather_object(handles, (bytes(handle.raw), self.nbytes), group=group)
            self.peer_ptrs = []
            for r, (h, nb) in enumerate(handles):
                if nb != self.nbytes:
                    raise ValueError("SharedBuffer sizes differ between ranks ({} vs {})".format(nb, self.nbytes))
                if r == self.rank:
                    self.peer_ptrs.append(self._ptr)
                    continue
                q = C.c_void_p()
                _lib.check(L.col_ipc_open(C.byref(q), C.create_string_buffer(h, 64)), "col_ipc_open")
                self.peer_ptrs.append(q.value)
            self.local = torch.as_tensor(_RawCuda(self._ptr, self.nbytes), device=self.device)
            # device-side table of the ranks' base addresses (u64 per rank), for kernels
            self.table = torch.tensor(self.peer_ptrs, dtype=torch.int64, device=self.device)

    def peer(self, r):
        """uint8 tensor view (on THIS device) of rank r's buffer."""
        return self.local if r == self.rank else torch.as_tensor(_RawCuda(self.peer_ptrs[r], self.nbytes),
                                                                  device=self.device)

    def close(self):
        if getattr(self, "_ptr", None) is None:
            return
        L = _lib.lib()
        with torch.cuda.device(self.device):
            torch.cuda.synchronize()
            for r, q in enumerate(self.peer_ptrs):
                if r != self.rank:
                    L.col_ipc_close(q)
            dist.barrier(group=self.group)      # nobody still maps it
            L.col_ipc_free(self._ptr)
        self._ptr = None
