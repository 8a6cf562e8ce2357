"""Symmetric (peer-mapped) device buffers for the in-kernel NVLink data plane.

Every rank allocates the same set of buffers with the native allocator, exports a CUDA IPC handle for each,
and imports every peer's handles (exchanged once over the gloo control group).  The result per buffer is a
device-resident table of W pointers (own buffer at index `rank`) that the kernels of csrc/comm.cu index by
rank.  The reference has no communication layer to compare with.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List

import torch

from ..ops import native as N
from .env import DistEnv


class _RawCuda:
    """Minimal __cuda_array_interface__ holder so torch can view memory it did not allocate."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


class SymmetricBuffers:
    def __init__(self, env: DistEnv, sizes: Dict[str, int]):
        """sizes: name -> bytes.  Collective over env.ctrl."""
        self.env = env
        self.local_ptr: Dict[str, int] = {}
        self.peer_ptr: Dict[str, List[int]] = {}
        self.tables: Dict[str, torch.Tensor] = {}
        self._holders = []
        handles = {}
        for name, nbytes in sizes.items():
            ptr = N.symm_alloc(nbytes)
            self.local_ptr[name] = ptr
            handles[name] = N.ipc_export(ptr) if env.distributed else b""
        everyone = env.gather_objects(handles)
        for name in sizes:
            ptrs = []
            for r in range(env.world_size):
                ptrs.append(self.local_ptr[name] if r == env.rank else N.ipc_import(everyone[r][name]))
            self.peer_ptr[name] = ptrs
            self.tables[name] = torch.tensor(ptrs, dtype=torch.int64, device=env.device)
        self.sizes = dict(sizes)
        env.barrier()

    def close(self) -> None:
        """Unmap the peers' buffers and free the own ones.  Collective: every rank has stopped using every buffer
        (the caller synchronised its device) before anything is released; tensors handed out by `tensor()` must not
        be touched afterwards."""
        if not self.local_ptr:
            return
        env = self.env
        env.barrier()
        for name, ptrs in self.peer_ptr.items():
            for r, p in enumerate(ptrs):
                if r != env.rank:
                    N.lib().ftl_ipc_close(ctypes.c_void_p(p))
        env.barrier()            # nobody maps my memory any more
        for p in self.local_ptr.values():
            N.lib().ftl_symm_free(ctypes.c_void_p(p))
        self.local_ptr, self.peer_ptr, self.tables, self._holders = {}, {}, {}, []

    def tensor(self, name: str, dtype: torch.dtype) -> torch.Tensor:
        """Local buffer `name` as a torch tensor (no copy, not owned by torch's allocator)."""
        holder = _RawCuda(self.local_ptr[name], self.sizes[name])
        self._holders.append(holder)
        t = torch.as_tensor(holder, device=self.env.device)
        return t.view(dtype)

    def table(self, name: str) -> torch.Tensor:
        return self.tables[name]
