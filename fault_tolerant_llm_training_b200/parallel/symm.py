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


class VmmSymmetricBuffers(SymmetricBuffers):
    """The same heap built on the CUDA virtual-memory API (csrc/nvls.cu), optionally with NVSwitch multicast.

    Every rank creates its buffers with cuMemCreate and exports one POSIX file descriptor per buffer; descriptors
    cannot travel through torch.distributed, so each rank serves them on a unix-domain socket (SCM_RIGHTS) whose
    path is agreed over the gloo control group, and fetches its peers' descriptors the same way.  Imported
    allocations are mapped with cuMemMap, which yields the same per-buffer pointer tables as the CUDA-IPC heap.
    For the names in `multicast`, rank 0 additionally creates a multicast object over all W devices; every rank adds
    its device, binds its own allocation and maps the object: `mc_ptr[name]` is an address at which a store reaches
    all W copies and a `multimem.ld_reduce` returns their sum (used by the NVLS kernels, FTL_COLLECTIVE=nvls)."""

    def __init__(self, env: DistEnv, sizes: Dict[str, int], multicast=()):
        import os
        import socket
        import threading
        import uuid

        self.env = env
        W, r = env.world_size, env.rank
        vmm_ok, mc_ok = N.vmm_supported()
        if not vmm_ok:
            raise N.NativeError("the CUDA virtual-memory API with file-descriptor handles is not available on this device")
        multicast = [n for n in multicast if W > 1]
        if multicast and not mc_ok:
            raise N.NativeError("this GPU / driver does not support multicast objects (NVLS); use FTL_COLLECTIVE=p2p")
        names = list(sizes)
        self._req = dict(sizes)     # bytes asked for; the mappings are rounded up to the allocation granularity
        self.sizes = {n: N.vmm_round(sizes[n], W if n in multicast else 1) for n in names}
        self.local_ptr, self.peer_ptr, self.tables, self._holders = {}, {}, {}, []
        self.mc_ptr: Dict[str, int] = {}
        self._handles: Dict[str, List[int]] = {n: [0] * W for n in names}   # allocation handle per (name, rank)
        self._mc_handles: Dict[str, int] = {}
        my_fds = []
        for n in names:
            ptr, h, fd = N.vmm_alloc(self.sizes[n])
            self.local_ptr[n] = ptr
            self._handles[n][r] = h
            my_fds.append(fd)
        mc_fds = []
        if r == 0:
            for n in multicast:
                h, fd = N.mc_create(W, self.sizes[n])
                self._mc_handles[n] = h
                mc_fds.append(fd)
        peer_fds: Dict[int, List[int]] = {}
        if W > 1:
            tag = env.gather_objects(uuid.uuid4().hex[:12] if r == 0 else None)[0]
            path = lambda q: f"/tmp/ftl_vmm_{tag}_{q}.sock"
            srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            srv.bind(path(r))
            srv.listen(W)
            send_list = my_fds + mc_fds   # rank 0 also hands out the multicast objects

            def serve():
                for _ in range(W - 1):
                    conn, _ = srv.accept()
                    with conn:
                        socket.send_fds(conn, [b"f"], send_list)

            th = threading.Thread(target=serve, daemon=True)
            th.start()
            env.barrier()        # every socket is listening
            for q in range(W):
                if q == r:
                    continue
                want = len(names) + (len(multicast) if q == 0 else 0)
                with socket.socket(socket.AF_UNIX, socket.SOCK_STREAM) as c:
                    c.connect(path(q))
                    _, fds, _, _ = socket.recv_fds(c, 16, want)
                if len(fds) != want:
                    raise N.NativeError(f"rank {r}: expected {want} file descriptors from rank {q}, got {len(fds)}")
                peer_fds[q] = list(fds)
            th.join(timeout=60)
            srv.close()
            env.barrier()
            try:
                os.unlink(path(r))
            except OSError:
                pass
        for i, n in enumerate(names):
            ptrs = []
            for q in range(W):
                if q == r:
                    ptrs.append(self.local_ptr[n])
                else:
                    p, h = N.vmm_import(peer_fds[q][i], self.sizes[n])
                    self._handles[n][q] = h
                    ptrs.append(p)
            self.peer_ptr[n] = ptrs
            self.tables[n] = torch.tensor(ptrs, dtype=torch.int64, device=env.device)
        if multicast:
            for j, n in enumerate(multicast):
                if r != 0:
                    self._mc_handles[n] = N.mc_import(peer_fds[0][len(names) + j])
                N.mc_add_device(self._mc_handles[n])
            env.barrier()        # every device is part of every multicast object before anyone binds
            for n in multicast:
                self.mc_ptr[n] = N.mc_bind_map(self._mc_handles[n], self._handles[n][r], self.sizes[n])
        for fd in my_fds + mc_fds + [f for fds in peer_fds.values() for f in fds]:
            N.lib().ftl_fd_close(fd)
        torch.cuda.synchronize()
        env.barrier()

    def tensor(self, name: str, dtype: torch.dtype) -> torch.Tensor:
        holder = _RawCuda(self.local_ptr[name], self._req[name])
        self._holders.append(holder)
        return torch.as_tensor(holder, device=self.env.device).view(dtype)

    def close(self) -> None:
        if not self.local_ptr:
            return
        env = self.env
        env.barrier()
        for n, p in self.mc_ptr.items():
            N.vmm_unmap(p, self.sizes[n], self._mc_handles.get(n, 0))
        for n, ptrs in self.peer_ptr.items():
            for q, p in enumerate(ptrs):
                if q != env.rank:
                    N.vmm_unmap(p, self.sizes[n], self._handles[n][q])
        env.barrier()
        for n, p in self.local_ptr.items():
            N.vmm_unmap(p, self.sizes[n], self._handles[n][env.rank])
        self.local_ptr, self.peer_ptr, self.tables, self.mc_ptr, self._holders = {}, {}, {}, {}, []


def make_symmetric(env: DistEnv, sizes: Dict[str, int], multicast=()) -> SymmetricBuffers:
    """FTL_SYMM=ipc (default): cudaMalloc + CUDA IPC.  FTL_SYMM=vmm, or any multicast request: the virtual-memory heap."""
    import os
    if multicast or os.environ.get("FTL_SYMM", "ipc") == "vmm":
        return VmmSymmetricBuffers(env, sizes, multicast)
    return SymmetricBuffers(env, sizes)
