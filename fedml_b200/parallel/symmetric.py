"""Symmetric (peer-visible) device memory for the one-process-per-GPU runtime.

Every rank allocates one region of the same size; after ``rendezvous`` each rank knows the
virtual address at which every peer's region is mapped in *its own* address space, so kernels can
load/store peer memory directly over NVLink/NVSwitch.  Two providers:

  1. ``torch.distributed._symmetric_memory`` (CUDA VMM + fabric/posix handle exchange; also yields an
     NVLS multicast address when the platform supports it);
  2. CUDA IPC through this package's native library (``fb_dev_malloc`` / ``fb_ipc_*`` in csrc/host.cu),
     used when (1) fails — e.g. restricted containers.

The reference's counterpart is ``ProcessGroupManager`` (cross_silo/hierarchical/process_group_manager.py:30-34),
which only creates an NCCL/Gloo group; it has no notion of peer-mapped memory.
"""
from __future__ import annotations

import ctypes as C
import logging
from typing import Dict, List, Optional, Tuple

import torch
import torch.distributed as dist


class _RawCudaBuffer:
    """Adapter exposing a raw device pointer through ``__cuda_array_interface__``."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3,
                                         "strides": None}


class SymmetricRegion:
    def __init__(self, nbytes: int, device: torch.device, group=None, provider: str = "auto"):
        self.device = torch.device(device)
        self.nbytes = (int(nbytes) + 255) // 256 * 256
        self.group = group
        self.world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        self.rank = dist.get_rank(group) if self.world > 1 else 0
        self.multicast_ptr = 0
        self._cursor = 0
        self._names: Dict[str, Tuple[int, int]] = {}
        self._keep = []
        self.provider = None
        if self.world == 1:
            self.local = torch.zeros(self.nbytes, dtype=torch.uint8, device=self.device)
            self.peer_ptrs = [self.local.data_ptr()]
            self.provider = "local"
            return
        errors = []
        order = ["torch", "ipc"] if provider == "auto" else [provider]
        # every rank must take the same branch: agree on success collectively
        for prov in order:
            ok, err = True, ""
            try:
                getattr(self, f"_init_{prov}")()
            except Exception as exc:  # noqa: BLE001
                ok, err = False, f"{type(exc).__name__}: {exc}"
            flags = [None] * self.world
            dist.all_gather_object(flags, ok, group=self.group)
            if all(flags):
                self.provider = prov
                break
            errors.append(f"{prov}: {err or 'failed on a peer rank'}")
        if self.provider is None:
            raise RuntimeError("no symmetric-memory provider worked: " + " | ".join(errors))
        if self.rank == 0:
            logging.info("symmetric memory via %s (%d B/rank, multicast=%s)", self.provider, self.nbytes,
                         bool(self.multicast_ptr))

    # ------------------------------------------------------------------ providers
    def _init_torch(self):
        import torch.distributed._symmetric_memory as symm
        grp = self.group if self.group is not None else dist.group.WORLD
        t = symm.empty(self.nbytes, dtype=torch.uint8, device=self.device)
        hdl = symm.rendezvous(t, grp)
        t.zero_()
        self.local = t
        self.peer_ptrs = [int(p) for p in hdl.buffer_ptrs]
        try:
            self.multicast_ptr = int(hdl.multicast_ptr) if hdl.has_multicast_support(self.device.type, self.device.index) \
                else 0
        except Exception:  # noqa: BLE001
            try:
                self.multicast_ptr = int(hdl.multicast_ptr or 0)
            except Exception:  # noqa: BLE001
                self.multicast_ptr = 0
        self._keep.append(hdl)

    def _init_ipc(self):
        from ..ops.native import check, lib
        L = lib()
        torch.cuda.set_device(self.device)
        base = C.c_void_p()
        check(L.fb_dev_malloc(C.byref(base), C.c_int64(self.nbytes)), "fb_dev_malloc")
        hb = int(L.fb_ipc_handle_bytes())
        handle = (C.c_ubyte * hb)()
        check(L.fb_ipc_get_handle(base, handle), "fb_ipc_get_handle")
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(handle), group=self.group)
        ptrs = []
        for q, h in enumerate(handles):
            if q == self.rank:
                ptrs.append(int(base.value)); continue
            buf = (C.c_ubyte * hb).from_buffer_copy(h)
            out = C.c_void_p()
            check(L.fb_ipc_open_handle(buf, C.byref(out)), "fb_ipc_open_handle")
            ptrs.append(int(out.value))
        self.peer_ptrs = ptrs
        self.local = torch.as_tensor(_RawCudaBuffer(int(base.value), self.nbytes), device=self.device)
        self._keep.append(base)

    # ------------------------------------------------------------------ carving
    def carve(self, name: str, nbytes: int, align: int = 256) -> int:
        off = (self._cursor + align - 1) // align * align
        if off + nbytes > self.nbytes:
            raise RuntimeError(f"symmetric region exhausted carving {name!r}")
        self._cursor = off + nbytes
        self._names[name] = (off, nbytes)
        return off

    def tensor(self, name: str, dtype=torch.float32) -> torch.Tensor:
        off, nbytes = self._names[name]
        return self.local[off:off + nbytes].view(dtype)

    def peer_ptr(self, name: str, q: int) -> int:
        return self.peer_ptrs[q] + self._names[name][0]

    def peer_table(self, name: str) -> torch.Tensor:
        """Device int64 tensor [G]: address of ``name`` on every rank, as mapped into this process."""
        return torch.tensor([self.peer_ptr(name, q) for q in range(self.world)], dtype=torch.int64).to(self.device)

    def barrier(self):
        if self.world > 1:
            torch.cuda.synchronize(self.device)
            dist.barrier(group=self.group)
