"""Symmetric buffers and teams: the plumbing under the fused collectives.

The reference gets its communication from torch (train.py:58 `init_process_group('nccl', 'env://')`, the DDP
reducer, torch.cuda.comm).  Here torch.distributed is only used to exchange 64-byte IPC handles; the data
path is kernels reading and writing peer memory over NVLink (csrc/comm/fused_allreduce.cu).

Two ways to form a team:
  * `ProcessTeam`  - one process per GPU (DDP, pipeline): buffers are cudaMalloc'ed by the native library,
                     exported with CUDA IPC and imported by every peer.
  * `DeviceTeam`   - one process, several GPUs (DataParallel): ordinary tensors + peer access.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import torch
import torch.distributed as dist

from ..ops import native as N


class _RawCuda:
    """Adapter exposing a raw device pointer through __cuda_array_interface__ so torch can view it."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def _bind():
    L = N.lib()
    if getattr(L, "_dpt_symm_bound", False):
        return L
    vp = C.c_void_p
    L.dpt_symm_alloc.argtypes = [C.c_size_t, C.POINTER(vp)]
    L.dpt_symm_alloc.restype = C.c_int
    L.dpt_symm_free.argtypes = [vp]
    L.dpt_symm_free.restype = C.c_int
    L.dpt_ipc_export.argtypes = [vp, C.c_char_p]
    L.dpt_ipc_export.restype = C.c_int
    L.dpt_ipc_import.argtypes = [C.c_char_p, C.POINTER(vp)]
    L.dpt_ipc_import.restype = C.c_int
    L.dpt_ipc_close.argtypes = [vp]
    L.dpt_ipc_close.restype = C.c_int
    L._dpt_symm_bound = True
    return L


class SymmetricBuffer:
    """`nbytes` of zeroed device memory on every rank of `group`, mapped into every rank's address space."""

    def __init__(self, nbytes: int, device: torch.device, group: Optional[dist.ProcessGroup] = None):
        L = _bind()
        self.nbytes = int(nbytes)
        self.device = torch.device(device)
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        with torch.cuda.device(self.device):
            p = C.c_void_p()
            N.check(L.dpt_symm_alloc(self.nbytes, C.byref(p)))
            self.local_ptr = p.value
            self.ptrs: List[int] = [0] * self.world
            self.ptrs[self.rank] = self.local_ptr
            self._imported: List[int] = []
            if self.world > 1:
                handle = C.create_string_buffer(64)
                N.check(L.dpt_ipc_export(self.local_ptr, handle))
                mine = torch.tensor(list(handle.raw), dtype=torch.uint8, device=self.device)
                allh = torch.empty(self.world * 64, dtype=torch.uint8, device=self.device)
                dist.all_gather_into_tensor(allh, mine, group=group)
                allh = allh.cpu().numpy().tobytes()
                for r in range(self.world):
                    if r == self.rank:
                        continue
                    q = C.c_void_p()
                    N.check(L.dpt_ipc_import(allh[64 * r:64 * (r + 1)], C.byref(q)))
                    self.ptrs[r] = q.value
                    self._imported.append(q.value)
                dist.barrier(group=group)
        self._raw = torch.as_tensor(_RawCuda(self.local_ptr, self.nbytes), device=self.device)

    def tensor(self, dtype: torch.dtype, numel: Optional[int] = None, offset_bytes: int = 0) -> torch.Tensor:
        """A view of this rank's copy."""
        esz = torch.empty((), dtype=dtype).element_size()
        numel = (self.nbytes - offset_bytes) // esz if numel is None else numel
        return self._raw[offset_bytes:offset_bytes + numel * esz].view(dtype)

    def close(self) -> None:
        L = _bind()
        for q in self._imported:
            L.dpt_ipc_close(q)
        self._imported = []
        if self.local_ptr:
            self._raw = None
            L.dpt_symm_free(self.local_ptr)
            self.local_ptr = 0


def make_team_struct(rank: int, world: int, pad_ptrs: Sequence[int], epoch_ptr: int) -> N.Team:
    t = N.Team()
    t.rank, t.world = rank, world
    for r in range(world):
        t.pads[r] = pad_ptrs[r]
    t.epoch = epoch_ptr
    return t


class ProcessTeam:
    """All ranks of a process group (one GPU each) as a native team."""

    def __init__(self, device: torch.device, group: Optional[dist.ProcessGroup] = None):
        self.device = torch.device(device)
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        if self.world > N.MAX_PEERS:
            raise ValueError(f"teams are limited to {N.MAX_PEERS} ranks (one NVSwitch domain)")
        self.pad = SymmetricBuffer(N.PAD_BYTES, self.device, group)
        self.epoch = torch.zeros(N.EPOCH_BYTES // 4, dtype=torch.int32, device=self.device)
        self.struct = make_team_struct(self.rank, self.world, self.pad.ptrs, self.epoch.data_ptr())

    def buffer(self, nbytes: int) -> SymmetricBuffer:
        return SymmetricBuffer(nbytes, self.device, self.group)


class DeviceTeam:
    """Several devices of ONE process as a native team (peer access instead of IPC)."""

    def __init__(self, devices: Sequence[int]):
        self.devices = [int(d) for d in devices]
        self.world = len(self.devices)
        if self.world > N.MAX_PEERS:
            raise ValueError(f"teams are limited to {N.MAX_PEERS} devices")
        L = N.lib()
        for a in self.devices:
            for b in self.devices:
                if a != b:
                    N.check(L.dpt_enable_peer_access(a, b))
        self.pads = [torch.zeros(N.PAD_BYTES // 4, dtype=torch.int32, device=f"cuda:{d}") for d in self.devices]
        self.epochs = [torch.zeros(N.EPOCH_BYTES // 4, dtype=torch.int32, device=f"cuda:{d}") for d in self.devices]
        pad_ptrs = [p.data_ptr() for p in self.pads]
        self.structs = [make_team_struct(r, self.world, pad_ptrs, self.epochs[r].data_ptr()) for r in range(self.world)]
