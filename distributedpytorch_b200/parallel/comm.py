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
import os
import atexit
import shutil
import socket
import tempfile
import struct
from typing import Dict, List, Optional, Sequence

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


def _bind_vmm():
    L = N.lib()
    if getattr(L, "_dpt_vmm_bound", False):
        return L
    vp, ull, sz, ip = C.c_void_p, C.c_ulonglong, C.c_size_t, C.POINTER(C.c_int)
    L.dpt_vmm_query.argtypes = [C.c_int, ip, ip]
    L.dpt_vmm_round_size.argtypes = [C.c_int, sz, C.c_int, C.POINTER(sz)]
    L.dpt_vmm_alloc.argtypes = [C.c_int, sz, C.POINTER(vp), C.POINTER(ull), ip]
    L.dpt_vmm_import.argtypes = [C.c_int, C.c_int, sz, C.POINTER(vp), C.POINTER(ull)]
    L.dpt_vmm_map_local.argtypes = [C.c_int, ull, sz, C.POINTER(vp)]
    L.dpt_vmm_grant.argtypes = [vp, sz, C.c_int]
    L.dpt_vmm_unmap.argtypes = [vp, sz]
    L.dpt_vmm_release.argtypes = [ull]
    L.dpt_close_fd.argtypes = [C.c_int]
    L.dpt_mc_create.argtypes = [C.c_int, sz, C.POINTER(ull), ip]
    L.dpt_mc_import.argtypes = [C.c_int, C.POINTER(ull)]
    L.dpt_mc_add_device.argtypes = [ull, C.c_int]
    L.dpt_mc_bind.argtypes = [ull, C.c_int, ull, sz]
    L.dpt_mc_map.argtypes = [ull, C.c_int, sz, C.POINTER(vp)]
    for f in ("query", "round_size", "alloc", "import", "map_local", "grant", "unmap", "release"):
        getattr(L, "dpt_vmm_" + f).restype = C.c_int
    for f in ("create", "import", "add_device", "bind", "map"):
        getattr(L, "dpt_mc_" + f).restype = C.c_int
    L._dpt_vmm_bound = True
    return L


def multicast_supported(device: torch.device) -> bool:
    """True if `device` can do NVSwitch multicast / in-fabric reduction (NVLS) through the VMM API."""
    L = _bind_vmm()
    vmm, mc = C.c_int(0), C.c_int(0)
    N.check(L.dpt_vmm_query(torch.device(device).index or 0, C.byref(vmm), C.byref(mc)))
    return bool(mc.value)


_fd_round = 0
_fd_dir: Optional[str] = None


def _fd_directory(rank: int, group) -> str:
    """A directory only this user can enter (mkdtemp, mode 0700), created by rank 0 of the group and announced to the
    other ranks through the process group: the socket paths are unpredictable, cannot be pre-created by another local
    user, and two jobs of the same user never share them."""
    global _fd_dir
    if _fd_dir is None:
        box = [tempfile.mkdtemp(prefix="dpt_fd_") if rank == 0 else None]
        dist.broadcast_object_list(box, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        _fd_dir = box[0]
        if rank == 0:
            atexit.register(shutil.rmtree, _fd_dir, True)
    return _fd_dir


def _peer_uid(conn: socket.socket) -> int:
    pid, uid, gid = struct.unpack("3i", conn.getsockopt(socket.SOL_SOCKET, socket.SO_PEERCRED, struct.calcsize("3i")))
    return uid


def _exchange_fds(rank: int, world: int, sends: Dict[int, int], n_expect: int, group) -> Dict[int, int]:
    """Pass file descriptors between the ranks of one node over AF_UNIX sockets (SCM_RIGHTS): `sends` maps a
    destination rank to the fd it gets; returns {source rank: received fd} with `n_expect` entries.  Every rank of the
    group must call this the same number of times (the socket names are numbered per call).  A connection is only
    honoured if the kernel says the peer runs under this uid (SO_PEERCRED), names a rank in range, and that rank has
    not delivered already: what arrives is mapped as peer GPU memory, so it must come from a rank of this job."""
    global _fd_round
    _fd_round += 1
    base = os.path.join(_fd_directory(rank, group), f"r{_fd_round}")
    mine = f"{base}_{rank}.sock"
    srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    srv.bind(mine)
    srv.listen(world)
    dist.barrier(group=group)  # every listener exists before anyone connects
    for dst, fd in sends.items():
        c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        c.connect(f"{base}_{dst}.sock")
        socket.send_fds(c, [struct.pack("i", rank)], [fd])
        c.close()
    got: Dict[int, int] = {}
    srv.settimeout(120)
    while len(got) < n_expect:
        conn, _ = srv.accept()
        try:
            if _peer_uid(conn) != os.getuid():
                continue
            msg, fds, _, _ = socket.recv_fds(conn, 4, 1)
            src = struct.unpack("i", msg)[0] if len(msg) == 4 else -1
            if not fds or not 0 <= src < world or src == rank or src in got:
                for f in fds:
                    os.close(f)
                continue
            got[src] = fds[0]
        finally:
            conn.close()
    dist.barrier(group=group)
    srv.close()
    os.unlink(mine)
    return got


class MulticastBuffer:
    """A SymmetricBuffer on CUDA VMM allocations with an NVLS multicast mapping on top: `ptrs[r]` is rank r's copy
    (unicast, as in SymmetricBuffer), `mc_ptr` addresses all copies at once - `multimem.st` there writes every
    rank's copy, `multimem.ld_reduce` returns the sum over ranks, added inside the NVSwitch (csrc/comm/vmm.cu)."""

    def __init__(self, nbytes: int, device: torch.device, group: Optional[dist.ProcessGroup] = None):
        L = _bind_vmm()
        self.device = torch.device(device)
        self.group = group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        dev = self.device.index or 0
        with torch.cuda.device(self.device):
            sz = C.c_size_t()
            N.check(L.dpt_vmm_round_size(dev, int(nbytes), self.world, C.byref(sz)))
            self.nbytes, self.alloc_bytes = int(nbytes), sz.value
            p, h, fd = C.c_void_p(), C.c_ulonglong(), C.c_int(-1)
            N.check(L.dpt_vmm_alloc(dev, self.alloc_bytes, C.byref(p), C.byref(h), C.byref(fd)))
            self.local_ptr, self._handle = p.value, h.value
            self.ptrs: List[int] = [0] * self.world
            self.ptrs[self.rank] = self.local_ptr
            self._peer_handles: List[int] = []
            peers = [r for r in range(self.world) if r != self.rank]
            got = _exchange_fds(self.rank, self.world, {r: fd.value for r in peers}, len(peers), group)
            L.dpt_close_fd(fd.value)
            for r, pfd in sorted(got.items()):
                q, qh = C.c_void_p(), C.c_ulonglong()
                N.check(L.dpt_vmm_import(dev, pfd, self.alloc_bytes, C.byref(q), C.byref(qh)))
                L.dpt_close_fd(pfd)
                self.ptrs[r] = q.value
                self._peer_handles.append(qh.value)
            # the multicast object: created by rank 0, imported by the others, every device added before anyone binds
            mc = C.c_ulonglong()
            if self.rank == 0:
                mfd = C.c_int(-1)
                N.check(L.dpt_mc_create(self.world, self.alloc_bytes, C.byref(mc), C.byref(mfd)))
                _exchange_fds(0, self.world, {r: mfd.value for r in peers}, 0, group)
                L.dpt_close_fd(mfd.value)
            else:
                got = _exchange_fds(self.rank, self.world, {}, 1, group)
                N.check(L.dpt_mc_import(got[0], C.byref(mc)))
                L.dpt_close_fd(got[0])
            self._mc = mc.value
            N.check(L.dpt_mc_add_device(self._mc, dev))
            dist.barrier(group=group)
            N.check(L.dpt_mc_bind(self._mc, dev, self._handle, self.alloc_bytes))
            dist.barrier(group=group)
            m = C.c_void_p()
            N.check(L.dpt_mc_map(self._mc, dev, self.alloc_bytes, C.byref(m)))
            self.mc_ptr = m.value
            torch.cuda.synchronize()
            dist.barrier(group=group)
        self._raw = torch.as_tensor(_RawCuda(self.local_ptr, self.nbytes), device=self.device)

    tensor = SymmetricBuffer.tensor

    def close(self) -> None:
        """Unmap and release everything this rank mapped: the multicast view and handle, the peers' allocations, then its
        own (a collective in spirit: every rank should call it, after its last kernel on the buffer has finished).
        Without it the mappings simply live until the process exits."""
        if self.local_ptr == 0:
            return
        L = _bind_vmm()
        self._raw = None
        with torch.cuda.device(self.device):
            torch.cuda.synchronize()
            if dist.is_initialized():
                dist.barrier(group=self.group)  # nobody unmaps while a peer's kernel may still touch the memory
            L.dpt_vmm_unmap(self.mc_ptr, self.alloc_bytes)
            L.dpt_vmm_release(self._mc)
            for r, q in enumerate(self.ptrs):
                if r != self.rank and q:
                    L.dpt_vmm_unmap(q, self.alloc_bytes)
            for h in self._peer_handles:
                L.dpt_vmm_release(h)
            L.dpt_vmm_unmap(self.local_ptr, self.alloc_bytes)
            L.dpt_vmm_release(self._handle)
        self.local_ptr, self.mc_ptr, self._mc, self._handle, self._peer_handles = 0, 0, 0, 0, []
        self.ptrs = [0] * self.world


class DeviceMulticastBuffer:
    """MulticastBuffer for the devices of ONE process (DataParallel): one VMM allocation per device, each readable and
    writable by every device of the team under a single address (`ptrs[r]`), all bound to one multicast object that every
    device maps (`mc_ptrs[r]` is device r's view).  No file descriptors travel: the handles are used in place."""

    def __init__(self, nbytes: int, devices: Sequence[int]):
        prev = torch.cuda.current_device()  # the native calls below switch the current device
        try:
            self._build(nbytes, devices)
        finally:
            torch.cuda.set_device(prev)

    def _build(self, nbytes: int, devices: Sequence[int]) -> None:
        L = _bind_vmm()
        self.devices = [int(d) for d in devices]
        self.world = len(self.devices)
        sz = C.c_size_t()
        N.check(L.dpt_vmm_round_size(self.devices[0], int(nbytes), self.world, C.byref(sz)))
        self.nbytes, self.alloc_bytes = int(nbytes), sz.value
        self.ptrs: List[int] = []
        self._handles: List[int] = []
        for d in self.devices:
            p, h, fd = C.c_void_p(), C.c_ulonglong(), C.c_int(-1)
            N.check(L.dpt_vmm_alloc(d, self.alloc_bytes, C.byref(p), C.byref(h), C.byref(fd)))
            L.dpt_close_fd(fd.value)
            self.ptrs.append(p.value)
            self._handles.append(h.value)
            for e in self.devices:
                if e != d:
                    N.check(L.dpt_vmm_grant(p.value, self.alloc_bytes, e))
        mc, mfd = C.c_ulonglong(), C.c_int(-1)
        N.check(L.dpt_mc_create(self.world, self.alloc_bytes, C.byref(mc), C.byref(mfd)))
        L.dpt_close_fd(mfd.value)
        self._mc = mc.value
        for d in self.devices:  # every device joins before anyone binds
            N.check(L.dpt_mc_add_device(self._mc, d))
        for d, h in zip(self.devices, self._handles):
            N.check(L.dpt_mc_bind(self._mc, d, h, self.alloc_bytes))
        self.mc_ptrs: List[int] = []
        for d in self.devices:
            m = C.c_void_p()
            N.check(L.dpt_mc_map(self._mc, d, self.alloc_bytes, C.byref(m)))
            self.mc_ptrs.append(m.value)
        self._raw = []
        for d, p in zip(self.devices, self.ptrs):
            with torch.cuda.device(d):
                torch.cuda.synchronize()
                self._raw.append(torch.as_tensor(_RawCuda(p, self.nbytes), device=f"cuda:{d}"))

    def tensor(self, r: int, dtype: torch.dtype, numel: Optional[int] = None) -> torch.Tensor:
        """A view of device r's copy."""
        esz = torch.empty((), dtype=dtype).element_size()
        numel = self.nbytes // esz if numel is None else numel
        return self._raw[r][:numel * esz].view(dtype)

    def close(self) -> None:
        if not self.ptrs:
            return
        L = _bind_vmm()
        self._raw = []
        for d in self.devices:
            torch.cuda.synchronize(d)
        for m in self.mc_ptrs:
            L.dpt_vmm_unmap(m, self.alloc_bytes)
        L.dpt_vmm_release(self._mc)
        for p, h in zip(self.ptrs, self._handles):
            L.dpt_vmm_unmap(p, self.alloc_bytes)
            L.dpt_vmm_release(h)
        self.ptrs, self.mc_ptrs, self._handles, self._mc = [], [], [], 0


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

    def mc_buffer(self, nbytes: int) -> "MulticastBuffer":
        """A symmetric buffer that also has an NVLS multicast address (needs multicast_supported(device))."""
        return MulticastBuffer(nbytes, self.device, self.group)


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

    def multicast_supported(self) -> bool:
        return self.world >= 2 and all(multicast_supported(torch.device("cuda", d)) for d in self.devices)

    def mc_buffer(self, nbytes: int) -> DeviceMulticastBuffer:
        """One buffer per device plus an NVLS multicast address over all of them (needs multicast_supported())."""
        return DeviceMulticastBuffer(nbytes, self.devices)
