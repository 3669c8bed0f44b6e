"""Peer-mapped device memory for the row-sharded engine: one slab per rank, exported with CUDA IPC and mapped by
every other rank of the box, so kernels can store into a peer's staging buffers over NVLink 5 / NVSwitch
(include/ctr_b200.h, "peer memory").  torch.distributed only carries the 64-byte handles."""
from __future__ import annotations

import ctypes as C
from typing import List, Sequence

import torch

from ._lib import IPC_HANDLE_BYTES, MAX_PEERS, Handle, PeerRoute


class _Region:
    """a window of the own slab, visible to torch through the CUDA array interface"""

    def __init__(self, owner, addr: int, shape, typestr: str):
        self._owner = owner
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (addr, False), "version": 2,
                                         "strides": None}


class PeerSlab:
    def __init__(self, h: Handle, nbytes: int, comm, device: torch.device):
        """comm: deep_learning_b200.comm.DistComm (one process per GPU: CUDA IPC) or LoopbackComm (ranks are threads of
        one process on one GPU: the slabs share an address space and their addresses are handed over as they are)"""
        world, rank = comm.world, comm.rank
        if world > MAX_PEERS:
            raise ValueError("PeerSlab: at most %d ranks" % MAX_PEERS)
        self.h, self.world, self.rank, self.nbytes, self.device = h, world, rank, int(nbytes), device
        self.loopback = bool(comm.is_loopback)
        own, handle = C.c_void_p(), C.create_string_buffer(IPC_HANDLE_BYTES)
        h.invoke("ctr_peer_alloc", self.nbytes, C.byref(own), handle)
        self.own = own.value
        self.ptrs: List[int] = []
        if self.loopback:
            self.ptrs = [int(p) for p in comm.all_gather_object(self.own)]
        else:
            handles: List[bytes] = comm.all_gather_object(bytes(handle.raw))
            for r in range(world):
                if r == rank:
                    self.ptrs.append(self.own)
                    continue
                p = C.c_void_p()
                h.invoke("ctr_peer_open", C.create_string_buffer(handles[r], IPC_HANDLE_BYTES), C.byref(p))
                self.ptrs.append(p.value)
        comm.barrier()                   # nobody stores into a slab before everyone has zeroed and mapped theirs
        self._closed = False

    def tensor(self, offset: int, shape: Sequence[int], dtype=torch.float32) -> torch.Tensor:
        typestr = {torch.float32: "<f4", torch.int32: "<i4", torch.uint8: "|u1"}[dtype]
        return torch.as_tensor(_Region(self, self.own + int(offset), shape, typestr), device=self.device)

    def route(self, seg_end: Sequence[int], dst_row: Sequence[int], dst_off: int, dyn: int = 0) -> PeerRoute:
        """placement of a routed list in the peers' buffers at byte offset dst_off; `dyn` (device address of a route block
        written by ctr_peer_plan_routes) makes the kernels take segment ends / destination rows from the device"""
        rt = PeerRoute()
        rt.world, rt.rank, rt.dst_off = self.world, self.rank, int(dst_off)
        for p in range(self.world):
            rt.slab[p], rt.seg_end[p], rt.dst_row[p] = self.ptrs[p], int(seg_end[p]), int(dst_row[p])
        rt.dyn = int(dyn) if dyn else None
        return rt

    def close(self) -> None:
        if self._closed:
            return
        self._closed = True
        torch.cuda.synchronize(self.device)
        for r, p in enumerate(self.ptrs):
            if r != self.rank and not self.loopback:
                self.h.invoke("ctr_peer_close", p)
        self.h.invoke("ctr_peer_free", self.own)
