"""Set-up plumbing of the row-sharded engine: who the ranks are and how they swap a few small host objects.

Two carriers with one interface:

* `DistComm`     -- torch.distributed (NCCL on the GPUs, gloo in the CPU tests): one process per GPU, the production
                    arrangement (the reference itself is single-process).
* `LoopbackComm` -- N ranks as N Python threads of ONE process on ONE GPU.  Every rank is a full ShardedCtrEngine with
                    its own shard, slab, streams and id plans; the slabs are plain device allocations whose addresses
                    are handed over directly (same address space: no CUDA IPC), and a barrier is a host rendezvous plus
                    stream-to-stream event waits instead of the spinning `ctr_peer_barrier` kernel (kernels of
                    different ranks that wait on one another must not share a GPU: nothing guarantees co-residency).
                    Everything else -- routes, fused gather/push and finalize/push kernels, owner-side merge and Adam,
                    the dense-gradient reduction out of the slabs, id-plan prefetch -- is the production code path.
                    This is what lets a 1-GPU box run the world-2/4/8 parity check of the step
                    (tests/test_gpu_sharded_loopback.py).
"""
from __future__ import annotations

import threading
from typing import Any, Callable, List, Optional

import torch


class DistComm:
    is_loopback = False

    def __init__(self, group=None):
        import torch.distributed as dist
        if not dist.is_initialized():
            raise RuntimeError("ShardedCtrEngine needs an initialised torch.distributed process group")
        self.dist, self.group = dist, group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)

    def all_gather_object(self, obj: Any) -> List[Any]:
        out = [None] * self.world
        self.dist.all_gather_object(out, obj, group=self.group)
        return out

    def broadcast_object(self, obj: Any, src: int = 0) -> Any:
        box = [obj]
        self.dist.broadcast_object_list(box, src=src, group=self.group)
        return box[0]

    def barrier(self) -> None:
        self.dist.barrier(self.group)

    def all_gather_tensor(self, t: torch.Tensor) -> List[torch.Tensor]:
        parts = [torch.empty_like(t) for _ in range(self.world)]
        self.dist.all_gather(parts, t.contiguous(), group=self.group)
        return parts


class LoopbackWorld:
    """shared state of N loopback ranks; `comm(rank)` gives rank `rank` its view"""

    def __init__(self, world: int, timeout_s: float = 120.0):
        self.world = int(world)
        self.timeout = timeout_s
        self._bar = threading.Barrier(self.world)
        self._slots: List[Any] = [None] * self.world
        self._events: List[Optional[torch.cuda.Event]] = [None] * self.world

    def comm(self, rank: int) -> "LoopbackComm":
        return LoopbackComm(self, rank)

    def run(self, fn: Callable[["LoopbackComm"], Any], device: int = 0) -> List[Any]:
        """fn(comm) on one thread per rank; returns the results, re-raises the first failure (a failing rank breaks the
        rendezvous so the others do not hang)"""
        out: List[Any] = [None] * self.world
        err: List[Optional[BaseException]] = [None] * self.world

        def body(r: int) -> None:
            try:
                torch.cuda.set_device(device)
                out[r] = fn(self.comm(r))
            except BaseException as e:          # noqa: BLE001 -- reported to the caller below
                err[r] = e
                self._bar.abort()
        th = [threading.Thread(target=body, args=(r,), daemon=True) for r in range(self.world)]
        for t in th:
            t.start()
        for t in th:
            t.join()
        real = [e for e in err if e is not None and not isinstance(e, threading.BrokenBarrierError)]
        if real or any(err):
            raise (real or [e for e in err if e is not None])[0]
        return out


class LoopbackComm:
    is_loopback = True
    group = None

    def __init__(self, w: LoopbackWorld, rank: int):
        self.w, self.world, self.rank = w, w.world, int(rank)

    def _wait(self) -> None:
        self.w._bar.wait(timeout=self.w.timeout)

    def all_gather_object(self, obj: Any) -> List[Any]:
        self.w._slots[self.rank] = obj
        self._wait()
        out = list(self.w._slots)
        self._wait()                 # nobody overwrites a slot before everyone has read
        return out

    def broadcast_object(self, obj: Any, src: int = 0) -> Any:
        return self.all_gather_object(obj)[src]

    def barrier(self) -> None:
        self._wait()

    def all_gather_tensor(self, t: torch.Tensor) -> List[torch.Tensor]:
        torch.cuda.current_stream().synchronize()
        return self.all_gather_object(t.contiguous().clone())

    def device_barrier(self, stream: int) -> None:
        """What `ctr_peer_barrier` guarantees, without a spinning kernel: everything the ranks enqueued on their
        `stream`s before this point completes before anything they enqueue on them afterwards."""
        ts = torch.cuda.ExternalStream(stream) if stream else torch.cuda.current_stream()
        ev = torch.cuda.Event()
        ev.record(ts)
        self.w._events[self.rank] = ev
        self._wait()
        for r in range(self.world):
            if r != self.rank:
                ts.wait_event(self.w._events[r])
        self._wait()                 # every wait is enqueued before any rank records its next event


def as_comm(group):
    """a torch.distributed process group (or None for the default group) or a LoopbackComm -> comm object"""
    if isinstance(group, (LoopbackComm, DistComm)):
        return group
    return DistComm(group)
