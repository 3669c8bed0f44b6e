"""ShardedCtrEngine: one process per GPU; embedding tables row-sharded (owner = id mod N), samples data-parallel,
dense tower replicated with one sum-allreduce of its gradients (schedule in shard.py).

The arithmetic is the single-GPU engine's: the fused gather+FM kernel simply reads the rows it needs from the
staging buffer filled by the vector all-to-all (indexed by each lookup's position in this rank's unique list)
instead of from the table, and the Adam update runs on the owner after the gradient all-to-all.  The loss is the
mean over the GLOBAL batch, so N ranks on N shards of a batch reproduce one GPU on the concatenated batch.

Everything that depends only on the batch ids -- sort / de-dup / per-owner counts, the counts and key all-to-alls
and the owner-side sort of the requested keys -- is an "id plan".  Given `next_batch`, the plan of step i+1 is
built on a side stream underneath the backward pass of step i, so the critical path of a step holds only what
depends on the parameters: owner gather -> vector all-to-all -> compute -> gradient all-to-all -> owner Adam.
"""
from __future__ import annotations

from typing import Dict, Optional

import ctypes as C
import os

import numpy as np
import torch
import torch.distributed as dist

from . import shard as _shard
from .comm import as_comm
from ._lib import AdamHP, EmbedCfg, PLAN_GRAD, PLAN_NREQ, PLAN_U, PLAN_VEC, PLAN_WORDS, ptr
from .peer import PeerSlab
from .engine import CtrEngine, _ALIGN, _ru
from .spec import ModelSpec


class _IdPlan:
    """Every id-derived structure of one batch (requester side and owner side)."""

    def __init__(self, eng: "ShardedCtrEngine", slot: int = 0):
        self.slot = slot
        s, dev, B = eng.spec, eng.dev, eng.Bmax
        n = eng.n_max
        u32 = dict(dtype=torch.int32, device=dev)     # raw 32-bit storage; the library reads it as uint32
        self.ids = torch.zeros(B, max(s.ids_ld, 1), dtype=eng.id_torch, device=dev)
        self.keys, self.perm, self.uidx, self.uniq, self.inv = (torch.zeros(max(n, 1), **u32) for _ in range(5))
        self.seg_start = torch.zeros(n + 2, **u32)
        self.n_unique = torch.zeros(1, **u32)
        self.counts = torch.zeros(eng.world, **u32)
        self.plan_dev = torch.zeros(PLAN_WORDS, **u32)     # routes / n_req / U as ctr_peer_plan_routes leaves them
        self.counts_host = torch.zeros(eng.world * eng.world + 2, dtype=torch.int32).pin_memory()   # matrix + 2 status words
        self.route_v = self.route_g = self.route_k = None
        self.ws_sort = torch.zeros(_ru(max(eng.h.query("ctr_sort_ws_bytes", n), 4 * (n // 2048 + 8)), 256),
                                   dtype=torch.uint8, device=dev)
        self.ev_local, self.ev_ready, self.ev_free = torch.cuda.Event(), torch.cuda.Event(), torch.cuda.Event()
        self.cap = 0
        self.B = self.U = self.n_req = 0
        self.send_counts = self.recv_counts = None
        self.batch_id = None
        self.on_side = False
        self.in_flight = False

    def ensure_owner(self, eng: "ShardedCtrEngine", n_req: int) -> None:
        if n_req <= self.cap:
            return
        cap = max(int(n_req * 1.25), 1024)
        u32 = dict(dtype=torch.int32, device=eng.dev)
        if eng.exchange_mode == "peer":       # peers store their requests straight into this rank's slab
            self.req = eng.slab.tensor(eng._req_off[self.slot], (eng._req_cap,), torch.int32)
        else:
            self.req = torch.zeros(cap, **u32)
        if eng.train_mode:
            self.o_keys, self.o_idx, self.o_uniq, self.o_uidx = (torch.zeros(cap, **u32) for _ in range(4))
            self.o_seg = torch.zeros(cap + 2, **u32)
            self.o_nuniq = torch.zeros(1, **u32)
            self.o_ws = torch.zeros(_ru(4 * (cap // 2048 + 8), 256), dtype=torch.uint8, device=eng.dev)
        self.cap = cap

    def bind_device_routes(self, eng: "ShardedCtrEngine") -> None:
        """peer mode: the three routes of this plan slot, built ONCE -- slab pointers and buffer offsets are static, segment
        ends and destination rows come from the plan block on the device (no host read of the split sizes)"""
        base = self.plan_dev.data_ptr()
        z = [0] * eng.world
        self.route_v = eng.slab.route(z, z, eng._rows_off, dyn=base + 4 * PLAN_VEC)
        self.route_g = eng.slab.route(z, z, eng._g_off, dyn=base + 4 * PLAN_GRAD)
        self.route_k = eng.slab.route(z, z, eng._req_off[self.slot], dyn=base + 4 * PLAN_GRAD)   # keys travel like the gradients
        self.ensure_owner(eng, eng._req_cap)


class ShardedCtrEngine(CtrEngine):
    def __init__(self, spec: ModelSpec, max_batch: int, group=None, **kw):
        # `group`: a torch.distributed process group (None = the default one), or a comm.LoopbackComm (N ranks as N threads
        # of one process on one GPU: the arrangement the 1-GPU test box uses to run the world-N step)
        self.comm = as_comm(group)
        world, rank = self.comm.world, self.comm.rank
        self.group = None if self.comm.is_loopback else group
        # the id plans of step i+1 are built under step i on a side stream: their collectives get a communicator of their
        # own, or they would queue in front of step i's vector / gradient exchanges (one communicator = one FIFO)
        self.group_plan = self.group
        if not self.comm.is_loopback and world > 1 and os.environ.get("CTR_EXCHANGE", "peer") == "nccl":
            ranks = dist.get_process_group_ranks(group) if group is not None else list(range(dist.get_world_size()))
            self.group_plan = dist.new_group(ranks=ranks)
        self.plan = _shard.ShardPlan(spec.table_rows, world)
        self.wire = spec.grad_stride          # floats per row on the wire (k [+1], rounded up to 4): vectors and gradients
        super().__init__(spec, max_batch, world=world, rank=rank, **kw)
        s, B, dev = spec, self.Bmax, self.dev
        f32 = dict(dtype=torch.float32, device=dev)
        self.n_extra = s.C if (s.use_fm and s.fm_cont) else 0
        self.n_max = B * s.ids_ld + self.n_extra
        if not self.train_mode:
            self.side = torch.cuda.Stream(device=dev)
        self._ids_external = True
        self.cur: Optional[_IdPlan] = None
        # "peer": owners / requesters store rows straight into the peer's staging buffer over NVLink (fused kernels);
        # "nccl": gather -> all_to_all_single -> consume
        self.exchange_mode = os.environ.get("CTR_EXCHANGE", "peer" if world > 1 else "nccl")
        if self.exchange_mode not in ("peer", "nccl"):
            raise ValueError("CTR_EXCHANGE must be peer or nccl")
        if self.comm.is_loopback:
            self.exchange_mode = "peer"       # the peer-memory path is the one under test; there is no NCCL between threads
        # peer mode plans its routes on the device (no host read in the step); CTR_PLAN=host keeps the earlier host-side
        # planning (one host wait per step for the split sizes) as an A/B and a fallback
        self.device_plan = self.exchange_mode == "peer" and os.environ.get("CTR_PLAN", "device") == "device"
        if self.exchange_mode != "peer":
            # torch.distributed collectives run on torch's CURRENT stream, not on the engine's sparse stream: the
            # overlapped backward order is only valid where the exchange is a kernel on the stream it is given
            self.bwd_overlap = False
        # CTR_BWD_ORDER=split (peer-memory mode, opt-in): gradient rows are pushed BEFORE the weight-gradient GEMMs and applied
        # behind them (CtrEngine._backward_split).  Parity-green on real peers, but no faster: 1.1007 against 1.0968 ms at
        # N = 2 -- the barrier behind the GEMMs still costs 24 us, so it is not rank skew (profiles/r02m_bwd_split_n2.txt)
        self.bwd_split = (self.exchange_mode == "peer" and self.train_mode and not self.bwd_overlap and
                          os.environ.get("CTR_BWD_ORDER", "serial") == "split")
        self.peer_status = self.status[:2]      # [0] step barriers, [1] id-plan barriers: words of the engine's status array
        self._epoch, self._pepoch, self._sepoch, self._rows_dirty = 0, 0, 0, False
        if self.exchange_mode == "peer":
            # slab layout: arrival words of the step barriers | of the id-plan barriers | count matrix x 2 plan slots |
            # received rows | received gradients | requested keys x 2 plan slots | dense gradients (read by every peer)
            row_bytes, W = self.wire * 4, world
            # three independent sets of arrival words: step barriers on the main stream, id-plan barriers on the side
            # stream, gradient barriers on the sparse stream (they can be in flight at the same time)
            self._flags_off, self._pflags_off, self._sflags_off, self._mat_off = 0, 64, 128, 192
            self._rows_off = _ru(self._mat_off + 2 * W * W * 4, 256)
            self._g_off = self._rows_off + _ru((self.n_max + 1) * row_bytes, 256)
            self._g_cap = self.n_max if self.train_mode else 0
            self._req_cap = max(self.n_max, 1)
            r0 = self._g_off + _ru(max(self._g_cap, 1) * row_bytes, 256)
            self._req_off = [r0, r0 + _ru(self._req_cap * 4, 256)]
            self._dense_off = self._req_off[1] + _ru(self._req_cap * 4, 256)
            n_dense = self.n_dense + _ALIGN
            # last 256 bytes: this rank's share of the step's loss, two slots of 4 floats (read by every peer, early loss)
            self._loss_off = self._dense_off + _ru((n_dense if self.train_mode else 1) * 4, 256)
            self.slab = PeerSlab(self.h, self._loss_off + 256, self.comm, dev)
            self.rows_recv = self.slab.tensor(self._rows_off, (self.n_max + 1, self.wire))
            self.g_recv = self.slab.tensor(self._g_off, (max(self._g_cap, 1), self.wire)) if self.train_mode else None
            self._rt0 = self.slab.route([0] * W, [0] * W, 0)
            self._rt_counts = [self.slab.route([0] * W, [0] * W, self._mat_off + (sl * W * W + rank * W) * 4) for sl in range(2)]
            self._mat_view = [self.slab.tensor(self._mat_off + sl * W * W * 4, (W * W,), torch.int32) for sl in range(2)]
            if self.train_mode:
                # the dense gradients live in the slab: after the gradient barrier every rank sums all replicas' buffers
                # in rank order straight out of peer memory (no collective, bitwise identical results on every rank)
                self.grad = self.slab.tensor(self._dense_off, (n_dense,))
                self.grad_red = torch.zeros(n_dense, **f32)
                self._rt_dense = self.slab.route([0] * W, [0] * W, self._dense_off)
                self.ev_bar, self.ev_red = torch.cuda.Event(), torch.cuda.Event()
                self._loss_slab = self.slab.tensor(self._loss_off, (8,))
                self._rt_loss = [self.slab.route([0] * W, [0] * W, self._loss_off + 16 * sl) for sl in range(2)]
                self._loss_red = torch.zeros(4, **f32)
                self._loss_step = 0
        else:
            self.rows_recv = torch.zeros(self.n_max + 1, self.wire, **f32)
        self.plans = [_IdPlan(self, 0), _IdPlan(self, 1)]
        if self.device_plan:
            for pl in self.plans:
                pl.bind_device_routes(self)
        self.cont_buf = torch.zeros(max(s.C, 1), self.wire, **f32)
        self.grads_send = (torch.zeros(max(self.n_max, 1), self.wire, **f32)
                           if self.train_mode and self.exchange_mode == "nccl" else None)
        self._row_cap = 0
        self._ensure_rows(self.n_max)

    @property
    def U(self) -> int:
        """unique rows of the current batch on this rank (padding run excluded).  Peer mode: only the device knows it
        during the step; this reads it back (a host wait -- for tests and reports, never in the step)"""
        if self.cur is None:
            return 0
        return int(self.cur.plan_dev[PLAN_U].item()) if self.cur.U is None else int(self.cur.U)

    @property
    def n_req(self) -> int:
        """rows the peers requested of this rank for the current batch"""
        if self.cur is None:
            return 0
        return int(self.cur.plan_dev[PLAN_NREQ].item()) if self.cur.n_req is None else int(self.cur.n_req)

    def _ensure_rows(self, n_req: int) -> None:
        """owner-side float staging (vectors out, gradients in) grows with the largest request seen"""
        if n_req <= self._row_cap:
            return
        cap = max(int(n_req * 1.25), 1024)
        f32 = dict(dtype=torch.float32, device=self.dev)
        peer = self.exchange_mode == "peer"
        if not peer:
            self.rows_send = torch.zeros(cap, self.wire, **f32)
        if self.train_mode:
            if not peer:
                self.g_recv = torch.zeros(cap, self.wire, **f32)
            if not peer or os.environ.get("CTR_OWNER_UPDATE", "fused") != "fused":
                self.g_own = torch.zeros(cap, self.wire, **f32)
        self._row_cap = cap

    # ---- layout hooks ------------------------------------------------------------------------------------------
    def global_batch(self, B: int) -> int:
        return B * self.world        # every rank feeds the same batch size (drop_remainder, utils/data_loader.py:33)

    def _cfg(self, B: int) -> EmbedCfg:
        cfg = super()._cfg(B)
        cfg.id_bytes = 4             # lookups index the staging buffer by their position in the unique list
        cfg.table_rows = 0           # positions are valid by construction; the fed ids are range-checked by ctr_sort_ids
        cfg.zero_row0 = 0            # padding ids share one all-zero staging row instead
        cfg.table_stride = self.wire
        return cfg

    def _cont_rows(self):
        s = self.spec
        return self.cont_buf if (s.use_fm and s.fm_cont) else None

    def _assemble_rows(self, shard_rows: torch.Tensor) -> np.ndarray:
        parts = self.comm.all_gather_tensor(shard_rows)
        R = self.spec.table_rows
        full = np.zeros((R,) + tuple(shard_rows.shape[1:]), dtype=np.float32)
        for o, t in enumerate(parts):
            cnt = len(range(o, R, self.world))
            full[o::self.world] = t[:cnt].cpu().numpy()
        return full

    # ---- id plans ----------------------------------------------------------------------------------------------
    def _build_local(self, pl: _IdPlan, B: int, stream: torch.cuda.Stream) -> None:
        """enqueue on `stream`: sort / de-dup / per-owner counts / inverse map, then the counts all-to-all and an
        async copy of (send, recv) counts to pinned host memory.  No host synchronisation."""
        s, h, sp = self.spec, self.h, self.plan
        st = stream.cuda_stream
        n = B * s.ids_ld + self.n_extra
        ws, wsn = ptr(pl.ws_sort), pl.ws_sort.numel()
        h.call("ctr_sort_ids", st, ptr(pl.ids), self.id_bytes, B, s.ids_ld, 0, s.ids_ld, 0, int(s.zero_row0),
               s.cont_base, self.n_extra, self.world, sp.local_bits, sp.key_bits, int(s.table_rows), ptr(self.status),
               ptr(pl.keys), ptr(pl.perm), ws, wsn)
        h.call("ctr_unique_segments", st, ptr(pl.keys), n, ptr(pl.uniq), ptr(pl.seg_start), ptr(pl.uidx),
               ptr(pl.n_unique), ws, wsn, None)
        h.call("ctr_shard_counts", st, ptr(pl.uniq), ptr(pl.n_unique), self.world, sp.local_bits, ptr(pl.counts))
        h.call("ctr_invert_perm", st, ptr(pl.perm), ptr(pl.uidx), n, ptr(pl.inv))
        W = self.world
        if self.device_plan:
            # exchange #1a without a collective: my counts row goes into every peer's matrix; one barrier later the whole
            # matrix is in the own slab and ONE small kernel turns it into this rank's routes, n_req and U -- on the device.
            # The host never reads the split sizes: nothing in the step waits for the GPU, so the host can run ahead
            h.call("ctr_peer_bcast", st, ptr(pl.counts), W, C.byref(self._rt_counts[pl.slot]))
            self._plan_barrier(st)
            h.call("ctr_peer_plan_routes", st, ptr(self._mat_view[pl.slot]), W, self.rank, self._req_cap,
                   self._g_cap if self.train_mode else self._req_cap, ptr(pl.plan_dev), ptr(self.status))
            # exchange #1b: every requester stores its unique keys into the owner's request list
            h.call("ctr_peer_push_rows", st, ptr(pl.uniq), n, 1, C.byref(pl.route_k))
            self._plan_barrier(st)
            if self.train_mode:
                # every peer's slice is already ascending: a stable merge (equal rows keep peer order, so the owner's
                # sum is order-fixed) replaces a radix sort of the request list
                cap = self._req_cap
                h.call("ctr_merge_sorted_segments", st, ptr(pl.req), cap, None, W, ptr(pl.o_keys), ptr(pl.o_idx),
                       pl.plan_dev.data_ptr() + 4 * PLAN_VEC)
                h.call("ctr_unique_segments", st, ptr(pl.o_keys), cap, ptr(pl.o_uniq), ptr(pl.o_seg), ptr(pl.o_uidx),
                       ptr(pl.o_nuniq), ptr(pl.o_ws), pl.o_ws.numel(), pl.plan_dev.data_ptr() + 4 * PLAN_NREQ)
            pl.ev_ready.record(stream)
            pl.B, pl.U, pl.n_req = B, None, None
            return
        # host-side planning (CTR_PLAN=host, or CTR_EXCHANGE=nccl: NCCL wants the split sizes on the host)
        with torch.cuda.stream(stream):
            if self.exchange_mode == "peer":
                h.call("ctr_peer_bcast", st, ptr(pl.counts), W, C.byref(self._rt_counts[pl.slot]))
                self._plan_barrier(st)
                pl.counts_host[:W * W].copy_(self._mat_view[pl.slot], non_blocking=True)
            else:
                mat = _shard.gather_counts(pl.counts, self.group_plan)     # exchange #1a: every rank learns every split
                pl.counts_host[:W * W].copy_(mat.reshape(-1), non_blocking=True)
            pl.counts_host[W * W:].copy_(self.peer_status, non_blocking=True)
            pl.ev_local.record(stream)
        pl.B = B

    def _build_exchange(self, pl: _IdPlan, stream: torch.cuda.Stream) -> None:
        """host-side planning only (CTR_PLAN=host / CTR_EXCHANGE=nccl; the default peer-memory path plans its routes on the
        device, see _build_local): the host reads the split sizes, then on `stream`: the key exchange and the owner-side
        merge + run structure"""
        h, sp = self.h, self.plan
        if self.device_plan:
            return
        if self.host_prof is not None:
            import time
            t0 = time.perf_counter()
            pl.ev_local.synchronize()
            self.host_prof["sync_s"] += time.perf_counter() - t0
        else:
            pl.ev_local.synchronize()
        flat, W = pl.counts_host.tolist(), self.world
        if flat[W * W] or flat[W * W + 1]:
            raise RuntimeError("peer barrier timed out waiting for rank %d" % ((flat[W * W] or flat[W * W + 1]) - 1))
        mat = [flat[p * W:(p + 1) * W] for p in range(W)]
        pl.send_counts, pl.recv_counts, vec, grad = _shard.peer_routes(mat, self.rank)
        pl.U, pl.n_req = sum(pl.send_counts), sum(pl.recv_counts)
        pl.ensure_owner(self, pl.n_req)
        if self.exchange_mode == "peer":
            worst = max(sum(mat[p][o] for p in range(W)) for o in range(W))
            if self.train_mode and worst > self._g_cap:      # the same verdict on every rank (all hold the matrix)
                raise RuntimeError("peer exchange: %d rows requested of one owner, receive buffer holds %d" % (worst, self._g_cap))
            pl.route_v = self.slab.route(vec[0], vec[1], self._rows_off)
            pl.route_g = self.slab.route(grad[0], grad[1], self._g_off)
            pl.route_k = self.slab.route(grad[0], grad[1], self._req_off[pl.slot])     # keys travel like the gradients will
            if max(pl.n_req, pl.U) > self._req_cap:
                raise RuntimeError("peer exchange: %d keys requested of this owner, request buffer holds %d" % (pl.n_req, self._req_cap))
        n_req, st = pl.n_req, stream.cuda_stream
        with torch.cuda.stream(stream):
            if self.exchange_mode == "peer":
                # exchange #1b without a collective: every requester stores its keys into the owner's request list
                h.call("ctr_peer_push_rows", st, ptr(pl.uniq), pl.U, 1, C.byref(pl.route_k))
                self._plan_barrier(st)
            else:
                _shard.exchange(pl.uniq[:pl.U], pl.send_counts, pl.recv_counts, out=pl.req, group=self.group_plan)   # #1b
            if self.train_mode and n_req:
                ws, wsn = ptr(pl.o_ws), pl.o_ws.numel()
                # every peer's slice is already ascending: a stable merge (equal rows keep peer order, so the owner's
                # sum is order-fixed) replaces the radix sort of the request list
                ends = (C.c_uint32 * self.world)(*vec[0])
                h.call("ctr_merge_sorted_segments", st, ptr(pl.req), n_req, ends, self.world, ptr(pl.o_keys), ptr(pl.o_idx), None)
                h.call("ctr_unique_segments", st, ptr(pl.o_keys), n_req, ptr(pl.o_uniq), ptr(pl.o_seg), ptr(pl.o_uidx),
                       ptr(pl.o_nuniq), ws, wsn, None)
            pl.ev_ready.record(stream)

    def _activate(self, pl: _IdPlan) -> None:
        self.cur = pl
        self.ids, self.keys, self.perm, self.uidx, self.uniq = pl.ids, pl.keys, pl.perm, pl.uidx, pl.uniq
        self.seg_start, self.n_unique = pl.seg_start, pl.n_unique

    def _other_plan(self) -> _IdPlan:
        return self.plans[1] if self.cur is self.plans[0] else self.plans[0]

    def load_batch(self, batch: Dict, non_blocking: bool = True, skip_ids: bool = False) -> int:
        main = self._main_stream()
        pl = next((c for c in self.plans if c.in_flight and c.batch_id == id(batch)), None)
        if pl is not None:                         # prefetched during the previous step
            B = super().load_batch(batch, non_blocking, skip_ids=True)     # (flips the input sets if it was uploaded too)
            if self._pending_exchange is pl:       # (only if the previous step never reached its sparse update)
                self._build_exchange(pl, self.side)
                self._pending_exchange = None
            pl.in_flight = False
        else:
            pl = self._other_plan() if self.cur is not None else self.plans[0]
            if pl.in_flight:                       # a prefetched plan nobody claimed: let it drain, then reuse it
                if self._pending_exchange is pl:
                    self._build_exchange(pl, self.side)
                    self._pending_exchange = None
                main.wait_event(pl.ev_ready)
                pl.in_flight = False
            B = super().load_batch(batch, non_blocking, skip_ids=True)
            cate = batch["cate_feats"]
            if not isinstance(cate, torch.Tensor):
                cate = torch.as_tensor(np.asarray(cate))
            if self.spec.ids_ld:
                pl.ids[:B].copy_(cate.to(self.id_torch) if cate.dtype != self.id_torch else cate, non_blocking=non_blocking)
            self._build_local(pl, B, main)
            self._build_exchange(pl, main)
            pl.on_side = False
        self._activate(pl)
        return B

    _pending_exchange: Optional[_IdPlan] = None
    host_prof = None       # tests/host_timing.py: seconds the host spent blocked on the split-size read

    def _prefetch_next(self, next_batch: Optional[Dict], ev_call) -> None:
        """start the id plan of the next batch on the side stream (upload ids, sort, de-dup, counts all-to-all)"""
        super()._prefetch_next(next_batch, ev_call)         # non-id inputs: upload into the other input set
        if next_batch is None or not self.spec.ids_ld:
            return
        pl = self._other_plan()
        if pl.in_flight:
            return
        cate = next_batch["cate_feats"]
        if not isinstance(cate, torch.Tensor):
            cate = torch.as_tensor(np.asarray(cate))
        B = int(cate.shape[0])
        self.side.wait_event(pl.ev_free) if pl.batch_id is not None else None
        self.side.wait_event(ev_call)             # a device-resident batch may still be in production on the main stream
        with torch.cuda.stream(self.side):
            pl.ids[:B].copy_(cate.to(self.id_torch) if cate.dtype != self.id_torch else cate, non_blocking=True)
        self._build_local(pl, B, self.side)
        pl.batch_id, pl.in_flight, pl.on_side = id(next_batch), True, True
        pl.batch_obj = next_batch             # keeps the id from being recycled while the plan is pending
        self._pending_exchange = None if self.device_plan else pl

    def _end_step(self) -> None:
        super()._end_step()
        if self.cur is not None:
            self.cur.ev_free.record(self._main_stream())
            if self.cur.batch_id is None:
                self.cur.batch_id = -1             # marks ev_free as recorded

    # ---- forward: owner gather -> vectors all-to-all -> fused gather+FM -------------------------------------------
    def _launch_sort(self, B: int) -> None:
        self._sort_wide(B, self._stream())      # embedding ids are planned in load_batch / prefetched

    def _embed_forward(self, B: int) -> None:
        s, st, h, sp, pl = self.spec, self._stream(), self.h, self.plan, self.cur
        main = self._main_stream()
        if pl.on_side:
            main.wait_event(pl.ev_ready)
        self._mark("wait_plan")
        if self.exchange_mode == "peer":
            # owners read their [v | w] rows and store them straight into the requester's staging buffer, in the
            # requester's unique-list order; the barrier is the only other launch of exchange #2.  The request list's
            # length and the placement come from the plan block on the device (the launch is sized by the capacity)
            if self._rows_dirty:                 # no gradient barrier since the last forward (predict): peers may still read
                self._barrier(st)
            n_list = self._req_cap if self.device_plan else pl.n_req
            if n_list:
                h.call("ctr_gather_rows_push", st, ptr(self.table), s.table_stride, ptr(pl.req), n_list, sp.local_bits,
                       s.row_width, self.wire, C.byref(pl.route_v))
            self._mark("owner_gather")
            self._barrier(st)
            self._rows_dirty = True
            self._mark("a2a_vectors")
            # the run of padding ids (key 0xFFFFFFFF), if any, is unique index U: an all-zero staging row
            if self.device_plan:
                h.call("ctr_zero_row_at", st, ptr(self.rows_recv), self.wire, self.wire, pl.plan_dev.data_ptr() + 4 * PLAN_U)
            else:
                self.rows_recv[pl.U].zero_()
        else:
            U, n_req = pl.U, pl.n_req
            self._ensure_rows(n_req)
            # owners gather [v | w] rows; all-to-all #2 returns them in the requester's unique-list order
            if n_req:
                h.call("ctr_gather_rows", st, ptr(self.table), s.table_stride, ptr(pl.req), None, n_req, sp.local_bits,
                       s.row_width, ptr(self.rows_send), self.wire)
            self._mark("owner_gather")
            _shard.exchange(self.rows_send[:n_req], pl.recv_counts, pl.send_counts, out=self.rows_recv, group=self.group)
            self._mark("a2a_vectors")
            self.rows_recv[U].zero_()    # the run of padding ids (key 0xFFFFFFFF), if any, is unique index U
        if s.use_fm and s.fm_cont and s.C:
            h.call("ctr_gather_rows", st, ptr(self.rows_recv), self.wire, ptr(pl.inv[B * s.ids_ld:]), None, s.C, 0,
                   s.row_width, ptr(self.cont_buf), self.wire)
        cfg = self._cfg(B)
        if s.Fm:
            h.call("ctr_pooled_gather_fwd", st, cfg, ptr(self.rows_recv), ptr(pl.inv), ptr(self.multi_start), None,
                   ptr(self.x0), ptr(self.pooled_first), ptr(self.inv_cnt_emb), ptr(self.inv_cnt_first))
        h.call("ctr_fm_gather_fwd", st, cfg, ptr(self.rows_recv), ptr(pl.inv), 0, ptr(self.cont), ptr(self.vec),
               ptr(self._cont_rows()), ptr(self.pooled_first), ptr(self.x0), ptr(self.fm), ptr(self.sum_e))

    # ---- backward: per-unique gradients -> all-to-all #3 -> owner sum in peer order -> Adam ------------------------
    def _sparse_update(self, B: int, n: int, hp: AdamHP) -> None:
        self._sparse_push(B, n)
        self._sparse_apply(B, n, hp)

    def _sparse_push(self, B: int, n: int) -> None:
        s, st, h, pl = self.spec, self._stream(), self.h, self.cur
        # The second half of the next batch's id plan (key all-to-all, owner-side sort) needs its split sizes on the
        # host.  By now the whole forward + dense backward of THIS step is queued on the main stream, so the host wait
        # does not starve the GPU.  The position is fixed on purpose: the plan collectives run on their own
        # communicator, and two communicators are only safe if every rank issues all collectives in one global order
        # (polling for "sizes ready" made that order rank-dependent and deadlocked).
        if self._pending_exchange is not None:
            self._build_exchange(self._pending_exchange, self.side)
            self._pending_exchange = None
        if self.exchange_mode == "peer":
            # one gradient row per unique key, summed and stored straight into its owner's receive buffer (exchange #3)
            h.call("ctr_embed_bwd_finalize_push", st, ptr(pl.seg_start), ptr(pl.n_unique), n, s.grad_width, self.wire,
                   ptr(self.part), C.byref(pl.route_g))
            if not self.bwd_split:
                self._mark("finalize")

    def _sparse_apply(self, B: int, n: int, hp: AdamHP) -> None:
        s, st, h, sp, pl = self.spec, self._stream(), self.h, self.plan, self.cur
        U, n_req = pl.U, pl.n_req
        if self.exchange_mode == "peer":
            if self.device_plan:
                n_req = self._req_cap        # the device knows the real count (*o_nuniq); this only bounds the launch
            self._barrier(st)
            self._rows_dirty = False
            if not self._on_sparse:
                self._launch_peer_reduce()       # the dense gradients of every replica were complete before this barrier
            # (overlapped backward: the weight gradients are still being computed on the main stream; their reduction
            # gets a barrier of its own behind them, _reduce_dense_grads)
        else:
            h.call("ctr_embed_bwd_finalize", st, ptr(pl.seg_start), ptr(pl.n_unique), n, s.grad_width, self.wire,
                   ptr(self.part), ptr(self.grads_send))
            self._mark("finalize")
            _shard.exchange(self.grads_send[:U], pl.send_counts, pl.recv_counts, out=self.g_recv, group=self.group)
        self._mark("a2a_grads")
        if not n_req:
            return
        self._ensure_rows(n_req)
        if os.environ.get("CTR_OWNER_UPDATE", "fused") == "fused":
            # sum of the peers' contributions (fixed peer order) + TF1 Adam on the owned row, one launch
            h.call("ctr_adam_rows_gathered", st, ptr(pl.o_uniq), ptr(pl.o_nuniq), n_req, sp.local_bits, ptr(self.g_recv),
                   ptr(pl.o_idx), ptr(pl.o_seg), self.wire, s.row_width, ptr(self.table), ptr(self.m_table), ptr(self.v_table),
                   s.table_stride, self.rows_local, ptr(self.touched), self._fm_w2(), s.k, hp)
            return
        h.call("ctr_segment_sum_rows", st, ptr(pl.o_seg), ptr(pl.o_nuniq), n_req, ptr(pl.o_idx), ptr(self.g_recv),
               self.wire, s.grad_width, ptr(self.g_own), self.wire)
        h.call("ctr_adam_rows", st, ptr(pl.o_uniq), ptr(pl.o_nuniq), n_req, sp.local_bits, ptr(self.g_own), None,
               self.wire, s.row_width, ptr(self.table), ptr(self.m_table), ptr(self.v_table), s.table_stride,
               self.rows_local, ptr(self.touched), None, 0, self._fm_w2(), s.k, hp)

    def _barrier(self, st: int, third_set: bool = False) -> None:
        """all ranks: everything stored into peer slabs so far is visible, and every peer is at the same point.
        third_set: the barrier of the early loss on the loss stream (the words the overlapped backward's sparse stream uses;
        the two are never active in the same engine)"""
        if self._on_sparse or third_set:          # its own words and epoch: can be in flight beside a step barrier
            self._sepoch += 1
            flags, epoch = self._sflags_off, self._sepoch
        else:
            self._epoch += 1
            flags, epoch = self._flags_off, self._epoch
        if self.comm.is_loopback:
            self.comm.device_barrier(st)
            return
        self.h.call("ctr_peer_barrier", st, C.byref(self.cur.route_v), flags, epoch, ptr(self.peer_status),
                    int(os.environ.get("CTR_PEER_TIMEOUT_MS", 10000)))

    def align_ranks(self) -> None:
        """One device-side rank barrier on the current stream (peer-memory mode; a no-op otherwise).  For timing harnesses:
        processes leave a HOST barrier milliseconds apart, so the early ranks' first step would otherwise wait for the latest
        rank's host inside the timed window; after this call every rank's stream is at the same instant."""
        if self.exchange_mode != "peer" or self.cur is None:
            return
        self._pin_stream()
        try:
            self._barrier(self._stream())
        finally:
            self._unpin_stream()

    _dense_work = None

    def _begin_dense_reduce(self) -> None:
        # dlogit already carries 1/B_global, so a plain SUM over ranks is the global-batch gradient; the loss share rides
        # in the tail slot of the same buffer.  Issued as soon as the last weight gradient is queued: NCCL's stream runs it
        # underneath the sparse backward (partials, gradient exchange, row Adam), which touches neither buffer.
        self.grad[self.n_dense:self.n_dense + 1].copy_(self.loss)
        if self.exchange_mode != "peer":
            self._dense_work = dist.all_reduce(self.grad, group=self.group, async_op=True)

    def _plan_barrier(self, st: int) -> None:
        """barrier of the id-plan exchanges: its own arrival words and epoch, since it runs on another stream than the step's"""
        self._pepoch += 1
        if self.comm.is_loopback:
            self.comm.device_barrier(st)
            return
        self.h.call("ctr_peer_barrier", st, C.byref(self._rt0), self._pflags_off, self._pepoch, ptr(self.peer_status[1:]),
                    int(os.environ.get("CTR_PEER_TIMEOUT_MS", 10000)))

    def _launch_peer_reduce(self) -> None:
        """after a step barrier: every replica's dense gradients are visible -> sum them on the aux stream"""
        main = self._main_stream()
        self.ev_bar.record(main)
        self.aux.wait_event(self.ev_bar)
        self.h.call("ctr_peer_reduce_sum", self.aux.cuda_stream, C.byref(self._rt_dense), self.n_dense + _ALIGN, ptr(self.grad_red))
        self.ev_red.record(self.aux)
        self._red_pending = True

    _red_pending = False

    def _dense_grad(self) -> torch.Tensor:
        return self.grad_red if (self.exchange_mode == "peer" and self.train_mode) else self.grad

    def _reduce_dense_grads(self) -> None:
        if self.exchange_mode == "peer":
            if not self._red_pending:            # no sparse side this step: the barrier has not happened yet
                self._barrier(self._stream())
                self._launch_peer_reduce()
            self._main_stream().wait_event(self.ev_red)
            self._red_pending = False
            return
        if self._dense_work is not None:
            self._dense_work.wait()          # stream-side wait: the main stream continues behind the collective
            self._dense_work = None

    @property
    def loss_is_early(self) -> bool:
        """Peer-memory mode: every rank leaves its share of the loss in its slab right behind the head kernel, and the
        loss stream sums the shares out of peer memory in rank order (the dense-gradient mechanism on 16 bytes) -- so the
        host gets the global loss after the forward pass, as on one GPU.  Otherwise the loss comes back with the gradient
        all-reduce, after the backward.  CTR_EARLY_LOSS=0 switches it off."""
        return (self.exchange_mode == "peer" and self.train_mode and self.world > 1 and not self.bwd_overlap and
                CtrEngine.loss_is_early.fget(self) and os.environ.get("CTR_EARLY_LOSS", "1") != "0")

    def _global_loss(self) -> torch.Tensor:
        if not self.loss_is_early:
            return self._dense_grad()[self.n_dense:self.n_dense + 1].clone()
        # runs on torch's current stream = the loss stream (_publish_loss), behind the head kernel
        ls = torch.cuda.current_stream(self.dev)
        slot = self._loss_step & 1
        self._loss_step += 1
        self._loss_slab[4 * slot:4 * slot + 1].copy_(self.loss)
        self._barrier(ls.cuda_stream, third_set=True)
        self.h.call("ctr_peer_reduce_sum", ls.cuda_stream, C.byref(self._rt_loss[slot]), 4, ptr(self._loss_red))
        return self._loss_red[0:1].clone()

    def predict(self, batch: Optional[Dict] = None) -> torch.Tensor:
        out = super().predict(batch)
        self._end_step()
        return out
