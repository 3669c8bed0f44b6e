"""ShardedCtrEngine: one process per GPU; embedding tables row-sharded (owner = id mod N), samples data-parallel,
dense tower replicated with one sum-allreduce of its gradients (schedule in shard.py).

The arithmetic is the single-GPU engine's: the fused gather+FM kernel simply reads the rows it needs from the
staging buffer filled by the vector all-to-all (indexed by each lookup's position in this rank's unique list)
instead of from the table, and the Adam update runs on the owner after the gradient all-to-all.  The loss is the
mean over the GLOBAL batch, so N ranks on N shards of a batch reproduce one GPU on the concatenated batch.
"""
from __future__ import annotations

from typing import Dict, Optional

import numpy as np
import torch
import torch.distributed as dist

from . import shard as _shard
from ._lib import AdamHP, EmbedCfg, ptr
from .engine import CtrEngine, _ru
from .spec import ModelSpec


class ShardedCtrEngine(CtrEngine):
    def __init__(self, spec: ModelSpec, max_batch: int, group=None, **kw):
        if not dist.is_initialized():
            raise RuntimeError("ShardedCtrEngine needs an initialised torch.distributed process group")
        self.group = group
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        self.plan = _shard.ShardPlan(spec.table_rows, world)
        self.wire = spec.grad_stride          # floats per row on the wire (k [+1], rounded up to 4): vectors and gradients
        super().__init__(spec, max_batch, world=world, rank=rank, **kw)
        s, B, dev = spec, self.Bmax, self.dev
        u32 = dict(dtype=torch.int32, device=dev)
        f32 = dict(dtype=torch.float32, device=dev)
        if not self.train_mode:      # the de-dup structure is needed by the forward exchange, training or not
            self.n_extra = s.C if (s.use_fm and s.fm_cont) else 0
            n = B * s.ids_ld + self.n_extra
            self.keys, self.perm, self.uidx, self.uniq = (torch.zeros(max(n, 1), **u32) for _ in range(4))
            self.seg_start = torch.zeros(n + 2, **u32)
            self.n_unique = torch.zeros(1, **u32)
            self.ws_sort = torch.zeros(_ru(max(self.h.query("ctr_sort_ws_bytes", n), 4 * (n // 2048 + 8)), 256),
                                       dtype=torch.uint8, device=dev)
        n = B * s.ids_ld + self.n_extra
        self.n_max = n
        self.inv = torch.zeros(max(n, 1), **u32)
        self.counts = torch.zeros(world, **u32)
        self.rows_recv = torch.zeros(n + 1, self.wire, **f32)
        self.cont_buf = torch.zeros(max(s.C, 1), self.wire, **f32)
        self.grads_send = torch.zeros(max(n, 1), self.wire, **f32) if self.train_mode else None
        self._cap = 0
        self._ensure_owner(n)
        self.send_counts = self.recv_counts = None
        self.n_req = self.U = 0
        self.exchange_rows_step = 0          # rows this rank sent+received in the last step (reported by bench.py)

    # ---- owner-side buffers grow with the largest request seen ------------------------------------------------
    def _ensure_owner(self, n_req: int) -> None:
        if n_req <= self._cap:
            return
        cap = max(int(n_req * 1.25), 1024)
        dev = self.dev
        u32 = dict(dtype=torch.int32, device=dev)
        f32 = dict(dtype=torch.float32, device=dev)
        self.req = torch.zeros(cap, **u32)
        self.rows_send = torch.zeros(cap, self.wire, **f32)
        if self.train_mode:
            self.g_recv = torch.zeros(cap, self.wire, **f32)
            self.g_own = torch.zeros(cap, self.wire, **f32)
            self.o_keys, self.o_idx, self.o_uniq, self.o_uidx = (torch.zeros(cap, **u32) for _ in range(4))
            self.o_seg = torch.zeros(cap + 2, **u32)
            self.o_nuniq = torch.zeros(1, **u32)
            self.o_ws = torch.zeros(_ru(max(self.h.query("ctr_sort_ws_bytes", cap), 4 * (cap // 2048 + 8)), 256),
                                    dtype=torch.uint8, device=dev)
        self._cap = cap

    # ---- layout hooks ------------------------------------------------------------------------------------------
    def global_batch(self, B: int) -> int:
        return B * self.world        # every rank feeds the same batch size (drop_remainder, utils/data_loader.py:33)

    def _cfg(self, B: int) -> EmbedCfg:
        cfg = super()._cfg(B)
        cfg.id_bytes = 4             # lookups index the staging buffer by their position in the unique list
        cfg.zero_row0 = 0            # padding ids share one all-zero staging row instead
        cfg.table_stride = self.wire
        return cfg

    def _cont_rows(self):
        s = self.spec
        return self.cont_buf if (s.use_fm and s.fm_cont) else None

    def _assemble_rows(self, shard_rows: torch.Tensor) -> np.ndarray:
        parts = [torch.empty_like(shard_rows) for _ in range(self.world)]
        dist.all_gather(parts, shard_rows.contiguous(), group=self.group)
        R = self.spec.table_rows
        full = np.zeros((R,) + tuple(shard_rows.shape[1:]), dtype=np.float32)
        for o, t in enumerate(parts):
            cnt = len(range(o, R, self.world))
            full[o::self.world] = t[:cnt].cpu().numpy()
        return full

    # ---- forward: de-dup -> ids all-to-all -> owner gather -> vectors all-to-all -> fused gather+FM ---------------
    def _launch_sort(self, B: int) -> None:
        self._sort_wide(B, self._stream())      # embedding ids are sorted inside _embed_forward (predict needs it too)

    def _embed_forward(self, B: int) -> None:
        s, st, h, pl = self.spec, self._stream(), self.h, self.plan
        n = B * s.ids_ld + self.n_extra
        ws, wsn = ptr(self.ws_sort), self.ws_sort.numel()
        h.call("ctr_sort_ids", st, ptr(self.ids), self.id_bytes, B, s.ids_ld, 0, s.ids_ld, 0, int(s.zero_row0),
               s.cont_base, self.n_extra, self.world, pl.local_bits, pl.key_bits, ptr(self.keys), ptr(self.perm), ws, wsn)
        h.call("ctr_unique_segments", st, ptr(self.keys), n, ptr(self.uniq), ptr(self.seg_start), ptr(self.uidx),
               ptr(self.n_unique), ws, wsn)
        h.call("ctr_shard_counts", st, ptr(self.uniq), ptr(self.n_unique), self.world, pl.local_bits, ptr(self.counts))
        h.call("ctr_invert_perm", st, ptr(self.perm), ptr(self.uidx), n, ptr(self.inv))
        # all-to-all #1: counts (one host read per step: NCCL wants split sizes on the host), then unique keys
        self.send_counts, self.recv_counts = _shard.split_sizes(self.counts, self.group)
        U, n_req = sum(self.send_counts), sum(self.recv_counts)
        self.U, self.n_req = U, n_req
        self._ensure_owner(n_req)
        _shard.exchange(self.uniq[:U], self.send_counts, self.recv_counts, out=self.req, group=self.group)
        # owners gather [v | w] rows; all-to-all #2 returns them in the requester's unique-list order
        if n_req:
            h.call("ctr_gather_rows", st, ptr(self.table), s.table_stride, ptr(self.req), None, n_req, pl.local_bits,
                   s.row_width, ptr(self.rows_send), self.wire)
        _shard.exchange(self.rows_send[:n_req], self.recv_counts, self.send_counts, out=self.rows_recv, group=self.group)
        self.rows_recv[U].zero_()    # the run of padding ids (key 0xFFFFFFFF), if any, is unique index U
        self.exchange_rows_step = U + n_req
        if s.use_fm and s.fm_cont and s.C:
            h.call("ctr_gather_rows", st, ptr(self.rows_recv), self.wire, ptr(self.inv[B * s.ids_ld:]), None, s.C, 0,
                   s.row_width, ptr(self.cont_buf), self.wire)
        cfg = self._cfg(B)
        if s.Fm:
            h.call("ctr_pooled_gather_fwd", st, cfg, ptr(self.rows_recv), ptr(self.inv), ptr(self.multi_start), None,
                   ptr(self.x0), ptr(self.pooled_first), ptr(self.inv_cnt_emb), ptr(self.inv_cnt_first))
        h.call("ctr_fm_gather_fwd", st, cfg, ptr(self.rows_recv), ptr(self.inv), 0, ptr(self.cont), ptr(self.vec),
               ptr(self._cont_rows()), ptr(self.pooled_first), ptr(self.x0), ptr(self.fm), ptr(self.sum_e))

    # ---- backward: per-unique gradients -> all-to-all #3 -> owner sum in peer order -> Adam ------------------------
    def _sparse_update(self, B: int, n: int, hp: AdamHP) -> None:
        s, st, h, pl = self.spec, self._stream(), self.h, self.plan
        U, n_req = self.U, self.n_req
        h.call("ctr_embed_bwd_finalize", st, ptr(self.seg_start), ptr(self.n_unique), n, s.row_width, self.wire,
               ptr(self.part), ptr(self.grads_send))
        _shard.exchange(self.grads_send[:U], self.send_counts, self.recv_counts, out=self.g_recv, group=self.group)
        if not n_req:
            return
        self.o_keys[:n_req].copy_(self.req[:n_req])
        torch.arange(n_req, out=self.o_idx[:n_req])
        ws, wsn = ptr(self.o_ws), self.o_ws.numel()
        # stable sort by local row: equal rows keep peer order, so the sum below is order-fixed
        h.call("ctr_sort_pairs", st, ptr(self.o_keys), ptr(self.o_idx), n_req, pl.local_bits, ws, wsn)
        h.call("ctr_unique_segments", st, ptr(self.o_keys), n_req, ptr(self.o_uniq), ptr(self.o_seg), ptr(self.o_uidx),
               ptr(self.o_nuniq), ws, wsn)
        h.call("ctr_segment_sum_rows", st, ptr(self.o_seg), ptr(self.o_nuniq), n_req, ptr(self.o_idx), ptr(self.g_recv),
               self.wire, s.row_width, ptr(self.g_own), self.wire)
        h.call("ctr_adam_rows", st, ptr(self.o_uniq), ptr(self.o_nuniq), n_req, pl.local_bits, ptr(self.g_own), None,
               self.wire, s.row_width, ptr(self.table), ptr(self.m_table), ptr(self.v_table), s.table_stride,
               self.rows_local, ptr(self.touched), None, 0, hp)

    def _reduce_dense_grads(self) -> None:
        # dlogit already carries 1/B_global, so a plain SUM over ranks is the global-batch gradient; the loss
        # share rides in the tail slot of the same buffer
        self.grad[self.n_dense:self.n_dense + 1].copy_(self.loss)
        dist.all_reduce(self.grad, group=self.group)

    def _global_loss(self) -> torch.Tensor:
        return self.grad[self.n_dense:self.n_dense + 1].clone()
