"""CtrEngine: device state and the step schedule of one model on one GPU.

PyTorch is plumbing here (device memory, streams, H2D/D2H copies); every arithmetic step of the
hot path is a call into libctr_b200.so through `_lib.Handle.call`.
"""
from __future__ import annotations

import ctypes as C
import math
import os
from typing import Dict, List, Optional

import numpy as np
import torch

from . import _lib
from . import shard as _shard
from ._lib import DenseLayout, MAX_LAYERS, AdamHP, EmbedCfg, ptr

# where the side-stream id sort starts: behind the gather (default; it then shares the machine with the forward GEMMs and
# the head kernel, which it slows from 64 to 95 us) or behind the head kernel (shares it with the backward GEMMs)
_SORT_AFTER_HEAD = os.environ.get("CTR_SORT_AT", "embed") == "head"
# order of the backward pass.  "serial" (default): wgrad/dgrad layer by layer, the sparse side (partials, [gradient
# exchange], row Adam) behind the last GEMM.  "overlap" (opt-in): the data gradients of all layers first, then the sparse
# side on its own stream UNDER the weight-gradient GEMMs.  Same kernels, same arithmetic, parity-green in both orders.
# Measured on one B200 (round 2, B = 64K): overlap 1.052 ms/step against 1.022 serial -- the sparse side is fully hidden
# (12 us of waiting left) but the three weight-gradient GEMMs take 364 us with it underneath against ~200 us alone: the
# interference (HBM bandwidth, SM slots) costs more than the 169 us the sparse kernels take by themselves.
_BWD_OVERLAP = os.environ.get("CTR_BWD_ORDER", "serial") == "overlap"
from .spec import ModelSpec

ADAM_BETA1, ADAM_BETA2, ADAM_EPS = 0.9, 0.999, 1e-8   # tf.train.AdamOptimizer defaults (models/deepfm.py:161)
def _loss_stream_min_batch() -> int:
    """below it a step is launch-bound: the loss stream's events cost more than they hide (CTR_LOSS_STREAM_MIN_BATCH: tests
    lower it to run the loss-stream paths on small batches)"""
    return int(os.environ.get("CTR_LOSS_STREAM_MIN_BATCH", 4096))


FTRL_INIT_ACCUM = 0.1                                 # tf.train.FtrlOptimizer default (test_model/wdl.py:136)
PREC = {"fp32": 0, "tf32": 1}
_ALIGN = 64   # floats; keeps every dense parameter 256-byte aligned inside the flat buffer


def _ru(a: int, b: int) -> int:
    return (a + b - 1) // b * b


class CtrEngine:
    def __init__(self, spec: ModelSpec, max_batch: int, device: Optional[torch.device] = None,
                 learning_rate: float = 0.001, decay_steps: int = 10000000, decay_rate: float = 0.9,
                 precision: str = "fp32", adam_mode: str = "tf_dense", train: bool = True,
                 world: int = 1, rank: int = 0):
        if not torch.cuda.is_available():
            raise _lib.CtrError("CtrEngine needs a CUDA device: the hot path has no CPU fallback")
        assert adam_mode in ("tf_dense", "lazy") and precision in PREC
        self.spec, self.Bmax = spec, int(max_batch)
        self.world, self.rank = int(world), int(rank)
        # table rows held by this rank: row id lives on rank id % world at local row id // world
        self.rows_local = (spec.table_rows + self.world - 1) // self.world
        self.dev = torch.device(device if device is not None else "cuda:%d" % torch.cuda.current_device())
        self.h = _lib.Handle(self.dev.index if self.dev.index is not None else torch.cuda.current_device())
        self.lr0, self.decay_steps, self.decay_rate = learning_rate, decay_steps, decay_rate
        self.prec, self.adam_mode, self.train_mode = PREC[precision], adam_mode, train
        if self.prec == 1 and not self.h.has_tcgen05:
            raise _lib.CtrError("precision='tf32' needs an sm_100 device")
        self.global_step, self.beta1_power, self.beta2_power = 0, ADAM_BETA1, ADAM_BETA2
        s, B, dev = spec, self.Bmax, self.dev
        # device status words (include/ctr_b200.h CTR_STATUS_*): [step barrier, plan barrier, bad id, spare].  Kernels set
        # them, the optimizer kernels refuse to move weights while one is set, check_status() turns them into an exception
        self.status = torch.zeros(_lib.STATUS_WORDS, dtype=torch.int32, device=dev)
        self._status_host = torch.zeros(_lib.STATUS_WORDS, dtype=torch.int32).pin_memory()
        self._status_ev = torch.cuda.Event()
        self._status_pending = False
        self._steps_done = self._steps_verified = self._steps_at_snap = 0     # train steps enqueued / known to be applied
        f32 = dict(dtype=torch.float32, device=dev)
        self.id_torch = torch.int64 if s.id_dtype == "int64" else torch.int32
        self.id_bytes = 8 if s.id_dtype == "int64" else 4

        # ---- parameters ---------------------------------------------------------------------
        if (s.frozen_rows or s.weighted_pool) and self.world > 1:
            raise NotImplementedError("%s: frozen rows / weighted pooling are single-GPU only" % s.alg)
        # frozen rows (a reference tf.constant looked up like a table, models/dnn_multi_textline.py:46-48) sit behind the
        # trainable rows: every update kernel is told `rows_local` and never touches them
        self.table = torch.zeros(self.rows_local + s.frozen_rows, s.table_stride, **f32)
        self.frozen_off = torch.from_numpy(s.frozen_slot_offset()).to(dev, self.id_torch) if s.frozen_rows else None
        self.dense_shapes = s.dense_param_shapes()
        # flat buffer ordered by regulariser group so Adam needs at most three launches
        order = {"none": 0, "l2": 1, "l1": 2}
        grp = lambda n: s.reg.get(n, "none") if s.l2_reg > 0 else "none"
        names = sorted(range(len(self.dense_shapes)), key=lambda i: (order[grp(self.dense_shapes[i][0])], i))
        self.off: Dict[str, int] = {}
        self.groups: List[List[int]] = [[0, 0], [0, 0], [0, 0]]   # [start, end) per reg group
        cur = 0
        for gi in range(3):
            self.groups[gi][0] = cur
            for i in names:
                n, shp = self.dense_shapes[i]
                if order[grp(n)] != gi:
                    continue
                self.off[n] = cur
                cur += _ru(int(np.prod(shp)), _ALIGN)
            self.groups[gi][1] = cur
        self.n_dense = cur
        self.theta = torch.zeros(cur, **f32)
        self.shape = dict(self.dense_shapes)
        # K-major copies of the tower weights for the tcgen05 path
        self.wt = [torch.zeros(h_, _ru(k_, 4), **f32) for (k_, h_) in self._layer_dims()] if self.prec == 1 else None
        # tf32-rounded copies [K, N] the TF32 data-gradient GEMM reads
        self.wr = [torch.zeros(k_, h_, **f32) for (k_, h_) in self._layer_dims()] if (self.prec == 1 and train) else None

        # ---- batch + activations ----------------------------------------------------------------
        # two sets of input buffers: the batch of step i+1 is uploaded (copy stream) while step i computes
        self._in = []
        for _ in range(2):
            self._in.append(dict(ids=torch.zeros(B, max(s.ids_ld, 1), dtype=self.id_torch, device=dev),
                                 cont=torch.zeros(B, max(s.C, 1), **f32), vec=torch.zeros(B, max(s.V, 1), **f32),
                                 label=torch.zeros(B, **f32),
                                 sw=torch.zeros(B, max(s.M, 1), **f32) if s.weighted_pool else None,
                                 wide=torch.zeros(B, max(s.wide_fields, 1), dtype=self.id_torch, device=dev) if s.head == "wdl" else None))
        self._in_free = [torch.cuda.Event(), torch.cuda.Event()]
        self._in_used = [False, False]
        self.copy_stream = torch.cuda.Stream(device=dev)
        self._pf_event, self._pf_id, self._pf_B = torch.cuda.Event(), None, 0
        self._ids_external = False          # sharded engine: ids live in its id plans
        self._use_inputs(0)
        self.x0 = torch.zeros(B, s.x0_ld, **f32)
        # hidden activations carry pad columns; column H holds 1.0 for good, so the weight-gradient GEMM of the next
        # layer yields the bias gradient as an extra row (no GEMM ever reads or writes past column H-1 otherwise).  The row
        # pitch is a multiple of 128 bytes: with H + 4 floats (1616 B at H = 400) every other row starts in the middle of a
        # 32-byte sector and the epilogue's stores cost 6 % of the GEMM (62.3 vs 66.0 us, profiles/r02j_gemm_bound.txt)
        self.act_ld = [_ru(h_ + 1, 32) for h_ in s.hidden]
        self.acts = [torch.zeros(B, ld, **f32) for ld in self.act_ld]
        for a, h_ in zip(self.acts, s.hidden):
            a[:, h_] = 1.0
        # TF32 training: the forward pass of every hidden layer but the last also leaves its ReLU sign bits, and the data
        # gradient of the next layer masks with those instead of re-reading the fp32 activation (ctr_mlp_fwd_bits);
        # CTR_RELU_BITS=0 keeps the fp32 mask for A/B
        self.relu_bits = None
        if self.prec == 1 and train and os.environ.get("CTR_RELU_BITS", "1") != "0":
            self.relu_bits = [torch.zeros(max(self.h.query("ctr_mlp_relu_bits_bytes", B, h_) // 4, 1), dtype=torch.int32, device=dev)
                              for h_ in s.hidden[:-1]]
        self.fm = torch.zeros(B, s.fm_ld, **f32)
        self.sum_e = torch.zeros(B, s.k, **f32)
        self.logit, self.score, self.dlogit = (torch.zeros(B, **f32) for _ in range(3))
        self.loss = torch.zeros(1, **f32)
        self.reg_terms = torch.zeros(4, **f32)
        self.pooled_first = torch.zeros(B, max(s.Fm, 1), **f32)
        self.inv_cnt_emb = torch.zeros(B, max(s.Fm, 1), **f32)
        self.inv_cnt_first = torch.zeros(B, max(s.Fm, 1), **f32)
        self.multi_start = torch.from_numpy(s.multi_start()).to(dev)
        self.field_of_slot = torch.from_numpy(s.field_of_slot()).to(dev)
        self.wide_sum = torch.zeros(B, **f32) if s.head == "wdl" else None

        ws_bytes = max(self.h.query("ctr_colsum_ws_bytes", B, max(s.hidden + [s.fm_ld])), 4 * (B // 8 + 8), 1 << 16,
                       self.h.query("ctr_head_ws_bytes", B, s.n_fm, s.hidden[-1]))
        if train:
            self.m_table, self.v_table = torch.zeros_like(self.table), torch.zeros_like(self.table)
            self.m, self.v = (torch.zeros(cur, **f32) for _ in range(2))
            if s.optimizer == "ftrl":
                # FtrlOptimizer's slots: m holds `accum` (initial_accumulator_value 0.1), v holds `linear` (0); the row
                # kernels are its sparse apply, so the step never sweeps the whole table whatever adam_mode says
                self.m_table.fill_(FTRL_INIT_ACCUM)
                self.m.fill_(FTRL_INIT_ACCUM)
            self.grad = torch.zeros(cur + _ALIGN, **f32)     # tail slot [cur] carries the loss share through the allreduce
            maxh = max(s.hidden)
            # one dZ per layer: the data gradients of ALL layers run first (so dX0 exists early and the sparse side can
            # start), the weight gradients read these buffers afterwards (CTR_BWD_ORDER=serial needs only two)
            self.dz = [torch.zeros(B, maxh, **f32) for _ in range(max(2, len(s.hidden)))]
            self.dx0_ld = _ru(max(s.E, 4), 4)
            self.dx0 = torch.zeros(B, self.dx0_ld, **f32)
            self.n_extra = s.C if (s.use_fm and s.fm_cont) else 0
            n = B * s.ids_ld + self.n_extra
            self.n_sort = n
            u32 = dict(dtype=torch.int32, device=dev)   # raw 32-bit storage; the library treats it as uint32
            self.keys, self.perm, self.uidx, self.uniq = (torch.zeros(max(n, 1), **u32) for _ in range(4))
            self.seg_start = torch.zeros(n + 2, **u32)
            self.n_unique = torch.zeros(1, **u32)
            self.part = torch.zeros((n + n // 32 + 2) * s.grad_stride, **f32)
            self.cont_grad = torch.zeros(max(s.C, 1), s.grad_stride, **f32)
            self.touched = torch.zeros(self.rows_local, dtype=torch.uint8, device=dev) if adam_mode == "tf_dense" else None
            cfg = self._cfg(B)
            # the id sort depends only on the batch: it runs on a side stream underneath the forward pass
            self.side = torch.cuda.Stream(device=dev)
            self.ev_batch, self.ev_sorted = torch.cuda.Event(), torch.cuda.Event()
            self.aux = torch.cuda.Stream(device=dev)
            self.ev_head, self.ev_cont = torch.cuda.Event(), torch.cuda.Event()
            # the sparse backward (partials -> [exchange] -> row Adam) runs on its own stream underneath the weight-gradient
            # GEMMs; it has its own workspace (the GEMMs' split-K partials live in self.ws at the same time)
            self.sparse_stream = torch.cuda.Stream(device=dev)
            self.ev_dx0, self.ev_sparse = torch.cuda.Event(), torch.cuda.Event()
            self.ws_sparse = torch.zeros(_ru(max(self.h.query("ctr_adam_rows_ws_bytes", n), 4 * (n // 2048 + 8), 4096), 256),
                                         dtype=torch.uint8, device=dev)
            self.ws_cont = torch.zeros(_ru(max(self.h.query("ctr_cont_bwd_ws_bytes", cfg), 256), 256), dtype=torch.uint8, device=dev)
            self.ws_sort = torch.zeros(_ru(max(self.h.query("ctr_sort_ws_bytes", n), 4 * (n // 2048 + 8)), 256),
                                       dtype=torch.uint8, device=dev)
            ws_bytes = max(ws_bytes, self.h.query("ctr_cont_bwd_ws_bytes", cfg),
                           self.h.query("ctr_adam_rows_ws_bytes", n),
                           4 * (n // 2048 + 8))
            for (k_, h_) in self._layer_dims():
                ws_bytes = max(ws_bytes, self.h.query("ctr_mlp_bwd_weight_ws_bytes", self.prec, B, k_, h_))
            if s.head == "wdl":
                # wide ids get their own sorted run structure (gradient of the scalar gather, models/wdl.py:240-246)
                nw = B * s.wide_fields
                self.w_sq = torch.zeros(nw + s.hidden[-1] + 1, **f32) if s.gathered_l2 else None
                self.w_keys, self.w_perm, self.w_uidx, self.w_uniq = (torch.zeros(max(nw, 1), **u32) for _ in range(4))
                self.w_seg_start = torch.zeros(nw + 2, **u32)
                self.w_n_unique = torch.zeros(1, **u32)
                self.ws_sort_w = torch.zeros(_ru(max(self.h.query("ctr_sort_ws_bytes", nw), 4 * (nw // 2048 + 8)), 256),
                                             dtype=torch.uint8, device=dev)
                self.ev_sorted_w = torch.cuda.Event()
        self.ws = torch.zeros(_ru(ws_bytes, 256), dtype=torch.uint8, device=dev)
        self.last_batch = 0

    # ------------------------------------------------------------------------------------------
    def _layer_dims(self):
        dims, prev = [], self.spec.D0
        for h_ in self.spec.hidden:
            dims.append((prev, h_))
            prev = h_
        return dims

    def _cfg(self, B: int) -> EmbedCfg:
        s = self.spec
        return EmbedCfg(B=B, id_bytes=self.id_bytes, ids_ld=max(s.ids_ld, 1) if s.ids_ld == 0 else s.ids_ld, Fc=s.Fc,
                        Fm=s.Fm, M=s.M, C=s.C, V=s.V, k=s.k, has_first=int(s.has_first), use_fm=int(s.use_fm),
                        fm_cont=int(s.fm_cont), zero_row0=int(s.zero_row0), count_rule=0, combiner=0,
                        table_stride=s.table_stride, x0_ld=s.x0_ld, fm_ld=s.fm_ld, grad_stride=s.grad_stride,
                        round_tf32=int(self.prec == 1), reserved0=0, table_rows=self._id_limit(),
                        status=self.status.data_ptr())

    def _id_limit(self) -> int:
        """ids K1/K2 may be fed: rows of the (single-GPU) table incl. the frozen rows behind it"""
        return int(self.spec.table_rows + self.spec.frozen_rows)

    # ---- fault reporting ------------------------------------------------------------------------------------------
    def _snapshot_status(self) -> None:
        """queue a copy of the status words to pinned host memory behind the step (no host wait)"""
        self._status_host.copy_(self.status, non_blocking=True)
        self._status_ev.record(self._main_stream())
        self._status_pending = True
        self._steps_at_snap = self._steps_done

    def check_status(self, wait: bool = True) -> None:
        """Raise if a kernel reported a fault: an id outside its table (tf.nn.embedding_lookup raises
        InvalidArgumentError there, models/deepfm.py:78,85,98) or a peer barrier that timed out.  The weights were not
        touched by the faulty step (the optimizer kernels check the same words).  Called by the fit / predict loops right
        after they read a loss or score (already synchronised, so it costs nothing), by get_weights and by the
        checkpoint writers; wait=False only looks if the snapshot has already landed."""
        if not self._status_pending:
            self._snapshot_status()
        if not wait and not self._status_ev.query():
            return
        self._status_ev.synchronize()
        self._status_pending = False
        st = self._status_host.tolist()
        if not any(st[:4]):
            self._steps_verified = self._steps_at_snap
            return
        # the status words are sticky and stop every optimizer kernel: no step since the last clean snapshot was applied,
        # so the host-side step counters go back by as many steps
        lost = self._steps_done - self._steps_verified
        self.global_step -= lost
        self.beta1_power /= ADAM_BETA1 ** lost
        self.beta2_power /= ADAM_BETA2 ** lost
        self._steps_verified = self._steps_done
        self.status.zero_()
        if st[2]:
            raise _lib.CtrError("id out of range in column %d of cate_feats / wide_feats (table has %d rows): the step "
                                "was not applied" % (st[2] - 1, self._id_limit()))
        if st[3]:
            raise _lib.CtrError("row exchange: owner %d would receive more rows than its buffers hold (capacity = one rank's "
                                "lookups per step): the step was not applied" % (st[3] - 1))
        raise _lib.CtrError("peer barrier timed out waiting for rank %d: the step was not applied" % ((st[0] or st[1]) - 1))

    # the stream a step runs on is looked up ONCE per public call (torch.cuda.current_stream costs ~7 us of host time, and a
    # step asked for it ~20 times: the multi-GPU step is close to host-bound)
    _main_cached = None
    _st_cached = None

    def _main_stream(self) -> torch.cuda.Stream:
        return self._main_cached if self._main_cached is not None else torch.cuda.current_stream(self.dev)

    def _stream(self) -> int:
        return self._st_cached if self._st_cached is not None else torch.cuda.current_stream(self.dev).cuda_stream

    def _pin_stream(self) -> None:
        self._main_cached = torch.cuda.current_stream(self.dev)
        self._st_cached = self._main_cached.cuda_stream

    def _unpin_stream(self) -> None:
        self._main_cached = self._st_cached = None

    host_prof = None       # tests/host_timing.py
    _pf_obj = None
    _sort_B = None         # batch size whose id sort forward() still has to enqueue (train steps only)
    phase_marks = None     # list of (name, event) when a timing script asks for the main-stream phase picture

    def _mark(self, name: str) -> None:
        if self.phase_marks is not None:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record(self._main_stream())
            self.phase_marks.append((name, ev))

    @property
    def _vw(self) -> Dict:
        d = self.__dict__.get("_vw_cache")
        if d is None:
            d = self.__dict__["_vw_cache"] = {}
        return d

    def view(self, name: str, buf: Optional[torch.Tensor] = None) -> torch.Tensor:
        """flat-buffer view of a dense parameter (internal layout)"""
        buf = self.theta if buf is None else buf
        rng = self._vw.get(name)
        if rng is None:
            o = self.off[name]
            rng = self._vw[name] = (o, o + int(np.prod(self.shape[name])))
        return buf[rng[0]:rng[1]]

    def _w(self, i: int, buf=None):
        return self.view("deep_%d" % i, buf)

    def _wd(self, i: int):
        """weights as read by the data-gradient GEMM (tf32-rounded copy on the TF32 path)"""
        return self.wr[i] if self.wr is not None else self._w(i)

    def _b(self, i: int, buf=None):
        return self.view("deep_bias_%d" % i, buf)

    def _head_parts(self, buf=None):
        """(w_fm, w_deep, bias) views of the head parameters"""
        s = self.spec
        H = s.hidden[-1]
        if s.head == "deepfm":
            w = self.view(s.head_w_name, buf)
            return w[:s.n_fm], w[s.n_fm:s.n_fm + H], self.view(s.head_b_name, buf)
        if s.head == "dnn":
            return None, self.view(s.head_w_name, buf), self.view(s.head_b_name, buf)
        # quirk Q4: in models/wdl.py:224-225 the deep weights alias rows W..W+H-1; test_model/wdl.py:110 puts them behind
        w = self.view("wdl_weights", buf)
        off = s.wide_fields if s.deep_row_offset is None else s.deep_row_offset
        return None, w[off:off + H], self.view("wdl_bias", buf)

    # ---- parameter injection / read-back (reference names and shapes) ------------------------------
    def set_weights(self, p: Dict[str, np.ndarray]) -> None:
        """Load variables named and shaped as in the reference model file (e.g. feats_emb [R,k],
        feats [R,1], deep_0 [D0,H0] in [cont|vector|emb] row order, deep_fm_weight [F+k+H,1])."""
        self._load(p, self.table, self.theta)
        self._refresh_transposes()

    def _load(self, p: Dict[str, np.ndarray], table: torch.Tensor, flat: torch.Tensor) -> None:
        s, st = self.spec, self._stream()
        t = lambda a: torch.as_tensor(np.ascontiguousarray(np.asarray(a, dtype=np.float32))).to(self.dev)
        emb_np = np.asarray(p[s.emb_name], dtype=np.float32).reshape(s.table_rows, s.k)[self.rank::self.world]
        emb = t(emb_np)
        first = t(np.asarray(p[s.first_name], dtype=np.float32).reshape(s.table_rows)[self.rank::self.world]) if s.has_first else None
        self.h.call("ctr_table_pack", st, ptr(emb), ptr(first), emb.shape[0], s.k, s.table_stride, ptr(table))
        if s.frozen_rows and table is self.table:
            fz = t(np.asarray(p[s.frozen_name], dtype=np.float32).reshape(s.frozen_rows, s.k))
            self.h.call("ctr_table_pack", st, ptr(fz), None, s.frozen_rows, s.k, s.table_stride, ptr(table[self.rows_local:]))
        for name, shp in self.dense_shapes:
            a = np.asarray(p[name], dtype=np.float32).reshape(shp)
            if name == "deep_0":
                a = a[s.deep0_perm(), :]
            if s.head == "deepfm" and name == s.head_w_name:
                perm = np.concatenate([s.head_fm_perm(), s.n_fm + np.arange(s.hidden[-1])])
                a = a[perm, :]
            self.view(name, flat).copy_(t(a).reshape(-1))
        self._main_stream().synchronize()

    def set_optimizer_state(self, m: Dict[str, np.ndarray], v: Dict[str, np.ndarray], global_step: int,
                            beta1_power: float, beta2_power: float) -> None:
        """Adam slots named and shaped like the variables they belong to (what tf.train.Saver stores,
        models/deepfm_multi.py:269-278,308-310), plus the step counters."""
        self._load(m, self.m_table, self.m)
        self._load(v, self.v_table, self.v)
        self.global_step, self.beta1_power, self.beta2_power = int(global_step), float(beta1_power), float(beta2_power)

    def set_dense_weights(self, p: Dict[str, np.ndarray]) -> None:
        """dense variables only (reference names/shapes); the table is left as is"""
        s = self.spec
        for name, shp in self.dense_shapes:
            a = np.asarray(p[name], dtype=np.float32).reshape(shp)
            if name == "deep_0":
                a = a[s.deep0_perm(), :]
            if s.head == "deepfm" and name == s.head_w_name:
                a = a[np.concatenate([s.head_fm_perm(), s.n_fm + np.arange(s.hidden[-1])]), :]
            self.view(name).copy_(torch.from_numpy(np.ascontiguousarray(a)).reshape(-1).to(self.dev))
        self._refresh_transposes()

    def init_table_on_device(self, seed: int = 2019, chunk_rows: int = 8 << 20) -> None:
        """reference initial distributions drawn directly in HBM (tables too large to stage on the
        host): embeddings ~ N(0, 0.01^2) (xavier-uniform for weight_mat), first-order ~ U(0,1)"""
        s = self.spec
        g = torch.Generator(device=self.dev); g.manual_seed(seed)
        lim = math.sqrt(6.0 / (s.table_rows + s.k))
        g.manual_seed(seed + 7919 * self.rank)
        for r0 in range(0, self.rows_local, chunk_rows):
            r1 = min(self.rows_local, r0 + chunk_rows)
            blk = self.table[r0:r1]
            if s.emb_name == "weight_mat":
                blk[:, :s.k] = (torch.rand(r1 - r0, s.k, generator=g, device=self.dev) * 2 - 1) * lim
            else:
                blk[:, :s.k] = torch.randn(r1 - r0, s.k, generator=g, device=self.dev) * 0.01
            if s.has_first:
                blk[:, s.k] = torch.rand(r1 - r0, generator=g, device=self.dev)

    def get_weights(self) -> Dict[str, np.ndarray]:
        self.check_status()
        return self._dump(self.table, self.theta)

    def get_optimizer_state(self):
        """(m, v) dicts named/shaped like the variables (the Adam slots tf.train.Saver would store)"""
        return self._dump(self.m_table, self.m), self._dump(self.v_table, self.v)

    def _dump(self, table: torch.Tensor, flat: torch.Tensor) -> Dict[str, np.ndarray]:
        s, st = self.spec, self._stream()
        out: Dict[str, np.ndarray] = {}
        emb = torch.empty(self.rows_local, s.k, dtype=torch.float32, device=self.dev)
        first = torch.empty(self.rows_local, dtype=torch.float32, device=self.dev) if s.has_first else None
        self.h.call("ctr_table_unpack", st, ptr(table), self.rows_local, s.k, s.table_stride, ptr(emb), ptr(first))
        out[s.emb_name] = self._assemble_rows(emb)
        if s.frozen_rows:
            fz = torch.zeros(s.frozen_rows, s.k, dtype=torch.float32, device=self.dev)
            if table is self.table:      # the Adam slots of a constant are zero
                self.h.call("ctr_table_unpack", st, ptr(table[self.rows_local:]), s.frozen_rows, s.k, s.table_stride, ptr(fz), None)
            out[s.frozen_name] = fz.cpu().numpy()
        if s.has_first:
            out[s.first_name] = self._assemble_rows(first.reshape(-1, 1))
        for name, shp in self.dense_shapes:
            a = self.view(name, flat).cpu().numpy().reshape(shp).copy()
            if name == "deep_0":
                inv = np.empty_like(s.deep0_perm()); inv[s.deep0_perm()] = np.arange(len(inv))
                a = a[inv, :]
            if s.head == "deepfm" and name == s.head_w_name:
                perm = np.concatenate([s.head_fm_perm(), s.n_fm + np.arange(s.hidden[-1])])
                inv = np.empty_like(perm); inv[perm] = np.arange(len(perm))
                a = a[inv, :]
            out[name] = a
        return out

    # ---- per-shard checkpoint (the reference's tf.train.Saver stores every variable with its Adam
    # slots and the step counters, models/deepfm_multi.py:269-278,308-310) ---------------------------------------------------
    @staticmethod
    def shard_file(path: str, rank: int, world: int) -> str:
        return os.path.join(path, "shard-%05d-of-%05d.npz" % (rank, world))

    def save_shard(self, path: str) -> str:
        """Write THIS rank's state: its rows of the table (device layout, no all-gather), their Adam slots, the replicated
        dense buffer and the step counters.  Every rank calls it; the files of one checkpoint are
        `shard-<rank>-of-<world>.npz`."""
        s = self.spec
        os.makedirs(path, exist_ok=True)
        self.check_status()          # a faulted step never reaches a checkpoint
        self._main_stream().synchronize()
        R = self.rows_local
        blob = {"meta": np.asarray([self.world, self.rank, s.table_rows, s.k, s.table_stride, self.global_step, self.n_dense,
                                    int(self.train_mode)], dtype=np.int64),
                "powers": np.asarray([self.beta1_power, self.beta2_power], dtype=np.float64),
                "table": self.table[:R].cpu().numpy(), "theta": self.theta.cpu().numpy()}
        if s.frozen_rows:
            blob["frozen"] = self.table[R:].cpu().numpy()
        if self.train_mode:
            blob.update({"m_table": self.m_table[:R].cpu().numpy(), "v_table": self.v_table[:R].cpu().numpy(),
                         "m": self.m.cpu().numpy(), "v": self.v.cpu().numpy()})
        f = self.shard_file(path, self.rank, self.world)
        tmp = f + ".tmp.npz"
        np.savez(tmp, **blob)
        os.replace(tmp, f)           # a crash mid-write never leaves a half-written shard under the final name
        return f

    def load_shard(self, path: str) -> None:
        """Restore from `save_shard` files.  Same world size: each rank reads its own file.  Different world size (train
        on 8, serve on 1): each rank walks all old shards and keeps the rows it owns now (row id = local * W_old + r_old,
        new owner id % W_new, new local row id // W_new)."""
        s = self.spec
        files = sorted(f for f in os.listdir(path) if f.startswith("shard-") and f.endswith(".npz") and ".tmp" not in f)
        if not files:
            raise FileNotFoundError("no shard-*.npz under %s" % path)
        w_old = int(files[0].split("-of-")[1].split(".")[0])
        if len(files) != w_old:
            raise ValueError("checkpoint %s is incomplete: %d of %d shards" % (path, len(files), w_old))
        R = self.rows_local
        mine = [self.rank] if w_old == self.world else range(w_old)
        first = True
        for r_old in mine:
            z = np.load(self.shard_file(path, r_old, w_old))
            meta = z["meta"]
            if int(meta[2]) != s.table_rows or int(meta[3]) != s.k or int(meta[4]) != s.table_stride or int(meta[6]) != self.n_dense:
                raise ValueError("checkpoint %s was written for another model (rows %d k %d)" % (path, meta[2], meta[3]))
            has_opt = bool(meta[7]) and self.train_mode
            pairs = [("table", self.table)] + ([("m_table", self.m_table), ("v_table", self.v_table)] if has_opt else [])
            if w_old == self.world:
                for name, dst in pairs:
                    dst[:R].copy_(torch.from_numpy(z[name]))
            else:
                keep, loc_np = _shard.reshard_rows(z["table"].shape[0], w_old, r_old, self.world, self.rank, s.table_rows)
                loc = torch.from_numpy(loc_np).to(self.dev)
                for name, dst in pairs:
                    dst[loc] = torch.from_numpy(z[name][keep]).to(self.dev)
            if first:
                self.theta.copy_(torch.from_numpy(z["theta"]))
                if s.frozen_rows:
                    self.table[R:].copy_(torch.from_numpy(z["frozen"]))
                if has_opt:
                    self.m.copy_(torch.from_numpy(z["m"])); self.v.copy_(torch.from_numpy(z["v"]))
                self.global_step = int(meta[5])
                self.beta1_power, self.beta2_power = float(z["powers"][0]), float(z["powers"][1])
                first = False
        self._refresh_transposes()
        self._main_stream().synchronize()

    def _assemble_rows(self, shard: torch.Tensor) -> np.ndarray:
        """full [table_rows, w] host array from this rank's shard (one GPU: the shard IS the table)"""
        return shard.cpu().numpy()

    def _dense_layout(self) -> DenseLayout:
        """positions of the regulariser groups and of the tower weights in the flat buffer (built once)"""
        if self._lay is None:
            s, lay = self.spec, DenseLayout()
            for gi in range(3):
                lay.group_end[gi] = self.groups[gi][1]
            lay.reg_scale = float(s.l2_reg)
            dims = self._layer_dims()
            if self.wt is not None and len(dims) > MAX_LAYERS:
                raise ValueError("at most %d tower layers on the TF32 path" % MAX_LAYERS)
            lay.n_layers = len(dims) if self.wt is not None else 0
            for i, (k_, h_) in enumerate(dims[:lay.n_layers]):
                lay.w_off[i], lay.K[i], lay.N[i], lay.ld_kmajor[i] = self.off["deep_%d" % i], k_, h_, _ru(k_, 4)
                lay.w_kmajor[i] = self.wt[i].data_ptr()
                lay.w_rounded[i] = self.wr[i].data_ptr() if self.wr is not None else None
            self._lay = lay
        return self._lay

    _lay = None

    def _refresh_transposes(self) -> None:
        if self.wt is None:
            return
        st = self._stream()
        for i, (k_, h_) in enumerate(self._layer_dims()):
            self.h.call("ctr_transpose", st, ptr(self._w(i)), k_, h_, ptr(self.wt[i]), _ru(k_, 4), 1)
            if self.wr is not None:
                self.h.call("ctr_round_tf32_copy", st, ptr(self._w(i)), ptr(self.wr[i]), k_ * h_)

    # ---- batch upload ------------------------------------------------------------------------------
    def _use_inputs(self, i: int) -> None:
        self._cur_in = i
        t = self._in[i]
        self.ids, self.cont, self.vec, self.label, self.wide = t["ids"], t["cont"], t["vec"], t["label"], t["wide"]
        self.sw = t["sw"]

    def load_batch(self, batch: Dict, non_blocking: bool = True, skip_ids: bool = False) -> int:
        """Make `batch` (host or device arrays/tensors, reference key names) the current device batch.  Returns B
        (may be < max_batch for the short last batch, models/deepfm.py:178).  A batch announced as `next_batch`
        of the previous train_step is already on the device: the buffer sets just flip."""
        if self._pf_id is not None and self._pf_id == id(batch):
            self._main_stream().wait_event(self._pf_event)
            self._use_inputs(1 - self._cur_in)
            self._pf_id = self._pf_obj = None
            self.last_batch = self._pf_B
            return self._pf_B
        self._pf_id = self._pf_obj = None
        B = self._copy_inputs(batch, self._in[self._cur_in], non_blocking, skip_ids)
        self.last_batch = B
        return B

    def _copy_inputs(self, batch: Dict, tgt: Dict, non_blocking: bool, skip_ids: bool) -> int:
        s = self.spec

        def as_t(x, dtype):
            if isinstance(x, torch.Tensor):
                return x if x.dtype == dtype else x.to(dtype)
            return torch.as_tensor(np.asarray(x)).to(dtype)
        cate = as_t(batch["cate_feats"], self.id_torch)
        B = int(cate.shape[0])
        if B > self.Bmax:
            raise ValueError("batch of %d exceeds max_batch %d" % (B, self.Bmax))
        if s.ids_ld:
            if cate.shape[1] != s.ids_ld:
                raise ValueError("cate_feats has %d columns, model expects %d" % (cate.shape[1], s.ids_ld))
            if not skip_ids:
                tgt["ids"][:B].copy_(cate, non_blocking=non_blocking)
                if self.frozen_off is not None:      # 'tag' slots address the frozen rows behind the trainable table
                    tgt["ids"][:B] += self.frozen_off
        if s.weighted_pool:
            sw = as_t(batch["multi_feats_value"] if "multi_feats_value" in batch else batch["mul_cat_feats_value"], torch.float32)
            if sw.shape[1] != s.M:
                raise ValueError("multi_feats_value has %d columns, model expects %d" % (sw.shape[1], s.M))
            tgt["sw"][:B].copy_(sw, non_blocking=non_blocking)
        cont = as_t(batch["cont_feats"], torch.float32)
        if s.alg == "dnn" and "vector_feats" in batch and batch["vector_feats"] is not None:
            vec = as_t(batch["vector_feats"], torch.float32)
            if vec.shape[1] and cont.shape[1] + vec.shape[1] == s.C:
                cont = torch.cat([cont, vec.to(cont.device)], dim=1)
        if s.C:
            if cont.shape[1] != s.C:
                raise ValueError("cont_feats has %d columns, model expects %d" % (cont.shape[1], s.C))
            tgt["cont"][:B].copy_(cont, non_blocking=non_blocking)
        if s.V:
            tgt["vec"][:B].copy_(as_t(batch["vector_feats"], torch.float32), non_blocking=non_blocking)
        if s.head == "wdl":
            tgt["wide"][:B].copy_(as_t(batch["wide_feats"], self.id_torch), non_blocking=non_blocking)
        lab = batch.get("label", batch.get("labels"))
        if lab is not None:
            tgt["label"][:B].copy_(as_t(lab, torch.float32).reshape(-1), non_blocking=non_blocking)
        return B

    # ---- forward -------------------------------------------------------------------------------------
    def _embed_forward(self, B: int) -> None:
        """K2 + K1: rows of the (local) table -> X0, fm, sum_e"""
        s, st, h = self.spec, self._stream(), self.h
        cfg = self._cfg(B)
        if s.Fm:
            h.call("ctr_pooled_gather_fwd", st, cfg, ptr(self.table), ptr(self.ids), ptr(self.multi_start), ptr(self.sw),
                   ptr(self.x0), ptr(self.pooled_first), ptr(self.inv_cnt_emb), ptr(self.inv_cnt_first))
        h.call("ctr_fm_gather_fwd", st, cfg, ptr(self.table), ptr(self.ids), 0, ptr(self.cont), ptr(self.vec),
               ptr(self._cont_rows()), ptr(self.pooled_first), ptr(self.x0), ptr(self.fm), ptr(self.sum_e))

    def _cont_rows(self):
        """the C table rows of the continuous fields (models/deepfm.py:63-66), contiguous"""
        s = self.spec
        return self.table[s.cont_base:] if (s.use_fm and s.fm_cont) else None

    def forward(self, B: int, with_loss: bool) -> None:
        s, st, h = self.spec, self._stream(), self.h
        self._embed_forward(B)
        if self._sort_B is not None and not _SORT_AFTER_HEAD:
            # the id sort starts BEHIND the gather: K1 is the HBM-bound kernel of the step and a co-running radix sort took
            # 20 % of its bandwidth; under the tensor-core GEMMs that follow the sort is close to free, and its result is
            # only needed after the MLP backward
            self._launch_sort(self._sort_B)
            self._sort_B = None
        self._mark("embed")
        x, ldx = self.x0, s.x0_ld
        for i, (k_, h_) in enumerate(self._layer_dims()):
            if self.relu_bits is not None and self.train_mode and i < len(self.relu_bits):
                h.call("ctr_mlp_fwd_bits", st, B, k_, h_, ptr(x), ldx, ptr(self.wt[i]), ptr(self._b(i)), 1.0, 0,
                       ptr(self.acts[i]), self.act_ld[i], ptr(self.relu_bits[i]))
            else:
                h.call("ctr_mlp_fwd", st, self.prec, B, k_, h_, ptr(x), ldx, ptr(self._w(i)),
                       ptr(self.wt[i]) if self.wt is not None else None, ptr(self._b(i)), 1, 1.0, 0, ptr(self.acts[i]), self.act_ld[i])
            x, ldx = self.acts[i], self.act_ld[i]
        w_fm, w_deep, bias = self._head_parts()
        self._mark("mlp_fwd")
        extra = None
        if s.head == "wdl":
            h.call("ctr_scalar_gather_sum", st, ptr(self.view("wdl_weights")), ptr(self.wide), self.id_bytes, B,
                   s.wide_fields, s.wide_rows + s.hidden[-1], ptr(self.status), ptr(self.wide_sum))
            extra = self.wide_sum
        lab = self.label if with_loss else None
        g_fm = g_deep = g_bias = None
        if with_loss and self.train_mode:
            # head-weight gradients d w = sum_b dlogit[b] * feats_input[b, :] (models/deepfm.py:130,143) come out of the
            # same kernel; for wdl the deep block aliases rows W..W+H-1 of wdl_weights (quirk Q4), zeroed first so the
            # wide scatter can add on top
            if s.head == "wdl":
                self.view("wdl_weights", self.grad).zero_()
            g_fm, g_deep, g_bias = self._head_parts(self.grad)
        h.call("ctr_head_fwd", st, B, self.global_batch(B), ptr(self.fm), s.fm_ld, s.n_fm, ptr(w_fm), ptr(x), ldx, s.hidden[-1], ptr(w_deep),
               ptr(extra), ptr(bias), ptr(lab), ptr(self.logit), ptr(self.score),
               ptr(self.dlogit) if with_loss else None,
               ptr(self.dz[0]) if (with_loss and self.train_mode) else None, s.hidden[-1], int(self.prec == 1),
               ptr(self.loss) if with_loss else None, ptr(g_fm), ptr(g_deep), ptr(g_bias), ptr(self.ws), self.ws.numel(),
               _lib.HEAD_SCE_ON_SCORE if s.loss == "sce_on_score" else 0)
        if self._sort_B is not None:        # CTR_SORT_AT=head: the sort runs under the backward GEMMs instead (A/B knob)
            self._launch_sort(self._sort_B)
            self._sort_B = None

    def global_batch(self, B: int) -> int:
        """samples the loss is a mean over (all ranks' batches when sharded)"""
        return B

    def _reg_loss(self, st: Optional[int] = None, ws: Optional[torch.Tensor] = None) -> torch.Tensor:
        """regulariser terms of the reported loss (models/deepfm.py:155-156, models/wdl.py:268-275,
        models/dnn.py:88-90): one deterministic reduction per regulariser group of the flat buffer.  `st` / `ws`: another
        stream than the step's needs a workspace of its own (the torch arithmetic follows torch's current stream)."""
        s = self.spec
        st = self._stream() if st is None else st
        ws = self.ws if ws is None else ws
        total = None
        for gi, mode, scale in ((1, 1, 0.5 * s.l2_reg), (2, 2, s.l2_reg)):
            a, b = self.groups[gi]
            if b > a:
                out = self.reg_terms[gi:gi + 1]
                self.h.call("ctr_reduce", st, ptr(self.theta[a:b]), b - a, mode, ptr(out), ptr(ws), ws.numel())
                total = out * scale if total is None else total + out * scale
        return total

    # ---- backward + update -------------------------------------------------------------------------
    def _adam_hp(self) -> AdamHP:
        """TF1 Adam step size (adam.py): lr from exponential_decay(staircase) on the PRE-increment
        global_step (models/deepfm.py:159-162), then lr_t = lr*sqrt(1-b2^t)/(1-b1^t)."""
        lr = self.lr0 * self.decay_rate ** math.floor(self.global_step / self.decay_steps)
        if self.spec.optimizer == "ftrl":      # test_model/wdl.py:132-136: FtrlOptimizer(learning_rate=<decayed lr>)
            return AdamHP(lr_t=lr, guard=self.status.data_ptr(), kind=AdamHP.FTRL)
        lr_t = lr * math.sqrt(1.0 - self.beta2_power) / (1.0 - self.beta1_power)
        return AdamHP(lr_t=lr_t, beta1=ADAM_BETA1, beta2=ADAM_BETA2, eps=ADAM_EPS, guard=self.status.data_ptr())

    def backward_and_update(self, B: int) -> None:
        s, st, h = self.spec, self._stream(), self.h
        cfg = self._cfg(B)
        L = len(s.hidden)
        dims = self._layer_dims()
        H = s.hidden[-1]
        w_fm, w_deep, _ = self._head_parts()
        ws, wsn = ptr(self.ws), self.ws.numel()
        # head-weight gradients were written by ctr_head_fwd
        if s.head == "wdl":
            self._wdl_wide_grad(B)
        if s.ids_ld and self.n_extra:
            # gradient of the C continuous-field rows: needs only dlogit / sum_e / cont, so it runs on its own stream
            # underneath the tower's backward GEMMs (26 us off the critical path at config 4), with its own workspace
            main = self._main_stream()
            self.ev_head.record(main)
            self.aux.wait_event(self.ev_head)
            h.call("ctr_fm_cont_bwd", self.aux.cuda_stream, cfg, ptr(self.cont), ptr(self._cont_rows()), ptr(self.sum_e),
                   ptr(self.dlogit), ptr(w_fm), ptr(self.cont_grad), ptr(self.ws_cont), self.ws_cont.numel())
            self.ev_cont.record(self.aux)
        # dz[0] already holds the gradient through the last ReLU (written by ctr_head_fwd)
        if self.bwd_overlap and s.ids_ld:
            self._backward_overlapped(B, cfg, dims, w_fm)
            return
        if self.bwd_split and s.ids_ld:
            self._backward_split(B, cfg, dims, w_fm)
            return
        cur = 0
        for i in range(L - 1, -1, -1):
            k_, h_ = dims[i]
            x, ldx = (self.acts[i - 1], self.act_ld[i - 1]) if i > 0 else (self.x0, s.x0_ld)
            h.call("ctr_mlp_bwd_weight", st, self.prec, B, k_, h_, ptr(x), ldx, ptr(self.dz[cur]), h_,
                   ptr(self._w(i, self.grad)), ptr(self._b(i, self.grad)), ws, wsn, 1)
            if i > 0:
                self._dgrad_hidden(st, B, i, k_, h_, self.dz[cur], self.dz[cur ^ 1])
                cur ^= 1
            elif s.E:
                # only the embedding columns of X0 carry parameters: canonical order puts them first
                h.call("ctr_mlp_bwd_data", st, self.prec, B, s.E, h_, ptr(self.dz[cur]), h_, ptr(self._wd(0)), None, 0,
                       1.0, ptr(self.dx0), self.dx0_ld)
        hp = self._adam_hp()
        self._mark("mlp_bwd")
        self._begin_dense_reduce()          # every dense gradient is final: the reduction runs under the sparse side
        # ---- sparse side: sort -> runs -> partials -> Adam on touched rows ------------------------------
        if s.ids_ld:
            n = B * s.ids_ld + self.n_extra
            self._main_stream().wait_event(self.ev_sorted)
            self._mark("wait_sort")
            if self.n_extra:
                self._main_stream().wait_event(self.ev_cont)
            h.call("ctr_embed_bwd_partials_weighted", st, cfg, ptr(self.keys), ptr(self.perm), ptr(self.uidx), n,
                   ptr(self.field_of_slot), ptr(self.x0), ptr(self.dx0), self.dx0_ld, ptr(self.sum_e), ptr(self.dlogit),
                   ptr(w_fm), ptr(self.inv_cnt_emb), ptr(self.inv_cnt_first), ptr(self.cont_grad), ptr(self.sw),
                   ptr(self.part))
            self._mark("partials")
            self._sparse_update(B, n, hp)
            self._mark("sparse_update")
        self._reduce_dense_grads()
        self._mark("dense_reduce")
        if self.adam_mode == "tf_dense" and s.optimizer != "ftrl":
            # TF1 Adam moves EVERY row each step (m, v decay; theta -= lr_t*m/(sqrt(v)+eps)) -- exact mode
            h.call("ctr_adam_dense", st, ptr(self.table), ptr(self.m_table), ptr(self.v_table), None,
                   self.rows_local * s.table_stride, 0.0, 0.0, ptr(self.touched), s.table_stride, hp)
        # ---- dense side: Adam over the flat buffer + the GEMM-side weight copies, one launch ---------------------------
        self._before_dense_update()
        h.call("ctr_dense_update", st, ptr(self.theta), ptr(self.m), ptr(self.v), ptr(self._dense_grad()), C.byref(self._dense_layout()), hp)
        self.beta1_power *= ADAM_BETA1
        self.beta2_power *= ADAM_BETA2
        self.global_step += 1
        self._steps_done += 1

    def _dgrad_hidden(self, st: int, B: int, i: int, k_: int, h_: int, dz_in, dz_out) -> None:
        """dZ of layer i-1 = (dZ of layer i @ W_i^T) masked by the ReLU of layer i-1's output (models/deepfm.py:121-126
        differentiated): sign bits left by the forward pass when there are any, else the fp32 activation itself"""
        if self.relu_bits is not None:
            self.h.call("ctr_mlp_bwd_data_bits", st, B, k_, h_, ptr(dz_in), h_, ptr(self._wd(i)), ptr(self.relu_bits[i - 1]), 1.0,
                        ptr(dz_out), k_)
        else:
            self.h.call("ctr_mlp_bwd_data", st, self.prec, B, k_, h_, ptr(dz_in), h_, ptr(self._wd(i)), ptr(self.acts[i - 1]),
                        self.act_ld[i - 1], 1.0, ptr(dz_out), k_)

    class _OnStream:
        """while active, every helper of the engine enqueues on `stream` instead of the step's main stream"""

        def __init__(self, eng, stream):
            self.eng, self.stream = eng, stream

        def __enter__(self):
            e = self.eng
            self.saved = (e._main_cached, e._st_cached, e._on_sparse)
            e._main_cached, e._st_cached, e._on_sparse = self.stream, self.stream.cuda_stream, True

        def __exit__(self, *exc):
            e = self.eng
            e._main_cached, e._st_cached, e._on_sparse = self.saved
            return False

    _on_sparse = False
    bwd_overlap = _BWD_OVERLAP      # (the sharded engine turns it off outside the peer-memory exchange)

    bwd_split = False         # CTR_BWD_ORDER=split, sharded peer-memory engines only (opt-in: measured no faster)

    def _backward_split(self, B: int, cfg, dims, w_fm) -> None:
        """Backward of the sharded step, ONE stream, reordered:  dgrad L-1 .. dgrad 0 (-> dX0) | partials | gradient rows
        stored into their owners' buffers | wgrad L-1 .. wgrad 0 | barrier | owner-side row update | dense update.
        The idea: the stores over NVLink and the ranks' skew drain under the three weight-gradient GEMMs.  Measured at N = 2:
        no gain (1.1007 vs 1.0968 ms), the barrier still costs 24 us behind the GEMMs -- kept as an opt-in A/B."""
        s, st, h = self.spec, self._stream(), self.h
        L = len(s.hidden)
        ws, wsn = ptr(self.ws), self.ws.numel()
        main = self._main_stream()
        for i in range(L - 1, 0, -1):                      # dz[j] is dZ of layer L-1-j
            j = L - 1 - i
            k_, h_ = dims[i]
            self._dgrad_hidden(st, B, i, k_, h_, self.dz[j], self.dz[j + 1])
        if s.E:
            h.call("ctr_mlp_bwd_data", st, self.prec, B, s.E, dims[0][1], ptr(self.dz[L - 1]), dims[0][1], ptr(self._wd(0)), None, 0,
                   1.0, ptr(self.dx0), self.dx0_ld)
        hp = self._adam_hp()
        self._mark("mlp_bwd_data")
        n = B * s.ids_ld + self.n_extra
        main.wait_event(self.ev_sorted)
        if self.n_extra:
            main.wait_event(self.ev_cont)
        h.call("ctr_embed_bwd_partials_weighted", st, cfg, ptr(self.keys), ptr(self.perm), ptr(self.uidx), n,
               ptr(self.field_of_slot), ptr(self.x0), ptr(self.dx0), self.dx0_ld, ptr(self.sum_e), ptr(self.dlogit),
               ptr(w_fm), ptr(self.inv_cnt_emb), ptr(self.inv_cnt_first), ptr(self.cont_grad), ptr(self.sw),
               ptr(self.part))
        self._mark("partials")
        self._sparse_push(B, n)
        self._mark("finalize")
        for i in range(L - 1, -1, -1):
            k_, h_ = dims[i]
            x, ldx = (self.acts[i - 1], self.act_ld[i - 1]) if i > 0 else (self.x0, s.x0_ld)
            h.call("ctr_mlp_bwd_weight", st, self.prec, B, k_, h_, ptr(x), ldx, ptr(self.dz[L - 1 - i]), h_,
                   ptr(self._w(i, self.grad)), ptr(self._b(i, self.grad)), ws, wsn, 1)
        self._mark("mlp_bwd_weight")
        self._begin_dense_reduce()
        self._sparse_apply(B, n, hp)
        self._mark("sparse_update")
        self._reduce_dense_grads()
        self._mark("dense_reduce")
        if self.adam_mode == "tf_dense" and s.optimizer != "ftrl":
            h.call("ctr_adam_dense", st, ptr(self.table), ptr(self.m_table), ptr(self.v_table), None,
                   self.rows_local * s.table_stride, 0.0, 0.0, ptr(self.touched), s.table_stride, hp)
        self._before_dense_update()
        h.call("ctr_dense_update", st, ptr(self.theta), ptr(self.m), ptr(self.v), ptr(self._dense_grad()), C.byref(self._dense_layout()), hp)
        self.beta1_power *= ADAM_BETA1
        self.beta2_power *= ADAM_BETA2
        self.global_step += 1
        self._steps_done += 1

    def _sparse_push(self, B: int, n: int) -> None:
        """first half of _sparse_update where it splits (sharded engines): gradient rows -> their owners"""

    def _sparse_apply(self, B: int, n: int, hp: AdamHP) -> None:
        """second half: the row update"""
        self._sparse_update(B, n, hp)

    def _backward_overlapped(self, B: int, cfg, dims, w_fm) -> None:
        """Backward with the sparse side underneath the weight-gradient GEMMs.
        main stream   : dgrad L-1 .. dgrad 0 (-> dX0) | wgrad L-1 .. wgrad 0 | dense reduce + dense update | wait(sparse)
        sparse stream :                  wait(dX0, sort, cont) -> partials -> [finalize/push, barrier] -> row Adam
        The weight gradients do not depend on the data gradient of their own layer, only on dZ of that layer, which the
        dgrad chain has already produced (one dZ buffer per layer)."""
        s, st, h = self.spec, self._stream(), self.h
        L = len(s.hidden)
        ws, wsn = ptr(self.ws), self.ws.numel()
        main = self._main_stream()
        for i in range(L - 1, 0, -1):                      # dz[j] is dZ of layer L-1-j
            j = L - 1 - i
            k_, h_ = dims[i]
            self._dgrad_hidden(st, B, i, k_, h_, self.dz[j], self.dz[j + 1])
        if s.E:      # only the embedding columns of X0 carry parameters: canonical order puts them first
            h.call("ctr_mlp_bwd_data", st, self.prec, B, s.E, dims[0][1], ptr(self.dz[L - 1]), dims[0][1], ptr(self._wd(0)), None, 0,
                   1.0, ptr(self.dx0), self.dx0_ld)
        self.ev_dx0.record(main)
        hp = self._adam_hp()
        self._mark("mlp_bwd_data")
        # ---- sparse side, on its own stream ---------------------------------------------------------------------
        sp = self.sparse_stream
        sp.wait_event(self.ev_dx0)
        sp.wait_event(self.ev_sorted)
        if self.n_extra:
            sp.wait_event(self.ev_cont)
        n = B * s.ids_ld + self.n_extra
        with self._OnStream(self, sp):
            self.h.call("ctr_embed_bwd_partials_weighted", sp.cuda_stream, cfg, ptr(self.keys), ptr(self.perm), ptr(self.uidx), n,
                        ptr(self.field_of_slot), ptr(self.x0), ptr(self.dx0), self.dx0_ld, ptr(self.sum_e), ptr(self.dlogit),
                        ptr(w_fm), ptr(self.inv_cnt_emb), ptr(self.inv_cnt_first), ptr(self.cont_grad), ptr(self.sw),
                        ptr(self.part))
            self._sparse_update(B, n, hp)
            if self.adam_mode == "tf_dense" and s.optimizer != "ftrl":
                # TF1 Adam moves EVERY row each step (m, v decay; theta -= lr_t*m/(sqrt(v)+eps)) -- exact mode
                h.call("ctr_adam_dense", sp.cuda_stream, ptr(self.table), ptr(self.m_table), ptr(self.v_table), None,
                       self.rows_local * s.table_stride, 0.0, 0.0, ptr(self.touched), s.table_stride, hp)
        self.ev_sparse.record(sp)
        # ---- weight gradients on the main stream, the sparse side running underneath ----------------------------------
        for i in range(L - 1, -1, -1):
            k_, h_ = dims[i]
            x, ldx = (self.acts[i - 1], self.act_ld[i - 1]) if i > 0 else (self.x0, s.x0_ld)
            h.call("ctr_mlp_bwd_weight", st, self.prec, B, k_, h_, ptr(x), ldx, ptr(self.dz[L - 1 - i]), h_,
                   ptr(self._w(i, self.grad)), ptr(self._b(i, self.grad)), ws, wsn, 1)
        self._mark("mlp_bwd_weight")
        self._begin_dense_reduce()
        self._reduce_dense_grads()
        self._mark("dense_reduce")
        self._before_dense_update()
        h.call("ctr_dense_update", st, ptr(self.theta), ptr(self.m), ptr(self.v), ptr(self._dense_grad()), C.byref(self._dense_layout()), hp)
        # the table is final (and, sharded, every peer past its gradient barrier) before anything that follows on the main
        # stream: the next step's gather, get_weights, a checkpoint
        main.wait_event(self.ev_sparse)
        self._mark("sparse_done")
        self.beta1_power *= ADAM_BETA1
        self.beta2_power *= ADAM_BETA2
        self.global_step += 1
        self._steps_done += 1

    def _sparse_update(self, B: int, n: int, hp: AdamHP) -> None:
        """per-run partial gradients (self.part) -> TF1 Adam on the touched rows of the local table"""
        s, st = self.spec, self._stream()
        self.h.call("ctr_adam_rows", st, ptr(self.uniq), ptr(self.n_unique), n, 0, ptr(self.part), ptr(self.seg_start),
                    s.grad_stride, s.row_width, ptr(self.table), ptr(self.m_table), ptr(self.v_table), s.table_stride,
                    self.rows_local, ptr(self.touched), ptr(self.ws_sparse), self.ws_sparse.numel(), self._fm_w2(), s.k, hp)

    def _fm_w2(self):
        """device pointer of the k second-order head weights (the e-term of the FM gradient is applied in ctr_adam_rows)"""
        s = self.spec
        if not s.use_fm:
            return None
        w_fm = self._head_parts()[0]
        return w_fm.data_ptr() + 4 * (s.Fc + s.Fm + s.C)

    def _dense_grad(self) -> torch.Tensor:
        """the buffer the dense update reads (sharded engines: the sum over ranks)"""
        return self.grad

    def _begin_dense_reduce(self) -> None:
        """data-parallel hook: start summing the dense gradients over ranks (no-op on one GPU)"""

    def _reduce_dense_grads(self) -> None:
        """data-parallel hook: the dense gradients are summed over ranks from here on (no-op on one GPU)"""

    def _launch_sort(self, B: int) -> None:
        """sort ids -> (keys, perm), run-length structure -> (uniq, seg_start, uidx, n_unique); enqueued on
        the side stream right behind the gather (K1) so it overlaps the MLP"""
        s, h = self.spec, self.h
        if not s.ids_ld:
            return
        main = self._main_stream()
        self.ev_batch.record(main)
        self.side.wait_event(self.ev_batch)
        n = B * s.ids_ld + self.n_extra
        key_bits = max(1, int(s.table_rows + s.frozen_rows - 1).bit_length())
        st, ws, wsn = self.side.cuda_stream, ptr(self.ws_sort), self.ws_sort.numel()
        h.call("ctr_sort_ids", st, ptr(self.ids), self.id_bytes, B, s.ids_ld, 0, s.ids_ld, 0, int(s.zero_row0),
               s.cont_base, self.n_extra, 1, 0, key_bits, self._id_limit(), ptr(self.status), ptr(self.keys), ptr(self.perm),
               ws, wsn)
        h.call("ctr_unique_segments", st, ptr(self.keys), n, ptr(self.uniq), ptr(self.seg_start), ptr(self.uidx),
               ptr(self.n_unique), ws, wsn, None)
        self.ev_sorted.record(self.side)
        self._sort_wide(B, st)

    def _sort_wide(self, B: int, st: int) -> None:
        """run structure of the wide ids (gradient of the scalar gather of models/wdl.py:240-246)"""
        s, h = self.spec, self.h
        if s.head == "wdl" and s.wide_fields and self.train_mode:
            nw = B * s.wide_fields
            rows = s.wide_rows + s.hidden[-1]
            ws, wsn = ptr(self.ws_sort_w), self.ws_sort_w.numel()
            h.call("ctr_sort_ids", st, ptr(self.wide), self.id_bytes, B, s.wide_fields, 0, s.wide_fields, 0, 0, 0, 0, 1, 0,
                   max(1, int(rows - 1).bit_length()), rows, ptr(self.status), ptr(self.w_keys), ptr(self.w_perm), ws, wsn)
            h.call("ctr_unique_segments", st, ptr(self.w_keys), nw, ptr(self.w_uniq), ptr(self.w_seg_start),
                   ptr(self.w_uidx), ptr(self.w_n_unique), ws, wsn, None)
            self.ev_sorted_w.record(self.side if st == self.side.cuda_stream else self._main_stream())

    _ws_wide = None

    def _wdl_wide_grad(self, B: int) -> None:
        """d wdl_weights[id] += sum of dlogit over the wide lookups of id (added on top of the deep-projection
        rows W..W+H-1 that the colsum already wrote: quirk Q4 makes the two alias, models/wdl.py:224-228)"""
        s = self.spec
        if not s.wide_fields:
            return
        self._main_stream().wait_event(self.ev_sorted_w)
        if self._ws_wide is None:       # piece / run lists of the long-run path (zeroed once; every call leaves the counters at zero)
            self._ws_wide = torch.zeros(self.h.query("ctr_scalar_segment_add_ws_bytes"), dtype=torch.uint8, device=self.dev)
        self.h.call("ctr_scalar_segment_add", self._stream(), ptr(self.w_uniq), ptr(self.w_seg_start), ptr(self.w_n_unique),
                    B * s.wide_fields, ptr(self.w_perm), ptr(self.dlogit), s.wide_fields, s.wide_rows + s.hidden[-1],
                    ptr(self.view("wdl_weights", self.grad)), ptr(self._ws_wide), self._ws_wide.numel())
        self._gathered_l2_term(B)

    def _gathered_l2_term(self, B: int) -> None:
        """quirk Q6 (models/wdl_textline.py:264-269,284-286): l2 over the gathered [B, W+H] weights; the deep rows
        W..W+H-1 are looked up once per sample.  Gradient goes on top of the wdl_weights gradient, the local share of the
        loss term into self.loss BEFORE the dense reduction, so the sum over ranks is the term of the global batch."""
        s = self.spec
        if not (s.gathered_l2 and s.l2_reg > 0):
            return
        nw, H = B * s.wide_fields, s.hidden[-1]
        self.h.call("ctr_scalar_gathered_l2", self._stream(), ptr(self.w_uniq), ptr(self.w_seg_start), ptr(self.w_n_unique),
                    nw, ptr(self.view("wdl_weights")), s.l2_reg, s.wide_rows + H, s.wide_fields, H, float(B),
                    ptr(self.view("wdl_weights", self.grad)), ptr(self.w_sq))
        self.h.call("ctr_reduce", self._stream(), ptr(self.w_sq), nw + H, 0, ptr(self.reg_terms[3:4]), ptr(self.ws),
                    self.ws.numel())
        self.loss.add_(self.reg_terms[3:4], alpha=0.5 * s.l2_reg)

    def eval_loss(self, batch: Optional[Dict] = None) -> torch.Tensor:
        """`sess.run([label, out, loss])` without the optimizer (the evaluation half of the textline fit loops,
        models/wdl_textline.py:313, models/dnn_multi_textline.py:185): the loss of `batch` under the current weights,
        scores in self.score.  One GPU."""
        if not self.train_mode or self.world > 1:
            raise _lib.CtrError("eval_loss needs a single-GPU training engine")
        s = self.spec
        B = self.load_batch(batch) if batch is not None else self.last_batch
        self.forward(B, with_loss=True)
        if s.gathered_l2 and s.l2_reg > 0:
            self._sort_wide(B, self._stream())
            self._gathered_l2_term(B)
        reg = self._reg_loss()
        loss = self.loss.clone()
        return loss if reg is None else loss + reg

    # ---- public step API -------------------------------------------------------------------------------
    def train_step(self, batch: Optional[Dict] = None, next_batch: Optional[Dict] = None) -> torch.Tensor:
        self._pin_stream()
        try:
            return self._train_step(batch, next_batch)
        finally:
            self._unpin_stream()

    def _train_step(self, batch: Optional[Dict] = None, next_batch: Optional[Dict] = None) -> torch.Tensor:
        """One `sess.run([loss, optimizer])` (models/deepfm.py:181).  Returns the loss as a device
        scalar tensor (log-loss + regulariser); call .item() to read it.  `next_batch` (optional) is the batch the
        NEXT call will train on: the input pipeline knows it one step ahead (the reference's tf.data iterator
        prefetches the same way, utils/data_loader.py:29-40), and the sharded engine uses it to take the id
        de-duplication and id exchange of step i+1 off the critical path of step i."""
        if not self.train_mode:
            raise _lib.CtrError("engine was created with train=False")
        if self._status_pending:
            self.check_status(wait=False)      # a fault of an earlier step whose snapshot has landed: raise now
        self._mark("start")
        ev_call = None
        if next_batch is not None:
            # whatever produced `next_batch` was enqueued before this call: work on other streams that reads it waits
            # for this point of the main stream and no later one
            ev_call = torch.cuda.Event()
            ev_call.record(self._main_stream())
        B = self.load_batch(batch) if batch is not None else self.last_batch
        self._sort_B = B                    # enqueued by forward(), right behind the gather
        self._mark("load")
        self.forward(B, with_loss=True)
        self._mark("head")
        # queued from the host only now, with the whole forward already in the main stream's queue (the launches of the
        # prefetch would otherwise sit between this step's upload and its first kernel); it still STARTS early, since
        # it waits for ev_call and not for the forward
        self._prefetch_next(next_batch, ev_call)
        out = None
        if self.loss_is_early:
            # the value is final once the head kernel has run: the regulariser sums, the additions and the copy to the host
            # go to the loss stream now, off the step's critical path (six small launches), and a caller that reads every
            # step's loss (last_loss()) waits for the forward pass only and queues the next step under the backward
            out = self._publish_loss(None, early=True)
            self.backward_and_update(B)
            self._main_stream().wait_event(self._loss_dev_done)     # the returned tensor is safe on the caller's stream
        else:
            # the loss itself arrives late (sharded: with the gradient all-reduce), but its regulariser terms depend on the
            # parameters only: they are summed on the loss stream now, off the step's main stream (small batches: the step
            # is launch-bound and the extra stream costs more host time than it hides)
            if self.Bmax >= _loss_stream_min_batch():
                reg = self._reg_loss_on_loss_stream()
                self.backward_and_update(B)
                self._main_stream().wait_event(self._loss_dev_done)
            else:
                reg = self._reg_loss()
                self.backward_and_update(B)
            out = self._publish_loss(reg)
        self._end_step()
        self._mark("update")
        return out

    @property
    def loss_is_early(self) -> bool:
        """sharded engines: the global loss comes back with the gradient all-reduce, after the backward; the gathered-l2
        models (quirk Q6) add a loss term that the backward computes"""
        return not (self.spec.gathered_l2 and self.spec.l2_reg > 0) and self.Bmax >= _loss_stream_min_batch()

    def _loss_stream_init(self) -> None:
        if self._loss_host is None:
            self._loss_host = torch.zeros(2, dtype=torch.float32).pin_memory()
            self._loss_stream = torch.cuda.Stream(device=self.dev)
            self._loss_events = [torch.cuda.Event(), torch.cuda.Event()]
            self._loss_made, self._loss_dev_done = torch.cuda.Event(), torch.cuda.Event()
            self._ws_reg = torch.zeros(1 << 16, dtype=torch.uint8, device=self.dev)

    def _reg_loss_on_loss_stream(self):
        """regulariser sums on the loss stream, behind everything queued on the main stream so far; the dense update of this
        step waits for them (_before_dense_update), and so must whoever reads the result on another stream"""
        self._loss_stream_init()
        main, ls = self._main_stream(), self._loss_stream
        self._loss_made.record(main)
        ls.wait_event(self._loss_made)
        with torch.cuda.stream(ls):
            reg = self._reg_loss(ls.cuda_stream, self._ws_reg)
        self._loss_dev_done.record(ls)
        self._loss_dev_pending = True
        if reg is not None:
            reg.record_stream(main)
        return reg

    def _publish_loss(self, reg, early: bool = False) -> torch.Tensor:
        main = self._main_stream()
        if not early:
            # the value exists only after the backward: nothing to hide, so no extra stream work (a launch-bound small-batch
            # step would pay ~30 us of host time for it); last_loss() reads the tensor
            loss = self._global_loss()
            if reg is not None:
                loss = loss + reg
            self._loss_tensor = loss
            return loss
        self._loss_tensor = None
        self._loss_stream_init()
        ls = self._loss_stream
        self._loss_slot ^= 1
        # everything on the loss stream, behind the head kernel.  The regulariser reads the dense parameters: the dense
        # update of this step waits for _loss_dev_done (backward_and_update)
        self._loss_made.record(main)
        ls.wait_event(self._loss_made)
        with torch.cuda.stream(ls):
            reg = self._reg_loss(ls.cuda_stream, self._ws_reg)
            loss = self._global_loss()
            if reg is not None:
                loss = loss + reg
            self._loss_host[self._loss_slot:self._loss_slot + 1].copy_(loss, non_blocking=True)
        self._loss_dev_done.record(ls)
        self._loss_dev_pending = True
        loss.record_stream(main)
        self._loss_events[self._loss_slot].record(ls)
        return loss

    _loss_host = None
    _loss_slot = 0
    _loss_dev_pending = False
    _loss_tensor = None        # late-loss steps: the returned tensor itself (last_loss() reads it)

    def _before_dense_update(self) -> None:
        """the loss stream's regulariser sums read the dense parameters the dense update rewrites"""
        if self._loss_dev_pending:
            self._main_stream().wait_event(self._loss_dev_done)
            self._loss_dev_pending = False

    def last_loss(self) -> float:
        """The loss of the latest train_step as a host float (what `sess.run([loss, optimizer])` hands back,
        models/deepfm.py:181).  Waits for that value's device-to-host copy only -- on one GPU the loss is final after the
        forward pass -- and not for the rest of the step, unlike `.item()` on the returned tensor."""
        if self._loss_tensor is not None:
            return float(self._loss_tensor.item())
        if self._loss_host is None:
            raise _lib.CtrError("last_loss: no train_step has run")
        self._loss_events[self._loss_slot].synchronize()
        return float(self._loss_host[self._loss_slot])

    def _prefetch_next(self, next_batch: Optional[Dict], ev_call) -> None:
        """work for the next step that runs under this one: upload the next batch into the other input set"""
        if next_batch is None:
            return
        other = 1 - self._cur_in
        cs = self.copy_stream
        if self._in_used[other]:
            cs.wait_event(self._in_free[other])        # the step that last read that set has finished
        cs.wait_event(ev_call)                         # a device-resident batch may still be in production upstream
        with torch.cuda.stream(cs):
            self._pf_B = self._copy_inputs(next_batch, self._in[other], True, self._ids_external)
        self._pf_event.record(cs)
        self._pf_id = id(next_batch)
        self._pf_obj = next_batch          # keeps the object alive, so its id cannot be recycled while the upload is pending

    def _end_step(self) -> None:
        self._in_free[self._cur_in].record(self._main_stream())
        self._in_used[self._cur_in] = True
        self._snapshot_status()            # read by the next public call (or check_status) without a host wait

    def _global_loss(self) -> torch.Tensor:
        """log-loss of the global batch (sharded engines read it back from the gradient allreduce)"""
        return self.loss.clone()

    def predict(self, batch: Optional[Dict] = None) -> torch.Tensor:
        """`sess.run("score:0")` (models/deepfm.py:250): scores [B] on the device."""
        B = self.load_batch(batch) if batch is not None else self.last_batch
        self.forward(B, with_loss=False)
        return self.score[:B]

    _pred_graphs = None

    def predict_graphed(self, batch: Dict) -> torch.Tensor:
        """`predict` with the forward pass replayed from a CUDA graph (one graph per batch size, captured on first use):
        for small serving batches the ~10 launches of a predict are host time, not GPU time (config 5 at B = 1K: 95 us per
        call launched one by one).  Same kernels, same results; single-GPU engines (the sharded forward plans ids on the
        host side of the step).  The batch is copied into input set 0 outside the graph."""
        if self.world > 1:
            raise _lib.CtrError("predict_graphed is single-GPU")
        self._use_inputs(0)
        self._pf_id = self._pf_obj = None
        B = self.load_batch(batch)
        if self._pred_graphs is None:
            self._pred_graphs = {}
        g = self._pred_graphs.get(B)
        if g is None:
            self.forward(B, with_loss=False)               # warm-up: one-time attribute settings happen outside the capture
            torch.cuda.current_stream(self.dev).synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self.forward(B, with_loss=False)
            self._pred_graphs[B] = g
        g.replay()
        return self.score[:B]

    def auc(self, scores: torch.Tensor, labels: torch.Tensor) -> float:
        """roc_auc_score(labels, scores) on the device (models/deepfm.py:231)."""
        n = int(scores.numel())
        sc = scores.reshape(-1).to(self.dev, torch.float32).contiguous()
        lb = labels.reshape(-1).to(self.dev, torch.float32).contiguous()
        wsn = self.h.query("ctr_auc_ws_bytes", n)
        ws = torch.empty(wsn, dtype=torch.uint8, device=self.dev)
        out = torch.zeros(1, dtype=torch.float64, device=self.dev)
        self.h.call("ctr_auc", self._stream(), ptr(sc), ptr(lb), n, ptr(out), ptr(ws), wsn)
        return float(out.item())
