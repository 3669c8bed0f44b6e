"""SplitFmEngine: models/deepfm_pipeline.py on one GPU.

That model looks one categorical feature up in TWO rows of the same table: the FM part at `id + C`
(models/deepfm_pipeline.py:89), the deep part at the raw `id` (:120); the C continuous fields own rows 0..C-1 and row 0 is
zeroed (:83-86), so continuous field 0 has no weight.  No kernel is new -- the step is composed from the existing
entry points:

  forward   K1 twice: pass A (use_fm, id_offset = C) -> fm / sum_e; pass B (no FM, raw ids) -> the tower input X0
  backward  two sorted lookup lists (A: id + C plus the continuous rows 1..C-1; B: raw ids, id 0 dropped), two partials
            passes (A against an all-zero dX0: FM gradient only; B: dX0 only), one gradient row per unique key of each
            list (ctr_embed_bwd_finalize), then the owner-side machinery of the sharded engine on ONE GPU: the two
            sorted key lists are merged (ctr_merge_sorted_segments, "world" = 2) and ctr_adam_rows_gathered sums the
            contributions of a row hit through both lists BEFORE TF1 Adam -- one Adam step per row, as TensorFlow applies
            the summed IndexedSlices of one variable.

Parity: tests/test_gpu_parity.py::test_deepfm_pipeline_split_fm (K1 fast path and general kernel, TF-exact and lazy Adam,
padding ids and ids that alias the continuous-field rows), 1e-5 against the CPU restatement; golden vectors in
tests/golden/deepfm_pipeline_tiny.npz.
"""
from __future__ import annotations

import ctypes as C

import torch

from ._lib import EmbedCfg, ptr
from .engine import ADAM_BETA1, ADAM_BETA2, CtrEngine, _ru
from .spec import ModelSpec


class SplitFmEngine(CtrEngine):
    def __init__(self, spec: ModelSpec, max_batch: int, **kw):
        if not spec.split_fm:
            raise ValueError("SplitFmEngine is for split_fm specs (deepfm_pipeline)")
        if kw.get("world", 1) != 1:
            raise NotImplementedError("deepfm_pipeline: single GPU only")
        super().__init__(spec, max_batch, **kw)
        s, B, dev = spec, self.Bmax, self.dev
        f32 = dict(dtype=torch.float32, device=dev)
        u32 = dict(dtype=torch.int32, device=dev)
        self.cont_buf = torch.zeros(max(s.C, 1), s.table_stride, **f32)
        self.x0_a = torch.zeros(B, s.x0_ld, **f32)                      # pass A's tower tile: written, never read
        self.fm_b, self.sum_e_b = torch.zeros(B, s.fm_ld, **f32), torch.zeros(B, s.k, **f32)
        if self.train_mode:
            self.nA_cap, self.nB_cap = B * s.Fc + s.C, B * s.Fc
            nT = self.nA_cap + self.nB_cap
            self.dx0_zero = torch.zeros(B, self.dx0_ld, **f32)          # never written
            self.keys_b, self.perm_b, self.uidx_b = (torch.zeros(max(self.nB_cap, 1), **u32) for _ in range(3))
            self.seg_b = torch.zeros(self.nB_cap + 2, **u32)
            self.nuniq_b = torch.zeros(1, **u32)
            self.part_b = torch.zeros((self.nB_cap + self.nB_cap // 32 + 2) * s.grad_stride, **f32)
            self.ukeys = torch.zeros(nT, **u32)                         # [unique keys of A | of B], 0xFFFFFFFF behind each
            self.gcat = torch.zeros(nT, s.grad_stride, **f32)           # one gradient row per unique key, same positions
            self.m_keys, self.m_idx, self.o_uniq, self.o_uidx = (torch.zeros(nT, **u32) for _ in range(4))
            self.o_seg = torch.zeros(nT + 2, **u32)
            self.o_nuniq = torch.zeros(1, **u32)
            wsn = _ru(max(self.h.query("ctr_sort_ws_bytes", nT), 4 * (nT // 2048 + 8)), 256)
            self.ws_b = torch.zeros(wsn, dtype=torch.uint8, device=dev)     # list B's sort / scans (side stream)
            self.ws_m = torch.zeros(wsn, dtype=torch.uint8, device=dev)     # merged list's scan (main stream)
            self.ends = (C.c_uint32 * 2)(self.nA_cap, nT)

    # ---- the two configurations -------------------------------------------------------------------------------------
    def _cfg_b(self, B: int) -> EmbedCfg:
        cfg = self._cfg(B)
        cfg.use_fm, cfg.fm_cont = 0, 0
        return cfg

    def _cont_rows(self):
        return self.cont_buf

    # ---- forward ------------------------------------------------------------------------------------------------------
    def _embed_forward(self, B: int) -> None:
        s, st, h = self.spec, self._stream(), self.h
        if s.C:
            # rows 0..C-1 as the FM sees them: row 0 reads as zero (models/deepfm_pipeline.py:83-86)
            self.cont_buf[:s.C].copy_(self.table[:s.C])
            self.cont_buf[0].zero_()
        # pass A: FM over [cont rows | rows id + C]
        h.call("ctr_fm_gather_fwd", st, self._cfg(B), ptr(self.table), ptr(self.ids), s.C, ptr(self.cont), ptr(self.vec),
               ptr(self.cont_buf), None, ptr(self.x0_a), ptr(self.fm), ptr(self.sum_e))
        # pass B: tower input [emb of the RAW ids | cont | vector]
        h.call("ctr_fm_gather_fwd", st, self._cfg_b(B), ptr(self.table), ptr(self.ids), 0, ptr(self.cont), ptr(self.vec),
               None, None, ptr(self.x0), ptr(self.fm_b), ptr(self.sum_e_b))

    def _launch_sort(self, B: int) -> None:
        s, h = self.spec, self.h
        main = self._main_stream()
        self.ev_batch.record(main)
        self.side.wait_event(self.ev_batch)
        st = self.side.cuda_stream
        key_bits = max(1, int(s.table_rows - 1).bit_length())
        nA, nB = B * s.Fc + max(s.C - 1, 0), B * s.Fc
        with torch.cuda.stream(self.side):
            self.ukeys.fill_(-1)                                         # 0xFFFFFFFF: sorts behind every real key
        ws, wsn = ptr(self.ws_sort), self.ws_sort.numel()
        # list A: FM lookups at id + C, then the continuous rows 1..C-1 (row 0 is cut out of the graph)
        h.call("ctr_sort_ids", st, ptr(self.ids), self.id_bytes, B, s.ids_ld, 0, s.Fc, s.C, 1, 1, max(s.C - 1, 0), 1, 0, key_bits,
               int(s.table_rows), ptr(self.status), ptr(self.keys), ptr(self.perm), ws, wsn)
        h.call("ctr_unique_segments", st, ptr(self.keys), nA, ptr(self.ukeys), ptr(self.seg_start), ptr(self.uidx),
               ptr(self.n_unique), ws, wsn, None)
        # list B: deep lookups at the raw id, id 0 dropped
        wsb, wsbn = ptr(self.ws_b), self.ws_b.numel()
        h.call("ctr_sort_ids", st, ptr(self.ids), self.id_bytes, B, s.ids_ld, 0, s.Fc, 0, 1, 0, 0, 1, 0, key_bits,
               int(s.table_rows), ptr(self.status), ptr(self.keys_b), ptr(self.perm_b), wsb, wsbn)
        h.call("ctr_unique_segments", st, ptr(self.keys_b), nB, ptr(self.ukeys[self.nA_cap:]), ptr(self.seg_b), ptr(self.uidx_b),
               ptr(self.nuniq_b), wsb, wsbn, None)
        self.ev_sorted.record(self.side)

    # ---- backward + update (dense part as CtrEngine.backward_and_update) ---------------------------------------------------
    def backward_and_update(self, B: int) -> None:
        s, st, h = self.spec, self._stream(), self.h
        cfg_a, cfg_b = self._cfg(B), self._cfg_b(B)
        L, dims = len(s.hidden), self._layer_dims()
        w_fm, _, _ = self._head_parts()
        ws, wsn = ptr(self.ws), self.ws.numel()
        main = self._main_stream()
        if s.C:
            self.ev_head.record(main)
            self.aux.wait_event(self.ev_head)
            h.call("ctr_fm_cont_bwd", self.aux.cuda_stream, cfg_a, ptr(self.cont), ptr(self.cont_buf), ptr(self.sum_e),
                   ptr(self.dlogit), ptr(w_fm), ptr(self.cont_grad), ptr(self.ws_cont), self.ws_cont.numel())
            self.ev_cont.record(self.aux)
        cur = 0
        for i in range(L - 1, -1, -1):
            k_, h_ = dims[i]
            x, ldx = (self.acts[i - 1], self.act_ld[i - 1]) if i > 0 else (self.x0, s.x0_ld)
            h.call("ctr_mlp_bwd_weight", st, self.prec, B, k_, h_, ptr(x), ldx, ptr(self.dz[cur]), h_,
                   ptr(self._w(i, self.grad)), ptr(self._b(i, self.grad)), ws, wsn, 1)
            if i > 0:
                h.call("ctr_mlp_bwd_data", st, self.prec, B, k_, h_, ptr(self.dz[cur]), h_, ptr(self._wd(i)), ptr(x), ldx, 1.0,
                       ptr(self.dz[cur ^ 1]), k_)
                cur ^= 1
            elif s.E:
                h.call("ctr_mlp_bwd_data", st, self.prec, B, s.E, h_, ptr(self.dz[cur]), h_, ptr(self._wd(0)), None, 0, 1.0,
                       ptr(self.dx0), self.dx0_ld)
        hp = self._adam_hp()
        # ---- sparse side ---------------------------------------------------------------------------------------------------
        gs, nT = s.grad_stride, self.nA_cap + self.nB_cap
        nA, nB = B * s.Fc + max(s.C - 1, 0), B * s.Fc
        main.wait_event(self.ev_sorted)
        if s.C:
            main.wait_event(self.ev_cont)
        cont_grad_1 = self.cont_grad[1:] if s.C > 1 else None            # extra entry c' is continuous row c' + 1
        # A: FM gradient only (dX0 = 0): g_v = dlogit * w2 * sum_e, g_w = dlogit * w1[field], D += dlogit
        h.call("ctr_embed_bwd_partials", st, cfg_a, ptr(self.keys), ptr(self.perm), ptr(self.uidx), nA, None, ptr(self.x0_a),
               ptr(self.dx0_zero), self.dx0_ld, ptr(self.sum_e), ptr(self.dlogit), ptr(w_fm), None, None, ptr(cont_grad_1),
               ptr(self.part))
        # B: tower gradient only
        h.call("ctr_embed_bwd_partials", st, cfg_b, ptr(self.keys_b), ptr(self.perm_b), ptr(self.uidx_b), nB, None, None,
               ptr(self.dx0), self.dx0_ld, None, None, None, None, None, None, ptr(self.part_b))
        h.call("ctr_embed_bwd_finalize", st, ptr(self.seg_start), ptr(self.n_unique), nA, s.grad_width, gs, ptr(self.part),
               ptr(self.gcat))
        h.call("ctr_embed_bwd_finalize", st, ptr(self.seg_b), ptr(self.nuniq_b), nB, s.grad_width, gs, ptr(self.part_b),
               ptr(self.gcat[self.nA_cap:]))
        # rows reached through both lists: merge the two sorted key lists, sum per row, ONE TF1 Adam step per row
        h.call("ctr_merge_sorted_segments", st, ptr(self.ukeys), nT, self.ends, 2, ptr(self.m_keys), ptr(self.m_idx), None)
        h.call("ctr_unique_segments", st, ptr(self.m_keys), nT, ptr(self.o_uniq), ptr(self.o_seg), ptr(self.o_uidx),
               ptr(self.o_nuniq), ptr(self.ws_m), self.ws_m.numel(), None)
        h.call("ctr_adam_rows_gathered", st, ptr(self.o_uniq), ptr(self.o_nuniq), nT, 0, ptr(self.gcat), ptr(self.m_idx),
               ptr(self.o_seg), gs, s.row_width, ptr(self.table), ptr(self.m_table), ptr(self.v_table), s.table_stride,
               self.rows_local, ptr(self.touched), self._fm_w2(), s.k, hp)
        if self.adam_mode == "tf_dense":
            h.call("ctr_adam_dense", st, ptr(self.table), ptr(self.m_table), ptr(self.v_table), None,
                   self.rows_local * s.table_stride, 0.0, 0.0, ptr(self.touched), s.table_stride, hp)
        self._before_dense_update()
        h.call("ctr_dense_update", st, ptr(self.theta), ptr(self.m), ptr(self.v), ptr(self.grad), C.byref(self._dense_layout()), hp)
        self.beta1_power *= ADAM_BETA1
        self.beta2_power *= ADAM_BETA2
        self.global_step += 1
        self._steps_done += 1
