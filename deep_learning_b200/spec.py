"""Canonical description of one reference model and the mapping between the reference's
variables (names/shapes of models/<alg>.py) and the device layouts of include/ctr_b200.h."""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import numpy as np


def _round_up(a: int, b: int) -> int:
    return (a + b - 1) // b * b


@dataclass
class ModelSpec:
    alg: str
    C: int                      # continuous fields fed to the tower
    V: int                      # vector_feats width
    Fc: int                     # single-valued categorical fields
    k: int                      # embedding_size
    hidden: List[int]
    table_rows: int
    emb_name: str               # reference variable holding the [R,k] table
    first_name: Optional[str]   # reference variable holding the [R,1] first-order table
    use_fm: bool = False
    fm_cont: bool = False
    zero_row0: bool = False
    cont_base: int = 0          # table row of continuous field 0 (fm_cont models)
    multi_ranges: List[Tuple[int, int]] = field(default_factory=list)   # slot ranges inside the multi block
    head: str = "dnn"           # 'deepfm' | 'dnn' | 'wdl'
    head_w_name: str = "deep_res"
    head_b_name: str = "deep_res_bias"
    first_order_ref: str = "cate_cont"   # order of the first-order block in the reference head weight
    l2_reg: float = 0.0
    reg: Dict[str, str] = field(default_factory=dict)   # reference variable -> 'l2' | 'l1'
    wide_fields: int = 0        # W (wdl)
    wide_rows: int = 0          # rows of wdl_weights minus H (wdl)
    id_dtype: str = "int32"
    gathered_l2: bool = False   # wdl_textline quirk Q6: the l2 term runs over the GATHERED wdl_weights [B, W+H]
    weighted_pool: bool = False  # multi-valued slots carry a weight (mul_cat_feats_value, models/dnn_multi_textline.py:96-101)
    frozen_name: Optional[str] = None     # reference constant appended to the table as never-updated rows ('tags_emb')
    frozen_rows: int = 0
    frozen_fields: Tuple[int, ...] = ()   # multi-valued fields that look up the frozen rows (feat_name == 'tag')
    optimizer: str = "adam"     # 'adam': tf.train.AdamOptimizer (every models/*.py) | 'ftrl': FtrlOptimizer (test_model/wdl.py:136)
    loss: str = "log_loss"      # 'log_loss' (models/deepfm.py:151) | 'sce_on_score' (test_model/wdl.py:119-121)
    deep_row_offset: Optional[int] = None   # wdl head: first row of the deep-projection weights inside wdl_weights.  None:
                                            # wide_fields, the aliasing of models/wdl.py:224-228 (quirk Q4);
                                            # test_model/wdl.py:110 puts them behind the wide ids (cate_feats_size)
    split_fm: bool = False      # quirk Q2 (models/deepfm_pipeline.py:89,120): the FM part reads row id + C, the deep part row id

    @property
    def Fm(self) -> int:
        return len(self.multi_ranges)

    @property
    def M(self) -> int:
        return self.multi_ranges[-1][1] if self.multi_ranges else 0

    @property
    def has_first(self) -> bool:
        return self.first_name is not None

    @property
    def ids_ld(self) -> int:
        return self.Fc + self.M

    @property
    def E(self) -> int:            # embedding columns of X0
        return (self.Fc + self.Fm) * self.k

    @property
    def D0(self) -> int:
        return self.E + self.C + self.V

    @property
    def x0_ld(self) -> int:
        return _round_up(self.D0 + 1, 4)       # >= one pad column: it holds the 1.0 of the fused bias gradient

    @property
    def n_fm(self) -> int:
        return (self.Fc + self.Fm + self.C + self.k) if self.use_fm else 0

    @property
    def fm_ld(self) -> int:
        return _round_up(max(self.n_fm, 1), 4)

    @property
    def table_stride(self) -> int:
        return _round_up(self.k + (1 if self.has_first else 0), 8)

    @property
    def grad_stride(self) -> int:
        return _round_up(self.k + (1 if self.has_first else 0), 4)

    @property
    def row_width(self) -> int:
        return self.k + (1 if self.has_first else 0)

    @property
    def grad_width(self) -> int:
        """columns of a gradient row that carry data: [0,k) embedding, [k] first-order, [k+1] D = sum of dlogit over the
        row's single-valued FM lookups (include/ctr_b200.h, ctr_embed_bwd_partials)"""
        return self.k + 2 if self.use_fm else self.row_width

    # ---- reference <-> canonical permutations -------------------------------------------------
    def deep0_perm(self) -> np.ndarray:
        """canonical X0 column i  ->  row of the reference deep_0 ([cont | vector | emb] order,
        models/deepfm.py:100, models/deepfm_multi.py:188-190, models/wdl.py:179, models/dnn.py:57)."""
        E, CV = self.E, self.C + self.V
        return np.concatenate([np.arange(E) + CV, np.arange(CV)]).astype(np.int64)

    def head_fm_perm(self) -> np.ndarray:
        """canonical fm column i -> row of the reference deep_fm_weight.
        canonical: [first single Fc | first multi Fm | first cont C | second k]
        deepfm.py:   [cate Fc | cont C | second k | deep H]                  (models/deepfm.py:70,130)
        deepfm_multi:[cont C | single Fc | multi Fm | second k | deep H]     (models/deepfm_multi.py:151,176,226)"""
        Fc, Fm, C, k = self.Fc, self.Fm, self.C, self.k
        if self.first_order_ref == "cate_cont":
            assert Fm == 0
            return np.arange(Fc + C + k, dtype=np.int64)
        single = C + np.arange(Fc)
        multi = C + Fc + np.arange(Fm)
        cont = np.arange(C)
        second = C + Fc + Fm + np.arange(k)
        return np.concatenate([single, multi, cont, second]).astype(np.int64)

    def frozen_slot_offset(self) -> np.ndarray:
        """per column of the id matrix: what to add to a raw id so it addresses the frozen rows appended behind the
        trainable table (slots of the `frozen_fields`), 0 elsewhere"""
        off = np.zeros(max(self.ids_ld, 1), dtype=np.int64)
        for j in self.frozen_fields:
            s0, s1 = self.multi_ranges[j]
            off[self.Fc + s0:self.Fc + s1] = self.table_rows
        return off

    def multi_start(self) -> np.ndarray:
        if not self.multi_ranges:
            return np.zeros(1, dtype=np.int32)
        return np.asarray([r[0] for r in self.multi_ranges] + [self.multi_ranges[-1][1]], dtype=np.int32)

    def field_of_slot(self) -> np.ndarray:
        out = np.zeros(max(self.M, 1), dtype=np.int32)
        for j, (s, e) in enumerate(self.multi_ranges):
            out[s:e] = j
        return out

    def dense_param_shapes(self) -> List[Tuple[str, Tuple[int, ...]]]:
        """trainable dense variables, reference names and shapes, in a fixed order"""
        out = []
        prev = self.D0
        for i, h in enumerate(self.hidden):
            out.append(("deep_%d" % i, (prev, h)))
            out.append(("deep_bias_%d" % i, (1, h)))
            prev = h
        if self.head == "deepfm":
            out.append((self.head_w_name, (self.n_fm + self.hidden[-1], 1)))
            out.append((self.head_b_name, (1,)))
        elif self.head == "dnn":
            out.append((self.head_w_name, (self.hidden[-1], 1)))
            out.append((self.head_b_name, (1, 1)))
        elif self.head == "wdl":
            out.append(("wdl_weights", (self.wide_rows + self.hidden[-1], 1)))
            out.append(("wdl_bias", (1,)))
        return out


def _ranges(hp) -> List[Tuple[int, int]]:
    return [(int(r[0]), int(r[1])) for r in getattr(hp, "multi_feats_range", [])]


def spec_from_hparams(alg: str, hp) -> ModelSpec:
    """Build the spec from the reference's hyper-parameter bag (attribute names of
    models/deepfm.py:17-33, models/deepfm_multi.py:17-39, models/wdl.py:19-39, models/dnn.py:16-30)."""
    H = [int(x) for x in hp.hidden_units]
    l2 = float(hp.l2_reg)
    k = int(hp.embedding_size)
    if alg == "deepfm":
        C, Fc = int(hp.cont_field_size), int(hp.cate_field_size)
        return ModelSpec(alg, C=C, V=int(hp.vector_feats_size), Fc=Fc, k=k, hidden=H,
                         table_rows=C + int(hp.cate_feats_size), emb_name="feats_emb", first_name="feats",
                         use_fm=True, fm_cont=True, zero_row0=False,
                         cont_base=Fc,    # quirk Q1: offset is cate_FIELD_size (models/deepfm.py:66)
                         head="deepfm", head_w_name="deep_fm_weight", head_b_name="deep_fm_bias",
                         first_order_ref="cate_cont", l2_reg=l2, reg={"deep_fm_weight": "l2"})
    if alg == "deepfm_multi":
        C, Fc = int(hp.cont_field_size), int(hp.cate_field_size)
        return ModelSpec(alg, C=C, V=int(hp.vector_feats_size), Fc=Fc, k=k, hidden=H,
                         table_rows=int(hp.cate_feats_size) + C, emb_name="feats_emb",
                         first_name="fm_first_order_emb", use_fm=True, fm_cont=True, zero_row0=True,
                         cont_base=int(hp.cate_feats_size),   # models/deepfm_multi.py:139
                         multi_ranges=_ranges(hp), head="deepfm", head_w_name="deep_fm_weight",
                         head_b_name="deep_fm_bias", first_order_ref="cont_single_multi", l2_reg=l2,
                         reg={"deep_fm_weight": "l2"}, id_dtype="int64")
    if alg == "dnn":
        # models/dnn.py:27: cont_field_size already includes the vector block
        C = int(hp.cont_field_size) + int(hp.vector_feats_size)
        return ModelSpec(alg, C=C, V=0, Fc=int(hp.cate_field_size), k=k, hidden=H,
                         table_rows=int(hp.cate_feats_size), emb_name="weight_mat", first_name=None,
                         head="dnn", l2_reg=l2, reg={"deep_%d" % i: "l1" for i in range(len(H))})   # quirk Q5
    if alg == "dnn_multi":
        return ModelSpec(alg, C=int(hp.cont_field_size), V=int(hp.vector_feats_size), Fc=int(hp.cate_field_size),
                         k=k, hidden=H, table_rows=int(hp.cate_feats_size), emb_name="feats_emb", first_name=None,
                         zero_row0=True, multi_ranges=_ranges(hp), head="dnn", l2_reg=l2,
                         reg={"deep_res": "l2"}, id_dtype="int64")
    # ---- the variants local_run_test.py dispatches: the same kernels, other field mixes
    if alg in ("deepfm_cate", "deepfm_multi_cate"):
        # models/deepfm_cate.py:72-171, models/deepfm_multi_cate.py:113-223: no continuous block, row-0-zero tables
        return ModelSpec(alg, C=0, V=int(hp.vector_feats_size), Fc=int(hp.cate_field_size), k=k, hidden=H,
                         table_rows=int(hp.cate_feats_size), emb_name="feats_emb", first_name="fm_first_order_emb",
                         use_fm=True, fm_cont=False, zero_row0=True, cont_base=int(hp.cate_feats_size),
                         multi_ranges=_ranges(hp) if alg == "deepfm_multi_cate" else [], head="deepfm",
                         head_w_name="deep_fm_weight", head_b_name="deep_fm_bias", first_order_ref="cont_single_multi",
                         l2_reg=l2, reg={"deep_fm_weight": "l2"}, id_dtype="int64")
    if alg in ("dnn_pipeline", "dnn_cate", "dnn_multi_cate"):
        # models/dnn_pipeline.py:68-122, models/dnn_cate.py:62-115, models/dnn_multi_cate.py:64-147
        C = int(hp.cont_field_size) if alg == "dnn_pipeline" else 0
        return ModelSpec(alg, C=C, V=int(hp.vector_feats_size), Fc=int(hp.cate_field_size), k=k, hidden=H,
                         table_rows=int(hp.cate_feats_size), emb_name="feats_emb", first_name=None, zero_row0=True,
                         multi_ranges=_ranges(hp) if alg == "dnn_multi_cate" else [], head="dnn", l2_reg=l2,
                         reg={"deep_res": "l2"}, id_dtype="int64")
    if alg == "deepfm_pipeline":
        # models/deepfm_pipeline.py:76-169: table rows [C continuous | cate_feats_size categorical], row 0 zeroed, FM ids
        # shifted by C while the deep part uses the raw ids (quirk Q2); head order [cont | cate | second | deep]
        C, Fc = int(hp.cont_field_size), int(hp.cate_field_size)
        return ModelSpec(alg, C=C, V=int(hp.vector_feats_size), Fc=Fc, k=k, hidden=H, table_rows=C + int(hp.cate_feats_size),
                         emb_name="feats_emb", first_name="fm_first_order_emb", use_fm=True, fm_cont=True, zero_row0=True,
                         cont_base=0, head="deepfm", head_w_name="deep_fm_weight", head_b_name="deep_fm_bias",
                         first_order_ref="cont_single_multi", l2_reg=l2, reg={"deep_fm_weight": "l2"}, id_dtype="int64",
                         split_fm=True)
    if alg == "dnn_multi_textline":
        # models/dnn_multi_textline.py:16-48,69-161: its own attribute spellings (cat_field_size, cat_index_size) are
        # accepted next to the canonical ones; row-0-zero `weight_mat`, frozen word2vec constant for the 'tag' field,
        # weighted pooling, L1 on the tower weights
        g = lambda a, b: int(getattr(hp, a, 0) or getattr(hp, b, 0))
        ranges = getattr(hp, "multi_feats_range", None) or getattr(hp, "mul_cat_index_range", [])
        frozen = tuple(j for j, r in enumerate(ranges) if len(r) > 2 and r[2] == "tag")
        w2v = int(getattr(hp, "word2vec_rows", 0))
        if frozen and w2v <= 0:
            raise ValueError("dnn_multi_textline: a 'tag' field needs word2vec_rows (rows of the frozen tags_emb table)")
        return ModelSpec(alg, C=int(hp.cont_field_size), V=0, Fc=g("cate_field_size", "cat_field_size"), k=k, hidden=H,
                         table_rows=g("cate_feats_size", "cat_index_size"), emb_name="weight_mat", first_name=None,
                         zero_row0=True, multi_ranges=[(int(r[0]), int(r[1])) for r in ranges], head="dnn", l2_reg=l2,
                         reg={"deep_%d" % i: "l1" for i in range(len(H))}, weighted_pool=True,
                         frozen_name="tags_emb" if frozen else None, frozen_rows=w2v if frozen else 0,
                         frozen_fields=frozen)
    if alg == "wdl_textline":
        # models/wdl_textline.py:43-50 (row-0-zero table), :278-294 (l2 on the gathered wdl_weights, quirk Q6)
        reg = {"deep_%d" % i: "l2" for i in range(len(H))}
        return ModelSpec(alg, C=int(hp.cont_field_size), V=0, Fc=int(hp.cate_field_size), k=k, hidden=H,
                         table_rows=int(hp.cate_index_size), emb_name="weight_mat", first_name=None, zero_row0=True,
                         head="wdl", head_w_name="wdl_weights", head_b_name="wdl_bias", l2_reg=l2, reg=reg,
                         wide_fields=int(hp.wide_field_size), wide_rows=int(hp.cate_index_size), gathered_l2=True)
    if alg == "wdl_ftrl":
        # test_model/wdl.py: ONE size for the table, the wide ids and the offset of the deep-projection
        # rows (:61-64, :101-111); FTRL (:136); cross-entropy on the sigmoided value (:119-121); l2 on wdl_weights + deep_i
        reg = {"deep_%d" % i: "l2" for i in range(len(H))}
        reg["wdl_weights"] = "l2"
        R = int(hp.cate_feats_size)
        return ModelSpec(alg, C=int(hp.cont_field_size), V=0, Fc=int(hp.cate_field_size), k=k, hidden=H,
                         table_rows=R, emb_name="weight_mat", first_name=None,
                         head="wdl", head_w_name="wdl_weights", head_b_name="wdl_bias", l2_reg=l2, reg=reg,
                         wide_fields=int(hp.wide_field_size), wide_rows=R, deep_row_offset=R,
                         optimizer="ftrl", loss="sce_on_score")
    if alg == "wdl":
        reg = {"deep_%d" % i: "l2" for i in range(len(H))}
        reg["wdl_weights"] = "l2"
        return ModelSpec(alg, C=int(hp.cont_field_size), V=0, Fc=int(hp.cate_field_size), k=k, hidden=H,
                         table_rows=int(hp.cate_index_size), emb_name="weight_mat", first_name=None,
                         head="wdl", head_w_name="wdl_weights", head_b_name="wdl_bias", l2_reg=l2, reg=reg,
                         wide_fields=int(hp.wide_field_size), wide_rows=int(hp.cate_index_size))
    raise ValueError("unknown alg_name %r" % alg)
