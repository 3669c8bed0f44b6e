"""Shared test helpers: model configurations (BASELINE.json configs at oracle-friendly sizes),
seeded batches and the oracle<->engine glue."""
from __future__ import annotations

import numpy as np
import torch

from oracle import ctr_oracle as O
from deep_learning_b200 import data as D


def hp_c1(**kw):
    """configs[0]: models/deepfm.py, 10 cont + 20 single-valued cate, k=8, MLP 256-128, batch 256"""
    hp = O.HParams(hidden_units=[256, 128], batch_size=256, embedding_size=8, cont_field_size=10,
                   cate_field_size=20, cate_feats_size=10000, l2_reg=1e-5, learning_rate=1e-3)
    for k, v in kw.items():
        setattr(hp, k, v)
    return hp


def hp_multi(B=512, **kw):
    """configs[1] at reduced batch: models/deepfm_multi.py, 13 cont + 20 single + 6 multi (top_n 16), k=16"""
    ranges, s = [], 0
    for j in range(6):
        ranges.append([s, s + 16, "m%d" % j]); s += 16
    hp = O.HParams(hidden_units=[512, 256, 128], batch_size=B, embedding_size=16, cont_field_size=13,
                   cate_field_size=20, cate_feats_size=20001, multi_feats_size=96, multi_field_size=6,
                   multi_feats_range=ranges, l2_reg=1e-5)
    for k, v in kw.items():
        setattr(hp, k, v)
    return hp


def hp_variant(alg, **kw):
    """the variants local_run_test.py dispatches, at oracle-friendly sizes.  `*_cate` models have no
    continuous block; a small vector block exercises the [vector | emb] deep input."""
    if alg in ("deepfm_multi_cate", "dnn_multi_cate"):
        return hp_multi(B=256, cont_field_size=0, vector_feats_size=5, **kw)
    cont = 7 if alg == "dnn_pipeline" else 0
    return hp_c1(cont_field_size=cont, vector_feats_size=5, embedding_size=16, **kw)


def hp_textline(alg, **kw):
    """models/wdl_textline.py and models/dnn_multi_textline.py at oracle-friendly sizes.  The multi
    block of dnn_multi_textline has a trainable field and a 'tag' field that reads the frozen word2vec rows."""
    if alg == "wdl_textline":
        return hp_c1(cate_index_size=10000, wide_field_size=20, embedding_size=16, **kw)
    hp = hp_c1(embedding_size=16, cate_feats_size=12001, multi_feats_size=12, multi_field_size=2,
               multi_feats_range=[[0, 5, "kw"], [5, 12, "tag"]], word2vec_rows=300, **kw)
    return hp


def hp_criteo(B=1024, rows=50000, **kw):
    """configs[2..4] shape at reduced batch/rows: 13 cont + 26 cate, k=16, MLP 400-400-400"""
    hp = O.HParams(hidden_units=[400, 400, 400], batch_size=B, embedding_size=16, cont_field_size=13,
                   cate_field_size=26, cate_feats_size=rows, cate_index_size=rows, wide_field_size=26, l2_reg=1e-5)
    for k, v in kw.items():
        setattr(hp, k, v)
    return hp


def make_batch(alg, hp, seed, B=None, dist="zipf", id_dtype=None):
    B = B or hp.batch_size
    if alg == "dnn_multi_textline":
        return make_textline_batch(hp, seed, B)
    multi = alg in ("deepfm_multi", "dnn_multi", "deepfm_multi_cate", "dnn_multi_cate")
    reserve = multi or alg in ("deepfm_cate", "dnn_cate", "dnn_pipeline", "wdl_textline", "deepfm_pipeline")   # row-0-zero models: id 0 is padding
    n_single_rows = (hp.cate_index_size if alg == "wdl_textline" else hp.cate_feats_size) - (1 if reserve else 0)
    if multi:
        Fm = hp.multi_field_size
        multi_rows = [max(2, n_single_rows // (4 * Fm))] * Fm
        sizes = D.field_sizes(n_single_rows - sum(multi_rows), hp.cate_field_size)
        tops = [int(r[1]) - int(r[0]) for r in hp.multi_feats_range]
    else:
        sizes, multi_rows, tops = D.field_sizes(n_single_rows, hp.cate_field_size), [], []
    if id_dtype is None:
        id_dtype = torch.int64 if reserve else torch.int32          # the tf.data generation parses int64 ids
    b = D.synthetic_batch(B, hp.cont_field_size, hp.cate_field_size, sizes, seed, V=hp.vector_feats_size,
                          multi_top_n=tops, multi_sizes=multi_rows, wide=hp.wide_field_size if alg in ("wdl", "wdl_textline", "wdl_ftrl") else 0,
                          dist=dist, reserve_zero=reserve, id_dtype=id_dtype)
    return b


def make_textline_batch(hp, seed, B=None, id_dtype=torch.int32):
    """dnn_multi_textline batch in the canonical key layout: cate_feats = [cat_feats | mul_cat_feats], weights in
    multi_feats_value.  Trainable multi fields draw ids from the main table (0 = padding), the 'tag' field from the
    word2vec rows (its padding id 0 reads the real row 0 of that table, as in the reference)."""
    B = B or hp.batch_size
    g = torch.Generator(); g.manual_seed(seed)
    Fc = hp.cate_field_size
    sizes = D.field_sizes(hp.cate_feats_size - 1, Fc)
    base = D.synthetic_batch(B, hp.cont_field_size, Fc, sizes, seed, dist="zipf", reserve_zero=True, id_dtype=id_dtype)
    cols = []
    for s0, s1, name in hp.multi_feats_range:
        tn = s1 - s0
        hi = hp.word2vec_rows if name == "tag" else min(hp.cate_feats_size, 400)
        ids = torch.randint(1, hi, (B, tn), generator=g)
        length = torch.randint(0, tn + 1, (B, 1), generator=g)
        cols.append(torch.where(torch.arange(tn).unsqueeze(0) < length, ids, torch.zeros_like(ids)))
    multi = torch.cat(cols, 1).to(id_dtype)
    base["cate_feats"] = torch.cat([base["cate_feats"], multi], 1).contiguous()
    base["multi_feats_value"] = torch.rand(B, multi.shape[1], generator=g) * 2.0
    return base


def oracle_batch(b, dtype):
    return {k: (v.to(dtype) if v.is_floating_point() else v.long()) for k, v in b.items()}


def rel_err(a, b):
    """max |a-b| / max(|b|) : relative to the tensor's scale (1e-5 FP32 / 2e-3 TF32 bars, BASELINE.json)"""
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b)) / max(float(np.max(np.abs(b))), 1e-30))
