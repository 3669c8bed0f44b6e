"""Generates the frozen vectors in this directory from the oracle (oracle/ctr_oracle.py).

    python tests/golden/make_golden.py

The reference ships no golden vectors and TensorFlow 1.x cannot be imported here, so these files do
NOT pin the oracle to the reference; they freeze the oracle's outputs on fixed inputs so that (1) any later edit of
the oracle that changes results is caught on CPU (tests/test_golden.py) and (2) the CUDA path is checked against
committed numbers, not only against whatever the oracle computes today (tests/test_gpu_golden.py).
One file per model: inputs (parameters by their reference variable names, one batch), the forward outputs and the
state after TWO train steps with TF1 dense Adam (step 2 is where lazy and dense Adam differ).
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

from oracle import ctr_oracle as O   # noqa: E402
from deep_learning_b200 import data as D   # noqa: E402


BASE = ("deepfm", "deepfm_multi", "wdl", "dnn", "dnn_multi")
VARIANTS = ("deepfm_cate", "deepfm_multi_cate", "dnn_pipeline", "dnn_cate", "dnn_multi_cate")
TEXTLINE = ("wdl_textline", "dnn_multi_textline", "deepfm_pipeline")
MULTI = ("deepfm_multi", "dnn_multi", "deepfm_multi_cate", "dnn_multi_cate")
ROW0 = MULTI + ("deepfm_cate", "dnn_cate", "dnn_pipeline", "wdl_textline", "deepfm_pipeline")          # row-0-zero tables: id 0 is padding


def tiny_hp(alg):
    kw = dict(hidden_units=[8, 4], batch_size=16, embedding_size=4, cont_field_size=3, cate_field_size=5,
              cate_feats_size=60, l2_reg=1e-3, learning_rate=1e-2)
    if alg in MULTI:
        kw.update(cate_feats_size=81, multi_feats_size=8, multi_field_size=2, multi_feats_range=[[0, 3, "m0"], [3, 8, "m1"]])
    if alg in O.NO_CONT:
        kw.update(cont_field_size=0)
    if alg in VARIANTS or alg == "deepfm_pipeline":
        kw.update(vector_feats_size=2)
    if alg in ("wdl", "wdl_textline"):
        kw.update(cate_index_size=60, wide_field_size=5)
    if alg == "wdl_ftrl":                  # test_model/wdl.py: one size for the table and the wide ids
        kw.update(wide_field_size=5)
    if alg == "dcn":                       # test_model/DCN.py: a dense feature vector only
        kw.update(deep_feature_size=7)
    if alg == "dnn_multi_textline":
        kw.update(cate_feats_size=81, multi_feats_size=8, multi_field_size=2, multi_feats_range=[[0, 3, "kw"], [3, 8, "tag"]],
                  word2vec_rows=12)
    return O.HParams(**kw)


def tiny_batch(alg, hp, seed):
    if alg == "deepfm_pipeline":
        b = _tiny_batch("deepfm_pipeline", hp, seed)
        b["cate_feats"][0, 0] = 0          # padding ids: the deep part reads the zero row, the FM part row C (quirk Q2)
        b["cate_feats"][3, 2] = 0
        return b
    return _tiny_batch(alg, hp, seed)


def _tiny_batch(alg, hp, seed):
    if alg == "dcn":
        g = torch.Generator(); g.manual_seed(seed)
        return {"features": torch.randn(hp.batch_size, hp.deep_feature_size, generator=g),
                "label": (torch.rand(hp.batch_size, 1, generator=g) < 0.3).float()}
    if alg == "dnn_multi_textline":
        import ctr_helpers as H
        return H.make_textline_batch(hp, seed)
    multi, row0 = alg in MULTI, alg in ROW0
    if multi:
        sizes, msizes, tops = D.field_sizes(60, 5), [10, 10], [3, 5]
    else:
        sizes, msizes, tops = D.field_sizes(hp.cate_feats_size - (1 if row0 else 0), 5), [], []
    return D.synthetic_batch(hp.batch_size, hp.cont_field_size, hp.cate_field_size, sizes, seed, V=hp.vector_feats_size,
                             multi_top_n=tops, multi_sizes=msizes, wide=hp.wide_field_size if alg in ("wdl", "wdl_textline", "wdl_ftrl") else 0,
                             dist="uniform", reserve_zero=row0, id_dtype=torch.int64 if row0 else torch.int32)


def main():
    only = sys.argv[1:]      # optional: regenerate just these models (the committed files of the others stay byte-identical)
    for alg in (only or BASE + VARIANTS + TEXTLINE):
        hp = tiny_hp(alg)
        p = O.init_params(alg, hp, seed=2019, dtype=torch.float64)
        # golden inputs are the fp32-representable values (what gets injected into the CUDA engine)
        p = {k: v.float().double() for k, v in p.items()}
        st = O.adam_init(p)
        out = {}
        for k, v in p.items():
            out["param/" + k] = v.numpy()
        batches = [tiny_batch(alg, hp, seed=11), tiny_batch(alg, hp, seed=12)]
        for i, b in enumerate(batches):
            for k, v in b.items():
                out["batch%d/%s" % (i, k)] = v.numpy()
        ob = {k: (v.double() if v.is_floating_point() else v.long()) for k, v in batches[0].items()}
        logit, score = O.FORWARD[alg](p, ob, hp)
        out["fwd/logit"], out["fwd/score"] = logit.numpy(), score.numpy()
        out["fwd/auc"] = np.float64(O.roc_auc(batches[0]["label"].numpy(), score.numpy()))
        for i, b in enumerate(batches):
            ob = {k: (v.double() if v.is_floating_point() else v.long()) for k, v in b.items()}
            loss, p = O.train_step(alg, p, st, ob, hp, adam_mode="tf_dense")
            out["step%d/loss" % (i + 1)] = np.float64(float(loss))
            for k, v in p.items():
                out["step%d/%s" % (i + 1, k)] = v.numpy()
        path = os.path.join(HERE, "%s_tiny.npz" % alg)
        np.savez_compressed(path, **out)
        print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
