"""CPU: the self-describing part of the serving export (deep_learning_b200/models/_base.py `save`): hyper-parameter bag,
model file name and the named-tensor signature of models/deepfm.py:193-201 / models/wdl.py:320-328."""
import json
import os
import types

import numpy as np
import pytest

import ctr_helpers as H
from deep_learning_b200.hparams import ModelParams
from deep_learning_b200.models import _base as B
from deep_learning_b200.spec import spec_from_hparams


def test_hparams_round_trip_through_json():
    mp = ModelParams(hidden_units=[400, 400], cate_field_size=26, multi_feats_range=[[0, 3, "kw"], [3, 8, "tag"]], l2_reg=1e-5)
    d = B.hparams_to_dict(mp)
    back = types.SimpleNamespace(**json.loads(json.dumps(d)))
    assert back.hidden_units == [400, 400] and back.multi_feats_range == [[0, 3, "kw"], [3, 8, "tag"]] and back.l2_reg == 1e-5
    # a launcher's own bag (local_run_textline.py:17-67 is a plain class): attributes that are not JSON stay behind
    ns = types.SimpleNamespace(hidden_units=(8, 4), batch_size=np.int64(16), handle=object(), _private=1, model_pb="x")
    d = B.hparams_to_dict(ns)
    assert d == {"hidden_units": [8, 4], "batch_size": 16, "model_pb": "x"}


@pytest.mark.parametrize("alg,extra", [("deepfm", set()), ("wdl", {"wide_feats"}), ("dnn_multi_textline", {"multi_feats_value"})])
def test_signature_names_the_reference_tensors(alg, extra):
    hp = H.hp_textline(alg) if alg == "dnn_multi_textline" else (H.hp_criteo(B=8, rows=100) if alg == "wdl" else H.hp_c1())
    sig = B.serving_signature(spec_from_hparams(alg, hp))
    assert set(sig["inputs"]) == {"cont_feats", "vector_feats", "cate_feats", "input_data_size"} | extra
    assert sig["inputs"]["cate_feats"]["tensor"] == "cate_feats:0" and sig["outputs"]["output"]["tensor"] == "score:0"
    assert sig["inputs"]["cate_feats"]["width"] == hp.cate_field_size + hp.multi_feats_size
    if alg == "wdl":
        assert sig["inputs"]["wide_feats"] == {"tensor": "wide_feats:0", "dtype": "int32", "width": hp.wide_field_size}


def test_meta_is_read_back_and_old_exports_say_what_they_lack(tmp_path):
    hp = H.hp_c1()
    meta = {"format": B.EXPORT_FORMAT, "alg": "deepfm", "generation": "A", "hparams": B.hparams_to_dict(hp),
            "signature": B.serving_signature(spec_from_hparams("deepfm", hp))}
    d = tmp_path / "pb"
    os.makedirs(d)
    np.savez(d / "model.npz", **{"w/x": np.zeros(2), "meta/json": np.array(json.dumps(meta))})
    got = B.read_export_meta(str(d))
    assert got["alg"] == "deepfm" and got["hparams"]["cate_feats_size"] == hp.cate_feats_size
    assert B.hparams_from_export(str(d)).hidden_units == hp.hidden_units
    old = tmp_path / "old"
    os.makedirs(old)
    np.savez(old / "model.npz", **{"w/x": np.zeros(2)})
    with pytest.raises(ValueError, match="variables-only export"):
        B.read_export_meta(str(old))
