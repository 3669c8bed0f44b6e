"""-m gpu: the serving export is self-describing, like the reference's SavedModel (models/deepfm.py:193-211,
models/deepfm_multi.py:312-334): a fresh process needs only the export directory -- module-level
`predict(data, model_pb)` with the reference's TWO arguments (local_run_test.py:82) and `session.run("score:0",
feed_dict)` by tensor NAME (models/deepfm.py:243-250, models/wdl.py:370-377)."""
import numpy as np
import pytest
import torch

import ctr_helpers as H

pytestmark = pytest.mark.gpu


def test_two_argument_predict_needs_only_the_export(tmp_path, capsys):
    import deep_learning_b200.models.deepfm_multi as alg_model
    hp = H.hp_multi(B=128)
    hp.model_pb = str(tmp_path / "pb")
    train = [H.make_batch("deepfm_multi", hp, seed=s) for s in range(3)]
    val = [H.make_batch("deepfm_multi", hp, seed=100 + s) for s in range(2)]
    m = alg_model.DeepModel(hp, iter(train), adam_mode="lazy")
    m.fit(2, iter(val))
    want = m._auc(val)
    del m
    capsys.readouterr()
    auc = alg_model.predict(val, hp.model_pb)                 # no hyper-parameter bag: read from the export
    assert "val of auc:" in capsys.readouterr().out
    assert abs(auc - want) < 1e-12


@pytest.mark.parametrize("alg", ["deepfm", "wdl"])
def test_scores_by_tensor_name_from_the_export_alone(alg, tmp_path):
    from deep_learning_b200 import data as D
    from deep_learning_b200.models._base import load_exported, read_export_meta
    import importlib
    alg_model = importlib.import_module("deep_learning_b200.models." + alg)
    hp = H.hp_c1(batch_size=128) if alg == "deepfm" else H.hp_c1(batch_size=128, cate_index_size=10000, wide_field_size=20)
    hp.model_pb = str(tmp_path / "pb")
    train = [D.to_pickled_batch(H.make_batch(alg, hp, seed=s)) for s in range(3)]
    b = H.make_batch(alg, hp, seed=77, B=50)
    m = alg_model.DeepModel(hp, adam_mode="lazy")
    m.fit(train, train[:1])
    want = m.predict_scores(D.to_pickled_batch(b))
    del m
    meta = read_export_meta(hp.model_pb)
    assert meta["alg"] == alg and meta["signature"]["outputs"]["output"]["tensor"] == "score:0"
    served = load_exported(hp.model_pb)
    feed = {"cont_feats:0": b["cont_feats"].numpy(), "cate_feats:0": b["cate_feats"].numpy(), "input_data_size:0": [50]}
    if alg == "wdl":
        feed["wide_feats:0"] = b["wide_feats"].numpy()
    got = served.run("score:0", feed)
    assert got.shape == (50, 1) and np.array_equal(got, want)             # same weights, same kernels: bit-identical
    with pytest.raises(KeyError, match="not an input"):
        served.run("score:0", {"no_such:0": [1]})
    with pytest.raises(KeyError, match="not an output"):
        served.run("logits:0", feed)
