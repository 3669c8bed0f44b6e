"""-m gpu: fault reporting of the CUDA path.  tf.nn.embedding_lookup raises InvalidArgumentError for an id outside
its table (call sites models/deepfm.py:78,85,98, models/wdl.py:240-246); here such a batch reads no memory out of
bounds, its loss is NaN, NO weight moves (the optimizer kernels check the engine's status words) and the next host
read raises, naming the column."""
import numpy as np
import pytest
import torch

import ctr_helpers as H
from oracle import ctr_oracle as O

pytestmark = pytest.mark.gpu


def _engine(alg, hp, **kw):
    from deep_learning_b200.engine import CtrEngine
    from deep_learning_b200.spec import spec_from_hparams
    p = O.init_params(alg, hp, dtype=torch.float64)
    eng = CtrEngine(spec_from_hparams(alg, hp), hp.batch_size, learning_rate=hp.learning_rate, precision="fp32", **kw)
    eng.set_weights({k: v.numpy() for k, v in p.items()})
    return eng


@pytest.mark.parametrize("alg,col,value", [
    ("deepfm", 3, 10 ** 6),            # K1 fast path, id past the table
    ("deepfm", 19, -5),                # negative id
    ("deepfm_multi", 25, 10 ** 6),     # K2 (pooled) slot
    ("wdl", 2, 10 ** 6),               # embedding id of the deep part
])
def test_bad_embedding_id_is_reported_and_moves_no_weight(alg, col, value):
    from deep_learning_b200._lib import CtrError
    hp = H.hp_multi(B=256) if alg == "deepfm_multi" else (H.hp_criteo(B=256, rows=5000) if alg == "wdl" else H.hp_c1())
    eng = _engine(alg, hp, adam_mode="tf_dense")
    good = H.make_batch(alg, hp, seed=7)
    loss0 = float(eng.train_step(good).item())
    assert np.isfinite(loss0)
    before = eng.get_weights()
    bad = {k: v.clone() for k, v in H.make_batch(alg, hp, seed=8).items()}
    bad["cate_feats"][5, col] = value
    loss = float(eng.train_step(bad).item())
    if alg != "wdl":          # the NaN row reaches the logit through the FM terms; a plain tower's ReLU turns NaN into 0
        assert np.isnan(loss)
    with pytest.raises(CtrError, match="column %d" % col):
        eng.check_status()
    after = eng.get_weights()               # status was cleared by the raise; nothing moved in the faulted step
    for k in before:
        assert np.array_equal(before[k], after[k]), k
    assert np.isfinite(float(eng.train_step(good).item()))
    eng.check_status()


def test_bad_wide_id_is_reported():
    from deep_learning_b200._lib import CtrError
    hp = H.hp_criteo(B=128, rows=5000)
    eng = _engine("wdl", hp, adam_mode="lazy")
    bad = {k: v.clone() for k, v in H.make_batch("wdl", hp, seed=3).items()}
    bad["wide_feats"][0, 4] = 10 ** 7
    scores = eng.predict(bad).cpu().numpy()
    assert np.isnan(scores[0]) and np.all(np.isfinite(scores[1:]))
    with pytest.raises(CtrError, match="column 4"):
        eng.check_status()


def test_fault_surfaces_at_the_next_step_without_an_explicit_check():
    from deep_learning_b200._lib import CtrError
    hp = H.hp_c1()
    eng = _engine("deepfm", hp, adam_mode="lazy")
    bad = {k: v.clone() for k, v in H.make_batch("deepfm", hp, seed=8).items()}
    bad["cate_feats"][0, 0] = 2 ** 30
    eng.train_step(bad)
    torch.cuda.synchronize()
    with pytest.raises(CtrError, match="id out of range"):
        eng.train_step(H.make_batch("deepfm", hp, seed=9))
