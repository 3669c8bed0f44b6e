"""-m gpu: the CUDA path against the COMMITTED golden vectors (tests/golden/*.npz): forward, AUC and two train steps
with TF1 dense Adam, for every model family.  FP32 path, 1e-5 relative (BASELINE.json)."""
import numpy as np
import pytest
import torch

import ctr_helpers as H
from golden.make_golden import tiny_hp
from test_golden import ALGS, load

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("alg", ALGS)
def test_cuda_matches_golden(alg):
    from deep_learning_b200.engine import CtrEngine
    from deep_learning_b200.spec import spec_from_hparams
    z, p, batches = load(alg)
    hp = tiny_hp(alg)
    eng = CtrEngine(spec_from_hparams(alg, hp), hp.batch_size, learning_rate=hp.learning_rate, precision="fp32",
                    adam_mode="tf_dense")
    eng.set_weights({k: v.numpy() for k, v in p.items()})
    score = eng.predict(batches[0]).cpu().numpy()
    assert H.rel_err(eng.logit[:hp.batch_size].cpu().numpy(), z["fwd/logit"].reshape(-1)) < 1e-5
    assert np.max(np.abs(score - z["fwd/score"].reshape(-1))) < 1e-5
    assert abs(eng.auc(torch.from_numpy(score), batches[0]["label"]) - float(z["fwd/auc"])) < 1e-12
    for i, b in enumerate(batches):
        loss = float(eng.train_step(b).item())
        ref = float(z["step%d/loss" % (i + 1)])
        assert abs(loss - ref) < 1e-5 * max(1.0, abs(ref)), (i, loss, ref)
        got = eng.get_weights()
        for k in p:
            assert H.rel_err(got[k].reshape(-1), z["step%d/%s" % (i + 1, k)].reshape(-1)) < 2e-5, (alg, i, k)
