"""-m gpu: the CUDA path against runs of the reference's own model files (tests/golden/ref_<model>.npz, see
tests/test_ref_golden.py for how they were made): same initial variables, the train steps and evaluations of the
reference's `fit` replayed in its order over its batches, FP32 path, TF1 dense Adam, FREE-RUNNING (no re-injection
between steps).  Bars: every step's loss, every evaluation output and every final variable 1e-5 relative to the
tensor's scale (BASELINE.json's FP32 tolerance, unrelaxed).

Beside the bar, the float32 floor: tests/golden/fp32floor_<model>.npz hold the SAME reference runs with the stand-in
computing in float32 (TF's arithmetic type; `make_ref_golden.py --fp32-floor`).  Their distance to the float64
fixtures is what float32 arithmetic costs by itself (1e-7 .. 8e-7 on the final variables); the CUDA path must stay
within FLOOR_FACTOR of it (or under ABS_FLOOR where the floor is only the final rounding), so a systematic error far
above float32 noise -- round 1's `1.f - 0.999f` in the Adam kernels was 55x the floor -- fails even when it is under
1e-5.  `tools/parity_report.py` prints the measured table (committed under profiles/)."""
import os

import numpy as np
import pytest
import torch

import ctr_helpers as H
from golden.make_golden import tiny_hp
from test_ref_golden import CUDA_ALGS as ALGS, HERE, LOGIT_EVAL          # the 13 files of models/ + test_model/wdl.py (DCN: oracle level only so far)

pytestmark = pytest.mark.gpu

BAR = 1e-5                # BASELINE.json, FP32 path
FLOOR_FACTOR = 8.0        # two float32 implementations with different summation orders: a few times the floor
ABS_FLOOR = 2e-6          # where the float32 floor is only the final rounding (1e-8 .. 1e-7)


def _batch(z, prefix):
    return {k.split("/", 1)[1]: torch.from_numpy(z[k]) for k in z.files if k.startswith(prefix + "/")}


def _rel(a, ref):
    a, ref = np.asarray(a, np.float64).reshape(-1), np.asarray(ref, np.float64).reshape(-1)
    return float(np.max(np.abs(a - ref)) / max(1.0, float(np.max(np.abs(ref)))))


def replay(alg, precision="fp32", adam_mode="tf_dense"):
    """-> [(what, err, floor)]: the CUDA path's and the float32 stand-in's distance to the float64 fixture"""
    from deep_learning_b200.engine import CtrEngine
    from deep_learning_b200.spec import spec_from_hparams
    z = np.load(os.path.join(HERE, "golden", "ref_%s.npz" % alg))
    zf = np.load(os.path.join(HERE, "golden", "fp32floor_%s.npz" % alg))
    hp = tiny_hp(alg)
    spec = spec_from_hparams(alg, hp)
    cls = CtrEngine
    if spec.split_fm:
        from deep_learning_b200.split_engine import SplitFmEngine as cls
    eng = cls(spec, hp.batch_size, learning_rate=hp.learning_rate, decay_steps=hp.learning_rate_decay_steps,
              decay_rate=hp.learning_rate_decay_rate, precision=precision, adam_mode=adam_mode)
    eng.set_weights({k[5:]: z[k].astype(np.float32) for k in z.files if k.startswith("init/")})
    family, events, n_val = str(z["family"]), str(z["events"]), int(z["n_val"])
    B = hp.batch_size
    t = e = pulled = 0
    ntb = len([k for k in z.files if k.startswith("train") and k.endswith("/label")])
    rows = []
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
            rows.append(("loss step %d" % t, abs(loss - ref) / max(1.0, abs(ref)),
                         abs(float(zf["losses"][t]) - ref) / max(1.0, abs(ref))))
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
            assert got.shape == ref.shape, (alg, e)
            rows.append(("eval %d" % e, _rel(got, ref), _rel(zf["eval%d" % e], ref)))
            e += 1
    got = eng.get_weights()
    for k in [k[6:] for k in z.files if k.startswith("final/")]:
        ref = z["final/" + k].reshape(-1)
        rows.append(("final " + k, H.rel_err(got[k].reshape(-1), ref), H.rel_err(zf["final/" + k].reshape(-1), ref)))
    return rows


@pytest.mark.parametrize("alg", ALGS)
def test_cuda_reproduces_the_reference_run(alg):
    for what, err, floor in replay(alg):
        assert err < BAR, (alg, what, err, floor)
        assert err <= max(FLOOR_FACTOR * floor, ABS_FLOOR), (alg, what, err, floor)
