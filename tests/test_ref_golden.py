"""The oracle against runs of the reference's own model files.

`tests/golden/ref_<model>.npz` hold what /root/reference/models/<model>.py computed -- unmodified, `fit` loop and
all -- when executed under the TensorFlow-1.x stand-in (`tests/golden/tf1_shim.py`, generator
`tests/golden/make_ref_golden.py`, run in the build container where /root/reference exists).  Here the oracle starts
from the same initial variables, replays the same order of train steps and evaluations over the same batches, and
must land on the same losses, evaluation outputs and final variables.  This is what ties the oracle's wiring (index
offsets, aliasing quirks, concatenation order, regularised weights, batch consumption of each `fit`) to the
reference's code rather than to a reading of it; the arithmetic of the individual `tf.*` symbols is still a
restatement (see the stand-in's header)."""
import glob
import os
import re

import numpy as np
import pytest
import torch

from oracle import ctr_oracle as O
from golden import make_golden as G

HERE = os.path.dirname(os.path.abspath(__file__))
FILES = sorted(f for f in glob.glob(os.path.join(HERE, "golden", "ref_*.npz")) if not f.endswith("ref_loader.npz"))
ALGS = [os.path.basename(f)[4:-4] for f in FILES]
# fixtures of models/*.py (the 13 files the product has a CUDA path and a drop-in module for).  The two test_model/ files
# (wdl_ftrl, dcn) are pinned at the ORACLE level only so far -- their CUDA path is not built, so the
# GPU replay (tests/test_ref_gpu.py) and the live surface checks (tests/test_ref_live.py) run over MODEL_ALGS
MODEL_ALGS = [a for a in ALGS if a not in O.TEST_MODEL_FILES]
CUDA_ALGS = MODEL_ALGS + [a for a in ALGS if a in ("wdl_ftrl",)]      # what tests/test_ref_gpu.py replays on the device

TOL = 1e-12          # float64 on both sides, same operation order up to summation order inside torch reductions

# what the reference's evaluation fetches: models/deepfm.py:229 runs `self.out`, which there is the LOGIT (:143; AUC is
# rank-invariant so the reference never notices); models/dnn.py:84,159 and models/wdl.py:264,355 run `self.out` after
# the sigmoid; the tf.data-fed files run "score:0" (models/deepfm_multi.py:376); the textline files fetch
# [label, out, loss] with `out` already the sigmoid (models/wdl_textline.py:278,312)
LOGIT_EVAL = ("deepfm",)


def _batch(z, prefix):
    b = {k.split("/", 1)[1]: torch.from_numpy(z[k]) for k in z.files if k.startswith(prefix + "/")}
    return {k: (v.double() if v.is_floating_point() else v.long()) for k, v in b.items()}


def _n_train_batches(z):
    return len([k for k in z.files if k.startswith("train") and k.endswith("/label")])


def test_every_model_file_of_the_reference_has_a_fixture():
    assert set(ALGS) >= set(O.FORWARD) - set(O.TEST_MODEL_FILES), "ref_*.npz must cover every models/*.py the oracle restates"


@pytest.mark.parametrize("alg", ALGS)
def test_oracle_reproduces_the_reference_run(alg):
    z = np.load(os.path.join(HERE, "golden", "ref_%s.npz" % alg))
    replay_with_oracle(alg, _Files(z), G.tiny_hp(alg))


class _Files(dict):
    """an .npz or the generator's dict behind one interface"""

    def __init__(self, z):
        super().__init__({k: z[k] for k in (z.files if hasattr(z, "files") else z)})
        self.files = list(self)


def replay_with_oracle(alg, z, hp):
    p = {k[5:]: torch.from_numpy(z[k]).double() for k in z.files if k.startswith("init/")}
    assert set(p) == set(O.param_shapes(alg, hp))
    st = O.optimizer_init(alg, p)          # Adam slots, or FTRL's for test_model/wdl.py
    family, events = str(z["family"]), str(z["events"])
    n_val = int(z["n_val"])
    t = e = pulled = 0
    ntb = _n_train_batches(z)
    assert events.count("T") == int(z["n_train"]) == len(z["losses"]) and events.count("E") == int(z["n_eval"])
    for ev in events:
        if family == "text":                      # one stream: every fetch of a model output pulls the next batch
            b = _batch(z, "train%d" % pulled)
            pulled += 1
        elif ev == "T":
            b = _batch(z, "train%d" % (t % ntb))  # generation A walks the same list once per epoch
        else:
            b = _batch(z, "val%d" % (e % n_val))
        if ev == "T":
            loss, p = O.train_step(alg, p, st, b, hp, adam_mode="tf_dense")
            assert abs(float(loss) - float(z["losses"][t])) < TOL, (alg, t)
            t += 1
        else:
            ref = z["eval%d" % e]
            if family == "text":
                loss, _, score = O.LOSS[alg](p, b, hp)
                got = np.concatenate([score.detach().numpy().ravel(), [float(loss)]])
            else:
                logit, score = O.FORWARD[alg](p, b, hp)[:2]
                got = (logit if alg in LOGIT_EVAL else score).detach().numpy()
            assert got.shape == ref.shape, (alg, e, got.shape, ref.shape)
            assert np.max(np.abs(got - ref)) < TOL, (alg, e)
            e += 1
    for k, v in p.items():
        assert float((v - torch.from_numpy(z["final/" + k]).double()).abs().max()) < TOL, (alg, k)
    if int(z["global_step"]) >= 0:
        assert int(z["global_step"]) == int(z["n_train"])


@pytest.mark.parametrize("alg", ALGS)
def test_reference_run_moves_every_variable_it_should(alg):
    """guards the fixture itself: a variable the run left untouched would make the final-state comparison vacuous"""
    z = np.load(os.path.join(HERE, "golden", "ref_%s.npz" % alg))
    for k in z.files:
        if k.startswith("init/") and k != "init/tags_emb":
            assert np.any(z[k] != z["final/" + k[5:]]), (alg, k)
    if "init/tags_emb" in z.files:
        assert np.array_equal(z["init/tags_emb"], z["final/tags_emb"])          # the word2vec rows are a constant


@pytest.mark.parametrize("alg", [a for a in ALGS if a not in ("wdl_textline", "dnn_multi_textline")])
def test_auc_printed_by_the_reference_fit(alg):
    """`fit` prints scikit-learn's roc_auc_score over the validation batches (models/deepfm.py:188-190,
    models/deepfm_multi.py:297-301); its last print is taken with the final variables"""
    z = np.load(os.path.join(HERE, "golden", "ref_%s.npz" % alg))
    hp = G.tiny_hp(alg)
    line = str(z["fit_stdout"]).splitlines()[-1]
    printed = float(re.search(r"auc[=:]\s*([0-9.]+)", line).group(1))
    p = {k[6:]: torch.from_numpy(z[k]).double() for k in z.files if k.startswith("final/")}
    labels, scores = [], []
    for i in range(int(z["n_val"])):
        b = _batch(z, "val%d" % i)
        labels.append(b["label"].numpy().ravel())
        scores.append(O.FORWARD[alg](p, b, hp)[1].detach().numpy().ravel())
    auc = O.roc_auc(np.concatenate(labels), np.concatenate(scores))
    digits = len(line.split("auc")[1].split()[0].split(".")[-1]) if "valid-auc" in line else 12
    assert abs(auc - printed) < 0.6 * 10 ** -min(digits, 5) if "valid-auc" in line else abs(auc - printed) < 1e-9, (line, auc)


class _StubEngine:
    """records what the product's fit loops ask of the engine; no arithmetic"""

    def __init__(self):
        self.log = []
        self.last_batch = 2
        self.label = torch.tensor([0.0, 1.0])
        self.score = torch.tensor([0.25, 0.75])

    def train_step(self, b, next_batch=None):
        self.log.append(("T", b["tag"]))
        return torch.tensor(0.5)

    def eval_loss(self, b):
        self.log.append(("E", b["tag"]))
        return torch.tensor(0.5)

    def predict(self, b):
        self.log.append(("E", b["tag"]))
        return torch.tensor([0.25, 0.75])

    def auc(self, scores, labels):
        return 0.5


@pytest.mark.parametrize("alg", ALGS)
def test_product_fit_loop_consumes_batches_like_the_reference_fit(alg, capsys):
    """host logic of deep_learning_b200/models/_base.py against the order in which the reference's `fit` issued train
    steps and evaluations (recorded in the fixture), driven with the generator's arguments"""
    from deep_learning_b200.models._base import BaseDeepModel
    z = np.load(os.path.join(HERE, "golden", "ref_%s.npz" % alg))
    family, events = str(z["family"]), str(z["events"])
    m = object.__new__(BaseDeepModel)
    m.engine = eng = _StubEngine()
    m._decode = lambda item: item
    m.epochs, m.batch_size, m.metric_type = int(z["epochs"]), 16, "auc"
    m.model_pb = m.save_model_checkpoint = m.restore_model_checkpoint = None
    m.model_restore = 0
    n_train = len([k for k in z.files if k.startswith("train") and k.endswith("/label")])
    train = [{"tag": ("train", i), "label": np.array([0.0, 1.0])} for i in range(n_train)]
    val = [{"tag": ("val", i), "label": np.array([0.0, 1.0])} for i in range(int(z["n_val"]))]
    if family == "feed":
        m._fit_a(train, val)
    elif family == "pipe":
        m.data_dict = iter(train)
        m._fit_b(1, iter(val))
    else:
        m.data_dict = iter(train)
        m._fit_c(2, 1)
    capsys.readouterr()
    assert "".join(k for k, _ in eng.log) == events
    t = e = pulled = 0
    for kind, tag in eng.log:
        if family == "text":
            assert tag == ("train", pulled)
            pulled += 1
        elif kind == "T":
            assert tag == ("train", t % n_train)
            t += 1
        else:
            assert tag == ("val", e % len(val))
            e += 1


class _OracleEngine:
    """the oracle behind the engine interface the product's fit loops use (test double: lets those loops run on CPU with
    real numbers, so what they PRINT can be held against what the reference's `fit` printed)"""

    def __init__(self, alg, hp, p):
        self.alg, self.hp, self.p, self.st = alg, hp, p, O.optimizer_init(alg, p)
        self.last_batch, self.label, self.score = 0, None, None

    @staticmethod
    def _ob(b):
        out = {}
        for k, v in b.items():
            if k == "tag":
                continue
            t = torch.as_tensor(v)
            out[k] = t.double() if t.is_floating_point() else t.long()
        return out

    def train_step(self, b, next_batch=None):
        loss, self.p = O.train_step(self.alg, self.p, self.st, self._ob(b), self.hp, adam_mode="tf_dense")
        return loss.detach()

    def predict(self, b):
        return O.FORWARD[self.alg](self.p, self._ob(b), self.hp)[1].detach().reshape(-1)

    def eval_loss(self, b):
        ob = self._ob(b)
        loss, _, score = O.LOSS[self.alg](self.p, ob, self.hp)
        self.last_batch, self.label, self.score = ob["label"].shape[0], ob["label"].reshape(-1), score.detach().reshape(-1)
        return loss.detach()

    def auc(self, scores, labels):
        return O.roc_auc(labels.numpy().reshape(-1), scores.numpy().reshape(-1))


def _progress_lines(text):
    """(iteration tag, auc, loss) of every progress line, whatever the generation's format"""
    out = []
    for line in text.splitlines():
        m = re.match(r"\[(\d+)\]\s*(?:valid-auc=|val_auc:|test auc:)([0-9.eE+-]+)\s*loss[=:]([0-9.eE+-]+)", line)
        if m:
            out.append((int(m.group(1)), float(m.group(2)), float(m.group(3))))
    return out


@pytest.mark.parametrize("alg", ALGS)
def test_product_fit_prints_what_the_reference_fit_printed(alg, capsys):
    """deep_learning_b200/models/_base.py `_fit_a/_fit_b/_fit_c` (running-loss bookkeeping, evaluation cadence, line
    formats) driven over the recorded batches with the oracle as engine: the progress lines carry the reference's
    iteration tags, AUCs and losses"""
    from deep_learning_b200.models._base import BaseDeepModel
    z = np.load(os.path.join(HERE, "golden", "ref_%s.npz" % alg))
    hp = G.tiny_hp(alg)
    family = str(z["family"])
    p = {k[5:]: torch.from_numpy(z[k]).double() for k in z.files if k.startswith("init/")}
    m = object.__new__(BaseDeepModel)
    m.engine = _OracleEngine(alg, hp, p)
    m._decode = lambda item: item
    m.epochs, m.batch_size, m.metric_type = int(z["epochs"]), hp.batch_size, "auc"
    m.model_pb = m.save_model_checkpoint = m.restore_model_checkpoint = None
    m.model_restore = 0
    n_train = len([k for k in z.files if k.startswith("train") and k.endswith("/label")])
    train = [{k.split("/", 1)[1]: z[k] for k in z.files if k.startswith("train%d/" % i)} for i in range(n_train)]
    val = [{k.split("/", 1)[1]: z[k] for k in z.files if k.startswith("val%d/" % i)} for i in range(int(z["n_val"]))]
    if family == "feed":
        m._fit_a(train, val)
    elif family == "pipe":
        m.data_dict = iter(train)
        m._fit_b(1, iter(val))
    else:
        m.data_dict = iter(train)
        m._fit_c(2, 1)
    got, want = _progress_lines(capsys.readouterr().out), _progress_lines(str(z["fit_stdout"]))
    assert want and len(got) == len(want), (got, want)
    for (gi, ga, gl), (wi, wa, wl) in zip(got, want):
        assert gi == wi
        tol = 6e-6 if family == "feed" else 1e-12          # generation A prints %.5f
        assert abs(ga - wa) <= tol and abs(gl - wl) <= tol, (alg, gi, (ga, gl), (wa, wl))


@pytest.mark.parametrize("alg", ALGS)
def test_reference_predict_after_export(alg, capsys):
    """the reference's predict (method or module function: reload the exported model, run "score:0" batch by batch, print
    scikit-learn's AUC; models/deepfm.py:233-255, models/deepfm_multi.py:383-418, models/wdl_textline.py:362-386) ran after
    `fit` when the fixture was made.  The oracle's scores under the final variables give that AUC, and the product's
    `_predict_a` prints the same line."""
    from deep_learning_b200.models._base import BaseDeepModel
    z = np.load(os.path.join(HERE, "golden", "ref_%s.npz" % alg))
    hp = G.tiny_hp(alg)
    line = str(z["predict_stdout"]).strip()
    assert re.fullmatch(r"val of auc:\d\.\d{5}", line), line
    p = {k[6:]: torch.from_numpy(z[k]).double() for k in z.files if k.startswith("final/")}
    names = [str(n) for n in z["predict_on"]]
    m = object.__new__(BaseDeepModel)
    m.engine = _OracleEngine(alg, hp, p)
    m._decode = lambda item: item
    m.model_pb = None
    batches = [{k.split("/", 1)[1]: z[k] for k in z.files if k.startswith(n + "/")} for n in names]
    auc = m._predict_a(batches)
    assert capsys.readouterr().out.strip().splitlines()[-1] == line
    assert abs(auc - float(line.split(":")[1])) < 6e-6
