"""CPU, build container only (skipped where /root/reference is absent): the reference's model files are run LIVE under
the TensorFlow-1.x stand-in at configurations other than the committed fixtures' -- one tower layer and three, no
regulariser, other embedding widths, batch sizes, continuous / vector widths -- and the oracle must reproduce each run
to 1e-12 (same replay as tests/test_ref_golden.py).  Widens the pin beyond one shape per model."""
import copy
import os
import sys

import numpy as np
import pytest

from oracle import ctr_oracle as O
from golden import make_golden as G
from test_ref_golden import MODEL_ALGS as ALGS, _Files, replay_with_oracle

pytestmark = pytest.mark.skipif(not os.path.isdir("/root/reference/models"), reason="needs /root/reference (build container)")


def _variant(alg, which):
    hp = copy.deepcopy(G.tiny_hp(alg))
    if which == "shallow_noreg":
        hp.hidden_units, hp.embedding_size, hp.l2_reg, hp.learning_rate, hp.batch_size = [6], 3, 0.0, 5e-3, 8
    else:
        hp.hidden_units, hp.embedding_size, hp.l2_reg, hp.batch_size = [5, 7, 3], 8, 1e-2, 24
        if alg not in O.NO_CONT:
            hp.cont_field_size = 4
        if hp.vector_feats_size:
            hp.vector_feats_size = 3
    return hp


@pytest.fixture(scope="module")
def generator():
    here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    sys.path.insert(0, here)
    import make_ref_golden as M
    M.S.install()
    yield M
    for k in [k for k in sys.modules if k == "tensorflow" or k.startswith("tensorflow.")]:
        if getattr(sys.modules[k], "__file__", None) is None:
            del sys.modules[k]


@pytest.mark.parametrize("which", ["shallow_noreg", "deep_wide"])
@pytest.mark.parametrize("alg", ALGS)
def test_oracle_reproduces_live_reference_runs(alg, which, generator, capsys):
    hp = _variant(alg, which)
    z = _Files(generator.run_reference(alg, hp))
    capsys.readouterr()
    assert int(z["n_train"]) >= 3
    replay_with_oracle(alg, z, hp)


@pytest.mark.parametrize("alg", ALGS)
def test_product_initial_distributions_follow_the_reference(alg, generator, capsys):
    """`models/_base.py::init_weights` against the variables the reference's own constructors initialise (drawn by the
    stand-in with the parameters the reference code passes: stddev / bounds / glorot scales).  Values cannot match (TF's
    generators are not reproducible); location, scale and support must."""
    from deep_learning_b200.models._base import init_weights
    from deep_learning_b200.spec import spec_from_hparams
    hp = copy.deepcopy(G.tiny_hp(alg))
    hp.hidden_units, hp.embedding_size = [48, 24], 8
    if alg in ("wdl", "wdl_textline"):
        hp.cate_index_size = 3000
    else:
        hp.cate_feats_size = 3000
    z = generator.run_reference(alg, hp)
    capsys.readouterr()
    mine = init_weights(spec_from_hparams(alg, hp), seed=7)
    ref = {k[5:]: np.asarray(v) for k, v in z.items() if k.startswith("init/") and k != "init/tags_emb"}
    assert set(ref) <= set(mine), sorted(set(ref) - set(mine))
    for name, r in ref.items():
        m = np.asarray(mine[name], dtype=np.float64)
        assert m.shape == r.shape, (alg, name, m.shape, r.shape)
        n = r.size
        if n < 400:
            continue                                     # biases of a few elements: shape only
        tol = 4.0 / np.sqrt(n)                            # ~4 sigma of the sampling error of a scale estimate
        assert abs(m.std() / r.std() - 1) < 1.5 * tol + 0.02, (alg, name, m.std(), r.std())
        assert abs(m.mean() - r.mean()) < 6 * r.std() / np.sqrt(n) + 1e-12, (alg, name, m.mean(), r.mean())
        # support: a uniform draw has hard bounds, a normal draw does not
        kurt = lambda a: float(np.mean((a - a.mean()) ** 4) / a.var() ** 2)
        assert abs(kurt(m) - kurt(r)) < 0.5, (alg, name, kurt(m), kurt(r))          # uniform 1.8 vs normal 3.0


@pytest.mark.parametrize("alg", ALGS)
def test_module_surface_matches_the_reference_module(alg, generator):
    """drop-in surface: every public callable of /root/reference/models/<alg>.py that a launcher can reach (DeepModel, its
    fit / evaluate / eval / predict / get_val_data, the module-level predict) exists in deep_learning_b200/models/<alg>.py
    with the same positional parameter names in the same order"""
    import importlib
    import inspect
    ref = generator.load_reference_model(alg)
    mine = importlib.import_module("deep_learning_b200.models." + alg)

    def params(fn):
        return [p.name for p in inspect.signature(fn).parameters.values()
                if p.kind in (p.POSITIONAL_ONLY, p.POSITIONAL_OR_KEYWORD) and p.name != "self"]

    assert hasattr(mine, "DeepModel")
    ref_init = params(ref.DeepModel.__init__)
    assert params(mine.DeepModel.__init__)[:len(ref_init)] == ref_init, (alg, params(mine.DeepModel.__init__), ref_init)
    for name in ("fit", "evaluate", "eval", "predict", "get_val_data"):
        if hasattr(ref.DeepModel, name):
            assert hasattr(mine.DeepModel, name), (alg, name)
            want, got = params(getattr(ref.DeepModel, name)), params(getattr(mine.DeepModel, name))
            assert got[:len(want)] == want, (alg, name, got, want)
    if hasattr(ref, "predict"):
        assert hasattr(mine, "predict"), alg
        want, got = params(ref.predict), params(mine.predict)
        assert got[:len(want)] == want, (alg, got, want)


def test_loader_surface_matches_the_reference_loaders(generator):
    """deep_learning_b200/loaders.py keeps the reference loader functions' names and leading positional parameters
    (utils/data_loader_load.py, utils/data_loader.py, utils/data_loader_textline.py); where two reference modules use
    one name for different things the product's second name is listed"""
    import importlib
    import inspect
    from deep_learning_b200 import loaders as L

    def params(fn):
        return [p.name for p in inspect.signature(fn).parameters.values()
                if p.kind in (p.POSITIONAL_ONLY, p.POSITIONAL_OR_KEYWORD)]

    RL = importlib.import_module("utils.data_loader_load")
    RP = importlib.import_module("utils.data_loader")
    RT = importlib.import_module("utils.data_loader_textline")
    pairs = [(RL.load_process, L.load_process), (RL.load_test, L.load_test), (RL.textline_process, L.textline_process),
             (RL.load_input_file, L.load_input_file), (RL.get_file_list, L.get_file_list),
             (RP.pipeline_process, L.pipeline_process), (RP.load_input_file, L.load_tfrecord_input_file),
             (RT.dnn_multi_process, L.dnn_multi_process)]
    for ref, mine in pairs:
        want, got = params(ref), params(mine)
        assert got[:len(want)] == want, (ref.__module__, ref.__name__, got, want)


def test_hyper_parameter_bag_carries_every_launcher_attribute():
    """every attribute the reference launchers' own ModelParams classes set (local_run*.py; parsed, not imported: the
    launchers import modules that do not exist) is a field of deep_learning_b200.hparams.ModelParams"""
    import ast
    import glob
    from deep_learning_b200.hparams import ModelParams
    mp, seen = ModelParams(), set()
    for f in sorted(glob.glob("/root/reference/local_run*.py")):
        for c in ast.walk(ast.parse(open(f).read())):
            if isinstance(c, ast.ClassDef) and c.name == "ModelParams":
                for n in ast.walk(c):
                    if isinstance(n, ast.Attribute) and isinstance(n.value, ast.Name) and n.value.id == "self" \
                            and isinstance(n.ctx, ast.Store):
                        seen.add(n.attr)
    assert len(seen) > 25
    assert not [a for a in sorted(seen) if not hasattr(mp, a)]
