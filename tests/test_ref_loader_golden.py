"""CPU: the oracle's and the product's input decoders against outputs of the reference's OWN plain-Python loaders.

`tests/golden/ref_loader.npz` holds small input files and the pickled batches `/root/reference/utils/data_loader_load.py`
produced from them (get_file_list / read_data / load_process / load_test, run as shipped in the build container by
`tests/golden/make_ref_loader_golden.py`; these functions are Python, the only `tf.*` they touch is local file
opening).  So, unlike the model arithmetic, this part of the path is pinned by the reference itself."""
import ast
import os
import pickle
import types

import numpy as np
import pytest

from oracle import loader_oracle as LO
from deep_learning_b200 import loaders as L

Z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_loader.npz"))
CASES = [str(c) for c in Z["cases"]]

KEYS = {   # what the models feed from each pickle, and the placeholder dtype it is converted to (models/deepfm.py:41-44)
    "": (("labels", np.float32), ("cont_feats", np.float32), ("vector_feats", np.float32), ("cate_feats", np.int32)),
    "tf_test": (("labels", np.float32), ("cont_feats_index", np.int32), ("cont_feats_value", np.float32),
                ("cate_feats_index", np.int32), ("cate_feats_value", np.int32)),
}


def _materialise(case, tmp_path):
    d = str(tmp_path) + "/"
    for k in Z.files:
        if k.startswith(case + "/file/"):
            with open(d + k.split("/file/")[1], "w") as f:
                f.write(str(Z[k]))
    mp = types.SimpleNamespace(**dict(ast.literal_eval(str(Z[case + "/mp"]))))
    ref = [pickle.loads(Z["%s/batch%d" % (case, i)].tobytes()) for i in range(int(Z[case + "/n_batches"]))]
    return d, mp, str(Z[case + "/kind"]), ref


@pytest.mark.parametrize("case", CASES)
def test_oracle_loader_equals_the_reference_output(case, tmp_path, capsys):
    d, mp, kind, ref = _materialise(case, tmp_path)
    files = sorted(d + f for f in os.listdir(d) if f[:4] == "part")
    got = [pickle.loads(b) for b in (LO.load_test if kind == "tf_test" else LO.load_process)(mp, files)]
    capsys.readouterr()
    assert got == ref                      # dict for dict, string for string


@pytest.mark.parametrize("case", CASES)
def test_native_loader_equals_the_reference_output(case, tmp_path, lib_built, capsys):
    d, mp, kind, ref = _materialise(case, tmp_path)
    files = sorted(L.get_file_list(d))
    got = (L.load_test if kind == "tf_test" else L.load_process)(mp, files, pin=False)
    capsys.readouterr()
    assert len(got) == len(ref)
    keys = list(KEYS[kind])
    if mp.alg_name == "wdl_textline":
        keys.append(("wide_feats", np.int32))
    if mp.alg_name == "dnn_multi":
        keys += [("mul_cate_feats", np.int32), ("mul_cate_feats_value", np.float32)]
    for g, r in zip(got, ref):
        for k, dt in keys:
            want = np.asarray(r[k], dtype=dt) if len(r[k]) and len(r[k][0]) else np.zeros((len(r["labels"]), 0), dt)
            assert np.array_equal(g[k].numpy(), want), (case, k)


# ---- tf.data text decoders: the reference's own `decode_libsvm` map functions (utils/data_loader_load.py:187-240,
# utils/data_loader_textline.py:4-36) and dataset chains, executed per line on the stand-in's eager string ops
TL_CASES = [str(c) for c in Z["textline_cases"]]
OWN_NAMES = {"cat_feats": "cate_feats", "mul_cat_feats": "mul_cate_feats", "mul_cat_feats_value": "mul_cate_feats_value"}


def _tl_materialise(case, tmp_path):
    d = str(tmp_path) + "/"
    files = []
    for k in Z.files:
        if k.startswith(case + "/file/"):
            files.append(d + k.split("/file/")[1])
            with open(files[-1], "w") as f:
                f.write(str(Z[k]))
    mp = types.SimpleNamespace(**dict(ast.literal_eval(str(Z[case + "/mp"]))))
    ref = []
    for i in range(int(Z[case + "/n_batches"])):
        pre = "%s/batch%d/" % (case, i)
        ref.append({k[len(pre):]: Z[k] for k in Z.files if k.startswith(pre)})
    return d, sorted(files), mp, str(Z[case + "/module"]), ref


@pytest.mark.parametrize("case", TL_CASES)
def test_oracle_textline_decoder_equals_the_reference_run(case, tmp_path):
    d, files, mp, module, ref = _tl_materialise(case, tmp_path)
    got = list(LO.textline_batches(mp, files, drop_remainder=(module == "load")))
    assert len(got) == len(ref)
    for g, r in zip(got, ref):
        for k, v in r.items():
            if v.ndim == 2:                    # 1-D components are the python 0 placeholders of unused columns (:192-197)
                assert np.array_equal(g[OWN_NAMES.get(k, k)], v), (case, k)


@pytest.mark.parametrize("case", TL_CASES)
def test_native_textline_decoder_equals_the_reference_run(case, tmp_path, lib_built):
    d, files, mp, module, ref = _tl_materialise(case, tmp_path)
    if module == "load":
        got = list(L.textline_process(mp, files, pin=False))
    else:
        got = list(L.dnn_multi_process(files[0], mp.batch_size, mp.epochs, mp=mp, pin=False))
    assert len(got) == len(ref)
    for g, r in zip(got, ref):
        for k, v in r.items():
            if v.ndim == 2:
                assert g[k].numpy().dtype.kind == v.dtype.kind and np.array_equal(g[k].numpy(), v), (case, k)
