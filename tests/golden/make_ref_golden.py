"""Runs the reference's OWN model files (/root/reference/models/*.py, unmodified, imported where they lie) under the
TensorFlow-1.x stand-in of `tf1_shim.py` and freezes what they compute into `tests/golden/ref_<model>.npz`.

    python tests/golden/make_ref_golden.py [model ...]          # build container only: needs /root/reference

What this pins and what it does not: the graph wiring (index tensors, offsets, concatenation order, which variables
are regularised, what `fit` fetches and in which order it consumes batches) is the reference's code executing; the
arithmetic of each `tf.*` symbol is the stand-in's torch-float64 restatement of TF's documented semantics, not
TensorFlow itself.  `tests/test_ref_golden.py` then requires the oracle to reproduce these files.

Per model: the variables right after `global_variables_initializer` (under the oracle's parameter names, with the
reference's variable name recorded beside each), the batches `fit` consumed, the loss of every train step, the
evaluation outputs `fit` fetched, and every variable after `fit` returned.
"""
import contextlib
import importlib.util
import io
import os
import pickle
import sys
import tempfile
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for _p in (ROOT, os.path.join(ROOT, "tests"), HERE, "/root/reference"):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import tf1_shim as S   # noqa: E402
import make_golden as G   # noqa: E402
from oracle import ctr_oracle as O   # noqa: E402

REF = "/root/reference"
PREDICT_MARK = "#### predict ####"
N_TRAIN = 3

FEED = ("deepfm", "dnn", "wdl", "wdl_ftrl", "dcn")                     # fit(train_data, val_data): pickled batches
MODELS_FEED = FEED[:3]                                                 # the ones under models/ (the other two: test_model/)
PIPE = ("deepfm_multi", "deepfm_cate", "deepfm_multi_cate", "deepfm_pipeline", "dnn_multi", "dnn_cate",
        "dnn_multi_cate", "dnn_pipeline")                              # fit(print_num_batch, predict_data): tf.data
TEXT = ("wdl_textline", "dnn_multi_textline")                          # fit(print_iter, num_batch_size): tf.data


def ref_args(alg, hp, tmp):
    a = types.SimpleNamespace(
        hidden_units=list(hp.hidden_units), epochs=1, batch_size=hp.batch_size, learning_rate=hp.learning_rate,
        model_pb=os.path.join(tmp, "pb"), learning_rate_decay_steps=hp.learning_rate_decay_steps,
        learning_rate_decay_rate=hp.learning_rate_decay_rate, l2_reg=hp.l2_reg, embedding_size=hp.embedding_size,
        cont_field_size=hp.cont_field_size, vector_feats_size=hp.vector_feats_size, cate_field_size=hp.cate_field_size,
        cate_feats_size=hp.cate_feats_size, cate_index_size=hp.cate_index_size, wide_field_size=hp.wide_field_size,
        multi_feats_size=hp.multi_feats_size, multi_field_size=hp.multi_field_size,
        multi_feats_range=[list(r) for r in hp.multi_feats_range], save_model_checkpoint=os.path.join(tmp, "ckpt"),
        restore_model_checkpoint=os.path.join(tmp, "ckpt"), model_restore=0, cate_feats_conf_path="",
        cat_field_size=hp.cate_field_size, cat_index_size=hp.cate_feats_size, cat_feat_conf_path="",
        word2vec_embedding_path="", deep_feature_size=hp.deep_feature_size)
    return a


def load_reference_model(alg):
    path = os.path.join(REF, O.TEST_MODEL_FILES.get(alg, os.path.join("models", alg + ".py")))
    spec = importlib.util.spec_from_file_location("reference_models_" + alg, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def np_batch(b):
    return {k: v.numpy() for k, v in b.items()}


def snapshot():
    return [(v.name, v.numpy()) for v in S.global_variables() if v.dtype.is_float]


def match_params(alg, hp, snap):
    """reference variables (creation order) -> oracle parameter names, by shape; same-shape variables keep their
    creation order, which is also the order of `param_shapes`"""
    shapes = O.param_shapes(alg, hp)
    left = list(snap)
    out, names = {}, {}
    for pname, shp in shapes.items():
        if pname == "tags_emb":
            continue
        hit = [i for i, (vn, val) in enumerate(left) if tuple(val.shape) == tuple(shp)]
        if not hit:
            raise AssertionError("%s: no reference variable of shape %s for %s (have %s)" % (
                alg, shp, pname, [(n, v.shape) for n, v in left]))
        vn, val = left.pop(hit[0])
        out[pname], names[pname] = val, vn
    if left:
        raise AssertionError("%s: reference variables without an oracle parameter: %s" % (
            alg, [(n, v.shape) for n, v in left]))
    return out, names


class Recorder:
    """wraps the stand-in's Session.run: remembers what each call fetched and returned"""

    def __init__(self):
        self.calls = []
        self._orig = S.Session.run
        self.init_snap = None

    def __enter__(self):
        rec = self

        def run(sess, fetches, feed_dict=None):
            out = rec._orig(sess, fetches, feed_dict)
            fl = fetches if isinstance(fetches, (list, tuple)) else [fetches]
            rec.calls.append(([getattr(f, "name", f) for f in fl], out, feed_dict))
            if rec.init_snap is None and any(str(getattr(f, "name", "")).startswith(("init", "group")) for f in fl):
                rec.init_snap = snapshot()
            return out

        S.Session.run = run
        return self

    def __exit__(self, *exc):
        S.Session.run = self._orig
        return False


def run_feed_model(alg, hp, mod, args, batches, val):
    """models/deepfm.py:164-214, models/dnn.py:98-146, models/wdl.py:287-341"""
    def pack(b):
        if alg == "dcn":                      # test_model/DCN.py:131-135
            return pickle.dumps({"features": b["features"], "labels": b["label"]})
        d = {"cont_feats": b["cont_feats"], "cate_feats": b["cate_feats"], "labels": b["label"]}
        if alg == "deepfm":
            d["vector_feats"] = b["vector_feats"]
        if alg == "dnn":                      # models/dnn.py:28: one float block holds [cont | vector]
            d["cont_feats"] = np.concatenate([b["cont_feats"], b["vector_feats"]], axis=1)
        if alg in ("wdl", "wdl_ftrl"):
            d["wide_feats"] = b["wide_feats"]
        return pickle.dumps(d)

    model = mod.DeepModel(args)
    with Recorder() as rec:
        model.fit([pack(b) for b in batches], [pack(b) for b in val])
    losses, evals = classify(rec)
    print(PREDICT_MARK)
    model.predict([pack(b) for b in val])          # models/deepfm.py:233-255: reload the exported model, score, print the AUC
    return rec.init_snap, losses, evals


def generate(alg):
    out = run_reference(alg, G.tiny_hp(alg))
    path = os.path.join(HERE, "ref_%s.npz" % alg)
    np.savez_compressed(path, **out)
    print("%-20s %6d bytes  steps=%d  losses=%s" % (alg, os.path.getsize(path), len(out["losses"]),
                                                     np.array2string(out["losses"], precision=6)))


def run_reference(alg, hp):
    """the reference's model file for `alg`, built with `hp` and run through its own `fit`: everything recorded"""
    S.reset_default_graph()
    np.random.seed(2019)                      # the reference draws the MLP weights from numpy's global generator
    mod = load_reference_model(alg)
    batches = [np_batch(G.tiny_batch(alg, hp, seed=21 + i)) for i in range(N_TRAIN)]
    val = [np_batch(G.tiny_batch(alg, hp, seed=31 + i)) for i in range(2)]
    log = io.StringIO()
    with tempfile.TemporaryDirectory() as tmp, contextlib.redirect_stdout(log):
        args = ref_args(alg, hp, tmp)
        if alg in FEED:
            args.epochs = 2                   # the running loss of generation A is never reset between epochs (:168-169)
            init, losses, evals = run_feed_model(alg, hp, mod, args, batches, val)
        elif alg in PIPE:
            init, losses, evals, batches, val = run_pipe_model(alg, hp, mod, args, batches, val)
        else:
            init, losses, evals, batches, val = run_text_model(alg, hp, mod, args, batches, val)
    p0, names = match_params(alg, hp, init)
    p1, _ = match_params(alg, hp, snapshot())
    fit_log, _, predict_log = log.getvalue().partition(PREDICT_MARK)
    out = {"n_train": np.int64(len(losses)), "n_val": np.int64(len(val)), "n_eval": np.int64(len(evals)),
           "fit_stdout": np.array("\n".join(l for l in fit_log.splitlines() if "auc" in l)),       # sklearn's AUC as `fit` printed it
           "predict_stdout": np.array("\n".join(l for l in predict_log.splitlines() if "auc" in l))}
    for k, v in getattr(run_text_model, "extra", {}).items():
        out["init/" + k] = out["final/" + k] = v
    run_text_model.extra = {}
    for k in p0:
        out["init/" + k], out["final/" + k], out["refname/" + k] = p0[k], p1[k], np.array(names[k])
    for i, b in enumerate(batches):
        for k, v in b.items():
            out["train%d/%s" % (i, k)] = v
    for i, b in enumerate(val):
        for k, v in b.items():
            out["val%d/%s" % (i, k)] = v
    out["losses"] = np.asarray(losses, dtype=np.float64)
    out["epochs"] = np.int64(2 if alg in FEED else 1)
    # the batches the reference's predict() scored after fit: the validation set, which the textline loop holds as the
    # two batches of the stream behind the train batches
    out["predict_on"] = np.array(["train%d" % (N_TRAIN + i) for i in range(2)] if alg in TEXT else ["val0", "val1"])
    out["events"] = np.array(classify.events)          # order of train steps (T) and evaluations (E) as `fit` issued them
    out["family"] = np.array("feed" if alg in FEED else "pipe" if alg in PIPE else "text")
    for i, e in enumerate(evals):
        out["eval%d" % i] = np.asarray(e, dtype=np.float64)
    gs = [v for v in S.global_variables() if not v.dtype.is_float and v.value.dim() == 0]
    out["global_step"] = np.int64(int(gs[0].value)) if gs else np.int64(-1)
    return out


def classify(rec):
    """train steps fetch the Adam op; evaluations are every other fetch of model outputs"""
    losses, evals = [], []
    rec.events = ""
    for names, out, feed in rec.calls:
        names = [str(n) for n in names]
        if any(n.startswith(("Adam", "Ftrl")) for n in names):
            losses.append(out[0])
            rec.events += "T"
        elif any(n.startswith(("init", "group", "IteratorGetNext")) for n in names):
            continue
        else:
            evals.append(out)
            rec.events += "E"
    classify.events = rec.events
    return losses, evals


def run_pipe_model(alg, hp, mod, args, batches, val):
    """fit(print_num_batch, predict_data) of the tf.data-fed files (models/deepfm_multi.py:262-335 and siblings):
    `data_dict` / `predict_data` are the component tensors of two one-shot iterators"""
    keys = {"label": (S.float32, [None, 1]), "cate_feats": (S.int64, [None, batches[0]["cate_feats"].shape[1]]),
            "vector_feats": (S.float32, [None, hp.vector_feats_size])}
    if alg not in O.NO_CONT:
        keys["cont_feats"] = (S.float32, [None, hp.cont_field_size])
    train_src, val_src = S.BatchSource(batches, keys), S.BatchSource(val, keys)
    model = mod.DeepModel(args, train_src.tensors)
    with Recorder() as rec:
        model.fit(1, val_src.tensors)                 # evaluate after every train step from the second on
    losses, evals = classify(rec)
    print(PREDICT_MARK)
    mod.predict(S.BatchSource(val, keys).tensors, args.model_pb)          # models/deepfm_multi.py:383-418
    return rec.init_snap, losses, evals, batches, val


def run_text_model(alg, hp, mod, args, batches, val):
    """fit(print_iter, num_batch_size) of models/wdl_textline.py:302-360 and models/dnn_multi_textline.py:173-231: every
    iteration pulls one batch for [label, out, loss] and, unless it printed, a SECOND batch for the train step"""
    allb = batches + val + [np_batch(G.tiny_batch(alg, hp, seed=41 + i)) for i in range(4)]
    if alg == "wdl_textline":
        keys = {"labels": (S.float32, [None, 1]), "cont_feats": (S.float32, [None, hp.cont_field_size]),
                "cate_feats": (S.int64, [None, hp.cate_field_size]), "wide_feats": (S.int64, [None, hp.wide_field_size])}
        ren = {"labels": "label", "cont_feats": "cont_feats", "cate_feats": "cate_feats", "wide_feats": "wide_feats"}
        feed = [{k: b[v] for k, v in ren.items()} for b in allb]
    else:
        Fs = hp.cate_field_size
        keys = {"labels": (S.float32, [None, 1]), "cont_feats": (S.float32, [None, hp.cont_field_size]),
                "cat_feats": (S.int64, [None, Fs]), "mul_cat_feats": (S.int64, [None, hp.multi_feats_size]),
                "mul_cat_feats_value": (S.float32, [None, hp.multi_feats_size])}
        feed = [{"labels": b["label"], "cont_feats": b["cont_feats"], "cat_feats": b["cate_feats"][:, :Fs],
                 "mul_cat_feats": b["cate_feats"][:, Fs:], "mul_cat_feats_value": b["multi_feats_value"]} for b in allb]
        # models/dnn_multi_textline.py:34 calls utils.my_utils.get_mul_index_range, which utils/my_utils.py does not
        # define (the parser exists only as a method, models/wdl_textline.py:82-110): supply its result
        import utils.my_utils as MU
        ranges = [[int(r[0]), int(r[1]), r[2]] for r in hp.multi_feats_range]
        MU.get_mul_index_range = lambda path: (ranges, ranges[-1][1])
        tags = O.init_params(alg, hp, seed=2019)["tags_emb"].numpy().astype(np.float32)
        args.word2vec_embedding_path = os.path.join(os.path.dirname(args.model_pb), "w2v.npy")
        np.save(args.word2vec_embedding_path, tags)
    src = S.BatchSource(feed, keys)
    model = mod.DeepModel(args, src.tensors)
    with Recorder() as rec:
        model.fit(2, 1)
    losses, evals = classify(rec)
    print(PREDICT_MARK)
    if alg == "wdl_textline":                      # models/wdl_textline.py:362-386: pickled batches
        mod.predict([pickle.dumps(f) for f in feed[len(batches):len(batches) + len(val)]], args.model_pb)
    else:                                          # models/dnn_multi_textline.py:233-256: plain dicts
        mod.predict(feed[len(batches):len(batches) + len(val)], args.model_pb)
    extra = {"tags_emb": tags} if alg == "dnn_multi_textline" else {}
    run_text_model.extra = extra
    # the stream as consumed: which batches were evaluated, which were trained on, is the reference loop's business
    return rec.init_snap, losses, [np.concatenate([np.asarray(e[1]).ravel(), [e[2]]]) for e in evals], allb[:src.n_pulled], []


def generate_fp32_floor(alg):
    """The SAME run with the stand-in computing in float32 (TF's arithmetic type), from the same initial variables and
    batches: how far float32 arithmetic itself lands from the float64 fixture.  tests/golden/fp32floor_<model>.npz holds
    only the results (losses, evaluations, final variables); tests/test_ref_gpu.py reports the CUDA path's distance to
    the float64 fixture next to this floor."""
    assert S.F is torch.float32, "run with TF1_SHIM_FLOAT=float32"
    out = run_reference(alg, G.tiny_hp(alg))
    z = np.load(os.path.join(HERE, "ref_%s.npz" % alg))
    for k in z.files:                          # same run: identical initial variables (after float32 rounding) and batches
        if k.startswith(("init/", "train", "val")) and z[k].dtype.kind in "fi":
            assert np.array_equal(np.asarray(out[k], dtype=np.float32), z[k].astype(np.float32)), (alg, k)
    assert str(out["events"]) == str(z["events"])
    keep = {k: np.asarray(v, dtype=np.float64) for k, v in out.items()
            if k.startswith(("final/", "eval")) or k == "losses"}
    path = os.path.join(HERE, "fp32floor_%s.npz" % alg)
    np.savez_compressed(path, **keep)
    worst = max(float(np.max(np.abs(keep[k] - z[k])) / max(float(np.max(np.abs(z[k]))), 1e-30))
                for k in keep if k.startswith("final/"))
    print("%-20s %6d bytes  worst final-variable deviation of float32 from the float64 fixture: %.2e" % (
        alg, os.path.getsize(path), worst))


def main():
    if not os.path.isdir(REF):
        raise SystemExit("needs /root/reference (build container only); the committed ref_*.npz are its outputs")
    S.install()
    argv = [a for a in sys.argv[1:] if a != "--fp32-floor"]
    if "--fp32-floor" in sys.argv[1:]:         # TF1_SHIM_FLOAT=float32 python tests/golden/make_ref_golden.py --fp32-floor
        for alg in (argv or FEED + PIPE + TEXT):
            generate_fp32_floor(alg)
        return
    for alg in (argv or FEED + PIPE + TEXT):
        generate(alg)


if __name__ == "__main__":
    main()
