"""Runs the reference's plain-Python input decoders (utils/data_loader_load.py: get_file_list, read_data, load_process,
load_test, reached through its own `load_input_file`) on small text files and freezes inputs and outputs into
`tests/golden/ref_loader.npz`.

    python tests/golden/make_ref_loader_golden.py          # build container only: needs /root/reference

These functions are Python, not TensorFlow: the only `tf.*` they touch is `tf.gfile.FastGFile` / `ListDirectory`
(local open / listdir in the stand-in), so the recorded pickles are the reference's own outputs byte for byte.
"""
import os
import sys
import tempfile
import types
import contextlib
import io

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
for _p in (HERE, "/root/reference"):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import tf1_shim as S   # noqa: E402


def fmt(x):
    return repr(float(np.float32(x)))


def lines(n, n_cont, n_cate, third=None, n_third=0, seed=0, awkward=False):
    g = np.random.RandomState(seed)
    out = []
    for _ in range(n):
        toks = [fmt(v) for v in g.rand(n_cont).astype(np.float32) * (1000.0 if awkward else 1.0)]
        if awkward:
            toks[:4] = ["1e-3", "-0.0", "3", "0.1000000014901161193847656250000001"]
        cols = [str(int(g.rand() < 0.3)), " ".join(toks), " ".join(str(v) for v in g.randint(0, 100000, size=n_cate))]
        if third == "wide":
            cols.append(" ".join(str(v) for v in g.randint(0, 5000, size=n_third)))
        if third == "libsvm":
            cols.append(" ".join("%d=%s" % (a, fmt(b)) for a, b in zip(g.randint(0, 900, size=n_third), g.rand(n_third) * 3)))
        out.append(",".join(cols))
    return out


CASES = {
    # name: (files {name: (lines, newline at end)}, mp attributes, read_file_type)
    "deepfm_two_files": ({"part-000": (lines(37, 9, 6, seed=1, awkward=True), True),
                          "part-001": (lines(12, 9, 6, seed=2), False), "_SUCCESS": ([], False)},
                         dict(alg_name="deepfm", batch_size=16, cont_field_size=9, vector_field_size=3, cate_field_size=6,
                              max_data_size=10 ** 9), ""),
    "wdl_textline_wide_max": ({"part-0": (lines(50, 4, 3, third="wide", n_third=5, seed=3), True)},
                              dict(alg_name="wdl_textline", batch_size=16, cont_field_size=4, vector_field_size=0,
                                   cate_field_size=3, wide_field_size=5, max_data_size=20), ""),
    "dnn_multi_libsvm": ({"part-0": (lines(20, 3, 4, third="libsvm", n_third=6, seed=5), True)},
                         dict(alg_name="dnn_multi", batch_size=8, cont_field_size=3, vector_field_size=0, cate_field_size=4,
                              multi_cate_field_size=6, max_data_size=10 ** 9), ""),
    "load_test_form": ({"part-0": (lines(40, 5, 4, seed=4), True)},
                       dict(alg_name="deepfm", batch_size=32, cont_field_size=5, vector_field_size=0, cate_field_size=4,
                            max_data_size=10 ** 9), "tf_test"),
}


def _tl(alg, third, n_third):
    return ({"part-a": (lines(29, 5, 4, third=third, n_third=n_third, seed=5), True),
             "part-b": (lines(14, 5, 4, third=third, n_third=n_third, seed=6), True)},
            dict(alg_name=alg, batch_size=8, epochs=2, cont_field_size=5, vector_field_size=2, cate_field_size=4,
                 wide_field_size=n_third, multi_feats_size=n_third))


# tf.data decoders (utils/data_loader_load.py:187-249 `textline_process`, utils/data_loader_textline.py:4-44): the map
# functions are the reference's, run per line on the stand-in's eager string ops
TEXTLINE_CASES = {
    "textline/wdl_textline": _tl("wdl_textline", "wide", 7) + ("load",),
    "textline/dnn_multi_textline": _tl("dnn_multi_textline", "libsvm", 6) + ("load",),
    "textline/deepfm_textline": _tl("deepfm_textline", None, 0) + ("load",),
    "textline/own_loader_short_batch": ({"part-999": (lines(21, 3, 2, third="libsvm", n_third=4, seed=7), True)},
                                        dict(alg_name="dnn_multi_textline", batch_size=8, epochs=1, cont_field_size=3,
                                             cate_field_size=2, multi_feats_size=4), "textline"),
}


def drain(tf, dd):
    keys = sorted(dd)
    got = []
    with tf.Session() as sess:
        try:
            while True:
                got.append(dict(zip(keys, sess.run([dd[k] for k in keys]))))
        except tf.errors.OutOfRangeError:
            pass
    return got


def main():
    if not os.path.isdir("/root/reference"):
        raise SystemExit("needs /root/reference (build container only)")
    tf = S.install()
    import utils.data_loader_load as RL
    import utils.data_loader_textline as RT
    out = {"cases": np.array(sorted(CASES)), "textline_cases": np.array(sorted(TEXTLINE_CASES))}
    for name, (files, mpd, module) in sorted(TEXTLINE_CASES.items()):
        with tempfile.TemporaryDirectory() as tmp:
            d = tmp + "/"
            for fn, (ls, nl) in files.items():
                txt = "\n".join(ls) + ("\n" if nl and ls else "")
                with open(d + fn, "w") as f:
                    f.write(txt)
                out["%s/file/%s" % (name, fn)] = np.array(txt)
            mp = types.SimpleNamespace(**mpd)
            with contextlib.redirect_stdout(io.StringIO()):
                if module == "load":
                    batches = drain(tf, RL.textline_process(mp, sorted(d + f for f in files)))
                else:
                    batches = drain(tf, RT.load_input_file(mp, d))
        out[name + "/mp"] = np.array(repr(sorted(mpd.items())))
        out[name + "/module"] = np.array(module)
        out[name + "/n_batches"] = np.int64(len(batches))
        for i, b in enumerate(batches):
            for k, v in b.items():
                out["%s/batch%d/%s" % (name, i, k)] = v.astype(np.float32) if v.dtype.kind == "f" else v
        print(name, len(batches), "batches", {k: v.shape for k, v in batches[0].items()})
    for name, (files, mpd, kind) in sorted(CASES.items()):
        with tempfile.TemporaryDirectory() as tmp:
            d = tmp + "/"
            for fn, (ls, nl) in files.items():
                txt = "\n".join(ls) + ("\n" if nl and ls else "")
                with open(d + fn, "w") as f:
                    f.write(txt)
                out["%s/file/%s" % (name, fn)] = np.array(txt)
            mp = types.SimpleNamespace(**mpd)
            with contextlib.redirect_stdout(io.StringIO()):
                listed = RL.get_file_list(d)
                # ListDirectory's order is the file system's; the launcher passes whatever it returns.  Sorted here so the
                # fixture does not depend on the directory hashing of this container.
                batches = (RL.load_test if kind == "tf_test" else RL.load_process)(mp, sorted(listed))
                via_entry = RL.load_input_file(mp, d, kind)
            assert sorted(os.path.basename(f) for f in listed) == sorted(k for k in files if k.startswith("part"))
            if len(listed) == 1:
                assert via_entry == batches          # the reference's own entry point gives the same pickles
        out[name + "/mp"] = np.array(repr(sorted(mpd.items())))
        out[name + "/kind"] = np.array(kind)
        out[name + "/n_batches"] = np.int64(len(batches))
        for i, b in enumerate(batches):
            out["%s/batch%d" % (name, i)] = np.frombuffer(b, dtype=np.uint8)
        print(name, len(batches), "batches")
    path = os.path.join(HERE, "ref_loader.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
