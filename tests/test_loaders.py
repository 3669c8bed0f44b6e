"""CPU: the native input decoders (csrc/host_loader.cu through deep_learning_b200/loaders.py) against the oracle
(oracle/loader_oracle.py: the reference's plain-Python load_process / load_test run as written, and restatements of
its tf.data decoders).  Ids, labels and shapes bit-exact; floats bit-exact (Python float() then a float32 cast)."""
import ctypes as C
import os
import pickle

import numpy as np
import pytest
import torch

from oracle import loader_oracle as LO
from deep_learning_b200 import loaders as L
from deep_learning_b200 import data as D
from deep_learning_b200.hparams import ModelParams


def _fmt(x):
    return repr(float(np.float32(x)))


def _write_lines(path, n, C_, Fc, third=None, n_third=0, seed=0, newline_at_end=True, awkward=False):
    g = np.random.RandomState(seed)
    lines = []
    for i in range(n):
        cont = g.rand(C_).astype(np.float32) * (1000.0 if awkward else 1.0)
        toks = [_fmt(v) for v in cont]
        if awkward and C_ >= 4:
            toks[0], toks[1], toks[2], toks[3] = "1e-3", "-0.0", "3", "0.1000000014901161193847656250000001"
        cate = g.randint(0, 100000, size=Fc)
        cols = [str(int(g.rand() < 0.3)), " ".join(toks), " ".join(str(v) for v in cate)]
        if third == "wide":
            cols.append(" ".join(str(v) for v in g.randint(0, 5000, size=n_third)))
        if third == "libsvm":
            cols.append(" ".join("%d=%s" % (a, _fmt(b)) for a, b in zip(g.randint(0, 900, size=n_third), g.rand(n_third) * 3)))
        lines.append(",".join(cols))
    txt = "\n".join(lines) + ("\n" if newline_at_end else "")
    with open(path, "w") as f:
        f.write(txt)


def _mp(**kw):
    mp = ModelParams(**kw)
    return mp


@pytest.fixture(scope="module")
def lib_ready(lib_built):
    return lib_built


def test_load_process_equals_the_reference_python(tmp_path, lib_ready):
    """utils/data_loader_load.py:61-139 run as written vs the native parser: two files, a short last batch, a last line
    without a newline, awkward float spellings, the vector block split off the tail of column 1"""
    d = str(tmp_path) + "/"
    _write_lines(d + "part-000", 700, 9, 6, seed=1, awkward=True)
    _write_lines(d + "part-001", 333, 9, 6, seed=2, newline_at_end=False)
    open(d + "_SUCCESS", "w").close()                       # not a "part" file: ignored by get_file_list
    mp = _mp(alg_name="deepfm", batch_size=256, cont_field_size=9, vector_field_size=3, cate_field_size=6)
    files = L.get_file_list(d)
    assert [os.path.basename(f) for f in files] == ["part-000", "part-001"]
    ref = [pickle.loads(b) for b in LO.load_process(mp, files)]
    got = L.load_process(mp, files, pin=False)
    assert len(got) == len(ref) == 5 and got[-1]["labels"].shape[0] == 1033 - 4 * 256
    for g_, r in zip(got, ref):
        assert np.array_equal(g_["cate_feats"].numpy(), LO.feed(r["cate_feats"], np.int32))
        assert np.array_equal(g_["cont_feats"].numpy(), LO.feed(r["cont_feats"], np.float32))
        assert np.array_equal(g_["vector_feats"].numpy(), LO.feed(r["vector_feats"], np.float32))
        assert np.array_equal(g_["labels"].numpy(), LO.feed(r["labels"], np.float32))
        assert g_["cont_feats"].shape[1] == 6 and g_["vector_feats"].shape[1] == 3
    # the reference wire format on request: decodes to the same arrays through the generation-A decoder
    pk = L.load_process(mp, files, as_pickle=True, pin=False)
    back = D.from_pickled_batch(pk[0])
    assert np.array_equal(back["cate_feats"], got[0]["cate_feats"].numpy())
    assert np.array_equal(back["cont_feats"], got[0]["cont_feats"].numpy())


def test_load_process_max_data_size_and_wide_column(tmp_path, lib_ready):
    d = str(tmp_path) + "/"
    _write_lines(d + "part-0", 50, 4, 3, third="wide", n_third=5, seed=3)
    mp = _mp(alg_name="wdl_textline", batch_size=16, cont_field_size=4, cate_field_size=3, wide_field_size=5)
    mp.max_data_size = 20
    ref = [pickle.loads(b) for b in LO.load_process(mp, [d + "part-0"])]
    got = L.load_process(mp, [d + "part-0"], pin=False)
    assert sum(len(r["labels"]) for r in ref) == 21 == sum(g_["labels"].shape[0] for g_ in got)   # :45 keeps max+1 lines
    for g_, r in zip(got, ref):
        assert np.array_equal(g_["wide_feats"].numpy(), LO.feed(r["wide_feats"], np.int32))
        assert np.array_equal(g_["cate_feats"].numpy(), LO.feed(r["cate_feats"], np.int32))


def test_load_test_index_value_form(tmp_path, lib_ready):
    d = str(tmp_path) + "/"
    _write_lines(d + "part-0", 40, 5, 4, seed=4)
    mp = _mp(alg_name="deepfm", batch_size=32, cont_field_size=5, vector_field_size=0, cate_field_size=4)
    ref = [pickle.loads(b) for b in LO.load_test(mp, [d + "part-0"])]
    got = L.load_test(mp, [d + "part-0"], pin=False)
    assert len(ref) == len(got) == 2
    for g_, r in zip(got, ref):
        for k, dt in (("labels", np.float32), ("cont_feats_index", np.int32), ("cont_feats_value", np.float32),
                      ("cate_feats_index", np.int32), ("cate_feats_value", np.int32)):
            assert np.array_equal(g_[k].numpy(), LO.feed(r[k], dt)), k


@pytest.mark.parametrize("alg,third,n_third", [("wdl_textline", "wide", 7), ("dnn_multi_textline", "libsvm", 6),
                                               ("deepfm_textline", None, 0)])
def test_textline_iterator(tmp_path, lib_ready, alg, third, n_third):
    """utils/data_loader_load.py:187-252: drop_remainder=True, repeat(epochs), batches run across file boundaries"""
    d = str(tmp_path) + "/"
    _write_lines(d + "part-a", 130, 5, 4, third=third, n_third=n_third, seed=5)
    _write_lines(d + "part-b", 75, 5, 4, third=third, n_third=n_third, seed=6)
    mp = _mp(alg_name=alg, batch_size=64, epochs=2, cont_field_size=5, vector_field_size=2, cate_field_size=4,
             wide_field_size=n_third, multi_feats_size=n_third)
    files = [d + "part-a", d + "part-b"]
    ref = list(LO.textline_batches(mp, files))
    got = list(L.textline_process(mp, files, pin=False))
    assert len(ref) == len(got) == 2 * (205 // 64)
    for g_, r in zip(got, ref):
        for k in r:
            if r[k].size or k in g_:
                assert np.array_equal(g_[k].numpy(), r[k]), k
    # small chunks: lines straddling chunk boundaries are carried over, results identical
    sc = L.schema_for(mp, "textline")
    blocks = L.LineBlocks(files, sc, chunk_bytes=997)
    small = list(L._rebatch(blocks, 64, True, sc.names, False))
    for a, b in zip(small, got[:3]):
        for k in a:
            assert torch.equal(a[k], b[k]), k


def test_dnn_multi_textline_loader_keeps_short_batch(tmp_path, lib_ready):
    """utils/data_loader_textline.py:39-44: batch() without drop_remainder; its own key names"""
    d = str(tmp_path) + "/"
    _write_lines(d + "part-999", 21, 3, 2, third="libsvm", n_third=4, seed=7)
    mp = _mp(alg_name="dnn_multi_textline", batch_size=8, epochs=1, cont_field_size=3, cate_field_size=2, multi_feats_size=4)
    got = list(L.dnn_multi_process(d + "part-999", 8, 1, mp=mp, pin=False))
    assert [b["labels"].shape[0] for b in got] == [8, 8, 5]
    assert set(got[0]) >= {"labels", "cont_feats", "cat_feats", "mul_cat_feats", "mul_cat_feats_value"}
    mp2 = _mp(alg_name="dnn_multi_textline", batch_size=8, epochs=1, cont_field_size=3, cate_field_size=2, multi_feats_size=4)
    ref = list(LO.textline_batches(mp2, [d + "part-999"], drop_remainder=False))
    for g_, r in zip(got, ref):
        assert np.array_equal(g_["mul_cat_feats"].numpy(), r["mul_cate_feats"])
        assert np.array_equal(g_["mul_cat_feats_value"].numpy(), r["mul_cate_feats_value"])
        assert np.array_equal(g_["cat_feats"].numpy(), r["cate_feats"])


def test_malformed_lines_are_reported(tmp_path, lib_ready):
    d = str(tmp_path) + "/"
    mp = _mp(alg_name="deepfm", batch_size=4, cont_field_size=2, vector_field_size=0, cate_field_size=2)
    for bad in ("1,0.5 0.25,3\n", "1,0.5,3 4\n", "1,0.5 x,3 4\n", "1,0.5 0.25,3 4.5\n", "1,0.5  0.25,3 4\n", "1\n"):
        with open(d + "part-0", "w") as f:
            f.write("0,0.5 0.25,3 4\n" + bad)
        with pytest.raises(ValueError, match="malformed line 1"):
            L.load_process(mp, [d + "part-0"], pin=False)
    with open(d + "part-0", "w") as f:
        f.write("0,0.5 0.25,3 4\n\n \n1, 7 1e2 ,3 4\r\n")          # blank lines skipped, CRLF tolerated, float() strips blanks
    try:
        got = L.load_process(mp, [d + "part-0"], pin=False)
    except ValueError:
        got = None
    # " 7" after the comma is the token "" then "7" for str.split(" "): float("") raises in the reference too
    assert got is None


def test_crc32c_known_answer_and_tfrecord_roundtrip(tmp_path, lib_ready):
    """RFC 3720 check value pins the CRC; Examples written by the oracle's writer (packed and unpacked list encodings,
    an extra key the spec does not name) are decoded by the native reader"""
    assert LO.crc32c(b"123456789") == 0xE3069283
    g = np.random.RandomState(8)
    n, C_, V, Fc, M = 150, 4, 3, 5, 6
    rows = [{"label": g.rand(1).round(), "cont_feats": g.rand(C_), "vector_feats": g.randn(V),
             "cate_feats": g.randint(0, 2 ** 40, size=Fc + M), "ignored": g.rand(2)} for _ in range(n)]
    rows[3]["cate_feats"][0] = -7                                   # negative int64: 10-byte varint
    d = str(tmp_path) + "/"
    LO.write_tfrecord(d + "part-r-00000", [LO.encode_example(r, packed=True) for r in rows[:80]])
    LO.write_tfrecord(d + "part-r-00001", [LO.encode_example(r, packed=False) for r in rows[80:]])
    mp = _mp(alg_name="deepfm_multi", batch_size=32, epochs=2, cont_field_size=C_, vector_feats_size=V, cate_field_size=Fc,
             multi_feats_size=M)
    files = L.get_file_list(d)
    ref = list(LO.pipeline_batches(mp, files, "train"))
    got = list(L.pipeline_process(mp, files, "train", pin=False, shuffle_seed=None))       # file order: comparable record by record
    assert len(ref) == len(got) == 2 * (n // 32)
    for g_, r in zip(got, ref):
        for k in ("label", "cont_feats", "vector_feats", "cate_feats"):
            assert np.array_equal(g_[k].numpy(), r[k]), k
        assert g_["cate_feats"].dtype == torch.int64 and g_["label"].shape == (32, 1)
    assert len(list(L.pipeline_process(mp, files, "predict", pin=False))) == n // 32       # no repeat outside training
    # tiny chunks: records straddle chunk boundaries
    small = list(L.TFRecordBatches(mp, files, "predict", pin=False, chunk_bytes=301))
    for a, b in zip(small, got):
        assert torch.equal(a["cate_feats"], b["cate_feats"]) and torch.equal(a["cont_feats"], b["cont_feats"])
    # *_cate generation: no cont_feats key in the spec (utils/data_loader.py:10-16)
    mpc = _mp(alg_name="deepfm_multi_cate", batch_size=32, epochs=1, cont_field_size=0, vector_feats_size=V, cate_field_size=Fc,
              multi_feats_size=M)
    gc = list(L.pipeline_process(mpc, files, "train", pin=False, shuffle_seed=None))
    assert gc[0]["cont_feats"].shape == (32, 0) and torch.equal(gc[0]["cate_feats"], got[0]["cate_feats"])
    # seeded shuffle keeps the multiset of records
    sh = list(L.pipeline_process(mp, files, "predict", pin=False, shuffle_seed=1))
    a = torch.cat([b["cate_feats"] for b in sh]); b_ = torch.cat([b["cate_feats"] for b in got[:n // 32]])
    assert not torch.equal(a, b_)
    # the DEFAULT for training is that shuffle (the reference never trains in file order, utils/data_loader.py:30): same
    # records, another order, and the second epoch differs from the first; scoring keeps file order.  (Shuffle comes BEFORE
    # batch(drop_remainder) there, so WHICH records fall into the dropped remainder changes with the order: an epoch is a
    # duplicate-free selection of (n // 32) * 32 of the file's records.)
    tr = list(L.pipeline_process(mp, files, "train", pin=False))
    e0 = torch.cat([b["cate_feats"] for b in tr[:n // 32]]); e1 = torch.cat([b["cate_feats"] for b in tr[n // 32:]])
    every = set(tuple(int(x) for x in r["cate_feats"]) for r in rows)
    for e in (e0, e1):
        recs = [tuple(r) for r in e.tolist()]
        assert len(recs) == (n // 32) * 32 == len(set(recs)) and set(recs) <= every
    assert not torch.equal(e0, b_) and not torch.equal(e0, e1)
    again = list(L.pipeline_process(mp, files, "train", pin=False))
    assert all(torch.equal(x["cate_feats"], y["cate_feats"]) for x, y in zip(tr, again))           # seeded: reproducible
    pr = list(L.pipeline_process(mp, files, "predict", pin=False))
    assert torch.equal(torch.cat([b["cate_feats"] for b in pr]), b_)


def test_tfrecord_corruption_is_detected(tmp_path, lib_ready):
    d = str(tmp_path) + "/"
    row = {"label": np.ones(1), "cont_feats": np.zeros(2), "vector_feats": np.zeros(1), "cate_feats": np.arange(3)}
    LO.write_tfrecord(d + "part-0", [LO.encode_example(row)] * 4)
    mp = _mp(alg_name="deepfm_multi", batch_size=2, epochs=1, cont_field_size=2, vector_feats_size=1, cate_field_size=3)
    assert len(list(L.pipeline_process(mp, [d + "part-0"], "train", pin=False))) == 2
    blob = bytearray(open(d + "part-0", "rb").read())
    blob[30] ^= 0x40
    open(d + "part-1", "wb").write(bytes(blob))
    with pytest.raises(ValueError, match="corrupt TFRecord"):
        list(L.pipeline_process(mp, [d + "part-1"], "train", pin=False))
    open(d + "part-2", "wb").write(bytes(open(d + "part-0", "rb").read()[:-5]))
    with pytest.raises(ValueError, match="truncated"):
        list(L.pipeline_process(mp, [d + "part-2"], "train", pin=False))
    mp_bad = _mp(alg_name="deepfm_multi", batch_size=2, epochs=1, cont_field_size=3, vector_feats_size=1, cate_field_size=3)
    with pytest.raises(ValueError):                                  # FixedLenFeature([3]) against 2 stored values
        list(L.pipeline_process(mp_bad, [d + "part-0"], "train", pin=False))


def _parse_one(line, n_cont, n_cate):
    sc = L._LineSchema(n_cont, 0, n_cate)
    blk, _ = L.LineBlocks([], sc, n_threads=1)._parse((line + "\n").encode())
    return blk


def test_number_spellings_follow_python(lib_ready):
    """the decoder accepts exactly the spellings the reference's conversion (Python float()/int() on str.split tokens, then
    a float32 / int32 cast) turns into finite numbers, with the same value bit for bit: exponents, signs, leading zeros,
    long mantissas (strtod path), denormals, values that overflow float32 to inf"""
    from hypothesis import given, settings, strategies as st

    spell = st.one_of(
        st.floats(allow_nan=False, allow_infinity=False, width=32).map(lambda x: repr(float(x))),
        st.floats(allow_nan=False, allow_infinity=False).map(lambda x: "%.17g" % x),
        st.floats(min_value=-1e6, max_value=1e6, allow_nan=False).map(lambda x: "%.4f" % x),
        st.floats(allow_nan=False, allow_infinity=False, width=32).map(lambda x: "%e" % x),
        st.integers(min_value=-10 ** 12, max_value=10 ** 12).map(str),
        st.sampled_from(["0", "-0", "+3", "007", "1e5", "1E-3", ".5", "5.", "1e39", "-1e-46", "0.1000000000000000055511151231257827",
                         "123456789012345678901234567890", "4.9406564584124654e-324"]))

    @settings(max_examples=300, deadline=None)
    @given(st.lists(spell, min_size=3, max_size=3), st.lists(st.integers(min_value=0, max_value=2 ** 31 - 1), min_size=2, max_size=2))
    def check(toks, ids):
        line = "1," + " ".join(toks) + "," + " ".join(str(i) for i in ids)
        blk = _parse_one(line, 3, 2)
        with np.errstate(over="ignore"):
            want = np.asarray([float(t) for t in toks], dtype=np.float64).astype(np.float32)
        assert np.array_equal(blk["cont"].numpy().reshape(-1).view(np.uint32), want.view(np.uint32)), toks
        assert blk["cate"].numpy().reshape(-1).tolist() == ids
    check()


@pytest.mark.parametrize("tok", ["", " ", "1e", "--1", "1.2.3", "0x10", "1,5", "abc", "1 2"])
def test_tokens_python_rejects_are_rejected(lib_ready, tok):
    try:
        float(tok)
        ok_py = True
    except ValueError:
        ok_py = False
    assert not ok_py or tok in ("1 2",)
    with pytest.raises(ValueError):
        _parse_one("1,0.5 %s 0.25,3 4" % tok, 3, 2)


def test_background_prefetch_keeps_order_and_surfaces_errors(tmp_path, lib_ready):
    d = str(tmp_path) + "/"
    _write_lines(d + "part-a", 300, 4, 3, seed=11)
    mp = _mp(alg_name="deepfm_textline", batch_size=32, epochs=2, cont_field_size=4, vector_field_size=0, cate_field_size=3)
    plain = list(L.textline_process(mp, [d + "part-a"], pin=False, prefetch=0))
    ahead = L.textline_process(mp, [d + "part-a"], pin=False, prefetch=3)
    for run in range(2):                                   # re-iterable
        got = list(ahead)
        assert len(got) == len(plain) == 2 * (300 // 32)
        for a, b in zip(got, plain):
            assert torch.equal(a["cate_feats"], b["cate_feats"]) and torch.equal(a["cont_feats"], b["cont_feats"])
    it = iter(ahead)                                        # abandoning the iterator must not hang the producer
    next(it)
    del it
    with open(d + "part-b", "w") as f:
        f.write("1,0.5 0.5 0.5 0.5,1 2 3\n" * 40 + "1,0.5 oops 0.5 0.5,1 2 3\n")
    bad = L.textline_process(mp, [d + "part-b"], pin=False, prefetch=2)
    with pytest.raises(ValueError, match="malformed"):
        list(bad)
    # the threaded decoder and the single-threaded one agree
    sc = L.schema_for(mp, "textline")
    buf = open(d + "part-a", "rb").read() * 20
    a, _ = L.LineBlocks([], sc, n_threads=1)._parse(buf)
    b, _ = L.LineBlocks([], sc, n_threads=7)._parse(buf)
    for k in a:
        assert torch.equal(a[k], b[k]), k


def _official_example_class():
    """tensorflow.Example built from its published schema (tensorflow/core/example/{example,feature}.proto) on the
    protobuf runtime that IS installed here: an independent implementation of the wire format"""
    from google.protobuf import descriptor_pb2, descriptor_pool, message_factory
    T = descriptor_pb2.FieldDescriptorProto
    f = descriptor_pb2.FileDescriptorProto(name="ctr_test_example.proto", package="ctr_test", syntax="proto3")

    def msg(name):
        m = f.message_type.add()
        m.name = name
        return m

    def field(m, name, num, typ, label=T.LABEL_OPTIONAL, type_name=None, oneof=None):
        x = m.field.add()
        x.name, x.number, x.type, x.label = name, num, typ, label
        if type_name:
            x.type_name = ".ctr_test." + type_name
        if oneof is not None:
            x.oneof_index = oneof
        return x

    field(msg("BytesList"), "value", 1, T.TYPE_BYTES, T.LABEL_REPEATED)
    field(msg("FloatList"), "value", 1, T.TYPE_FLOAT, T.LABEL_REPEATED)        # proto3: packed by default
    field(msg("Int64List"), "value", 1, T.TYPE_INT64, T.LABEL_REPEATED)
    feat = msg("Feature")
    feat.oneof_decl.add().name = "kind"
    field(feat, "bytes_list", 1, T.TYPE_MESSAGE, type_name="BytesList", oneof=0)
    field(feat, "float_list", 2, T.TYPE_MESSAGE, type_name="FloatList", oneof=0)
    field(feat, "int64_list", 3, T.TYPE_MESSAGE, type_name="Int64List", oneof=0)
    feats = msg("Features")
    entry = feats.nested_type.add()
    entry.name = "FeatureEntry"
    entry.options.map_entry = True
    field(entry, "key", 1, T.TYPE_STRING)
    field(entry, "value", 2, T.TYPE_MESSAGE, type_name="Feature")
    field(feats, "feature", 1, T.TYPE_MESSAGE, T.LABEL_REPEATED, type_name="Features.FeatureEntry")
    field(msg("Example"), "features", 1, T.TYPE_MESSAGE, type_name="Features")
    pool = descriptor_pool.DescriptorPool()
    pool.Add(f)
    return message_factory.GetMessageClass(pool.FindMessageTypeByName("ctr_test.Example"))


def test_tf_example_wire_format_against_the_protobuf_runtime(tmp_path, lib_ready):
    """the oracle's hand-written encoder / decoder and the native reader against google.protobuf itself"""
    Example = _official_example_class()
    g = np.random.RandomState(9)
    n, C_, V, Fc = 70, 3, 2, 5
    rows = [{"label": g.rand(1).round(), "cont_feats": g.rand(C_), "vector_feats": g.randn(V),
             "cate_feats": g.randint(-5, 2 ** 45, size=Fc)} for _ in range(n)]
    payloads = []
    for r in rows:
        ex = Example()
        for k, v in r.items():
            if k == "cate_feats":
                ex.features.feature[k].int64_list.value.extend(int(x) for x in v)
            else:
                ex.features.feature[k].float_list.value.extend(float(x) for x in v)
        payloads.append(ex.SerializeToString())
        # official bytes -> oracle decoder
        back = LO.decode_example(payloads[-1])
        assert np.array_equal(back["cate_feats"], r["cate_feats"])
        for k in ("label", "cont_feats", "vector_feats"):
            assert np.array_equal(back[k], np.asarray(r[k], dtype=np.float32)), k
        # oracle encoder (both list encodings) -> official parser
        for packed in (True, False):
            ex2 = Example.FromString(LO.encode_example(r, packed=packed))
            assert list(ex2.features.feature["cate_feats"].int64_list.value) == [int(x) for x in r["cate_feats"]]
            assert np.array_equal(np.asarray(ex2.features.feature["cont_feats"].float_list.value, dtype=np.float32),
                                  np.asarray(r["cont_feats"], dtype=np.float32))
            assert set(ex2.features.feature) == set(r)
    # official bytes -> record framing -> native reader
    d = str(tmp_path) + "/"
    LO.write_tfrecord(d + "part-r-00000", payloads)
    mp = _mp(alg_name="deepfm_pipeline", batch_size=32, epochs=1, cont_field_size=C_, vector_feats_size=V, cate_field_size=Fc)
    got = list(L.pipeline_process(mp, [d + "part-r-00000"], "predict", pin=False))
    assert len(got) == n // 32
    for j, b in enumerate(got):
        sl = rows[32 * j: 32 * j + 32]
        assert np.array_equal(b["cate_feats"].numpy(), np.stack([r["cate_feats"] for r in sl]))
        assert np.array_equal(b["cont_feats"].numpy(), np.stack([r["cont_feats"] for r in sl]).astype(np.float32))
        assert np.array_equal(b["vector_feats"].numpy(), np.stack([r["vector_feats"] for r in sl]).astype(np.float32))
        assert np.array_equal(b["label"].numpy(), np.stack([r["label"] for r in sl]).astype(np.float32))


def test_tfrecord_framing_against_tensorboard(tmp_path, lib_ready):
    """TensorBoard ships its own implementation of the TFRecord container (masked CRC-32C, length/CRC framing): its
    writer's files through the oracle and the native reader, the oracle writer's files through its CRC-checking reader"""
    rw = pytest.importorskip("tensorboard.summary.writer.record_writer")
    pw = pytest.importorskip("tensorboard.compat.tensorflow_stub.pywrap_tensorflow")
    for blob in (b"", b"a", b"123456789", bytes(range(256)) * 5):
        assert LO.crc32c(blob) == pw.crc32c(blob) and LO.masked_crc(blob) == rw.masked_crc32c(blob)
    g = np.random.RandomState(12)
    n, C_, V, Fc = 40, 2, 1, 3
    rows = [{"label": g.rand(1).round(), "cont_feats": g.rand(C_), "vector_feats": g.randn(V),
             "cate_feats": g.randint(0, 2 ** 33, size=Fc)} for _ in range(n)]
    payloads = [LO.encode_example(r) for r in rows]
    d = str(tmp_path) + "/"
    with open(d + "part-tb", "wb") as f:                       # written by TensorBoard's RecordWriter
        w = rw.RecordWriter(f)
        for p in payloads:
            w.write(p)
        w.flush()
    LO.write_tfrecord(d + "part-own", payloads)                 # written by the oracle
    assert open(d + "part-tb", "rb").read() == open(d + "part-own", "rb").read()
    rd = pw.PyRecordReader_New(d + "part-own")                  # TensorBoard's reader checks both CRCs of every record
    back = []
    while True:
        try:
            rd.GetNext()
        except Exception as e:
            assert type(e).__name__ == "OutOfRangeError", e
            break
        back.append(rd.record())
    assert back == payloads
    mp = _mp(alg_name="deepfm_pipeline", batch_size=8, epochs=1, cont_field_size=C_, vector_feats_size=V, cate_field_size=Fc)
    got = list(L.pipeline_process(mp, [d + "part-tb"], "predict", pin=False))
    assert len(got) == n // 8
    assert np.array_equal(torch.cat([b["cate_feats"] for b in got]).numpy(), np.stack([r["cate_feats"] for r in rows]))


def test_decimal_fast_path_boundaries(lib_ready):
    """the digit scanner's corners, each against Python float() + float32 cast: mantissas around 2^53 (the exact-division
    limit, beyond it std::from_chars), fraction lengths around the 8-digit blocks, 19 and 20 digits, leading zeros, the
    float32 rounding boundary of 0.1, signs, and a token that ends the column"""
    g = np.random.RandomState(3)
    toks = ["0.9007199254740991", "0.9007199254740992", "0.9007199254740993", "9007199254740993", "9007199254740992.5",
            "0.95488135218620300", "0.1234567", "0.12345678", "0.123456789", "0.1234567812345678", "0.12345678123456789",
            "1234567890123456789", "12345678901234567890", "0.0000000000000000000012345", "00000000.5", "-0.00000001",
            "0.100000001490116119384765625", "0.10000000149011611938476562500000001", "0.0999999940395355224609375",
            "+1.5", "-.5", "5.", "16777217", "16777216.999999999", "3.4028235677973366e38", "1e-45", "0.7e-45"]
    toks += [repr(float(np.float32(v))) for v in g.rand(200)] + ["%.19f" % v for v in g.rand(100)] + ["%.12f" % (v * 1e6) for v in g.rand(100)]
    for i in range(0, len(toks) - 2, 3):
        t3 = toks[i:i + 3]
        blk = _parse_one("0," + " ".join(t3) + ",7 8", 3, 2)
        with np.errstate(over="ignore"):
            want = np.asarray([float(t) for t in t3], dtype=np.float64).astype(np.float32)
        assert np.array_equal(blk["cont"].numpy().reshape(-1).view(np.uint32), want.view(np.uint32)), t3
        assert blk["cate"].numpy().reshape(-1).tolist() == [7, 8]


def test_tf_example_fuzz_against_the_protobuf_runtime(tmp_path, lib_ready):
    """random Examples serialised by google.protobuf (extra features of all three kinds, empty lists, values at the
    int64 / float32 extremes) through the native reader: same rows as the official parser gives back"""
    from hypothesis import given, settings, strategies as st
    Example = _official_example_class()
    C_, V, Fc = 3, 2, 4
    d = str(tmp_path) + "/"
    mp = _mp(alg_name="deepfm_pipeline", batch_size=1, epochs=1, cont_field_size=C_, vector_feats_size=V, cate_field_size=Fc)
    f32 = st.floats(width=32, allow_nan=False)
    i64 = st.integers(min_value=-2 ** 63, max_value=2 ** 63 - 1)
    counter = [0]

    @settings(max_examples=60, deadline=None)
    @given(st.lists(f32, min_size=C_, max_size=C_), st.lists(f32, min_size=V, max_size=V), st.lists(i64, min_size=Fc, max_size=Fc),
           st.sampled_from([0.0, 1.0]), st.lists(f32, max_size=3), st.lists(i64, max_size=3), st.lists(st.binary(max_size=5), max_size=2))
    def check(cont, vec, cate, label, extra_f, extra_i, extra_b):
        ex = Example()
        ex.features.feature["zz_extra_f"].float_list.value.extend(extra_f)
        ex.features.feature["cate_feats"].int64_list.value.extend(cate)
        ex.features.feature["aa_extra_i"].int64_list.value.extend(extra_i)
        ex.features.feature["label"].float_list.value.append(label)
        ex.features.feature["vector_feats"].float_list.value.extend(vec)
        ex.features.feature["extra_b"].bytes_list.value.extend(extra_b)
        ex.features.feature["cont_feats"].float_list.value.extend(cont)
        counter[0] += 1
        path = d + "part-%d" % counter[0]
        LO.write_tfrecord(path, [ex.SerializeToString()])
        (b,) = list(L.pipeline_process(mp, [path], "predict", pin=False))
        back = Example.FromString(ex.SerializeToString())
        assert b["cate_feats"].numpy().reshape(-1).tolist() == list(back.features.feature["cate_feats"].int64_list.value)
        want = np.asarray(list(back.features.feature["cont_feats"].float_list.value), dtype=np.float32)
        assert np.array_equal(b["cont_feats"].numpy().reshape(-1).view(np.uint32), want.view(np.uint32))
        want = np.asarray(list(back.features.feature["vector_feats"].float_list.value), dtype=np.float32)
        assert np.array_equal(b["vector_feats"].numpy().reshape(-1).view(np.uint32), want.view(np.uint32))
        assert float(b["label"].numpy().reshape(-1)[0]) == label
        os.remove(path)
    check()
