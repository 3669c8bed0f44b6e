"""CPU oracle for the input decoding on the upstream side of the boundary.  TEST INFRASTRUCTURE ONLY:
only tests/ may import this module; the product (deep_learning_b200/loaders.py + csrc/host_loader.cu) never does.

PARITY UNPINNED for the TF pieces: the reference ships no data files or fixtures, and `tf.data`, `tf.gfile`,
`tf.parse_single_example` live in TensorFlow 1.x, which cannot be imported here.
  * `load_process` / `load_test` are PLAIN PYTHON in the reference (str.split, list appends, pickle) and are restated
    here statement by statement with `open` in place of `tf.gfile.FastGFile`: for them this oracle IS the reference
    algorithm, run as-is.
  * `decode_libsvm` is a tf.data graph (string_split / string_to_number): restated from the documented op semantics.
  * TFRecord + tf.Example: third-party formats (TensorFlow `record_writer.cc` framing: u64 length, masked CRC-32C of the
    length, payload, masked CRC-32C of the payload; `example.proto` / `feature.proto` protobuf wire format), restated
    from their published definitions.  The CRC is pinned by the RFC 3720 known answer crc32c(b"123456789") = 0xE3069283.
"""
from __future__ import annotations

import pickle
import struct
from typing import Dict, Iterable, List, Sequence

import numpy as np


# ---- utils/data_loader_load.py, plain-Python half -----------------------------------------------------------------------
def get_libsvm_value(line):
    """utils/data_loader_load.py:12-21"""
    index_list, value_list = [], []
    for tok in line.split(" "):
        sp = tok.split("=")
        index_list.append(int(sp[0]))
        value_list.append(float(sp[1]))
    return index_list, value_list


def get_dense_value(line):
    """utils/data_loader_load.py:24-25: the tokens stay STRINGS; the cast happens when feed_dict converts them"""
    return line.split(" ")


def read_data(mp, file_dir_list):
    """utils/data_loader_load.py:28-58"""
    out, count = [], 0
    for path in file_dir_list:
        try:
            inf = open(path, "r")
        except OSError:
            continue
        stop = False
        for line in inf.readlines():
            if not line:
                continue
            out.append(line)
            count += 1
            if count > mp.max_data_size:
                stop = True
                break
        inf.close()
        if stop:
            break
    return out


def load_process(mp, file_dir_list) -> List[bytes]:
    """utils/data_loader_load.py:61-139 (the size checks of :62-92 print and exit(-1); here they raise)"""
    data = read_data(mp, file_dir_list)
    first = data[0].strip("\n").split(",")
    if len(first[1].split(" ")) != mp.cont_field_size or len(first[2].split(" ")) != mp.cate_field_size:
        raise ValueError("feature size is error!!!")
    out = []
    for i in range(0, len(data), mp.batch_size):
        labels, cont_l, cate_l, mul_l, mulv_l, wide_l, vec_l = [], [], [], [], [], [], []
        for line in data[i:i + mp.batch_size]:
            sp = line.strip("\n").split(",")
            labels.append([float(sp[0])])
            cont = get_dense_value(sp[1])
            cate = get_dense_value(sp[2])
            iv = mp.cont_field_size - mp.vector_field_size
            vec_l.append(cont[iv:])
            cont_l.append(cont[:iv])
            cate_l.append(cate)
            if mp.alg_name == "dnn_multi":
                a, b = get_libsvm_value(sp[3])
                mul_l.append(a); mulv_l.append(b)
            if mp.alg_name == "wide_deep" or mp.alg_name == "wdl_textline":
                wide_l.append(get_dense_value(sp[3]))
        out.append(pickle.dumps({"cont_feats": cont_l, "vector_feats": vec_l, "cate_feats": cate_l, "mul_cate_feats": mul_l,
                                 "mul_cate_feats_value": mulv_l, "wide_feats": wide_l, "labels": labels}))
    return out


def load_test(mp, file_dir_list) -> List[bytes]:
    """utils/data_loader_load.py:142-184"""
    data = read_data(mp, file_dir_list)
    out = []
    for i in range(0, len(data), mp.batch_size):
        d = {k: [] for k in ("labels", "cont_feats_index", "cont_feats_value", "cate_feats_index", "cate_feats_value",
                             "vector_feats_value")}
        for line in data[i:i + mp.batch_size]:
            sp = line.strip("\n").split(",")
            d["labels"].append([float(sp[0])])
            cont = get_dense_value(sp[1])
            iv = mp.cont_field_size - mp.vector_field_size
            d["cont_feats_index"].append([j for j in range(0, iv)])
            d["cate_feats_value"].append([j for j in range(0, mp.cate_field_size)])
            d["cate_feats_index"].append(get_dense_value(sp[2]))
            d["vector_feats_value"].append(cont[iv:])
            d["cont_feats_value"].append(cont[:iv])
        out.append(pickle.dumps(d))
    return out


def feed(values, dtype) -> np.ndarray:
    """what `feed_dict` does with the nested lists of strings / numbers: numpy converts them to the placeholder dtype
    (float32 via Python float = strtod then a narrowing cast; int32 via int)"""
    a = np.asarray(values)
    if a.dtype.kind in "US":
        flat = [float(x) if np.issubdtype(dtype, np.floating) else int(x) for x in a.reshape(-1)]
        return np.asarray(flat, dtype=dtype).reshape(a.shape)
    return a.astype(dtype)


def decode_libsvm(line: str, mp) -> Dict[str, np.ndarray]:
    """utils/data_loader_load.py:187-243 for ONE line (tf.string_split drops nothing here: single-space separated)"""
    cols = line.rstrip("\n").split(",")
    d = {"labels": np.asarray([float(cols[0])], dtype=np.float32)}
    cont = np.asarray([float(x) for x in cols[1].split(" ")], dtype=np.float32)
    vec = np.zeros(0, dtype=np.float32)
    if mp.alg_name == "deepfm_textline":
        s = mp.cont_field_size - mp.vector_field_size
        vec, cont = cont[s:], cont[:s]
    d["cont_feats"], d["vector_feats"] = cont, vec
    d["cate_feats"] = np.asarray([int(x) for x in cols[2].split(" ")], dtype=np.int32)
    if mp.alg_name == "dnn_multi_textline":
        pairs = [t.split("=") for t in cols[3].split(" ")]
        d["mul_cate_feats"] = np.asarray([int(a) for a, _ in pairs], dtype=np.int32)
        d["mul_cate_feats_value"] = np.asarray([float(b) for _, b in pairs], dtype=np.float32)
    if mp.alg_name == "wdl_textline":
        d["wide_feats_value"] = np.asarray([int(x) for x in cols[3].split(" ")], dtype=np.int32)
    return d


def textline_batches(mp, files, drop_remainder=True, epochs=None) -> Iterable[Dict[str, np.ndarray]]:
    """TextLineDataset(files).map(decode).batch(B, drop_remainder).repeat(epochs) (utils/data_loader_load.py:246-252)"""
    for _ in range(mp.epochs if epochs is None else epochs):
        rows = []
        for path in files:
            with open(path) as f:
                for line in f:
                    if line.strip():
                        rows.append(decode_libsvm(line, mp))
                    if len(rows) == mp.batch_size:
                        yield {k: np.stack([r[k] for r in rows]) for k in rows[0]}
                        rows = []
        if rows and not drop_remainder:
            yield {k: np.stack([r[k] for r in rows]) for k in rows[0]}


# ---- TFRecord + tf.Example (published formats) ----------------------------------------------------------------------------
def _crc_table():
    t = []
    for i in range(256):
        c = i
        for _ in range(8):
            c = (c >> 1) ^ 0x82F63B78 if c & 1 else c >> 1
        t.append(c)
    return t


_T = _crc_table()


def crc32c(b: bytes) -> int:
    c = 0xFFFFFFFF
    for x in b:
        c = (c >> 8) ^ _T[(c ^ x) & 0xFF]
    return c ^ 0xFFFFFFFF


def masked_crc(b: bytes) -> int:
    c = crc32c(b)
    return (((c >> 15) | (c << 17)) + 0xA282EAD8) & 0xFFFFFFFF


def _varint(v: int) -> bytes:
    v &= (1 << 64) - 1
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _ld(field: int, payload: bytes) -> bytes:
    return _varint((field << 3) | 2) + _varint(len(payload)) + payload


def encode_example(feats: Dict[str, np.ndarray], packed: bool = True) -> bytes:
    """tf.train.Example with float_list (field 2) / int64_list (field 3) features; map entries in sorted key order
    (what the deterministic Python serializer emits)"""
    entries = b""
    for key in sorted(feats):
        a = np.asarray(feats[key])
        if a.dtype.kind == "f":
            vals = a.astype("<f4").reshape(-1)
            body = _ld(1, vals.tobytes()) if packed else b"".join(_varint((1 << 3) | 5) + struct.pack("<f", float(x)) for x in vals)
            feature = _ld(2, body)
        else:
            vals = a.astype(np.int64).reshape(-1)
            body = _ld(1, b"".join(_varint(int(x)) for x in vals)) if packed else b"".join(_varint(1 << 3) + _varint(int(x)) for x in vals)
            feature = _ld(3, body)
        entries += _ld(1, _ld(1, key.encode()) + _ld(2, feature))
    return _ld(1, entries)


def write_tfrecord(path: str, payloads: Sequence[bytes]) -> None:
    with open(path, "wb") as f:
        for p in payloads:
            hdr = struct.pack("<Q", len(p))
            f.write(hdr + struct.pack("<I", masked_crc(hdr)) + p + struct.pack("<I", masked_crc(p)))


def pipeline_batches(mp, files, action_type="train") -> Iterable[Dict[str, np.ndarray]]:
    """utils/data_loader.py:7-40 without the (unseeded) shuffle: the FixedLenFeature spec applied record by record,
    batch(drop_remainder=True), repeat(epochs) for training.  `files` hold Examples written by `write_tfrecord`; the
    decoder below is independent of the C++ one (it walks the same wire format in Python)."""
    cate_alg = ("deepfm_cate", "dnn_cate", "dnn_multi_cate", "deepfm_multi_cate")
    keys = ["cate_feats", "label", "vector_feats"] + ([] if mp.alg_name in cate_alg else ["cont_feats"])
    for _ in range(mp.epochs if action_type == "train" else 1):
        rows = []
        for path in files:
            blob = open(path, "rb").read()
            pos = 0
            while pos < len(blob):
                (n,) = struct.unpack_from("<Q", blob, pos)
                assert struct.unpack_from("<I", blob, pos + 8)[0] == masked_crc(blob[pos:pos + 8])
                payload = blob[pos + 12:pos + 12 + n]
                assert struct.unpack_from("<I", blob, pos + 12 + n)[0] == masked_crc(payload)
                pos += 16 + n
                ex = decode_example(payload)
                rows.append({k: ex[k] for k in keys})
                if len(rows) == mp.batch_size:
                    yield {k: np.stack([r[k] for r in rows]) for k in keys}
                    rows = []


def _read_varint(b: bytes, p: int):
    v, sh = 0, 0
    while True:
        x = b[p]; p += 1
        v |= (x & 0x7F) << sh
        sh += 7
        if not x & 0x80:
            return v, p


def _fields(b: bytes):
    p = 0
    while p < len(b):
        tag, p = _read_varint(b, p)
        wt = tag & 7
        if wt == 0:
            v, p = _read_varint(b, p)
        elif wt == 2:
            n, p = _read_varint(b, p)
            v = b[p:p + n]; p += n
        elif wt == 5:
            v = b[p:p + 4]; p += 4
        elif wt == 1:
            v = b[p:p + 8]; p += 8
        else:
            raise ValueError("wire type %d" % wt)
        yield tag >> 3, wt, v


def decode_example(payload: bytes) -> Dict[str, np.ndarray]:
    out = {}
    for fn, wt, feats in _fields(payload):
        if fn != 1:
            continue
        for fn2, _, entry in _fields(feats):
            key, feat = None, None
            for fn3, _, v in _fields(entry):
                if fn3 == 1:
                    key = v.decode()
                elif fn3 == 2:
                    feat = v
            for kind, _, lst in _fields(feat):
                if kind == 2:
                    vals = []
                    for f4, wt4, v in _fields(lst):
                        vals += list(np.frombuffer(v, dtype="<f4")) if wt4 == 2 else [struct.unpack("<f", v)[0]]
                    out[key] = np.asarray(vals, dtype=np.float32)
                elif kind == 3:
                    vals = []
                    for f4, wt4, v in _fields(lst):
                        if wt4 == 2:
                            p = 0
                            while p < len(v):
                                x, p = _read_varint(v, p)
                                vals.append(x)
                        else:
                            vals.append(v)
                    out[key] = np.asarray(vals, dtype=np.uint64).astype(np.int64)
    return out
