"""Input pipeline on the upstream side of the boundary: the reference's three loader modules with
their function names and return shapes, decoding done by the native parsers of libctr_b200.so
(`ctr_parse_lines`, `ctr_parse_tfrecords`; include/ctr_b200.h) straight into pinned host tensors.

    utils/data_loader_load.py      load_input_file(mp, input_path, read_file_type)   -> list of batches | iterator
    utils/data_loader_textline.py  dnn_multi_process(filenames, batch_size, epochs)  -> iterator
    utils/data_loader.py           pipeline_process(mp, file_dir_list, action_type)  -> iterator

A batch is a dict of torch tensors under the reference's key names; every model module accepts it where the reference
passes pickled lists (generation A) or iterator tensors (generations B / textline).  `as_pickle=True` gives the
reference's wire format (`pickle.dumps` of Python lists) for code that insists on it.
"""
from __future__ import annotations

import ctypes as C
import os
import pickle
import queue
import threading
from typing import Dict, Iterable, Iterator, List, Optional, Sequence

import numpy as np
import torch

from . import _lib

CHUNK_BYTES = 32 << 20


class Prefetch:
    """Re-iterable wrapper: a background thread decodes up to `depth` batches ahead (tf.data's prefetch; the native decoders
    run without the GIL, so decoding overlaps the training loop's Python).  Exceptions of the producer surface in the
    consumer; abandoning the iterator stops the producer."""

    def __init__(self, source: Iterable, depth: int = 2):
        self.source, self.depth = source, max(1, int(depth))

    def __iter__(self):
        q: "queue.Queue" = queue.Queue(self.depth)
        stop = threading.Event()
        END, ERR = object(), object()

        def put(item) -> bool:
            while not stop.is_set():
                try:
                    q.put(item, timeout=0.1)
                    return True
                except queue.Full:
                    continue
            return False

        def run():
            try:
                for item in self.source:
                    if not put(item):
                        return
                put(END)
            except BaseException as e:      # noqa: BLE001 -- handed to the consumer
                put((ERR, e))

        th = threading.Thread(target=run, daemon=True, name="ctr-loader-prefetch")
        th.start()
        try:
            while True:
                item = q.get()
                if item is END:
                    return
                if isinstance(item, tuple) and len(item) == 2 and item[0] is ERR:
                    raise item[1]
                yield item
        finally:
            stop.set()


def _host(shape, dtype, pin: bool) -> torch.Tensor:
    t = torch.empty(shape, dtype=dtype)
    return t.pin_memory() if (pin and torch.cuda.is_available()) else t


def get_file_list(input_path: str) -> List[str]:
    """utils/data_loader_load.py:264-273 / utils/data_loader.py:49-58: entries whose name starts with "part";
    the path is `input_path + file` exactly as there (input_path carries its trailing separator)."""
    file_list = sorted(os.listdir(input_path))
    print("file_list_len:", len(file_list))
    return [input_path + f for f in file_list if f[:4] == "part"]


# ---- text lines ------------------------------------------------------------------------------------------------------
class _LineSchema:
    """what one line holds for a given alg_name (utils/data_loader_load.py:94-139,187-243)"""

    def __init__(self, n_cont: int, n_vec: int, n_cate: int, third_kind: int = 0, n_third: int = 0, id_dtype=torch.int32,
                 names: Optional[Dict[str, str]] = None):
        self.n_cont, self.n_vec, self.n_cate, self.third_kind, self.n_third = n_cont, n_vec, n_cate, third_kind, n_third
        self.id_dtype = id_dtype
        self.names = dict(label="labels", cont="cont_feats", vec="vector_feats", cate="cate_feats", third="wide_feats",
                          third_val="mul_cate_feats_value")
        if names:
            self.names.update(names)


def schema_for(mp, flavour: str = "load") -> _LineSchema:
    """flavour 'load': load_process (utils/data_loader_load.py:61-139); 'textline': decode_libsvm (:187-243);
    'dnn_multi_textline': utils/data_loader_textline.py:4-36"""
    alg = getattr(mp, "alg_name", "")
    n_cont_total = int(mp.cont_field_size)
    n_vec = int(getattr(mp, "vector_field_size", 0) or 0)
    if flavour == "dnn_multi_textline":
        return _LineSchema(n_cont_total, 0, int(getattr(mp, "cat_field_size", 0) or mp.cate_field_size), 2,
                           int(getattr(mp, "multi_feats_size", 0) or getattr(mp, "mul_cat_field_size", 0)),
                           names=dict(cate="cat_feats", third="mul_cat_feats", third_val="mul_cat_feats_value"))
    if flavour == "textline":
        # :203-206 splits the vector block off only for deepfm_textline
        nv = n_vec if alg == "deepfm_textline" else 0
        if alg == "dnn_multi_textline":
            return _LineSchema(n_cont_total, nv, int(mp.cate_field_size), 2, int(mp.multi_feats_size),
                               names=dict(third="mul_cate_feats", third_val="mul_cate_feats_value"))
        if alg == "wdl_textline":
            return _LineSchema(n_cont_total, nv, int(mp.cate_field_size), 1, int(mp.wide_field_size),
                               names=dict(third="wide_feats_value"))
        return _LineSchema(n_cont_total, nv, int(mp.cate_field_size))
    # load_process: the vector block is ALWAYS split off (:108-113); column 3 by alg_name (:115-127)
    if alg == "dnn_multi":
        return _LineSchema(n_cont_total, n_vec, int(mp.cate_field_size), 2, int(getattr(mp, "multi_cate_field_size", 0)),
                           names=dict(third="mul_cate_feats", third_val="mul_cate_feats_value"))
    if alg in ("wide_deep", "wdl_textline", "wdl"):
        # :125 tests "wide_deep" where :80 tests "wdl" (reference defect): plain "wdl" leaves wide_feats
        # empty there; here the column is parsed whenever the model has wide fields
        return _LineSchema(n_cont_total, n_vec, int(mp.cate_field_size), 1, int(mp.wide_field_size))
    return _LineSchema(n_cont_total, n_vec, int(mp.cate_field_size))


class LineBlocks:
    """files -> blocks of decoded rows (one block per chunk of bytes); a line never straddles two blocks"""

    def __init__(self, files: Sequence[str], sc: _LineSchema, n_threads: int = 0, chunk_bytes: int = CHUNK_BYTES,
                 max_rows: Optional[int] = None):
        self.files, self.sc, self.n_threads, self.chunk, self.max_rows = list(files), sc, n_threads, chunk_bytes, max_rows
        self.lib = _lib.load()

    def _parse(self, buf: bytes):
        sc = self.sc
        cap = buf.count(b"\n")
        idb = 8 if sc.id_dtype == torch.int64 else 4
        label = torch.empty(cap, 1, dtype=torch.float32)
        cont = torch.empty(cap, max(sc.n_cont, 1), dtype=torch.float32)
        cate = torch.empty(cap, max(sc.n_cate, 1), dtype=sc.id_dtype)
        tid = torch.empty(cap, max(sc.n_third, 1), dtype=sc.id_dtype)
        tval = torch.empty(cap, max(sc.n_third, 1), dtype=torch.float32)
        n, used, bad = C.c_int64(), C.c_int64(), C.c_int64()
        rc = self.lib.ctr_parse_lines(buf, len(buf), sc.n_cont, sc.n_cate, sc.third_kind, sc.n_third, idb, cap, self.n_threads,
                                      label.data_ptr(), cont.data_ptr(), cate.data_ptr(), tid.data_ptr(), tval.data_ptr(),
                                      C.byref(n), C.byref(used), C.byref(bad))
        if rc != 0:
            raise ValueError("feature size is error!!! (malformed line %d of the chunk; expected %d cont, %d cate, %d third)"
                             % (bad.value, sc.n_cont, sc.n_cate, sc.n_third))
        r = n.value
        nc = sc.n_cont - sc.n_vec
        blk = {"label": label[:r], "cont": cont[:r, :nc], "vec": cont[:r, nc:sc.n_cont], "cate": cate[:r, :sc.n_cate]}
        if sc.third_kind:
            blk["third"] = tid[:r, :sc.n_third]
            if sc.third_kind == 2:
                blk["third_val"] = tval[:r, :sc.n_third]
        return blk, used.value

    def __iter__(self):
        left = self.max_rows
        for path in self.files:
            try:
                f = open(path, "rb")
            except OSError:
                print("not file exception input file not find")     # utils/data_loader_load.py:37-39
                continue
            with f:
                rest = b""
                while True:
                    data = f.read(self.chunk)
                    buf = rest + data
                    if not data and buf and not buf.endswith(b"\n"):
                        buf += b"\n"                               # last line of a file without a newline
                    if not buf:
                        break
                    blk, used = self._parse(buf)
                    rest = buf[used:]
                    if left is not None:
                        if blk["label"].shape[0] > left:
                            blk = {k: v[:left] for k, v in blk.items()}
                        left -= blk["label"].shape[0]
                    if blk["label"].shape[0]:
                        yield blk
                    if left is not None and left <= 0:
                        return
                    if not data:
                        break


def _rebatch(blocks: Iterable[Dict[str, torch.Tensor]], B: int, drop_remainder: bool, names: Dict[str, str], pin: bool):
    """rows of consecutive blocks -> batches of exactly B rows (the last one short unless drop_remainder)"""
    cur, fill = None, 0
    for blk in blocks:
        n, at = blk["label"].shape[0], 0
        while at < n:
            if cur is None:
                cur = {k: _host((B,) + tuple(v.shape[1:]), v.dtype, pin) for k, v in blk.items()}
                fill = 0
            take = min(B - fill, n - at)
            for k, v in blk.items():
                cur[k][fill:fill + take] = v[at:at + take]
            fill += take
            at += take
            if fill == B:
                yield {names[k]: v for k, v in cur.items()}
                cur = None
    if cur is not None and fill and not drop_remainder:
        yield {names[k]: v[:fill] for k, v in cur.items()}


class TextLineBatches:
    """TextLineDataset(files).map(decode).batch(B[, drop_remainder]).repeat(epochs) as a re-iterable of batch dicts"""

    def __init__(self, files: Sequence[str], sc: _LineSchema, batch_size: int, epochs: int = 1, drop_remainder: bool = True,
                 pin: bool = True, n_threads: int = 0, max_rows: Optional[int] = None):
        self.files, self.sc, self.B, self.epochs, self.drop, self.pin = list(files), sc, int(batch_size), int(epochs), drop_remainder, pin
        self.n_threads, self.max_rows = n_threads, max_rows

    def __iter__(self) -> Iterator[Dict[str, torch.Tensor]]:
        for _ in range(self.epochs):
            yield from _rebatch(LineBlocks(self.files, self.sc, self.n_threads, max_rows=self.max_rows), self.B, self.drop,
                                self.sc.names, self.pin)


def _pickle_batch(b: Dict[str, torch.Tensor]) -> bytes:
    """the reference's generation-A wire format (utils/data_loader_load.py:130-137)"""
    g = lambda k: b[k].tolist() if k in b else []
    return pickle.dumps({"cont_feats": g("cont_feats"), "vector_feats": g("vector_feats"), "cate_feats": g("cate_feats"),
                         "mul_cate_feats": g("mul_cate_feats"), "mul_cate_feats_value": g("mul_cate_feats_value"),
                         "wide_feats": g("wide_feats"), "labels": g("labels")})


def load_process(mp, file_dir_list: Sequence[str], as_pickle: bool = False, pin: bool = True) -> List:
    """utils/data_loader_load.py:61-139: every line of every file (up to mp.max_data_size, :45), cut into batches of
    mp.batch_size, the last one short.  Returns the list the generation-A models' fit/predict take."""
    print("*****read data******")
    sc = schema_for(mp, "load")
    mx = getattr(mp, "max_data_size", None)
    mx = int(mx) + 1 if mx is not None else None       # :45 stops AFTER appending line number max_data_size + 1
    out = list(TextLineBatches(file_dir_list, sc, mp.batch_size, 1, drop_remainder=False, pin=pin, max_rows=mx))
    print("size of read data = ", sum(int(b["labels"].shape[0]) for b in out))
    return [_pickle_batch(b) for b in out] if as_pickle else out


def load_test(mp, file_dir_list: Sequence[str], pin: bool = True) -> List[Dict]:
    """utils/data_loader_load.py:142-184: the explicit index/value form (BASELINE.json's feat_index / feat_value / label):
    cont index = arange(C), cate value = arange(Fc) (sic, :156-157)"""
    sc = _LineSchema(int(mp.cont_field_size), int(getattr(mp, "vector_field_size", 0) or 0), int(mp.cate_field_size))
    out = []
    for b in TextLineBatches(file_dir_list, sc, mp.batch_size, 1, drop_remainder=False, pin=pin):
        n, C_, Fc = b["labels"].shape[0], b["cont_feats"].shape[1], b["cate_feats"].shape[1]
        out.append({"labels": b["labels"], "cont_feats_index": torch.arange(C_, dtype=torch.int32).expand(n, C_).contiguous(),
                    "cont_feats_value": b["cont_feats"], "cate_feats_index": b["cate_feats"],
                    "cate_feats_value": torch.arange(Fc, dtype=torch.int32).expand(n, Fc).contiguous(),
                    "vector_feats_value": b["vector_feats"]})
    return out


def textline_process(mp, filenames: Sequence[str], pin: bool = True, prefetch: int = 2):
    """utils/data_loader_load.py:246-252: batch(drop_remainder=True).repeat(epochs); `prefetch` batches are decoded ahead
    on a background thread (0: decode in the consumer's thread)"""
    it = TextLineBatches(filenames, schema_for(mp, "textline"), mp.batch_size, mp.epochs, drop_remainder=True, pin=pin)
    return Prefetch(it, prefetch) if prefetch else it


def dnn_multi_process(filenames, batch_size: int, epochs: int, mp=None, pin: bool = True) -> TextLineBatches:
    """utils/data_loader_textline.py:39-44: batch(batch_size).repeat(epochs) -- the last batch of an epoch may be short.
    The TF graph infers the column widths from the data; here they come from `mp`."""
    if mp is None:
        raise ValueError("dnn_multi_process(filenames, batch_size, epochs, mp=...): the field sizes are read from mp")
    files = [filenames] if isinstance(filenames, str) else list(filenames)
    return TextLineBatches(files, schema_for(mp, "dnn_multi_textline"), batch_size, epochs, drop_remainder=False, pin=pin)


def load_input_file(mp, input_path: str, read_file_type: str = "load"):
    """utils/data_loader_load.py:255-261"""
    files = get_file_list(input_path)
    if read_file_type == "textline":
        return textline_process(mp, files)
    if read_file_type == "tf_test":
        return load_test(mp, files)
    return load_process(mp, files)


# ---- TFRecord / tf.Example -----------------------------------------------------------------------------------------------
CATE_ALG = ("deepfm_cate", "dnn_cate", "dnn_multi_cate", "deepfm_multi_cate")      # utils/data_loader.py:8


class TFRecordBatches:
    """TFRecordDataset(files).map(parse_single_example).batch(B, drop_remainder=True)[.repeat(epochs)]
    (utils/data_loader.py:27-40).  The reference shuffles with a 10-batch buffer before batching (:30), from TF's
    unseeded generator, which nothing here can reproduce; what IS kept is that a training run does not see the file order:
    `shuffle_seed="default"` means a seeded shuffle inside each window of 10 batches for action_type == "train" (seed
    2019, the model files' random_seed, advanced by one per epoch) and file order for every other action type (scoring
    does not depend on the order, and a deterministic order keeps evaluations comparable).  `shuffle_seed=None` opts out
    (file order), an int picks the seed for any action type."""
    DEFAULT_SEED = 2019

    def __init__(self, mp, files: Sequence[str], action_type: str = "train", pin: bool = True, verify_crc: bool = True,
                 shuffle_seed="default", chunk_bytes: int = CHUNK_BYTES):
        if isinstance(shuffle_seed, str):
            if shuffle_seed != "default":
                raise ValueError("shuffle_seed: an int, None or 'default'")
            shuffle_seed = self.DEFAULT_SEED if action_type == "train" else None
        self.files, self.B, self.pin, self.crc, self.chunk = list(files), int(mp.batch_size), pin, int(verify_crc), chunk_bytes
        self.epochs = int(mp.epochs) if action_type == "train" else 1
        self.has_cont = getattr(mp, "alg_name", "") not in CATE_ALG
        self.n_cont = int(mp.cont_field_size) if self.has_cont else 0
        self.n_vec = int(mp.vector_feats_size)
        self.n_cate = int(mp.cate_field_size) + int(getattr(mp, "multi_feats_size", 0) or 0)
        self.seed = shuffle_seed
        self.lib = _lib.load()

    def _blocks(self):
        for path in self.files:
            with open(path, "rb") as f:
                rest = b""
                while True:
                    data = f.read(self.chunk)
                    buf = rest + data
                    if not buf:
                        break
                    cap = max(1, len(buf) // 16)
                    label = torch.empty(cap, 1, dtype=torch.float32)
                    cont = torch.empty(cap, max(self.n_cont, 1), dtype=torch.float32)
                    vec = torch.empty(cap, max(self.n_vec, 1), dtype=torch.float32)
                    cate = torch.empty(cap, self.n_cate, dtype=torch.int64)
                    n, used = C.c_int64(), C.c_int64()
                    rc = self.lib.ctr_parse_tfrecords(buf, len(buf), self.n_cont, self.n_vec, self.n_cate, self.crc, cap,
                                                      label.data_ptr(), cont.data_ptr(), vec.data_ptr(), cate.data_ptr(),
                                                      C.byref(n), C.byref(used))
                    if rc != 0:
                        raise ValueError("%s: corrupt TFRecord / Example does not match the feature spec (record %d of the chunk)"
                                         % (path, n.value))
                    r = n.value
                    rest = buf[used.value:]
                    if r:
                        yield {"label": label[:r], "cont": cont[:r, :self.n_cont], "vec": vec[:r, :self.n_vec], "cate": cate[:r]}
                    if not data:
                        if rest:
                            raise ValueError("%s: truncated TFRecord (%d trailing bytes)" % (path, len(rest)))
                        break

    def _shuffled(self, blocks, epoch: int = 0):
        g = torch.Generator(); g.manual_seed(self.seed + epoch)
        win = self.B * 10
        for big in _rebatch(blocks, win, False, {k: k for k in ("label", "cont", "vec", "cate")}, False):
            perm = torch.randperm(big["label"].shape[0], generator=g)
            yield {k: v[perm] for k, v in big.items()}

    def __iter__(self):
        names = {"label": "label", "cont": "cont_feats", "vec": "vector_feats", "cate": "cate_feats"}
        for epoch in range(self.epochs):
            blocks = self._blocks() if self.seed is None else self._shuffled(self._blocks(), epoch)
            for b in _rebatch(blocks, self.B, True, names, self.pin):
                if not self.has_cont:
                    b["cont_feats"] = b["cont_feats"][:, :0]
                yield b


def pipeline_process(mp, file_dir_list: Sequence[str], action_type: str, prefetch: int = 2, **kw):
    """utils/data_loader.py:27-40; `prefetch` batches are decoded ahead on a background thread (0: none)"""
    it = TFRecordBatches(mp, file_dir_list, action_type, **kw)
    return Prefetch(it, prefetch) if prefetch else it


def load_tfrecord_input_file(mp, input_path: str, action_type: str, **kw):
    """utils/data_loader.py:43-46 (`load_input_file` of that module)"""
    return pipeline_process(mp, get_file_list(input_path), action_type, **kw)
