"""Shared implementation behind the reference-shaped model modules (models/deepfm.py, ...).

Each module exposes `DeepModel` with the constructor / fit / evaluate|eval / predict signatures the
reference launchers call (local_run_textline.py:182-184, local_run_test.py:77-82); the TF session
argument is accepted and ignored.
"""
from __future__ import annotations

import dataclasses
import json
import math
import os
import sys
import time
import types
from typing import Dict, Iterable, List, Optional

import numpy as np
import torch

from .. import data as _data
from ..engine import CtrEngine
from ..spec import ModelSpec, spec_from_hparams


def init_weights(spec: ModelSpec, seed: int = 2019, with_table: bool = True) -> Dict[str, np.ndarray]:
    """Reference initial distributions: feats_emb ~ N(0, 0.01^2), first-order ~ U(0,1)
    (models/deepfm.py:58-60); xavier-uniform weight_mat (models/wdl.py:44-47); MLP W and b
    ~ N(0, 2/(fan_in+fan_out)) (models/deepfm.py:106-119); head bias ~ N(0,1) (models/deepfm.py:142)."""
    g = np.random.RandomState(seed)
    p: Dict[str, np.ndarray] = {}
    R, k = spec.table_rows, spec.k
    if not with_table:
        pass
    elif spec.emb_name == "weight_mat":
        lim = math.sqrt(6.0 / (R + k))
        p["weight_mat"] = g.uniform(-lim, lim, size=(R, k)).astype(np.float32)
    else:
        p[spec.emb_name] = g.normal(0.0, 0.01, size=(R, k)).astype(np.float32)
    if spec.frozen_rows and with_table:
        # stand-in for the word2vec table the reference loads from disk (models/dnn_multi_textline.py:46-48)
        p[spec.frozen_name] = g.normal(0.0, 0.1, size=(spec.frozen_rows, k)).astype(np.float32)
    if spec.first_name and with_table:
        p[spec.first_name] = g.uniform(0.0, 1.0, size=(R, 1)).astype(np.float32)
    shapes = dict(spec.dense_param_shapes())
    for name, shp in spec.dense_param_shapes():
        if name.startswith("deep_bias_"):
            w = shapes["deep_" + name.split("_")[-1]]
            std = math.sqrt(2.0 / (w[0] + w[1]))
        elif name in ("deep_fm_bias", "wdl_bias"):
            std = 1.0
        elif name == "deep_res_bias":
            std = math.sqrt(2.0 / (shapes["deep_res"][0] + 1))
        elif name == "wdl_weights":
            std = math.sqrt(2.0 / shp[0])
        else:
            std = math.sqrt(2.0 / (shp[0] + shp[1]))
        p[name] = g.normal(0.0, std, size=shp).astype(np.float32)
    return p


def _lookahead(items: Iterable):
    """(item, next item or None) pairs: the engine uploads / plans the next batch underneath the current step when it is
    told one step ahead what comes next (the reference's tf.data iterator prefetches the same way, utils/data_loader.py:29-40)"""
    it = iter(items)
    try:
        cur = next(it)
    except StopIteration:
        return
    for nxt in it:
        yield cur, nxt
        cur = nxt
    yield cur, None


EXPORT_FORMAT = 2          # 1: variables only (round 1); 2: + alg, hyper-parameters, serving signature


def hparams_to_dict(args) -> Dict:
    """the JSON-able attributes of a hyper-parameter bag (ModelParams, SimpleNamespace, a launcher's own class)"""
    src = dataclasses.asdict(args) if dataclasses.is_dataclass(args) else dict(vars(args))
    out = {}
    for k, v in src.items():
        if k.startswith("_"):
            continue
        if isinstance(v, (bool, int, float, str)) or v is None:
            out[k] = v
        elif isinstance(v, (list, tuple)):
            try:
                json.dumps(v)
                out[k] = list(v)
            except TypeError:
                pass
        elif isinstance(v, (np.integer, np.floating)):
            out[k] = v.item()
    return out


def serving_signature(spec) -> Dict:
    """The reference's signature_def (models/deepfm.py:193-201, models/wdl.py:320-328): tensor names, dtypes and
    per-sample widths of the inputs, and the output `score:0` [B, 1]."""
    ids = spec.id_dtype
    inputs = {"cont_feats": {"tensor": "cont_feats:0", "dtype": "float32", "width": spec.C},
              "vector_feats": {"tensor": "vector_feats:0", "dtype": "float32", "width": spec.V},
              "cate_feats": {"tensor": "cate_feats:0", "dtype": ids, "width": spec.ids_ld},
              "input_data_size": {"tensor": "input_data_size:0", "dtype": "int32", "width": None}}
    if spec.head == "wdl":
        inputs["wide_feats"] = {"tensor": "wide_feats:0", "dtype": ids, "width": spec.wide_fields}
    if spec.weighted_pool:
        inputs["multi_feats_value"] = {"tensor": "mul_cat_feats_value:0", "dtype": "float32", "width": spec.M}
    return {"method_name": "tensorflow/serving/predict", "inputs": inputs,
            "outputs": {"output": {"tensor": "score:0", "dtype": "float32", "width": 1}}}


def read_export_meta(path: str) -> Dict:
    """alg / hyper-parameters / signature of an export written by `save`"""
    z = np.load(os.path.join(path, "model.npz"))
    if "meta/json" not in z.files:
        raise ValueError("%s is a variables-only export (format 1): pass the hyper-parameter bag as args=mp" % path)
    return json.loads(str(z["meta/json"]))


def hparams_from_export(path: str):
    """the hyper-parameter bag an export was written with, as an attribute holder"""
    return types.SimpleNamespace(**read_export_meta(path)["hparams"])


class ExportedModel:
    """`tf.saved_model.loader.load(session, [SERVING], model_pb)` + `session.run("score:0", feed_dict)` by tensor NAME
    (models/deepfm.py:237-250, models/wdl.py:364-377): everything comes from the export directory."""

    def __init__(self, model_pb: str, max_batch: Optional[int] = None, **kw):
        import importlib
        meta = read_export_meta(model_pb)
        self.meta, self.signature = meta, meta["signature"]
        mod = importlib.import_module("deep_learning_b200.models." + meta["alg"])
        args = types.SimpleNamespace(**meta["hparams"])
        if max_batch:
            kw["max_batch"] = max_batch
        self.model = mod.DeepModel(args, **kw) if meta["generation"] == "A" else mod.DeepModel(args, None, **kw)
        self.model.restore(model_pb)
        self._by_tensor = {v["tensor"]: k for k, v in self.signature["inputs"].items()}

    def run(self, fetches, feed_dict: Dict) -> np.ndarray:
        """fetches: "score:0" (or the signature's output key); feed_dict keyed by tensor name ("cate_feats:0") or by
        signature key ("cate_feats").  Returns [B, 1] float32 like the reference's session.run."""
        out = self.signature["outputs"]["output"]["tensor"]
        if fetches not in (out, "output", "score"):
            raise KeyError("%r is not an output of this export (have %r)" % (fetches, out))
        batch = {}
        for k, v in feed_dict.items():
            key = self._by_tensor.get(k, k)
            if key not in self.signature["inputs"]:
                raise KeyError("%r is not an input of this export (have %s)" % (k, sorted(self._by_tensor)))
            batch[key] = v
        n = batch.pop("input_data_size", None)
        if "vector_feats" not in batch and self.model.spec.V == 0 and "cate_feats" in batch:
            batch["vector_feats"] = np.zeros((len(batch["cate_feats"]), 0), np.float32)
        scores = self.model.predict_scores(batch)
        return scores if n is None else scores[:int(np.asarray(n).reshape(-1)[0])]


def load_exported(model_pb: str, **kw) -> ExportedModel:
    return ExportedModel(model_pb, **kw)


class BaseDeepModel:
    ALG = ""
    GENERATION = "A"     # "A": feed_dict / pickled batches; "B": iterator of batch dicts

    def __init__(self, args, data_dict=None, *, precision: str = "fp32", adam_mode: str = "tf_dense",
                 device=None, weights: Optional[Dict[str, np.ndarray]] = None, max_batch: Optional[int] = None):
        self.args = args
        self.hidden_units = args.hidden_units
        self.epochs = getattr(args, "epochs", 1)
        self.batch_size = args.batch_size
        self.learning_rate = args.learning_rate
        self.model_pb = getattr(args, "model_pb", "")
        self.decay_steps = args.learning_rate_decay_steps
        self.decay_rate = args.learning_rate_decay_rate
        self.l2_reg = args.l2_reg
        self.metric_type = "auc"
        self.random_seed = 2019
        self.dropout_keep_deep = [1, 1, 1, 1, 1]
        self.dropout_keep_fm = [1, 1]
        self.embedding_size = args.embedding_size
        self.save_model_checkpoint = getattr(args, "save_model_checkpoint", "")
        self.restore_model_checkpoint = getattr(args, "restore_model_checkpoint", "")
        self.model_restore = getattr(args, "model_restore", 0)
        self.data_dict = data_dict
        self.spec = spec_from_hparams(self.ALG, args)
        eng_cls = CtrEngine
        if self.spec.split_fm:
            from ..split_engine import SplitFmEngine as eng_cls      # deepfm_pipeline: two-row lookup per feature (quirk Q2)
        self.engine = eng_cls(self.spec, max_batch or self.batch_size, device=device,
                                learning_rate=self.learning_rate, decay_steps=self.decay_steps,
                                decay_rate=self.decay_rate, precision=precision, adam_mode=adam_mode)
        self.engine.set_weights(weights if weights is not None else init_weights(self.spec, self.random_seed))

    # ---- batch decoding ---------------------------------------------------------------------------
    def _decode(self, item) -> Dict:
        if isinstance(item, (bytes, bytearray)):
            item = _data.from_pickled_batch(bytes(item), np.int64 if self.spec.id_dtype == "int64" else np.int32)
        return _data.canonical_batch(item, self.spec)

    # ---- persistence: the reference exports a SavedModel (.pb, models/deepfm.py:193-211) and, for the
    # tf.data generation, a Saver checkpoint with the Adam slots (models/deepfm_multi.py:308-310).
    def save(self, path: str, with_optimizer: bool = False) -> None:
        """The export is SELF-DESCRIBING, like the reference's SavedModel (models/deepfm.py:193-211): beside the variables
        it carries the model file's name, the hyper-parameter bag and the named-tensor serving signature, so
        `load_exported(path)` / module-level `predict(data, model_pb)` need nothing else (local_run_test.py:82)."""
        os.makedirs(path, exist_ok=True)
        blob = {"w/" + k: v for k, v in self.engine.get_weights().items()}
        blob["meta/json"] = np.array(json.dumps({"format": EXPORT_FORMAT, "alg": self.ALG, "generation": self.GENERATION,
                                                 "hparams": hparams_to_dict(self.args),
                                                 "signature": serving_signature(self.spec)}))
        if with_optimizer:
            e = self.engine
            blob.update({"opt/m_table": e.m_table.cpu().numpy(), "opt/v_table": e.v_table.cpu().numpy(),
                         "opt/m": e.m.cpu().numpy(), "opt/v": e.v.cpu().numpy(),
                         "opt/state": np.asarray([e.global_step, e.beta1_power, e.beta2_power], dtype=np.float64)})
        np.savez(os.path.join(path, "model.npz"), **blob)

    def restore(self, path: str) -> None:
        z = np.load(os.path.join(path, "model.npz"))
        self.engine.set_weights({k[2:]: z[k] for k in z.files if k.startswith("w/")})
        if "opt/state" in z.files:
            e = self.engine
            e.m_table.copy_(torch.from_numpy(z["opt/m_table"])); e.v_table.copy_(torch.from_numpy(z["opt/v_table"]))
            e.m.copy_(torch.from_numpy(z["opt/m"])); e.v.copy_(torch.from_numpy(z["opt/v"]))
            st = z["opt/state"]
            e.global_step, e.beta1_power, e.beta2_power = int(st[0]), float(st[1]), float(st[2])

    # ---- scoring ---------------------------------------------------------------------------------------
    def _score_batches(self, batches: Iterable):
        scores, labels = [], []
        for item in batches:
            b = self._decode(item)
            sc = self.engine.predict(b).clone()
            scores.append(sc)
            lab = b.get("label", b.get("labels"))
            labels.append(torch.as_tensor(np.asarray(lab) if not isinstance(lab, torch.Tensor) else lab)
                          .reshape(-1).to(sc.device, torch.float32))
        return torch.cat(scores), torch.cat(labels)

    def _auc(self, batches: Iterable) -> float:
        scores, labels = self._score_batches(batches)
        return self.engine.auc(scores, labels)

    def predict_scores(self, batch) -> np.ndarray:
        """`session.run("score:0", feed_dict)` for one batch: [B,1] host array."""
        out = self.engine.predict(self._decode(batch)).cpu().numpy().reshape(-1, 1)
        self.engine.check_status()        # an id outside its table raises here, as tf.nn.embedding_lookup does
        return out

    # ---- generation A: fit(train_data, val_data) (models/deepfm.py:164-214) -----------------------------
    def _fit_a(self, train_data: List, val_data: List) -> None:
        losses, num_samples = [], 0       # never reset between epochs, as in the reference (:168-169)
        for epoch in range(self.epochs):
            st = time.time()
            for b, nb in _lookahead(self._decode(item) for item in train_data):
                loss = self.engine.train_step(b, next_batch=nb)
                losses.append(loss)
                num_samples += self.batch_size
            total_loss = float(torch.stack(losses).sum().item() * self.batch_size / max(num_samples, 1))
            end_time = time.time()
            valid_metric = self._auc(val_data)
            print('[%s] valid-%s=%.5f\tloss=%.5f [%.1f s]' % (epoch + 1, self.metric_type, valid_metric, total_loss,
                                                              end_time - st))
            sys.stdout.flush()
        if self.model_pb:
            try:
                self.save(self.model_pb)
            except Exception as e:   # the reference swallows export failures (models/deepfm.py:212-214)
                print("Fail to export saved model, exception: {}".format(e))
                sys.stdout.flush()

    def _predict_a(self, data_val: List) -> float:
        if self.model_pb and os.path.exists(os.path.join(self.model_pb, "model.npz")):
            self.restore(self.model_pb)       # the reference reloads the exported graph (models/deepfm.py:237-238)
        auc = self._auc(data_val)
        print("val of auc:%.5f" % auc)
        sys.stdout.flush()
        return auc

    # ---- generation B: fit(print_num_batch, predict_data) (models/deepfm_multi.py:262-334) --------------
    def _fit_b(self, print_num_batch: int, predict_data: Iterable) -> None:
        if self.model_restore == 1:
            try:
                self.restore(self.restore_model_checkpoint)
                print("@_@~ Old Model Restored Successfully!")
            except Exception as e:
                print("=_=!! Error: There is no model checkpoint in %s" % self.restore_model_checkpoint)
                print(e)
                exit(-1)
        val_data = self.get_val_data(None, predict_data)
        print("Start of training")
        start_time = time.time()
        batch_count, i = 0, 0
        for b, nb in _lookahead(self._decode(item) for item in self.data_dict):   # the end of the iterator replaces tf.errors.OutOfRangeError
            loss = self.engine.train_step(b, next_batch=nb)
            if batch_count == print_num_batch:
                batch_end_time = time.time()
                auc = self.eval(None, val_data)
                print("[{}] val_auc:{}\t loss:{} time:{:.2f}s".format(i, auc, float(loss.item()),
                                                                     batch_end_time - start_time))
                sys.stdout.flush()
                start_time = time.time()
                batch_count = 0
            batch_count += 1
            i += 1
        print("--------------End of dataset-------------")
        if self.save_model_checkpoint:
            print("--------------save checkpoint model-------------")
            self.save(self.save_model_checkpoint, with_optimizer=True)
        if self.model_pb:
            print("--------------save pb model-------------")
            try:
                self.save(self.model_pb)
            except Exception as e:
                print("Fail to export saved model, exception: {}".format(e))
                sys.stdout.flush()

    # ---- textline generation: fit(print_iter, num_batch_size) (models/wdl_textline.py:302-357,
    # models/dnn_multi_textline.py:174-228) ---------------------------------------------------------------
    def _fit_c(self, print_iter: int, num_batch_size: int) -> None:
        """Every pass of the loop first EVALUATES one batch of the iterator (`sess.run([label, out, loss])`: the iterator
        advances, no update) and then, unless it just printed, TRAINS on the next one (`sess.run([loss, optimizer])`).
        Every `print_iter`-th evaluated batch is kept for the AUC; after `num_batch_size` kept batches the AUC of those
        batches is printed with sum(losses) / (count * print_iter)."""
        it = iter(self.data_dict)
        pred_list, label_list, losses, count = [], [], [], 0
        start_time = time.time()
        i = 0
        while True:
            try:
                b = self._decode(next(it))
                losses.append(self.engine.eval_loss(b))
                if i % print_iter == 0 and i > 0:
                    count += 1
                    B = self.engine.last_batch
                    label_list.append(self.engine.label[:B].clone())
                    pred_list.append(self.engine.score[:B].clone())
                if count == num_batch_size:
                    total_loss = float(torch.stack(losses).sum().item() / (count * print_iter))
                    auc = self.engine.auc(torch.cat(pred_list), torch.cat(label_list))
                    end_time = time.time()
                    print("[{}] test auc:{}\t loss:{} time:{:.2f}s".format(i, auc, total_loss, end_time - start_time))
                    sys.stdout.flush()
                    start_time = time.time()
                    pred_list, label_list, losses, count = [], [], [], 0
                else:
                    self.engine.train_step(self._decode(next(it)))
            except StopIteration:      # replaces tf.errors.OutOfRangeError
                print("--------------End of Dataset-------------")
                sys.stdout.flush()
                if self.model_pb:
                    try:
                        self.save(self.model_pb)
                    except Exception as e:
                        print("Fail to export saved model, exception: {}".format(e))
                        sys.stdout.flush()
                break
            i += 1

    @staticmethod
    def get_val_data(sess, data: Iterable) -> List[Dict]:
        """drain the validation iterator into host memory (models/deepfm_multi.py:336-361)"""
        return list(data)

    def eval(self, sess, val_data: List[Dict]) -> float:
        return self._auc(val_data)
