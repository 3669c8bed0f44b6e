"""B200 drop-in for the reference's test_model/DCN.py (generation A: pickled batches of dense features).

    m = DeepModel(parser); m.fit(train_data, val_data); m.predict(data_val)

`parser` carries what that file reads (:16-25): deep_feature_size, hidden_units, epochs, batch_size, learning_rate,
model_pb, learning_rate_decay_steps / _rate, l2_reg.  Batches: pickle.dumps({"features": [B][D], "labels": [B][1]}) (:131-135).
"""
import json
import os
import pickle
import sys

import numpy as np
import torch

from ..dcn_engine import DcnEngine
from ._base import BaseDeepModel, EXPORT_FORMAT, hparams_to_dict


class DeepModel(BaseDeepModel):
    ALG = "dcn"
    GENERATION = "A"

    def __init__(self, parser, *, device=None, weights=None, max_batch=None):
        self.args = parser
        self.deep_feature_size, self.hidden_units = int(parser.deep_feature_size), list(parser.hidden_units)
        self.epochs, self.batch_size = getattr(parser, "epochs", 1), int(parser.batch_size)
        self.learning_rate, self.model_pb = parser.learning_rate, getattr(parser, "model_pb", "")
        self.decay_steps, self.decay_rate = parser.learning_rate_decay_steps, parser.learning_rate_decay_rate
        self.l2_reg, self.metric_type, self.random_seed = parser.l2_reg, "auc", 2019
        self.cross_layer_num = 2                                             # test_model/DCN.py:28
        self.engine = DcnEngine(self.deep_feature_size, self.hidden_units, max_batch or self.batch_size, device=device,
                                learning_rate=self.learning_rate, decay_steps=self.decay_steps, decay_rate=self.decay_rate,
                                l2_reg=self.l2_reg, cross_layers=self.cross_layer_num)
        self.engine.set_weights(weights if weights is not None else self._init_weights())

    def _init_weights(self):
        """test_model/DCN.py:49-99: N(0, 2/(fan_in+fan_out)) for every matrix and bias, concat_bias = 0.01"""
        g, D, p, prev = np.random.RandomState(self.random_seed), self.deep_feature_size, {}, self.deep_feature_size
        for i, h in enumerate(self.hidden_units):
            std = np.sqrt(2.0 / (prev + h))
            p["deep_layer_%d" % i], p["deep_layer_bias_%d" % i] = g.normal(0, std, (prev, h)), g.normal(0, std, (1, h))
            prev = h
        for i in range(self.cross_layer_num):
            std = np.sqrt(2.0 / (2 * D))
            p["cross_layer_%d" % i], p["cross_bias_%d" % i] = g.normal(0, std, (D, 1)), g.normal(0, std, (D, 1))
        p["concat_projection"] = g.normal(0, np.sqrt(2.0 / (D + prev + 1)), (D + prev, 1))
        p["concat_bias"] = np.float32(0.01)
        return {k: np.asarray(v, dtype=np.float32) for k, v in p.items()}

    def _decode(self, item):
        if isinstance(item, (bytes, bytearray)):
            item = pickle.loads(bytes(item))
        return {"features": np.asarray(item["features"], dtype=np.float32), "label": np.asarray(item["labels"] if "labels" in item else item["label"], dtype=np.float32)}

    def save(self, path, with_optimizer=False):
        """self-describing like the other exports; signature of test_model/DCN.py:141-147, scored by "Sigmoid:0" (:190)"""
        os.makedirs(path, exist_ok=True)
        blob = {"w/" + k: v for k, v in self.engine.get_weights().items()}
        sig = {"method_name": "tensorflow/serving/predict",
               "inputs": {"feature_values": {"tensor": "feature_values:0", "dtype": "float32", "width": self.deep_feature_size},
                          "input_data_size": {"tensor": "input_data_size:0", "dtype": "int32", "width": None}},
               "outputs": {"output": {"tensor": "Sigmoid:0", "dtype": "float32", "width": 1}}}
        blob["meta/json"] = np.array(json.dumps({"format": EXPORT_FORMAT, "alg": self.ALG, "generation": self.GENERATION,
                                                 "hparams": hparams_to_dict(self.args), "signature": sig}))
        np.savez(os.path.join(path, "model.npz"), **blob)

    def restore(self, path):
        z = np.load(os.path.join(path, "model.npz"))
        self.engine.set_weights({k[2:]: z[k] for k in z.files if k.startswith("w/")})

    def fit(self, trian_data, val_data):
        """test_model/DCN.py:118-160"""
        self._fit_a(trian_data, val_data)

    def evaluate(self, sess, data_val):
        """test_model/DCN.py:162-175"""
        return self._auc(data_val)

    def predict(self, data_val):
        """test_model/DCN.py:177-197 (prints the AUC, no '---end---' line there)"""
        if self.model_pb and os.path.exists(os.path.join(self.model_pb, "model.npz")):
            self.restore(self.model_pb)
        auc = self._auc(data_val)
        print("val of auc:%.5f" % auc)
        sys.stdout.flush()
        return auc
