"""B200 drop-in for the reference's test_model/wdl.py (generation A: pickled batches): Wide&Deep with
the deep-projection rows BEHIND the wide ids (test_model/wdl.py:110), tf.train.FtrlOptimizer (:136) and
tf.nn.sigmoid_cross_entropy_with_logits applied to the already-sigmoided value (:119-121) -- reproduced as written.

    m = DeepModel(mp); m.fit(train_data, predict_data); m.predict(data_val)

`mp` carries the attributes that file reads (:17-32): hidden_units, epochs, batch_size, learning_rate, model_pb,
learning_rate_decay_steps / _rate, l2_reg, cont_field_size, cate_field_size, wide_field_size, cate_feats_size,
embedding_size.  Same kernels as models/wdl.py; the optimizer kernels run their FTRL element update
(ctr_adam_hp.kind = CTR_OPT_FTRL) and the head kernel its second loss (CTR_HEAD_SCE_ON_SCORE).
"""
from ._base import BaseDeepModel


class DeepModel(BaseDeepModel):
    ALG = "wdl_ftrl"
    GENERATION = "A"

    def __init__(self, args, **kw):
        kw.setdefault("adam_mode", "lazy")       # FTRL's table update is the sparse apply: looked-up rows only
        super().__init__(args, None, **kw)

    def fit(self, train_data, predict_data):
        """test_model/wdl.py:139-186"""
        self._fit_a(train_data, predict_data)

    def evaluate(self, sess, data_val):
        """test_model/wdl.py:188-203"""
        return self._auc(data_val)

    def predict(self, data_val):
        """test_model/wdl.py:205-228"""
        return self._predict_a(data_val)
