"""B200 drop-in for the reference's models/deepfm.py (generation A: feed_dict / pickled batches).

    m = DeepModel(mp); m.fit(train_data, val_data); m.predict(pred_data)     # local_run_textline.py:182-184
"""
from ._base import BaseDeepModel


class DeepModel(BaseDeepModel):
    ALG = "deepfm"
    GENERATION = "A"

    def __init__(self, args, **kw):
        super().__init__(args, None, **kw)

    def fit(self, train_data, val_data):
        """models/deepfm.py:164-214"""
        self._fit_a(train_data, val_data)

    def evaluate(self, sess, data_val):
        """models/deepfm.py:216-231 (ranks by logit; AUC is identical under the monotone sigmoid)"""
        return self._auc(data_val)

    def predict(self, data_val):
        """models/deepfm.py:233-255"""
        return self._predict_a(data_val)
