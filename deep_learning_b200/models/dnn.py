"""B200 drop-in for the reference's models/dnn.py (generation A: feed_dict / pickled batches).

    m = DeepModel(mp); m.fit(train_data, val_data); m.predict(pred_data)     # local_run_textline.py:182-184
"""
from ._base import BaseDeepModel


class DeepModel(BaseDeepModel):
    ALG = "dnn"
    GENERATION = "A"

    def __init__(self, args, **kw):
        super().__init__(args, None, **kw)

    def fit(self, train_data, val_data):
        """models/dnn.py:96-148"""
        self._fit_a(train_data, val_data)

    def evaluate(self, sess, data_val):
        """models/dnn.py:150-165"""
        return self._auc(data_val)

    def predict(self, data_val):
        """models/dnn.py:167-185"""
        return self._predict_a(data_val)
