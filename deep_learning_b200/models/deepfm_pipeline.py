"""B200 drop-in for the reference's models/deepfm_pipeline.py (generation B: iterator of batch dicts).

    m = DeepModel(mp, train_data); m.fit(print_num_batch, predict_data)      # local_run_test.py:77-79
    predict(predict_data, model_pb, args=mp)                                 # local_run_test.py:82

Quirk Q2 is reproduced: the FM part reads row id + cont_field_size, the deep part the raw id (split_engine.py).
"""
import sys

from ._base import BaseDeepModel, hparams_from_export


class DeepModel(BaseDeepModel):
    ALG = "deepfm_pipeline"
    GENERATION = "B"

    def __init__(self, args, data_dict, **kw):
        super().__init__(args, data_dict, **kw)

    def fit(self, print_num_batch, predict_data):
        """models/deepfm_pipeline.py:193-262"""
        self._fit_b(print_num_batch, predict_data)


def predict(predict_data, model_pb, args=None, **kw):
    """models/deepfm_pipeline.py:320-352"""
    if args is None:          # the export is self-describing: its own hyper-parameter bag (models/_base.py save)
        args = hparams_from_export(model_pb)
    m = DeepModel(args, None, **kw)
    m.restore(model_pb)
    auc = m._auc(list(predict_data))
    print("-----------end of data_set-----------")
    print("val of auc:%.5f" % auc)
    sys.stdout.flush()
    print('---end---')
    return auc
