"""B200 drop-in for the reference's models/deepfm_multi.py (generation B: iterator of batch dicts).

    m = DeepModel(mp, train_data); m.fit(print_num_batch, predict_data)      # local_run_test.py:77-79
    predict(predict_data, model_pb)                                          # local_run_test.py:82
"""
import sys

from ._base import BaseDeepModel, hparams_from_export


class DeepModel(BaseDeepModel):
    ALG = "deepfm_multi"
    GENERATION = "B"

    def __init__(self, args, data_dict, **kw):
        super().__init__(args, data_dict, **kw)

    def fit(self, print_num_batch, predict_data):
        """models/deepfm_multi.py:262-334"""
        self._fit_b(print_num_batch, predict_data)


def predict(predict_data, model_pb, args=None, **kw):
    """models/deepfm_multi.py:383-415.  The reference reloads the exported graph, which carries its own shapes; the
    export here carries the hyper-parameter bag too, so `args` is optional (an override)."""
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
