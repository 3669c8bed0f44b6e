"""B200 drop-in for the reference's models/dnn_pipeline.py (generation B: iterator of batch dicts).

    m = DeepModel(mp, train_data); m.fit(print_num_batch, predict_data)      # local_run_test.py:77-79
    predict(predict_data, model_pb)                                          # local_run_test.py:82
"""
import sys

from ._base import BaseDeepModel


class DeepModel(BaseDeepModel):
    ALG = "dnn_pipeline"
    GENERATION = "B"

    def __init__(self, args, data_dict, **kw):
        super().__init__(args, data_dict, **kw)

    def fit(self, print_num_batch, predict_data):
        """models/dnn_pipeline.py fit()"""
        self._fit_b(print_num_batch, predict_data)


def predict(predict_data, model_pb, args=None, **kw):
    """models/dnn_pipeline.py predict().  The reference reloads the exported graph, which carries its own shapes; the npz
    export here carries only variables, so the hyper-parameter bag is passed as `args`."""
    if args is None:
        raise ValueError("predict(predict_data, model_pb, args=mp): the hyper-parameter bag is required")
    m = DeepModel(args, None, **kw)
    m.restore(model_pb)
    auc = m._auc(list(predict_data))
    print("-----------end of data_set-----------")
    print("val of auc:%.5f" % auc)
    sys.stdout.flush()
    print('---end---')
    return auc
