"""B200 drop-in for the reference's models/wdl_textline.py (textline generation: iterator of batch dicts).

    m = DeepModel(mp, train_data); m.fit(print_iter=mp.print_iter, num_batch_size=mp.num_batch_size)
                                                                             # local_run_textline.py:186-189
    predict(pred_data, model_pb)                                             # local_run_textline.py:199-201

Differences from models/wdl.py that the engine reproduces: row-0-zero embedding table (:50) and the l2 term over the
GATHERED wdl_weights (:264-269 overwrite the variable with the lookup result before :284-286 regularise it).
"""
import sys

from ._base import BaseDeepModel, hparams_from_export


class DeepModel(BaseDeepModel):
    ALG = "wdl_textline"
    GENERATION = "C"

    def __init__(self, args, data_dict, **kw):
        super().__init__(args, data_dict, **kw)

    def fit(self, print_iter, num_batch_size):
        """models/wdl_textline.py:302-357"""
        self._fit_c(print_iter, num_batch_size)


def predict(data_out, model_pb, args=None, **kw):
    """models/wdl_textline.py:360-386: `data_out` is the pickled-batch list of utils/data_loader_load.py:94-139.  The
    export carries the hyper-parameter bag too, so `args` is optional (an override)."""
    if args is None:          # the export is self-describing: its own hyper-parameter bag (models/_base.py save)
        args = hparams_from_export(model_pb)
    m = DeepModel(args, None, **kw)
    m.restore(model_pb)
    auc = m._auc(list(data_out))
    print("val of auc:%.5f" % auc)
    sys.stdout.flush()
    print('---end---')
    return auc
