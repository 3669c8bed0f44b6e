"""B200 drop-in for the reference's models/dnn_multi_textline.py (textline generation: iterator of batch dicts with
`cat_feats`, `mul_cat_feats`, `mul_cat_feats_value`, `cont_feats`, `labels`; utils/data_loader_textline.py:30-34).

    m = DeepModel(mp, train_data); m.fit(print_iter=mp.print_iter, num_batch_size=mp.num_batch_size)
                                                                             # local_run_textline.py:186-189

The bag's attribute spellings of that file are accepted (cat_field_size, cat_index_size, word2vec_embedding_path);
`mul_cat_index_range` / `multi_feats_range` = [[start, end, feat_name], ...] (the reference calls a missing
my_utils.get_mul_index_range for it, models/dnn_multi_textline.py:34).  The field named 'tag' looks its ids up in the
frozen word2vec table (`np.load(word2vec_embedding_path)`, :46-48), whose width must equal embedding_size here.
"""
import sys

import numpy as np

from ._base import BaseDeepModel, hparams_from_export, init_weights


class DeepModel(BaseDeepModel):
    ALG = "dnn_multi_textline"
    GENERATION = "C"

    def __init__(self, args, data_dict, **kw):
        self._tags = None
        path = getattr(args, "word2vec_embedding_path", "")
        if path:
            self._tags = np.load(path).astype(np.float32)
            if self._tags.ndim != 2 or self._tags.shape[1] != int(args.embedding_size):
                raise ValueError("word2vec table %s is %s; its width must equal embedding_size=%d"
                                 % (path, self._tags.shape, int(args.embedding_size)))
            args.word2vec_rows = int(self._tags.shape[0])
        if kw.get("weights") is None:
            from ..spec import spec_from_hparams
            w = init_weights(spec_from_hparams(self.ALG, args), 2019)
            if self._tags is not None:
                w["tags_emb"] = self._tags
            kw["weights"] = w
        super().__init__(args, data_dict, **kw)

    def fit(self, print_iter, num_batch_size):
        """models/dnn_multi_textline.py:174-228"""
        self._fit_c(print_iter, num_batch_size)


def predict(data_out, model_pb, args=None, **kw):
    """models/dnn_multi_textline.py:231-256"""
    if args is None:          # the export is self-describing: its own hyper-parameter bag (models/_base.py save)
        args = hparams_from_export(model_pb)
    m = DeepModel(args, None, **kw)
    m.restore(model_pb)
    auc = m._auc(list(data_out))
    print("val of auc:%.5f" % auc)
    sys.stdout.flush()
    print('---end---')
    return auc
