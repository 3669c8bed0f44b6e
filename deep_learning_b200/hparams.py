"""The reference's hyper-parameter bag (`mp` / `args`): a plain attribute holder with the attribute names the
model files read (models/deepfm.py:17-33, models/deepfm_multi.py:17-39, models/wdl.py:19-39,
local_run_test.py:17-40, local_run_textline.py:17-67).  Launchers build one of these and hand it to
`DeepModel(mp, ...)`; defaults are the launchers' hard-coded values."""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List


@dataclass
class ModelParams:
    hidden_units: List[int] = field(default_factory=lambda: [512, 256, 128])   # local_run.py:28-35
    epochs: int = 1
    batch_size: int = 256
    learning_rate: float = 0.001
    learning_rate_decay_steps: int = 10000000
    learning_rate_decay_rate: float = 0.9
    l2_reg: float = 0.00001
    embedding_size: int = 8
    cont_field_size: int = 0
    vector_feats_size: int = 0
    cate_field_size: int = 0
    cate_feats_size: int = 0
    cate_index_size: int = 0            # models/wdl.py:33
    wide_field_size: int = 0            # models/wdl.py:39
    multi_feats_size: int = 0           # sum of top_n over the multi-valued fields
    multi_field_size: int = 0
    multi_feats_range: List[list] = field(default_factory=list)   # [[start, end, name], ...] utils/my_utils.py:60-65
    model_pb: str = ""
    save_model_checkpoint: str = ""
    restore_model_checkpoint: str = ""
    model_restore: int = 0
    cate_feats_conf_path: str = ""
    alg_name: str = "deepfm"
    action_type: str = "train"
    num_batch_size: int = 100           # print_num_batch (local_run_test.py:79)
    print_iter: int = 10                # textline fit loops (local_run_textline.py:189)
    word2vec_embedding_path: str = ""   # models/dnn_multi_textline.py:32
    word2vec_rows: int = 0              # rows of that table when the weights are injected instead of loaded
    max_data_size: int = 1 << 62        # utils/data_loader_load.py:45
    vector_field_size: int = 0          # utils/data_loader_load.py:109 (its spelling of vector_feats_size)
    # the rest of what the launchers' own ModelParams classes carry (local_run.py:17-38, local_run_test.py:17-45,
    # local_run_textline.py:17-67), so a launcher can fill this bag attribute for attribute
    dropout_keep_deep: List[float] = field(default_factory=lambda: [1, 1, 1, 1, 1])   # local_run.py:30 (never below 1 there)
    feat_conf_path: str = ""            # local_run.py:20
    cont_feats_conf_path: str = ""      # local_run_textline.py:60
    train_path: str = ""                # local_run.py:21
    predict_path: str = ""              # local_run.py:22
    pred_path: str = ""                 # local_run_textline.py:58
    multi_cate_field_size: int = 0      # local_run_textline.py:61; width of the libsvm column (utils/data_loader_load.py:73-78)
