"""CPU oracle for the CTR-zoo train/predict hot path.  TEST INFRASTRUCTURE ONLY.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s `cpu_baseline` / `--impl reference`
legs may import this module; the product (`deep_learning_b200/`) never does.

PARITY: the reference (RodJohn/deep_learning) ships no tests, golden vectors or fixtures for this
path, and its arithmetic lives in TensorFlow 1.x (un-vendored, unpinned, 1.13-1.15 by the .pyc
tags) which cannot be installed here (no wheel for py3.12, no network): against TensorFlow's own
kernels this file is PARITY UNPINNED.  What it IS pinned to: all 13 files of the reference's
`models/`, imported unmodified and run through their own `fit` under a TF1 graph-mode stand-in
(tests/golden/tf1_shim.py, generator tests/golden/make_ref_golden.py); this file reproduces the
recorded losses, evaluation outputs and final variables (tests/golden/ref_*.npz) to 1e-12
(tests/test_ref_golden.py).  The file RESTATES the TF graph of each model file op by op on
PyTorch-CPU tensors (`tf.nn.embedding_lookup` -> index, `tf.matmul` -> `@`, ...), takes gradients
with autograd on that same graph (what `Optimizer.minimize` does), and hand-writes TF1's Adam.
scikit-learn's `roc_auc_score` (its real implementation is present) pins `roc_auc`.

Every function cites the reference lines it follows (paths relative to /root/reference).
All functions take a `dtype`: float64 is the order-free truth used by the parity tests,
float32 matches TF's arithmetic type and is what the CPU baseline times.
"""
from __future__ import annotations

import math
import threading
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

Tensor = torch.Tensor

LOG_LOSS_EPS = 1e-7          # tf.losses.log_loss default epsilon (models/deepfm.py:151)
ADAM_BETA1, ADAM_BETA2, ADAM_EPS = 0.9, 0.999, 1e-8   # tf.train.AdamOptimizer defaults (models/deepfm.py:161)


# --------------------------------------------------------------------------------------
# hyper-parameter bag: attribute names are the reference's (models/deepfm.py:17-33,
# models/deepfm_multi.py:17-39, models/wdl.py:19-39, local_run_test.py:17-40)
# --------------------------------------------------------------------------------------
@dataclass
class HParams:
    hidden_units: List[int] = field(default_factory=lambda: [256, 128])
    epochs: int = 1
    batch_size: int = 256
    learning_rate: float = 0.001
    learning_rate_decay_steps: int = 10000000
    learning_rate_decay_rate: float = 0.9
    l2_reg: float = 0.00001
    embedding_size: int = 8
    cont_field_size: int = 10
    vector_feats_size: int = 0
    cate_field_size: int = 20
    cate_feats_size: int = 10000          # a.k.a. cate_index_size
    cate_index_size: int = 0              # wdl.py reads this name (wdl.py:33)
    wide_field_size: int = 0              # wdl.py:39
    multi_feats_size: int = 0             # sum of top_n over multi-valued fields (M)
    multi_field_size: int = 0             # number of multi-valued fields (Fm)
    multi_feats_range: List[list] = field(default_factory=list)   # [[start, end, name], ...] (utils/my_utils.py:60-65)
    model_pb: str = ""
    save_model_checkpoint: str = ""
    restore_model_checkpoint: str = ""
    model_restore: int = 0
    cate_feats_conf_path: str = ""
    word2vec_rows: int = 0                # rows of the frozen `tags_emb` constant (models/dnn_multi_textline.py:47-48)
    deep_feature_size: int = 0            # test_model/DCN.py:16: width of the dense feature vector


# --------------------------------------------------------------------------------------
# TF op restatements
# --------------------------------------------------------------------------------------
def embedding_lookup(table: Tensor, ids: Tensor) -> Tensor:
    """tf.nn.embedding_lookup(params, ids): rows of `params` at `ids`, shape ids.shape + [width]."""
    return table[ids.long()]


def zero_row0(table: Tensor) -> Tensor:
    """tf.concat((tf.zeros([1, w]), table[1:]), 0) -- models/deepfm_multi.py:130-133,
    models/dnn_multi.py:74.  Row 0 reads as zero and is cut out of the gradient graph."""
    return torch.cat((torch.zeros(1, table.shape[1], dtype=table.dtype), table[1:]), 0)


def nonzero_reduce_mean(emb: Tensor) -> Tensor:
    """models/deepfm_multi.py:83-88 (same arithmetic at models/dnn_multi.py:93-96).
    emb [B, top_n, w].  count = #{t : sum_w emb[b,t,:] != 0} (count_nonzero has no gradient),
    result = sum_t emb / count with tf.div_no_nan (0 where count == 0)."""
    axis_2_sum = emb.sum(dim=2)
    nonzero = (axis_2_sum != 0).sum(dim=1, keepdim=True).to(emb.dtype)   # non-differentiable
    multi_sum = emb.sum(dim=1)
    safe = torch.where(nonzero == 0, torch.ones_like(nonzero), nonzero)
    return torch.where(nonzero == 0, torch.zeros_like(multi_sum), multi_sum / safe)


# ReLU is continuous but its derivative is not: a pre-activation within rounding of zero can land on either
# side in a reduced-precision (TF32) run.  Tests that check GRADIENTS of such a run pin the branch each unit
# took (`with relu_branches(masks)`), so the comparison is same-branch and the tolerance measures rounding,
# not the jump of d relu at 0.  Forward values (logit / score / loss) are always compared with the true relu.
# per THREAD: the loopback multi-GPU tests run N ranks as N threads of one process, each calling this module on its own
# batch; a process-wide override would leak one rank's masks into another rank's forward
_RELU = threading.local()


class relu_branches:
    """context manager: masks[i] is a bool [B, H_i] tensor, True where unit (b, j) of layer i is active"""

    def __init__(self, masks):
        self.masks = masks

    def __enter__(self):
        self.prev = getattr(_RELU, "masks", None)
        _RELU.masks = self.masks
        return self

    def __exit__(self, *exc):
        _RELU.masks = self.prev
        return False


def mlp_tower(x: Tensor, p: Dict[str, Tensor], n_layers: int) -> Tensor:
    """models/deepfm.py:121-126: dropout(keep=1) is the identity; relu(x @ W_i + b_i)."""
    for i in range(n_layers):
        z = x @ p['deep_%d' % i] + p['deep_bias_%d' % i]
        masks = getattr(_RELU, "masks", None)
        x = torch.relu(z) if masks is None else z * masks[i].to(z.dtype)
    return x


def log_loss(labels: Tensor, predictions: Tensor, eps: float = LOG_LOSS_EPS) -> Tensor:
    """tf.losses.log_loss(labels, predictions) with default epsilon and SUM_BY_NONZERO_WEIGHTS
    (weights=1) == mean over all elements (models/deepfm.py:151-152; the extra reduce_mean is a
    no-op on a scalar)."""
    losses = -labels * torch.log(predictions + eps) - (1 - labels) * torch.log(1 - predictions + eps)
    return losses.mean()


def l2_regularizer(scale: float, w: Tensor) -> Tensor:
    """tf.contrib.layers.l2_regularizer(scale)(w) = scale * tf.nn.l2_loss(w) = scale * sum(w^2) / 2."""
    return scale * (w * w).sum() / 2


def l1_regularizer(scale: float, w: Tensor) -> Tensor:
    """tf.contrib.layers.l1_regularizer(scale)(w) = scale * sum(|w|)  (models/dnn.py:90)."""
    return scale * w.abs().sum()


def exponential_decay(lr0: float, global_step: int, decay_steps: int, decay_rate: float) -> float:
    """tf.train.exponential_decay(..., staircase=True) evaluated on the PRE-increment global_step
    (models/deepfm.py:159-162)."""
    return lr0 * decay_rate ** math.floor(global_step / decay_steps)


# --------------------------------------------------------------------------------------
# Model forward graphs.  Each returns (logit [B,1], score [B,1]).
# `batch` keys: label [B,1], cont_feats [B,C], vector_feats [B,V], cate_feats [B,Fc(+M)],
# wide_feats [B,W] -- the reference's batch-dict keys (utils/data_loader.py:12-24,
# utils/data_loader_load.py:130-137).
# --------------------------------------------------------------------------------------
def deepfm_forward(p: Dict[str, Tensor], batch: Dict[str, Tensor], hp: HParams):
    """models/deepfm.py:56-145."""
    C, Fc, k = hp.cont_field_size, hp.cate_field_size, hp.embedding_size
    cate, cont, vec = batch['cate_feats'].long(), batch['cont_feats'], batch['vector_feats']
    B = cont.shape[0]
    dt = cont.dtype
    # :63-75  NB offset is cate_FIELD_size (quirk Q1): cont rows alias categorical ids Fc..Fc+C-1
    cont_index_add = (torch.arange(C).unsqueeze(0).expand(B, C) + Fc)
    idx = torch.cat([cate, cont_index_add], dim=1)
    val = torch.cat([torch.ones(B, Fc, dtype=dt), cont], dim=1)
    F = Fc + C
    # :78-81 first order
    first = embedding_lookup(p['feats'], idx).reshape(B, F) * val
    # :85-93 second order
    e = embedding_lookup(p['feats_emb'], idx) * val.reshape(B, F, 1)
    sum_e = e.sum(1)
    second = 0.5 * (sum_e * sum_e - (e * e).sum(1))
    # :98-100 deep part
    cat_emb = embedding_lookup(p['feats_emb'], cate).reshape(B, Fc * k)
    dense_vector = torch.cat([cont, vec, cat_emb], dim=1)
    deep = mlp_tower(dense_vector, p, len(hp.hidden_units))
    # :128-145 head
    feats_input = torch.cat([first, second, deep], dim=1)
    out = feats_input @ p['deep_fm_weight'] + p['deep_fm_bias']
    return out, torch.sigmoid(out)


def deepfm_loss(p, batch, hp: HParams):
    """models/deepfm.py:147-156."""
    out, score = deepfm_forward(p, batch, hp)
    loss = log_loss(batch['label'], score)
    if hp.l2_reg > 0:
        loss = loss + l2_regularizer(hp.l2_reg, p['deep_fm_weight'])
    return loss, out, score


def _multi_cate_emb(p_first: Optional[Tensor], p_emb: Tensor, multi: Tensor, ranges):
    """models/deepfm_multi.py:90-121 / models/dnn_multi.py:83-104: loop over multi_feats_range,
    lookup, nonzero_reduce_mean, concat along axis 1."""
    firsts, embs = [], []
    for item in ranges:
        s, e = int(item[0]), int(item[1])
        if p_first is not None:
            firsts.append(nonzero_reduce_mean(embedding_lookup(p_first, multi[:, s:e])))
        embs.append(nonzero_reduce_mean(embedding_lookup(p_emb, multi[:, s:e])))
    all_first = torch.cat(firsts, dim=1) if firsts else None
    return all_first, torch.cat(embs, dim=1)


def deepfm_multi_forward(p, batch, hp: HParams):
    """models/deepfm_multi.py:123-243."""
    C, Fc, k, Fm = hp.cont_field_size, hp.cate_field_size, hp.embedding_size, hp.multi_field_size
    cate, cont, vec = batch['cate_feats'].long(), batch['cont_feats'], batch['vector_feats']
    B = cont.shape[0]
    single, multi = cate[:, :Fc], cate[:, Fc:]                          # :66-67
    feats_emb = zero_row0(p['feats_emb'])                               # :130-131
    fm_first = zero_row0(p['fm_first_order_emb'])                       # :132-133
    all_first_multi, all_multi = _multi_cate_emb(fm_first, feats_emb, multi, hp.multi_feats_range)
    cont_index_add = torch.arange(C).unsqueeze(0).expand(B, C) + hp.cate_feats_size   # :139
    # :143-152 first order  [cont*w | single w | pooled w]
    first_cont = embedding_lookup(fm_first, cont_index_add).reshape(B, C) * cont
    first_single = embedding_lookup(fm_first, single).reshape(B, Fc)
    first = torch.cat([first_cont, first_single, all_first_multi], dim=1)
    # :157-172 second order over [cont | single | pooled]
    second_cont = embedding_lookup(feats_emb, cont_index_add) * cont.reshape(B, C, 1)
    second_single = embedding_lookup(feats_emb, single)
    second_multi = all_multi.reshape(B, Fm, k)
    e = torch.cat([second_cont, second_single, second_multi], dim=1)
    sum_e = e.sum(1)
    second = 0.5 * (sum_e * sum_e - (e * e).sum(1))
    fm_res = torch.cat([first, second], dim=1)
    # :181-222 deep part
    single_emb = embedding_lookup(feats_emb, single).reshape(B, Fc * k)
    dense_vector = torch.cat([cont, vec, single_emb, all_multi], dim=1)
    deep = mlp_tower(dense_vector, p, len(hp.hidden_units))
    feats_input = torch.cat([fm_res, deep], dim=1)
    out = feats_input @ p['deep_fm_weight'] + p['deep_fm_bias']
    return out, torch.sigmoid(out)


def deepfm_multi_loss(p, batch, hp: HParams):
    """models/deepfm_multi.py:245-251."""
    out, score = deepfm_multi_forward(p, batch, hp)
    loss = log_loss(batch['label'], score)
    if hp.l2_reg > 0:
        loss = loss + l2_regularizer(hp.l2_reg, p['deep_fm_weight'])
    return loss, out, score


def wdl_forward(p, batch, hp: HParams):
    """models/wdl.py:101-253 with use_wide = use_deep = True."""
    C, Fc, k, W = hp.cont_field_size, hp.cate_field_size, hp.embedding_size, hp.wide_field_size
    cate, cont, wide = batch['cate_feats'].long(), batch['cont_feats'], batch['wide_feats'].long()
    B = cont.shape[0]
    H = hp.hidden_units[-1]
    cat_emb = embedding_lookup(p['weight_mat'], cate).reshape(B, Fc * k)       # :132-133
    dense_vector = torch.cat([cont, cat_emb], dim=1)                           # :179
    deep = mlp_tower(dense_vector, p, len(hp.hidden_units))                    # :207-212
    # :221-225 + :240-246.  NB offset is wide_feats_FIELD_size (quirk Q4): the deep projection
    # weights alias wide ids W..W+H-1 of the same [cate_index_size+H, 1] table.
    deep_index_add = torch.arange(H).unsqueeze(0).expand(B, H) + W
    idx = torch.cat([wide, deep_index_add], dim=1)
    val = torch.cat([torch.ones(B, W, dtype=cont.dtype), deep], dim=1)
    w = embedding_lookup(p['wdl_weights'], idx).reshape(B, W + H)
    wx = (w * val).sum(dim=1, keepdim=True)
    out = wx + p['wdl_bias']
    return out, torch.sigmoid(out)


def wdl_loss(p, batch, hp: HParams):
    """models/wdl.py:255-275."""
    out, score = wdl_forward(p, batch, hp)
    loss = log_loss(batch['label'], score)
    if hp.l2_reg > 0:
        loss = loss + l2_regularizer(hp.l2_reg, p['wdl_weights'])
        for i in range(len(hp.hidden_units)):
            loss = loss + l2_regularizer(hp.l2_reg, p['deep_%d' % i])
    return loss, out, score


def dnn_forward(p, batch, hp: HParams):
    """models/dnn.py:35-84.  cont_feats here is already [cont | vector] (dnn.py:27)."""
    Fc, k = hp.cate_field_size, hp.embedding_size
    cate, cont = batch['cate_feats'].long(), batch['cont_feats']
    if batch.get('vector_feats') is not None and batch['vector_feats'].shape[1] > 0:
        cont = torch.cat([cont, batch['vector_feats']], dim=1)
    B = cont.shape[0]
    cate_emb = embedding_lookup(p['weight_mat'], cate).reshape(B, Fc * k)      # :54-55
    dense_vector = torch.cat([cont, cate_emb], dim=1)                          # :57
    deep = mlp_tower(dense_vector, p, len(hp.hidden_units))
    out = deep @ p['deep_res'] + p['deep_res_bias']                            # :82
    return out, torch.sigmoid(out)


def dnn_loss(p, batch, hp: HParams):
    """models/dnn.py:84-90 -- NB the regulariser is L1 (quirk Q5)."""
    out, score = dnn_forward(p, batch, hp)
    loss = log_loss(batch['label'], score)
    if hp.l2_reg > 0:
        for i in range(len(hp.hidden_units)):
            loss = loss + l1_regularizer(hp.l2_reg, p['deep_%d' % i])
    return loss, out, score


def dnn_multi_forward(p, batch, hp: HParams):
    """models/dnn_multi.py:70-154."""
    Fc, k = hp.cate_field_size, hp.embedding_size
    cate, cont, vec = batch['cate_feats'].long(), batch['cont_feats'], batch['vector_feats']
    B = cont.shape[0]
    feats_emb = zero_row0(p['feats_emb'])                                      # :74
    cate_emb = embedding_lookup(feats_emb, cate[:, :Fc]).reshape(B, Fc * k)    # :79-81
    _, all_multi = _multi_cate_emb(None, feats_emb, cate[:, Fc:], hp.multi_feats_range)  # :83-104
    dense_vector = torch.cat([cont, vec, cate_emb, all_multi], dim=1)          # :107
    deep = mlp_tower(dense_vector, p, len(hp.hidden_units))
    out = deep @ p['deep_res'] + p['deep_res_bias']                            # :151
    return out, torch.sigmoid(out)


def dnn_multi_loss(p, batch, hp: HParams):
    """models/dnn_multi.py:156-162."""
    out, score = dnn_multi_forward(p, batch, hp)
    loss = log_loss(batch['label'], score)
    if hp.l2_reg > 0:
        loss = loss + l2_regularizer(hp.l2_reg, p['deep_res'])
    return loss, out, score


# ---- the variants local_run_test.py dispatches: restated from their own files, not derived from
# the functions above, so that the "same math, different field mix" claim is itself checked by the parity tests
def deepfm_cate_forward(p, batch, hp: HParams):
    """models/deepfm_cate.py:72-171: categorical fields only (value 1), row-0-zero tables, deep input [vector | emb]."""
    Fc, k = hp.cate_field_size, hp.embedding_size
    cate, vec = batch['cate_feats'].long()[:, :Fc], batch['vector_feats']
    B = cate.shape[0]
    feats_emb = zero_row0(p['feats_emb'])                                    # :79-80
    fm_first = zero_row0(p['fm_first_order_emb'])                            # :81-82
    first = embedding_lookup(fm_first, cate).reshape(B, Fc)                  # :90-92 (times ones)
    e = embedding_lookup(feats_emb, cate)                                    # :97-99
    sum_e = e.sum(1)
    second = 0.5 * (sum_e * sum_e - (e * e).sum(1))                          # :100-104
    fm_res = torch.cat([first, second], dim=1)                               # :109
    cat_emb = embedding_lookup(feats_emb, cate).reshape(B, Fc * k)           # :115-116
    dense_vector = torch.cat([vec, cat_emb], dim=1)                          # :118
    deep = mlp_tower(dense_vector, p, len(hp.hidden_units))                  # :143-148
    feats_input = torch.cat([fm_res, deep], dim=1)                           # :152-153
    out = feats_input @ p['deep_fm_weight'] + p['deep_fm_bias']              # :166
    return out, torch.sigmoid(out)


def deepfm_multi_cate_forward(p, batch, hp: HParams):
    """models/deepfm_multi_cate.py:113-223: single + pooled multi-valued categorical fields, no continuous block."""
    Fc, k, Fm = hp.cate_field_size, hp.embedding_size, hp.multi_field_size
    cate, vec = batch['cate_feats'].long(), batch['vector_feats']
    B = cate.shape[0]
    single, multi = cate[:, :Fc], cate[:, Fc:]
    feats_emb = zero_row0(p['feats_emb'])
    fm_first = zero_row0(p['fm_first_order_emb'])
    all_first_multi, all_multi = _multi_cate_emb(fm_first, feats_emb, multi, hp.multi_feats_range)   # :71-111
    first = torch.cat([embedding_lookup(fm_first, single).reshape(B, Fc), all_first_multi], dim=1)
    e = torch.cat([embedding_lookup(feats_emb, single), all_multi.reshape(B, Fm, k)], dim=1)
    sum_e = e.sum(1)
    second = 0.5 * (sum_e * sum_e - (e * e).sum(1))
    fm_res = torch.cat([first, second], dim=1)
    single_emb = embedding_lookup(feats_emb, single).reshape(B, Fc * k)
    dense_vector = torch.cat([vec, single_emb, all_multi], dim=1)            # :169
    deep = mlp_tower(dense_vector, p, len(hp.hidden_units))
    out = torch.cat([fm_res, deep], dim=1) @ p['deep_fm_weight'] + p['deep_fm_bias']
    return out, torch.sigmoid(out)


def _deepfm_variant_loss(fwd):
    def loss_fn(p, batch, hp: HParams):
        """log-loss + l2 on deep_fm_weight (models/deepfm_cate.py:174-180, models/deepfm_multi_cate.py:226-232)"""
        out, score = fwd(p, batch, hp)
        loss = log_loss(batch['label'], score)
        if hp.l2_reg > 0:
            loss = loss + l2_regularizer(hp.l2_reg, p['deep_fm_weight'])
        return loss, out, score
    return loss_fn


def dnn_pipeline_forward(p, batch, hp: HParams):
    """models/dnn_pipeline.py:68-122 (with cont_field_size = 0: models/dnn_cate.py:62-115)."""
    Fc, k = hp.cate_field_size, hp.embedding_size
    cate, cont, vec = batch['cate_feats'].long()[:, :Fc], batch['cont_feats'], batch['vector_feats']
    B = cate.shape[0]
    feats_emb = zero_row0(p['feats_emb'])                                    # dnn_pipeline.py:72-73
    cat_emb = embedding_lookup(feats_emb, cate).reshape(B, Fc * k)           # :78-79
    dense_vector = torch.cat([cont, vec, cat_emb], dim=1)                    # :82 ([vector | emb] in dnn_cate.py:75)
    deep = mlp_tower(dense_vector, p, len(hp.hidden_units))
    out = deep @ p['deep_res'] + p['deep_res_bias']                          # :119
    return out, torch.sigmoid(out)


def dnn_multi_cate_forward(p, batch, hp: HParams):
    """models/dnn_multi_cate.py:64-147."""
    Fc, k = hp.cate_field_size, hp.embedding_size
    cate, vec = batch['cate_feats'].long(), batch['vector_feats']
    B = cate.shape[0]
    feats_emb = zero_row0(p['feats_emb'])
    cate_emb = embedding_lookup(feats_emb, cate[:, :Fc]).reshape(B, Fc * k)
    _, all_multi = _multi_cate_emb(None, feats_emb, cate[:, Fc:], hp.multi_feats_range)
    dense_vector = torch.cat([vec, cate_emb, all_multi], dim=1)              # :103
    deep = mlp_tower(dense_vector, p, len(hp.hidden_units))
    out = deep @ p['deep_res'] + p['deep_res_bias']                          # :144
    return out, torch.sigmoid(out)


def _dnn_variant_loss(fwd):
    def loss_fn(p, batch, hp: HParams):
        """log-loss + l2 on deep_res (models/dnn_pipeline.py:125-131, dnn_cate.py:118-124, dnn_multi_cate.py:150-156)"""
        out, score = fwd(p, batch, hp)
        loss = log_loss(batch['label'], score)
        if hp.l2_reg > 0:
            loss = loss + l2_regularizer(hp.l2_reg, p['deep_res'])
        return loss, out, score
    return loss_fn


def deepfm_pipeline_forward(p, batch, hp: HParams):
    """models/deepfm_pipeline.py:76-169.  Table rows: [C continuous-field rows | cate_feats_size categorical rows], row 0
    zeroed (:83-86) -- so continuous field 0 has no weight at all.  Quirk Q2: the FM part looks categorical ids up at
    id + C (:89), the deep part at the RAW id (:120): one feature reads two different rows, and raw ids below C alias the
    continuous-field rows."""
    C, Fc, k = hp.cont_field_size, hp.cate_field_size, hp.embedding_size
    cate, cont, vec = batch['cate_feats'].long()[:, :Fc], batch['cont_feats'], batch['vector_feats']
    B = cate.shape[0]
    feats_emb = zero_row0(p['feats_emb'])                                      # :83-84
    first_tab = zero_row0(p['fm_first_order_emb'])                             # :85-86
    cont_index = torch.arange(C).unsqueeze(0).expand(B, C)                     # :58-61
    idx = torch.cat([cont_index, cate + C], dim=1)                             # :89-90
    val = torch.cat([cont, torch.ones(B, Fc, dtype=cont.dtype)], dim=1)        # :91
    F = C + Fc
    first_order = embedding_lookup(first_tab, idx).reshape(B, F) * val         # :95-97
    e = embedding_lookup(feats_emb, idx) * val.reshape(B, F, 1)                # :102-104
    sum_e = e.sum(dim=1)
    second_order = 0.5 * (sum_e * sum_e - (e * e).sum(dim=1))                  # :105-109
    fm_res = torch.cat([first_order, second_order], dim=1)                     # :114
    cat_emb = embedding_lookup(feats_emb, cate).reshape(B, Fc * k)             # :120-121 (raw ids)
    dense_vector = torch.cat([cont, vec, cat_emb], dim=1)                      # :123-124
    deep = mlp_tower(dense_vector, p, len(hp.hidden_units))                    # :149-154
    feats_input = torch.cat([fm_res, deep], dim=1)                             # :158
    out = feats_input @ p['deep_fm_weight'] + p['deep_fm_bias']                # :172
    return out, torch.sigmoid(out)


def wdl_textline_forward(p, batch, hp: HParams, with_gathered: bool = False):
    """models/wdl_textline.py:43-50 (row-0-zero embedding table), :133-235 (deep part), :237-276 (wide & deep).
    Same index construction as models/wdl.py (quirk Q4: deep weights alias wide ids W..W+H-1)."""
    C, Fc, k, W = hp.cont_field_size, hp.cate_field_size, hp.embedding_size, hp.wide_field_size
    cate, cont, wide = batch['cate_feats'].long(), batch['cont_feats'], batch['wide_feats'].long()
    B = cont.shape[0]
    H = hp.hidden_units[-1]
    embeddings = zero_row0(p['weight_mat'])                                    # :50
    cat_emb = embedding_lookup(embeddings, cate).reshape(B, Fc * k)            # :142-143
    dense_vector = torch.cat([cont, cat_emb], dim=1)                           # :190
    deep = mlp_tower(dense_vector, p, len(hp.hidden_units))                    # :215-221
    deep_index_add = torch.arange(H).unsqueeze(0).expand(B, H) + W             # :232
    idx = torch.cat([wide, deep_index_add], dim=1)                             # :262
    val = torch.cat([torch.ones(B, W, dtype=cont.dtype), deep], dim=1)         # :263
    gathered = embedding_lookup(p['wdl_weights'], idx).reshape(B, W + H)       # :264-269
    out = (gathered * val).sum(dim=1, keepdim=True) + p['wdl_bias']            # :271-273
    if with_gathered:
        return out, torch.sigmoid(out), gathered
    return out, torch.sigmoid(out)


def wdl_textline_loss(p, batch, hp: HParams):
    """models/wdl_textline.py:278-294.  Quirk Q6: `self.weights["wdl_weights"]` was overwritten at :264 by the
    GATHERED tensor [B, W+H], so the l2 term sums w^2 over every lookup of the batch (a row looked up c times counts
    c times, each deep row counts B times) instead of over the table."""
    out, score, gathered = wdl_textline_forward(p, batch, hp, with_gathered=True)
    loss = log_loss(batch['label'], score)
    if hp.l2_reg > 0:
        loss = loss + l2_regularizer(hp.l2_reg, gathered)
        for i in range(len(hp.hidden_units)):
            loss = loss + l2_regularizer(hp.l2_reg, p['deep_%d' % i])
    return loss, out, score


def dnn_multi_textline_forward(p, batch, hp: HParams):
    """models/dnn_multi_textline.py:37-48 (row-0-zero `weight_mat`, frozen `tags_emb` constant), :69-151.
    Batch keys are the canonical ones of this repo: cate_feats = [cat_feats | mul_cat_feats] (the reference feeds the two
    blocks as separate tensors, :55-57), multi_feats_value = mul_cat_feats_value.  Weighted pooling (:94-103): the count
    is taken on the UNWEIGHTED rows (row sum != 0), the sum on the weighted ones."""
    Fc, k = hp.cate_field_size, hp.embedding_size
    cate, cont, val = batch['cate_feats'].long(), batch['cont_feats'], batch['multi_feats_value']
    B = cont.shape[0]
    embeddings = zero_row0(p['weight_mat'])                                    # :44
    cat_emb = embedding_lookup(embeddings, cate[:, :Fc]).reshape(B, Fc * k)    # :73-74
    multi = cate[:, Fc:]
    pooled = []
    for item in hp.multi_feats_range:                                          # :77-110
        s, e, name = int(item[0]), int(item[1]), item[2]
        table = p['tags_emb'].detach() if name == 'tag' else embeddings        # :84-90: tf.constant has no gradient
        emb = embedding_lookup(table, multi[:, s:e])
        nonzero = (emb.sum(dim=2) != 0).sum(dim=1, keepdim=True).to(emb.dtype)  # :93-94
        wsum = (emb * val[:, s:e].unsqueeze(2)).sum(dim=1)                     # :96-101
        safe = torch.where(nonzero == 0, torch.ones_like(nonzero), nonzero)
        pooled.append(torch.where(nonzero == 0, torch.zeros_like(wsum), wsum / safe))   # :102 div_no_nan
    dense_vector = torch.cat([cont, cat_emb] + pooled, dim=1)                  # :113-114
    deep = mlp_tower(dense_vector, p, len(hp.hidden_units))
    out = deep @ p['deep_res'] + p['deep_res_bias']                            # :151
    return out, torch.sigmoid(out)


def dnn_multi_textline_loss(p, batch, hp: HParams):
    """models/dnn_multi_textline.py:153-161 -- L1 on the tower weights, as models/dnn.py (quirk Q5)."""
    out, score = dnn_multi_textline_forward(p, batch, hp)
    loss = log_loss(batch['label'], score)
    if hp.l2_reg > 0:
        for i in range(len(hp.hidden_units)):
            loss = loss + l1_regularizer(hp.l2_reg, p['deep_%d' % i])
    return loss, out, score


# --------------------------------------------------------------------------------------
# test_model/ extras: "models awaiting further development" that share the path
# --------------------------------------------------------------------------------------
def sigmoid_cross_entropy_with_logits(labels: Tensor, logits: Tensor) -> Tensor:
    """tf.nn.sigmoid_cross_entropy_with_logits (third-party TF 1.x, documented formula):
    max(x, 0) - x * z + log(1 + exp(-|x|)), element-wise."""
    return torch.clamp(logits, min=0) - logits * labels + torch.log1p(torch.exp(-torch.abs(logits)))


def wdl_ftrl_forward(p, batch, hp: HParams):
    """test_model/wdl.py:53-119: models/wdl.py with the deep-projection rows placed BEHIND the wide ids
    (`+ cate_feats_size`, :110 -- the un-aliased intent quirk Q4 of models/wdl.py misses)."""
    C, Fc, k, W = hp.cont_field_size, hp.cate_field_size, hp.embedding_size, hp.wide_field_size
    cate, cont, wide = batch['cate_feats'].long(), batch['cont_feats'], batch['wide_feats'].long()
    B = cont.shape[0]
    H = hp.hidden_units[-1]
    cat_emb = embedding_lookup(p['weight_mat'], cate).reshape(B, Fc * k)       # :67-68
    deep = mlp_tower(torch.cat([cont, cat_emb], dim=1), p, len(hp.hidden_units))   # :69, :93-97
    idx = torch.cat([wide, torch.arange(H).unsqueeze(0).expand(B, H) + hp.cate_feats_size], dim=1)    # :106-111
    val = torch.cat([torch.ones(B, W, dtype=cont.dtype), deep], dim=1)         # :112
    w = embedding_lookup(p['wdl_weights'], idx).reshape(B, W + H)              # :113-114
    out = (w * val).sum(dim=1, keepdim=True) + p['wdl_bias']                   # :115
    return out, torch.sigmoid(out)                                            # :119  (`self.out` IS the sigmoid from here on)


def wdl_ftrl_loss(p, batch, hp: HParams):
    """test_model/wdl.py:119-131: sigmoid_cross_entropy_with_logits fed the ALREADY-sigmoided value as `logits`
    (as written), mean over the batch, l2 on wdl_weights and every deep_i."""
    out, score = wdl_ftrl_forward(p, batch, hp)
    loss = sigmoid_cross_entropy_with_logits(batch['label'], score).mean()
    if hp.l2_reg > 0:
        loss = loss + l2_regularizer(hp.l2_reg, p['wdl_weights'])
        for i in range(len(hp.hidden_units)):
            loss = loss + l2_regularizer(hp.l2_reg, p['deep_%d' % i])
    return loss, out, score


DCN_CROSS_LAYERS = 2          # test_model/DCN.py:28


def dcn_forward(p, batch, hp: HParams):
    """test_model/DCN.py:43-103.  Dense features only; the deep tower has NO activation (:66-70: matmul + bias +
    dropout(1)); cross layer l: x_{l+1} = x0 * (x_l . w_l) + b_l + x_l (:84-86: tensordot(matmul(x0, x_l^T), w_l, 1)
    contracts the last axis, i.e. sum_j x0_i x_l_j w_j)."""
    x0 = batch['features']
    deep = x0
    for i in range(len(hp.hidden_units)):
        deep = deep @ p['deep_layer_%d' % i] + p['deep_layer_bias_%d' % i]
    xl = x0
    for i in range(DCN_CROSS_LAYERS):
        s = xl @ p['cross_layer_%d' % i]                                        # [B, 1]
        xl = x0 * s + p['cross_bias_%d' % i].reshape(1, -1) + xl
    out = torch.cat([xl, deep], dim=1) @ p['concat_projection'] + p['concat_bias']     # :101-102
    return out, torch.sigmoid(out)


def dcn_loss(p, batch, hp: HParams):
    """test_model/DCN.py:105-111: log_loss + l2 on concat_projection."""
    out, score = dcn_forward(p, batch, hp)
    loss = log_loss(batch['label'], score)
    if hp.l2_reg > 0:
        loss = loss + l2_regularizer(hp.l2_reg, p['concat_projection'])
    return loss, out, score


FORWARD = {'deepfm': deepfm_forward, 'deepfm_multi': deepfm_multi_forward, 'wdl': wdl_forward,
           'dnn': dnn_forward, 'dnn_multi': dnn_multi_forward,
           'deepfm_cate': deepfm_cate_forward, 'deepfm_multi_cate': deepfm_multi_cate_forward,
           'dnn_pipeline': dnn_pipeline_forward, 'dnn_cate': dnn_pipeline_forward, 'dnn_multi_cate': dnn_multi_cate_forward,
           'wdl_textline': wdl_textline_forward, 'dnn_multi_textline': dnn_multi_textline_forward,
           'deepfm_pipeline': deepfm_pipeline_forward, 'wdl_ftrl': wdl_ftrl_forward, 'dcn': dcn_forward}
LOSS = {'deepfm': deepfm_loss, 'deepfm_multi': deepfm_multi_loss, 'wdl': wdl_loss,
        'dnn': dnn_loss, 'dnn_multi': dnn_multi_loss,
        'deepfm_cate': _deepfm_variant_loss(deepfm_cate_forward),
        'deepfm_multi_cate': _deepfm_variant_loss(deepfm_multi_cate_forward),
        'dnn_pipeline': _dnn_variant_loss(dnn_pipeline_forward), 'dnn_cate': _dnn_variant_loss(dnn_pipeline_forward),
        'dnn_multi_cate': _dnn_variant_loss(dnn_multi_cate_forward),
        'wdl_textline': wdl_textline_loss, 'dnn_multi_textline': dnn_multi_textline_loss,
        'deepfm_pipeline': _deepfm_variant_loss(deepfm_pipeline_forward),     # models/deepfm_pipeline.py:177-183
        'wdl_ftrl': wdl_ftrl_loss, 'dcn': dcn_loss}
# which optimizer `.minimize` runs: tf.train.AdamOptimizer everywhere in models/; test_model/wdl.py:136 uses FTRL
OPTIMIZER = {'wdl_ftrl': 'ftrl'}
# the files of test_model/ (every other key of FORWARD is models/<key>.py)
TEST_MODEL_FILES = {'wdl_ftrl': 'test_model/wdl.py', 'dcn': 'test_model/DCN.py'}
# variants without a continuous block read cont_field_size as 0 (their model files never define it)
NO_CONT = ('deepfm_cate', 'deepfm_multi_cate', 'dnn_cate', 'dnn_multi_cate')


# --------------------------------------------------------------------------------------
# parameter shapes + distribution-matched initialisation (values are INJECTED into both
# the oracle and the CUDA model; the reference's own weights are not reproducible from
# seeds)
# --------------------------------------------------------------------------------------
def deep_input_size(alg: str, hp: HParams) -> int:
    k = hp.embedding_size
    if alg in ('deepfm', 'deepfm_pipeline'):
        return hp.cont_field_size + hp.vector_feats_size + hp.cate_field_size * k
    if alg in ('deepfm_multi', 'dnn_multi'):
        return hp.cont_field_size + hp.vector_feats_size + (hp.cate_field_size + hp.multi_field_size) * k
    if alg in ('deepfm_multi_cate', 'dnn_multi_cate'):
        return hp.vector_feats_size + (hp.cate_field_size + hp.multi_field_size) * k
    if alg in ('deepfm_cate', 'dnn_cate'):
        return hp.vector_feats_size + hp.cate_field_size * k
    if alg == 'dnn_pipeline':
        return hp.cont_field_size + hp.vector_feats_size + hp.cate_field_size * k
    if alg in ('wdl', 'wdl_textline'):
        return hp.cont_field_size + hp.cate_field_size * k
    if alg == 'dnn_multi_textline':
        return hp.cont_field_size + (hp.cate_field_size + hp.multi_field_size) * k
    if alg == 'dnn':
        return hp.cont_field_size + hp.vector_feats_size + hp.cate_field_size * k
    if alg == 'wdl_ftrl':
        return hp.cont_field_size + hp.cate_field_size * k
    if alg == 'dcn':
        return hp.deep_feature_size
    raise ValueError(alg)


def param_shapes(alg: str, hp: HParams) -> Dict[str, Tuple[int, ...]]:
    k, H = hp.embedding_size, hp.hidden_units
    s: Dict[str, Tuple[int, ...]] = {}
    d0 = deep_input_size(alg, hp)
    if alg == 'deepfm':                                   # models/deepfm.py:57-60
        R = hp.cont_field_size + hp.cate_feats_size
        s['feats_emb'] = (R, k); s['feats'] = (R, 1)
    elif alg == 'deepfm_pipeline':                        # models/deepfm_pipeline.py:77-81
        R = hp.cont_field_size + hp.cate_feats_size
        s['feats_emb'] = (R, k); s['fm_first_order_emb'] = (R, 1)
    elif alg == 'deepfm_multi':                           # models/deepfm_multi.py:124-128
        R = hp.cate_feats_size + hp.cont_field_size
        s['feats_emb'] = (R, k); s['fm_first_order_emb'] = (R, 1)
    elif alg in ('wdl', 'wdl_textline'):                  # models/wdl.py:43-47,231-236, models/wdl_textline.py:44-47,253-256
        s['weight_mat'] = (hp.cate_index_size, k)
        s['wdl_weights'] = (hp.cate_index_size + H[-1], 1); s['wdl_bias'] = (1,)
    elif alg == 'dnn':                                    # models/dnn.py:49-52
        s['weight_mat'] = (hp.cate_feats_size, k)
    elif alg == 'wdl_ftrl':                               # test_model/wdl.py:61-64,101-103
        s['weight_mat'] = (hp.cate_feats_size, k)
        s['wdl_weights'] = (hp.cate_feats_size + H[-1], 1); s['wdl_bias'] = (1,)
    if alg == 'dcn':                                      # test_model/DCN.py:49-99 (its own variable names)
        prev = d0
        for i, h in enumerate(H):
            s['deep_layer_%d' % i] = (prev, h); s['deep_layer_bias_%d' % i] = (1, h); prev = h
        for i in range(DCN_CROSS_LAYERS):
            s['cross_layer_%d' % i] = (d0, 1); s['cross_bias_%d' % i] = (d0, 1)
        s['concat_projection'] = (d0 + H[-1], 1); s['concat_bias'] = ()
        return s
    elif alg == 'dnn_multi_textline':                     # models/dnn_multi_textline.py:38-48 (tags_emb: frozen constant)
        s['weight_mat'] = (hp.cate_feats_size, k); s['tags_emb'] = (hp.word2vec_rows, k)
    elif alg in ('dnn_multi', 'dnn_pipeline', 'dnn_cate', 'dnn_multi_cate'):   # models/dnn_multi.py:71-73, dnn_pipeline.py:69-71
        s['feats_emb'] = (hp.cate_feats_size, k)
    elif alg in ('deepfm_cate', 'deepfm_multi_cate'):     # models/deepfm_cate.py:73-77
        s['feats_emb'] = (hp.cate_feats_size, k); s['fm_first_order_emb'] = (hp.cate_feats_size, 1)
    prev = d0
    for i, h in enumerate(H):
        s['deep_%d' % i] = (prev, h); s['deep_bias_%d' % i] = (1, h); prev = h
    if alg in ('deepfm', 'deepfm_multi', 'deepfm_pipeline'):
        F = hp.cont_field_size + hp.cate_field_size + (hp.multi_field_size if alg == 'deepfm_multi' else 0)
        s['deep_fm_weight'] = (F + k + H[-1], 1); s['deep_fm_bias'] = (1,)
    if alg in ('deepfm_cate', 'deepfm_multi_cate'):
        F = hp.cate_field_size + (hp.multi_field_size if alg == 'deepfm_multi_cate' else 0)
        s['deep_fm_weight'] = (F + k + H[-1], 1); s['deep_fm_bias'] = (1,)
    if alg in ('dnn', 'dnn_multi', 'dnn_pipeline', 'dnn_cate', 'dnn_multi_cate', 'dnn_multi_textline'):
        s['deep_res'] = (H[-1], 1); s['deep_res_bias'] = (1, 1)
    return s


def init_params(alg: str, hp: HParams, seed: int = 2019, dtype=torch.float64) -> Dict[str, Tensor]:
    """Distributions of SURVEY.md 8c-(9): feats_emb ~ N(0, 0.01^2); first-order ~ U(0,1)
    (models/deepfm.py:58-60); MLP W,b ~ N(0, 2/(fan_in+fan_out)) (:106-119); head bias ~ N(0,1)
    (:142); xavier-uniform for weight_mat (models/wdl.py:44-47)."""
    g = np.random.RandomState(seed)
    shapes = param_shapes(alg, hp)
    p: Dict[str, Tensor] = {}
    for name, shp in shapes.items():
        if name == 'feats_emb':
            a = g.normal(0.0, 0.01, size=shp)
        elif name in ('feats', 'fm_first_order_emb'):
            a = g.uniform(0.0, 1.0, size=shp)
        elif name == 'tags_emb':
            a = g.normal(0.0, 0.1, size=shp)
        elif name == 'weight_mat':
            lim = math.sqrt(6.0 / (shp[0] + shp[1])); a = g.uniform(-lim, lim, size=shp)
        elif name == 'wdl_weights':
            a = g.normal(0.0, math.sqrt(2.0 / shp[0]), size=shp)
        elif name in ('deep_fm_bias', 'wdl_bias'):
            a = g.normal(0.0, 1.0, size=shp)
        elif name == 'concat_bias':                       # test_model/DCN.py:99: tf.constant(0.01); drawn here so its gradient path is exercised
            a = g.normal(0.01, 0.1, size=shp)
        elif name.startswith('deep_layer_bias_'):
            w = shapes['deep_layer_%s' % name.split('_')[-1]]
            a = g.normal(0.0, math.sqrt(2.0 / (w[0] + w[1])), size=shp)
        elif name.startswith(('cross_layer_', 'cross_bias_')):
            a = g.normal(0.0, math.sqrt(2.0 / (2 * shp[0])), size=shp)
        elif name.startswith('deep_bias_'):
            w = shapes['deep_%s' % name.split('_')[-1]]
            a = g.normal(0.0, math.sqrt(2.0 / (w[0] + w[1])), size=shp)
        elif name == 'deep_res_bias':
            a = g.normal(0.0, math.sqrt(2.0 / (shapes['deep_res'][0] + 1)), size=shp)
        else:   # deep_i, deep_fm_weight, deep_res
            a = g.normal(0.0, math.sqrt(2.0 / (shp[0] + shp[1])), size=shp)
        # round to fp32 so fp64 oracle and fp32 device start from IDENTICAL values
        p[name] = torch.from_numpy(np.asarray(a, dtype=np.float32)).to(dtype)
    return p


# --------------------------------------------------------------------------------------
# gradients + TF1 Adam
# --------------------------------------------------------------------------------------
def loss_and_grads(alg: str, p: Dict[str, Tensor], batch: Dict[str, Tensor], hp: HParams):
    """What `AdamOptimizer.minimize` differentiates (models/deepfm.py:161-162): autodiff of the
    scalar loss w.r.t. every trainable variable.  Embedding-table gradients are returned DENSE
    [R, w]; rows hit several times in the batch hold the SUM of their per-lookup gradients, which
    is what TF's IndexedSlices de-duplication (unique + unsorted_segment_sum) feeds to Adam."""
    leaves = {n: t.detach().clone().requires_grad_(True) for n, t in p.items()}
    loss, out, score = LOSS[alg](leaves, batch, hp)
    names = list(leaves)
    gs = torch.autograd.grad(loss, [leaves[n] for n in names], allow_unused=True)
    grads = {n: (g if g is not None else torch.zeros_like(leaves[n])) for n, g in zip(names, gs)}
    return loss.detach(), out.detach(), score.detach(), grads


@dataclass
class AdamState:
    m: Dict[str, Tensor]
    v: Dict[str, Tensor]
    beta1_power: float = ADAM_BETA1     # TF initialises the accumulators to beta (adam.py _create_slots)
    beta2_power: float = ADAM_BETA2
    global_step: int = 0


def adam_init(p: Dict[str, Tensor]) -> AdamState:
    return AdamState({n: torch.zeros_like(t) for n, t in p.items()},
                     {n: torch.zeros_like(t) for n, t in p.items()})


def tf1_adam_step(p, grads, st: AdamState, hp: HParams, lazy_rows: Optional[Dict[str, Tensor]] = None):
    """tf.train.AdamOptimizer (TF 1.x adam.py) -- third-party, restated from its published algorithm:
        lr_t = lr * sqrt(1 - beta2^t) / (1 - beta1^t)
        m <- beta1*m + (1-beta1)*g ;  v <- beta2*v + (1-beta2)*g^2
        var <- var - lr_t * m / (sqrt(v) + eps)          (eps OUTSIDE the bias correction)
    applied to EVERY element of every variable each step (both _apply_dense and
    _apply_sparse_shared decay m,v and update var over the whole table); beta powers advance and
    global_step increments AFTER the apply; lr comes from exponential_decay on the pre-increment
    step (call site models/deepfm.py:159-162).

    `lazy_rows` (name -> bool mask [R]) restricts the update of the named tables to the masked
    rows: this is NOT the reference's behaviour, it restates the product's documented
    `adam_mode="lazy"` so that mode has an oracle too (identical to TF at step 1 from zero slots)."""
    lr = exponential_decay(hp.learning_rate, st.global_step, hp.learning_rate_decay_steps,
                           hp.learning_rate_decay_rate)
    lr_t = lr * math.sqrt(1 - st.beta2_power) / (1 - st.beta1_power)
    newp = {}
    for n, var in p.items():
        g = grads[n].to(var.dtype)
        m = ADAM_BETA1 * st.m[n] + (1 - ADAM_BETA1) * g
        v = ADAM_BETA2 * st.v[n] + (1 - ADAM_BETA2) * g * g
        nv = var - lr_t * m / (torch.sqrt(v) + ADAM_EPS)
        if lazy_rows is not None and n in lazy_rows:
            mask = lazy_rows[n].reshape(-1, *([1] * (var.dim() - 1)))
            m = torch.where(mask, m, st.m[n]); v = torch.where(mask, v, st.v[n])
            nv = torch.where(mask, nv, var)
        st.m[n], st.v[n], newp[n] = m, v, nv
    st.beta1_power *= ADAM_BETA1
    st.beta2_power *= ADAM_BETA2
    st.global_step += 1
    return newp


FTRL_INIT_ACCUM, FTRL_LR_POWER = 0.1, -0.5      # tf.train.FtrlOptimizer defaults (test_model/wdl.py:136): l1 = l2 = 0


def ftrl_init(p: Dict[str, Tensor]) -> AdamState:
    """FtrlOptimizer slots in the AdamState container: m <- `accum` (initial_accumulator_value 0.1), v <- `linear` (0)."""
    return AdamState({n: torch.full_like(t, FTRL_INIT_ACCUM) for n, t in p.items()},
                     {n: torch.zeros_like(t) for n, t in p.items()})


def tf1_ftrl_step(p, grads, st: AdamState, hp: HParams, sparse_rows: Dict[str, Tensor]):
    """tf.train.FtrlOptimizer (TF 1.x ftrl.py -> ApplyFtrl / SparseApplyFtrl kernels; third-party, restated from the
    published kernel) with its defaults lr_power = -0.5, l1 = l2 = 0:
        accum' = accum + g^2 ;  linear += g - (sqrt(accum') - sqrt(accum)) / lr * var
        var    = -linear / (sqrt(accum') / lr)        (|linear| > l1 = 0; else 0)
    Variables whose gradient is an IndexedSlices (the embedding table: its only consumer is embedding_lookup) take the
    SPARSE apply: only the looked-up rows change (`sparse_rows`: name -> row mask)."""
    lr = exponential_decay(hp.learning_rate, st.global_step, hp.learning_rate_decay_steps, hp.learning_rate_decay_rate)
    newp = {}
    for n, var in p.items():
        g = grads[n].to(var.dtype)
        accum, linear = st.m[n], st.v[n]
        accum_new = accum + g * g
        linear_new = linear + g - (torch.sqrt(accum_new) - torch.sqrt(accum)) / lr * var
        quadratic = torch.sqrt(accum_new) / lr
        nv = torch.where(linear_new.abs() > 0, -linear_new / quadratic, torch.zeros_like(var))
        if n in sparse_rows:
            mask = sparse_rows[n].reshape(-1, *([1] * (var.dim() - 1)))
            accum_new = torch.where(mask, accum_new, accum); linear_new = torch.where(mask, linear_new, linear)
            nv = torch.where(mask, nv, var)
        st.m[n], st.v[n], newp[n] = accum_new, linear_new, nv
    st.global_step += 1
    return newp


def touched_rows(alg: str, batch, hp: HParams) -> Dict[str, Tensor]:
    """Row masks for adam_mode='lazy': rows whose gradient is structurally non-zero this step."""
    if alg == 'dcn':
        return {}
    cate = batch['cate_feats'].long()
    shapes = param_shapes(alg, hp)
    out = {}

    def mask(R, ids, drop0=False):
        mk = torch.zeros(R, dtype=torch.bool)
        mk[ids.reshape(-1)] = True
        if drop0:
            mk[0] = False
        return mk
    if alg == 'deepfm':
        R = shapes['feats_emb'][0]
        mk = mask(R, cate)
        mk[hp.cate_field_size:hp.cate_field_size + hp.cont_field_size] = True
        out['feats_emb'] = mk; out['feats'] = mk
    elif alg == 'deepfm_multi':
        R = shapes['feats_emb'][0]
        mk = mask(R, cate, drop0=True)
        mk[hp.cate_feats_size:hp.cate_feats_size + hp.cont_field_size] = True
        out['feats_emb'] = mk; out['fm_first_order_emb'] = mk
    elif alg == 'deepfm_pipeline':
        C = hp.cont_field_size
        mk = mask(shapes['feats_emb'][0], torch.cat([cate.reshape(-1), cate.reshape(-1) + C]), drop0=True)
        mk[1:C] = True                                # continuous-field rows (row 0 is cut out of the graph)
        out['feats_emb'] = mk; out['fm_first_order_emb'] = mk
    elif alg in ('wdl', 'wdl_ftrl'):
        out['weight_mat'] = mask(shapes['weight_mat'][0], cate)
    elif alg == 'wdl_textline':
        out['weight_mat'] = mask(shapes['weight_mat'][0], cate, drop0=True)
    elif alg == 'dnn_multi_textline':
        Fc = hp.cate_field_size
        cols = [cate[:, :Fc].reshape(-1)] + [cate[:, Fc + int(r[0]):Fc + int(r[1])].reshape(-1)
                                             for r in hp.multi_feats_range if r[2] != 'tag']
        out['weight_mat'] = mask(shapes['weight_mat'][0], torch.cat(cols), drop0=True)
    elif alg == 'dnn':
        out['weight_mat'] = mask(shapes['weight_mat'][0], cate)
    elif alg in ('dnn_multi', 'dnn_pipeline', 'dnn_cate', 'dnn_multi_cate'):
        out['feats_emb'] = mask(shapes['feats_emb'][0], cate, drop0=True)
    elif alg in ('deepfm_cate', 'deepfm_multi_cate'):
        mk = mask(shapes['feats_emb'][0], cate, drop0=True)
        out['feats_emb'] = mk; out['fm_first_order_emb'] = mk
    return out


def train_step(alg, p, st: AdamState, batch, hp: HParams, adam_mode: str = 'tf_dense'):
    """One `sess.run([loss, optimizer])` (models/deepfm.py:181): returns (loss, new params)."""
    loss, out, score, grads = loss_and_grads(alg, p, batch, hp)
    if OPTIMIZER.get(alg) == 'ftrl':      # sparse apply for the table, dense for the rest -- whatever adam_mode says
        return loss, tf1_ftrl_step(p, grads, st, hp, touched_rows(alg, batch, hp))
    lazy = touched_rows(alg, batch, hp) if adam_mode == 'lazy' else None
    return loss, tf1_adam_step(p, grads, st, hp, lazy)


def optimizer_init(alg: str, p: Dict[str, Tensor]) -> AdamState:
    return ftrl_init(p) if OPTIMIZER.get(alg) == 'ftrl' else adam_init(p)


# --------------------------------------------------------------------------------------
# AUC -- restates sklearn.metrics.roc_auc_score (binary, no sample weights): ROC curve at
# distinct score thresholds, trapezoidal area.  Equivalent to Mann-Whitney U with ties = 1/2.
# Call sites: models/deepfm.py:231,253.
# --------------------------------------------------------------------------------------
def roc_auc(labels: np.ndarray, scores: np.ndarray) -> float:
    y = np.asarray(labels, dtype=np.float64).reshape(-1)
    s = np.asarray(scores).reshape(-1)
    order = np.argsort(-s, kind='mergesort')
    y, s = y[order], s[order]
    distinct = np.where(np.diff(s))[0]
    thr = np.r_[distinct, y.size - 1]
    tps = np.cumsum(y)[thr]
    fps = 1 + thr - tps
    tps = np.r_[0, tps]; fps = np.r_[0, fps]
    if fps[-1] <= 0 or tps[-1] <= 0:
        raise ValueError("Only one class present in y_true. ROC AUC score is not defined in that case.")
    fpr, tpr = fps / fps[-1], tps / tps[-1]
    return float(np.sum(np.diff(fpr) * (tpr[1:] + tpr[:-1]) / 2.0))


# --------------------------------------------------------------------------------------
# independent closed forms (NumPy) used by tests/ to guard the restatement itself
# --------------------------------------------------------------------------------------
def fm_second_order_pairwise(e: np.ndarray) -> np.ndarray:
    """sum_{i<j} e_i * e_j per embedding coordinate; equals 0.5((sum e)^2 - sum e^2). e [B,F,k]."""
    B, F, k = e.shape
    out = np.zeros((B, k), dtype=e.dtype)
    for i in range(F):
        for j in range(i + 1, F):
            out += e[:, i, :] * e[:, j, :]
    return out


def to_dtype(d: Dict[str, Tensor], dtype) -> Dict[str, Tensor]:
    return {n: (t.to(dtype) if t.is_floating_point() else t) for n, t in d.items()}
