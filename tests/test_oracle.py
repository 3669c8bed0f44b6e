"""CPU: the oracle (restatement of the reference TF graph) against hand-computed known answers,
independent closed forms and the one reference dependency present here (scikit-learn's AUC).
The reference ships no tests or golden vectors: parity is UNPINNED against
TensorFlow itself; these checks pin the restatement to the published TF op semantics."""
import math

import numpy as np
import pytest
import torch

from oracle import ctr_oracle as O
import ctr_helpers as H


def test_fm_known_answer_tiny():
    """B=1, Fc=2, C=1, k=2, hand computed (models/deepfm.py:63-93 incl. quirk Q1 offset = Fc)"""
    hp = O.HParams(hidden_units=[2], embedding_size=2, cont_field_size=1, cate_field_size=2, cate_feats_size=4, l2_reg=0.0)
    p = {n: torch.zeros(s, dtype=torch.float64) for n, s in O.param_shapes("deepfm", hp).items()}
    p["feats_emb"] = torch.tensor([[1., 2.], [3., 4.], [5., 6.], [7., 8.], [9., 10.]], dtype=torch.float64)
    p["feats"] = torch.tensor([[.1], [.2], [.3], [.4], [.5]], dtype=torch.float64)
    # head picks out first-order (3), second-order (2), deep (2)
    p["deep_fm_weight"] = torch.tensor([[1.], [10.], [100.], [1000.], [10000.], [0.], [0.]], dtype=torch.float64)
    b = {"cate_feats": torch.tensor([[0, 3]]), "cont_feats": torch.tensor([[0.5]], dtype=torch.float64),
         "vector_feats": torch.zeros(1, 0, dtype=torch.float64), "label": torch.ones(1, 1, dtype=torch.float64)}
    out, score = O.deepfm_forward(p, b, hp)
    # cont row id = 0 + Fc = 2 -> e = [1,2], [7,8], 0.5*[5,6]=[2.5,3]; first = [.1, .4, .3*.5]
    first = [0.1, 0.4, 0.15]
    s = np.array([10.5, 13.0]); q = np.array([1 + 49 + 6.25, 4 + 64 + 9.0])
    second = 0.5 * (s * s - q)
    want = first[0] + 10 * first[1] + 100 * first[2] + 1000 * second[0] + 10000 * second[1]
    assert abs(float(out) - want) < 1e-9
    assert abs(float(score) - 1 / (1 + math.exp(-want))) < 1e-12


def test_fm_identity_and_gradient_closed_form():
    rng = np.random.RandomState(0)
    e = rng.randn(5, 7, 3)
    second = 0.5 * (e.sum(1) ** 2 - (e ** 2).sum(1))
    assert np.allclose(second, O.fm_second_order_pairwise(e), atol=1e-12)
    # d second / d e_f = sum_e - e_f
    et = torch.tensor(e, requires_grad=True)
    g = torch.tensor(rng.randn(5, 3))
    s2 = 0.5 * (et.sum(1) ** 2 - (et ** 2).sum(1))
    (s2 * g).sum().backward()
    want = g.numpy()[:, None, :] * (e.sum(1)[:, None, :] - e)
    assert np.allclose(et.grad.numpy(), want, atol=1e-12)


def test_log_loss_matches_tf_formula():
    y = torch.tensor([[1.], [0.], [1.]], dtype=torch.float64); p = torch.tensor([[.9], [.2], [1e-9]], dtype=torch.float64)
    want = np.mean([-math.log(.9 + 1e-7), -math.log(.8 + 1e-7), -math.log(1e-9 + 1e-7)])
    assert abs(float(O.log_loss(y, p)) - want) < 1e-12
    assert abs(float(O.l2_regularizer(0.5, torch.tensor([3., 4.]))) - 0.5 * 25 / 2) < 1e-12
    assert abs(float(O.l1_regularizer(0.5, torch.tensor([-3., 4.]))) - 3.5) < 1e-12
    assert O.exponential_decay(0.1, 25, 10, 0.5) == 0.1 * 0.25


def test_nonzero_reduce_mean_semantics():
    """count rows whose SUM is non-zero (quirk Q3), div_no_nan, padding row 0 reads zero"""
    emb = torch.tensor([[[1., -1.], [2., 2.], [0., 0.]], [[0., 0.], [0., 0.], [0., 0.]]], dtype=torch.float64)
    r = O.nonzero_reduce_mean(emb)
    assert torch.allclose(r[0], torch.tensor([3., 1.], dtype=torch.float64))   # only one row counts
    assert torch.equal(r[1], torch.zeros(2, dtype=torch.float64))
    t = torch.arange(6, dtype=torch.float64).reshape(3, 2) + 1
    assert torch.equal(O.zero_row0(t)[0], torch.zeros(2, dtype=torch.float64)) and torch.equal(O.zero_row0(t)[1:], t[1:])


def test_tf1_adam_first_step_closed_form_and_epsilon_placement():
    hp = O.HParams()
    p = {"w": torch.tensor([1.0, -2.0, 3.0], dtype=torch.float64)}
    g = {"w": torch.tensor([1e-7, -0.5, 0.0], dtype=torch.float64)}
    st = O.adam_init(p)
    new = O.tf1_adam_step(p, g, st, hp)
    lr_t = hp.learning_rate * math.sqrt(1 - 0.999) / (1 - 0.9)
    for i in range(3):
        gi = float(g["w"][i])
        want = float(p["w"][i]) - lr_t * 0.1 * gi / (math.sqrt(0.001 * gi * gi) + 1e-8)
        assert abs(float(new["w"][i]) - want) < 1e-15
    assert st.global_step == 1 and abs(st.beta1_power - 0.81) < 1e-15
    # second step with zero gradient still moves the parameter (TF1 Adam is not lazy)
    new2 = O.tf1_adam_step(new, {"w": torch.zeros(3, dtype=torch.float64)}, st, hp)
    assert float((new2["w"] - new["w"]).abs()[1]) > 1e-5


def test_lazy_equals_dense_on_first_step_only():
    hp = H.hp_c1(batch_size=64)
    outs = {}
    for mode in ("tf_dense", "lazy"):
        p = O.init_params("deepfm", hp); st = O.adam_init(p)
        traj = []
        for s in range(2):
            b = H.oracle_batch(H.make_batch("deepfm", hp, seed=s), torch.float64)
            _, p = O.train_step("deepfm", p, st, b, hp, adam_mode=mode)
            traj.append(p["feats_emb"].clone())
        outs[mode] = traj
    assert torch.equal(outs["tf_dense"][0], outs["lazy"][0])
    assert not torch.equal(outs["tf_dense"][1], outs["lazy"][1])


@pytest.mark.parametrize("ties", [False, True])
def test_auc_matches_sklearn(ties):
    from sklearn.metrics import roc_auc_score
    rng = np.random.RandomState(3)
    y = (rng.rand(5000) < 0.3).astype(np.float32); s = rng.randn(5000).astype(np.float32)
    if ties:
        s = np.round(s, 1)
    assert abs(O.roc_auc(y, s) - roc_auc_score(y, s)) < 1e-12


@pytest.mark.parametrize("alg", ["deepfm", "deepfm_multi", "wdl", "dnn", "dnn_multi"])
def test_autograd_matches_finite_differences(alg):
    hp = {"deepfm": H.hp_c1(batch_size=16, hidden_units=[8, 4]), "dnn": H.hp_c1(batch_size=16, hidden_units=[8, 4]),
          "deepfm_multi": H.hp_multi(B=16, hidden_units=[8, 4]), "dnn_multi": H.hp_multi(B=16, hidden_units=[8, 4]),
          "wdl": H.hp_criteo(B=16, rows=500, hidden_units=[8, 4])}[alg]
    p = O.init_params(alg, hp)
    b = H.oracle_batch(H.make_batch(alg, hp, seed=1), torch.float64)
    loss, _, _, grads = O.loss_and_grads(alg, p, b, hp)
    rng = np.random.RandomState(0)
    for name in p:
        flat = p[name].reshape(-1)
        gflat = grads[name].reshape(-1)
        nz = torch.nonzero(gflat).reshape(-1)
        idx = int(nz[rng.randint(len(nz))]) if len(nz) else 0
        eps = 1e-6
        q = {k: v.clone() for k, v in p.items()}
        q[name].reshape(-1)[idx] += eps
        lp = float(O.LOSS[alg](q, b, hp)[0])
        q[name].reshape(-1)[idx] -= 2 * eps
        lm = float(O.LOSS[alg](q, b, hp)[0])
        fd = (lp - lm) / (2 * eps)
        assert abs(fd - float(gflat[idx])) < 1e-6 * max(1.0, abs(fd)) + 1e-9, (name, fd, float(gflat[idx]))


VARIANTS = ["deepfm_cate", "deepfm_multi_cate", "dnn_pipeline", "dnn_cate", "dnn_multi_cate"]


@pytest.mark.parametrize("alg", VARIANTS)
def test_variants_autograd_matches_finite_differences(alg):
    """the variants of SURVEY.md 8f N2, restated from their own files: gradient of every parameter block by central
    differences"""
    hp = H.hp_variant(alg, hidden_units=[8, 4])
    p = O.init_params(alg, hp)
    b = H.oracle_batch(H.make_batch(alg, hp, seed=1, B=16), torch.float64)
    loss, _, _, grads = O.loss_and_grads(alg, p, b, hp)
    rng = np.random.RandomState(0)
    for name in p:
        gflat = grads[name].reshape(-1)
        nz = torch.nonzero(gflat).reshape(-1)
        idx = int(nz[rng.randint(len(nz))]) if len(nz) else 0
        eps = 1e-6
        q = {k: v.clone() for k, v in p.items()}
        q[name].reshape(-1)[idx] += eps
        lp = float(O.LOSS[alg](q, b, hp)[0])
        q[name].reshape(-1)[idx] -= 2 * eps
        lm = float(O.LOSS[alg](q, b, hp)[0])
        fd = (lp - lm) / (2 * eps)
        assert abs(fd - float(gflat[idx])) < 1e-6 * max(1.0, abs(fd)) + 1e-9, (name, fd, float(gflat[idx]))


def test_variants_reduce_to_their_parents():
    """independent restatements agree where the graphs coincide: *_multi_cate == *_multi with no continuous block"""
    hp = H.hp_variant("deepfm_multi_cate", hidden_units=[8, 4])
    b = H.oracle_batch(H.make_batch("deepfm_multi_cate", hp, seed=3, B=32), torch.float64)
    for child, parent in (("deepfm_multi_cate", "deepfm_multi"), ("dnn_multi_cate", "dnn_multi")):
        p = O.init_params(child, hp)
        assert O.param_shapes(child, hp) == O.param_shapes(parent, hp)
        assert torch.equal(O.FORWARD[child](p, b, hp)[0], O.FORWARD[parent](p, b, hp)[0])


def test_data_formats_roundtrip():
    from deep_learning_b200 import data as D
    hp = H.hp_c1(batch_size=7)
    b = H.make_batch("deepfm", hp, seed=4)
    d = D.from_pickled_batch(D.to_pickled_batch(b))
    assert np.array_equal(d["cate_feats"], b["cate_feats"].numpy()) and np.allclose(d["cont_feats"], b["cont_feats"].numpy())
    assert d["label"].shape == (7, 1)
    sizes = D.field_sizes(10 ** 8, 26)
    assert sum(sizes) == 10 ** 8 and min(sizes) >= 2


# ---- textline models: models/wdl_textline.py, models/dnn_multi_textline.py -----------------------
@pytest.mark.parametrize("alg", ["wdl_textline", "dnn_multi_textline"])
def test_textline_autograd_matches_finite_differences(alg):
    hp = H.hp_textline(alg, hidden_units=[8, 4], l2_reg=1e-2)
    p = O.init_params(alg, hp)
    b = H.oracle_batch(H.make_batch(alg, hp, seed=1, B=16), torch.float64)
    loss, _, _, grads = O.loss_and_grads(alg, p, b, hp)
    rng = np.random.RandomState(0)
    for name in p:
        gflat = grads[name].reshape(-1)
        if name == "tags_emb":       # tf.constant: no gradient even though it shapes the output
            assert not gflat.any()
            continue
        nz = torch.nonzero(gflat).reshape(-1)
        idx = int(nz[rng.randint(len(nz))]) if len(nz) else 0
        eps = 1e-6
        q = {k: v.clone() for k, v in p.items()}
        q[name].reshape(-1)[idx] += eps
        lp = float(O.LOSS[alg](q, b, hp)[0])
        q[name].reshape(-1)[idx] -= 2 * eps
        lm = float(O.LOSS[alg](q, b, hp)[0])
        fd = (lp - lm) / (2 * eps)
        assert abs(fd - float(gflat[idx])) < 1e-6 * max(1.0, abs(fd)) + 1e-9, (name, fd, float(gflat[idx]))


def test_wdl_textline_l2_runs_over_the_gathered_weights():
    """quirk Q6 (models/wdl_textline.py:264-269,284-286): known answer for the regulariser.  With wide ids {5, 5} for
    sample 0 and {5, 9} for sample 1, W = 2, H = 4, the term is l2/2 * (3*w5^2 + w9^2 + B * sum_{r=2..5} w_r^2); row 5 is
    both a wide id and a deep row (quirk Q4), so it counts 3 + B times."""
    hp = O.HParams(hidden_units=[4], batch_size=2, embedding_size=4, cont_field_size=1, cate_field_size=1, cate_index_size=12,
                   wide_field_size=2, l2_reg=0.5)
    p = O.init_params("wdl_textline", hp)
    b = {"cate_feats": torch.tensor([[3], [0]]), "cont_feats": torch.rand(2, 1, dtype=torch.float64),
         "wide_feats": torch.tensor([[5, 5], [5, 9]]), "label": torch.tensor([[1.0], [0.0]], dtype=torch.float64)}
    loss, out, score = O.wdl_textline_loss(p, b, hp)
    w = p["wdl_weights"].reshape(-1)
    cnt = torch.zeros(16, dtype=torch.float64)
    cnt[5] += 3; cnt[9] += 1; cnt[2:6] += 2
    want = O.log_loss(b["label"], score) + 0.5 * hp.l2_reg * (cnt * w * w).sum() + O.l2_regularizer(hp.l2_reg, p["deep_0"])
    assert abs(float(loss) - float(want)) < 1e-14
    # id 0 reads the zero row: sample 1's embedding block is all zeros whatever weight_mat[0] holds
    q = {k: v.clone() for k, v in p.items()}
    q["weight_mat"][0] += 1.0
    assert torch.equal(O.wdl_textline_forward(q, b, hp)[0], out)
    # the table-wide l2 of models/wdl.py is a different number
    assert abs(float(O.wdl_loss(p, b, hp)[0]) - float(loss)) > 1e-6


def test_weighted_pooling_known_answer():
    """models/dnn_multi_textline.py:93-102: count on the UNWEIGHTED rows (row sum != 0), sum on the weighted ones.
    One trainable field [7, 0, 9] with weights [2, 5, 0.5]: id 0 is the zero row -> count 2, pooled = (2*r7 + 0.5*r9)/2.
    One 'tag' field [0, 1] with weights [1, 3]: row 0 of the frozen table is a REAL row -> count 2."""
    hp = O.HParams(hidden_units=[4], batch_size=1, embedding_size=2, cont_field_size=1, cate_field_size=1, cate_feats_size=12,
                   multi_feats_size=5, multi_field_size=2, multi_feats_range=[[0, 3, "kw"], [3, 5, "tag"]], word2vec_rows=3,
                   l2_reg=0.0)
    p = O.init_params("dnn_multi_textline", hp)
    p["deep_0"] = torch.eye(7, 4, dtype=torch.float64)            # tower passes the first 4 inputs through
    p["deep_bias_0"] = torch.full((1, 4), 10.0, dtype=torch.float64)   # keeps the ReLU open
    p["deep_res"] = torch.tensor([[0.0], [0.0], [0.0], [1.0]], dtype=torch.float64)
    p["deep_res_bias"] = torch.zeros(1, 1, dtype=torch.float64)
    b = {"cate_feats": torch.tensor([[4, 7, 0, 9, 0, 1]]), "cont_feats": torch.zeros(1, 1, dtype=torch.float64),
         "multi_feats_value": torch.tensor([[2.0, 5.0, 0.5, 1.0, 3.0]], dtype=torch.float64),
         "label": torch.ones(1, 1, dtype=torch.float64)}
    out, _ = O.dnn_multi_textline_forward(p, b, hp)
    W, T = p["weight_mat"], p["tags_emb"]
    pooled_kw = (2.0 * W[7] + 0.5 * W[9]) / 2.0
    # input order [cont(1) | cat_emb(2) | pooled kw(2) | pooled tag(2)]; output picks input 3 = pooled_kw[0] (+ bias 10)
    assert abs(float(out) - (float(pooled_kw[0]) + 10.0)) < 1e-14
    p["deep_res"] = torch.zeros(4, 1, dtype=torch.float64)
    p["deep_0"] = torch.zeros(7, 4, dtype=torch.float64); p["deep_0"][5, 0] = 1.0
    p["deep_res"][0, 0] = 1.0
    out, _ = O.dnn_multi_textline_forward(p, b, hp)
    pooled_tag = (1.0 * T[0] + 3.0 * T[1]) / 2.0
    assert abs(float(out) - (float(pooled_tag[0]) + 10.0)) < 1e-14


def test_textline_batch_keys_are_mapped():
    """utils/data_loader_textline.py:30-34 key names -> the canonical batch the engine reads"""
    from deep_learning_b200 import data as D
    from deep_learning_b200.spec import spec_from_hparams
    hp = H.hp_textline("dnn_multi_textline")
    spec = spec_from_hparams("dnn_multi_textline", hp)
    b = H.make_batch("dnn_multi_textline", hp, seed=3, B=8)
    Fc = hp.cate_field_size
    raw = {"labels": b["label"], "cont_feats": b["cont_feats"], "cat_feats": b["cate_feats"][:, :Fc],
           "mul_cat_feats": b["cate_feats"][:, Fc:], "mul_cat_feats_value": b["multi_feats_value"]}
    c = D.canonical_batch(raw, spec)
    assert torch.equal(c["cate_feats"], b["cate_feats"]) and torch.equal(c["multi_feats_value"], b["multi_feats_value"])
    assert torch.equal(c["label"], b["label"])
    assert spec.frozen_rows == 300 and spec.frozen_fields == (1,) and spec.weighted_pool
    off = spec.frozen_slot_offset()
    assert off[:Fc + 5].sum() == 0 and (off[Fc + 5:] == spec.table_rows).all()


def test_deepfm_pipeline_reads_two_rows_per_feature():
    """quirk Q2 (models/deepfm_pipeline.py:89,120) as a known answer: one sample, one categorical field with id 5, C = 2.
    FM row = 5 + 2 = 7, deep row = 5; continuous field 0 sits on the zeroed row 0 and contributes nothing."""
    hp = O.HParams(hidden_units=[2], batch_size=1, embedding_size=2, cont_field_size=2, cate_field_size=1, cate_feats_size=10,
                   vector_feats_size=0, l2_reg=0.0)
    p = O.init_params("deepfm_pipeline", hp)
    p["deep_0"] = torch.zeros(4, 2, dtype=torch.float64); p["deep_0"][2, 0] = 1.0; p["deep_0"][3, 1] = 1.0   # passes the embedding
    p["deep_bias_0"] = torch.full((1, 2), 5.0, dtype=torch.float64)
    w = torch.zeros(3 + 2 + 2, 1, dtype=torch.float64)
    p["deep_fm_bias"] = torch.zeros(1, dtype=torch.float64)
    b = {"cate_feats": torch.tensor([[5]]), "cont_feats": torch.tensor([[3.0, 0.5]], dtype=torch.float64),
         "vector_feats": torch.zeros(1, 0, dtype=torch.float64), "label": torch.ones(1, 1, dtype=torch.float64)}
    E, Wf = p["feats_emb"], p["fm_first_order_emb"]
    # first order: [cont0 * w[0] (zeroed) | cont1 * w[1] | w[7]]
    for j, want in enumerate([0.0, 0.5 * float(Wf[1]), float(Wf[7])]):
        w.zero_(); w[j] = 1.0; p["deep_fm_weight"] = w.clone()
        assert abs(float(O.deepfm_pipeline_forward(p, b, hp)[0]) - want) < 1e-14, j
    # second order over e = [0, 0.5 * E[1], E[7]]
    e = torch.stack([torch.zeros(2, dtype=torch.float64), 0.5 * E[1], E[7]])
    second = 0.5 * (e.sum(0) ** 2 - (e ** 2).sum(0))
    w.zero_(); w[3] = 1.0; p["deep_fm_weight"] = w.clone()
    assert abs(float(O.deepfm_pipeline_forward(p, b, hp)[0]) - float(second[0])) < 1e-14
    # deep part sees the RAW id's row
    w.zero_(); w[5] = 1.0; p["deep_fm_weight"] = w.clone()
    assert abs(float(O.deepfm_pipeline_forward(p, b, hp)[0]) - (float(E[5, 0]) + 5.0)) < 1e-14
    # a raw id below C aliases a continuous-field row; padding id 0 reads zeros in the deep part but row C in the FM part
    b0 = dict(b, cate_feats=torch.tensor([[0]]))
    assert abs(float(O.deepfm_pipeline_forward(p, b0, hp)[0]) - 5.0) < 1e-14
    w.zero_(); w[2] = 1.0; p["deep_fm_weight"] = w.clone()
    assert abs(float(O.deepfm_pipeline_forward(p, b0, hp)[0]) - float(Wf[2])) < 1e-14


def test_deepfm_pipeline_autograd_matches_finite_differences():
    hp = H.hp_c1(vector_feats_size=3, hidden_units=[8, 4], l2_reg=1e-2)
    p = O.init_params("deepfm_pipeline", hp)
    bt = H.make_batch("deepfm_pipeline", hp, seed=1, B=16)
    bt["cate_feats"][0, 0] = 0
    b = H.oracle_batch(bt, torch.float64)
    loss, _, _, grads = O.loss_and_grads("deepfm_pipeline", p, b, hp)
    assert not grads["feats_emb"][0].any() and not grads["fm_first_order_emb"][0].any()      # row 0 is cut out of the graph
    rng = np.random.RandomState(0)
    for name in p:
        gflat = grads[name].reshape(-1)
        nz = torch.nonzero(gflat).reshape(-1)
        idx = int(nz[rng.randint(len(nz))])
        eps = 1e-6
        q = {k: v.clone() for k, v in p.items()}
        q[name].reshape(-1)[idx] += eps
        lp = float(O.LOSS["deepfm_pipeline"](q, b, hp)[0])
        q[name].reshape(-1)[idx] -= 2 * eps
        lm = float(O.LOSS["deepfm_pipeline"](q, b, hp)[0])
        fd = (lp - lm) / (2 * eps)
        assert abs(fd - float(gflat[idx])) < 1e-6 * max(1.0, abs(fd)) + 1e-9, (name, fd, float(gflat[idx]))
    # the lazy-Adam row mask covers the support of the table gradient (a dead ReLU can zero a structurally touched row)
    tr = O.touched_rows("deepfm_pipeline", b, hp)
    support = (grads["feats_emb"].abs().sum(1) + grads["fm_first_order_emb"].abs().sum(1)) > 0
    assert not (support & ~tr["feats_emb"]).any() and not tr["feats_emb"][0]
    assert tr["feats_emb"][hp.cont_field_size]          # padding id 0 -> FM row C is touched


def test_tf1_adam_against_torch_optim_adam_over_ten_steps():
    """an independent implementation of the moment updates and both bias corrections: torch.optim.Adam computes
    theta -= lr/(1-b1^t) * m / (sqrt(v)/sqrt(1-b2^t) + eps), which is TF1's lr*sqrt(1-b2^t)/(1-b1^t) * m/(sqrt(v) + eps_tf)
    exactly when eps = eps_tf / sqrt(1-b2^t); with that per-step eps the two must agree over many steps, on gradients
    spanning the scale where eps matters (1e-9 .. 1)"""
    g = torch.Generator().manual_seed(5)
    p0 = {"w": torch.randn(7, 3, generator=g, dtype=torch.float64)}
    hp = O.HParams(learning_rate=1e-2)
    p, st = {k: v.clone() for k, v in p0.items()}, O.adam_init(p0)
    w = p0["w"].clone().requires_grad_(True)
    opt = torch.optim.Adam([w], lr=hp.learning_rate, betas=(O.ADAM_BETA1, O.ADAM_BETA2), eps=O.ADAM_EPS)
    for t in range(1, 11):
        grad = torch.randn(7, 3, generator=g, dtype=torch.float64) * torch.logspace(-9, 0, 21, dtype=torch.float64).reshape(7, 3)
        p = O.tf1_adam_step(p, {"w": grad}, st, hp)
        opt.param_groups[0]["eps"] = O.ADAM_EPS / math.sqrt(1 - O.ADAM_BETA2 ** t)
        w.grad = grad.clone()
        opt.step()
        assert float((p["w"] - w.detach()).abs().max()) < 1e-14, t


def test_relu_branch_override_is_per_thread():
    """the loopback multi-GPU tests run N ranks as N threads, each calling the oracle on its own batch: the masks one
    rank pins with `relu_branches` (global batch, 256 rows) must not reach another rank's 128-row forward"""
    import threading
    import ctr_helpers as H
    hp = H.hp_c1()
    p = O.init_params("deepfm", hp)
    b128 = H.oracle_batch(H.make_batch("deepfm", hp, seed=1, B=128), torch.float64)
    want = O.FORWARD["deepfm"](p, b128, hp)[0]
    masks = [torch.zeros(256, h, dtype=torch.bool) for h in hp.hidden_units]      # would zero the tower AND mis-size it
    go, done, box = threading.Event(), threading.Event(), {}

    def other():
        go.wait()
        try:
            box["out"] = O.FORWARD["deepfm"](p, b128, hp)[0]
        except Exception as e:           # noqa: BLE001 -- reported below
            box["err"] = e
        done.set()
    t = threading.Thread(target=other)
    t.start()
    with O.relu_branches(masks):
        go.set()
        done.wait()
    t.join()
    assert "err" not in box, box.get("err")
    assert torch.equal(box["out"], want)
