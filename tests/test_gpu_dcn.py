"""-m gpu: DCN (reference test_model/DCN.py) on the CUDA path -- the cross-network kernels against a
NumPy restatement, the whole step against the oracle, and the reference's own run of that file (tests/golden/ref_dcn.npz,
made under the TF1 stand-in) replayed free-running with the float32 floor beside it.  FP32, bars as tests/test_ref_gpu.py."""
import ctypes as C
import os

import numpy as np
import pytest
import torch

import ctr_helpers as H
from oracle import ctr_oracle as O
from golden.make_golden import tiny_hp
from test_ref_golden import HERE

pytestmark = pytest.mark.gpu


def _st():
    return torch.cuda.current_stream().cuda_stream


@pytest.mark.parametrize("B,D,L", [(1, 1, 1), (77, 7, 2), (513, 40, 2), (300, 128, 4)])
def test_cross_kernels_equal_the_closed_form(B, D, L):
    from deep_learning_b200 import _lib
    from deep_learning_b200._lib import CrossCfg, ptr
    hd = _lib.Handle(0)
    g = np.random.RandomState(B + D)
    x0 = g.standard_normal((B, D)); w = g.standard_normal((L, D)) * 0.3; b = g.standard_normal((L, D)) * 0.1
    dl = g.standard_normal(B) * 0.01; wp = g.standard_normal(D)
    # float64 closed form: forward, then the reverse walk (see csrc/cross_net.cu)
    xs = [x0]
    for l in range(L):
        s = xs[-1] @ w[l]
        xs.append(x0 * s[:, None] + b[l][None, :] + xs[-1])
    gw, gb = np.zeros((L, D)), np.zeros((L, D))
    gg = dl[:, None] * wp[None, :]
    for l in range(L - 1, -1, -1):
        t = (gg * x0).sum(1)
        gb[l] = gg.sum(0); gw[l] = (t[:, None] * xs[l]).sum(0)
        gg = gg + t[:, None] * w[l][None, :]
    ld = (D + 3) // 4 * 4
    t32 = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).cuda()
    x0_d = torch.zeros(B, ld, device="cuda"); x0_d[:, :D] = t32(x0)
    w_d, b_d, gw_d, gb_d = t32(w), t32(b), torch.full((L, D), 9.0, device="cuda"), torch.full((L, D), 9.0, device="cuda")
    out = torch.full((B, ld), 7.0, device="cuda")
    c = CrossCfg()
    c.B, c.D, c.n_layers, c.reserved, c.ldx = B, D, L, 0, ld
    for l in range(L):
        c.w[l], c.b[l], c.gw[l], c.gb[l] = w_d[l].data_ptr(), b_d[l].data_ptr(), gw_d[l].data_ptr(), gb_d[l].data_ptr()
    hd.call("ctr_cross_fwd", _st(), C.byref(c), ptr(x0_d), ptr(out), ld)
    wsn = hd.query("ctr_cross_bwd_ws_bytes", C.byref(c))
    ws = torch.zeros(wsn, dtype=torch.uint8, device="cuda")
    dl_d, wp_d = t32(dl), t32(wp)            # named: a temporary would be freed (and its memory reused) before the kernel runs
    hd.call("ctr_cross_bwd", _st(), C.byref(c), ptr(x0_d), ptr(dl_d), ptr(wp_d), ptr(ws), wsn)
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    assert H.rel_err(got[:, :D], xs[-1]) < 1e-5 and np.all(got[:, D:] == 7.0)
    assert H.rel_err(gw_d.cpu().numpy(), gw) < 2e-5 and H.rel_err(gb_d.cpu().numpy(), gb) < 2e-5
    # deterministic: a second launch gives the same bits
    gw2 = gw_d.clone()
    hd.call("ctr_cross_bwd", _st(), C.byref(c), ptr(x0_d), ptr(dl_d), ptr(wp_d), ptr(ws), wsn)
    torch.cuda.synchronize()
    assert torch.equal(gw2, gw_d)


def _engine(hp, p, B):
    from deep_learning_b200.dcn_engine import DcnEngine
    eng = DcnEngine(hp.deep_feature_size, hp.hidden_units, B, learning_rate=hp.learning_rate, decay_steps=hp.learning_rate_decay_steps,
                    decay_rate=hp.learning_rate_decay_rate, l2_reg=hp.l2_reg)
    eng.set_weights({k: v.numpy() for k, v in p.items()})
    return eng


def _batch(hp, seed, B):
    g = torch.Generator(); g.manual_seed(seed)
    return {"features": torch.randn(B, hp.deep_feature_size, generator=g), "label": (torch.rand(B, 1, generator=g) < 0.3).float()}


@pytest.mark.parametrize("D,hidden,B", [(40, [64, 32], 256), (13, [400, 400, 400], 1000)])
def test_dcn_steps_match_the_oracle(D, hidden, B):
    hp = O.HParams(hidden_units=hidden, batch_size=B, deep_feature_size=D, l2_reg=1e-3, learning_rate=1e-3)
    p = O.init_params("dcn", hp, seed=2019, dtype=torch.float64)
    st = O.adam_init(p)
    eng = _engine(hp, p, B)
    back = eng.get_weights()
    for n in p:
        assert np.array_equal(back[n].reshape(-1), p[n].float().numpy().reshape(-1)), n
    for step in range(3):
        b = _batch(hp, 100 + step, B if step else B - 3)          # a short batch too
        ob = H.oracle_batch(b, torch.float64)
        out_ref, score_ref = O.FORWARD["dcn"](p, ob, hp)
        score = eng.predict(b).cpu().numpy()
        assert np.max(np.abs(score - score_ref.numpy().reshape(-1))) < 1e-5
        loss_ref, p = O.train_step("dcn", p, st, ob, hp)
        loss = float(eng.train_step(b).item())
        assert abs(loss - float(loss_ref)) < 1e-5 * max(1.0, abs(float(loss_ref))), (step, loss, float(loss_ref))
        got = eng.get_weights()
        got_m, _ = eng.get_optimizer_state()
        for n in p:
            # the first moment is linear in the gradient: a tight check of the backward pass itself
            assert H.rel_err(got_m[n].reshape(-1), st.m[n].numpy().reshape(-1)) < 2e-5, ("m", n, step)
            assert H.rel_err(got[n].reshape(-1), p[n].numpy().reshape(-1)) < 2e-5, (n, step)


def test_dcn_reproduces_the_reference_run():
    """test_model/DCN.py, unmodified, ran under the TF1 stand-in: ref_dcn.npz (float64) and fp32floor_dcn.npz (float32)"""
    z = np.load(os.path.join(HERE, "golden", "ref_dcn.npz"))
    zf = np.load(os.path.join(HERE, "golden", "fp32floor_dcn.npz"))
    hp = tiny_hp("dcn")
    p = {k[5:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("init/")}
    eng = _engine(hp, p, hp.batch_size)
    ntb = len([k for k in z.files if k.startswith("train") and k.endswith("/label")])
    t = e = 0
    rows = []
    for ev in str(z["events"]):
        name = ("train%d" % (t % ntb)) if ev == "T" else ("val%d" % (e % int(z["n_val"])))
        b = {k.split("/", 1)[1]: torch.from_numpy(z[k]) for k in z.files if k.startswith(name + "/")}
        if ev == "T":
            loss, ref = float(eng.train_step(b).item()), float(z["losses"][t])
            rows.append(("loss %d" % t, abs(loss - ref) / max(1.0, abs(ref)), abs(float(zf["losses"][t]) - ref) / max(1.0, abs(ref))))
            t += 1
        else:
            got, ref = eng.predict(b).cpu().numpy(), z["eval%d" % e].reshape(-1)
            rows.append(("eval %d" % e, float(np.max(np.abs(got - ref))), float(np.max(np.abs(zf["eval%d" % e].reshape(-1) - ref)))))
            e += 1
    got = eng.get_weights()
    for k in [k[6:] for k in z.files if k.startswith("final/")]:
        ref = z["final/" + k].reshape(-1)
        rows.append(("final " + k, H.rel_err(got[k].reshape(-1), ref), H.rel_err(zf["final/" + k].reshape(-1), ref)))
    for what, err, floor in rows:
        assert err < 1e-5, (what, err, floor)
        assert err <= max(8.0 * floor, 2e-6), (what, err, floor)


def test_dcn_module_fit_export_predict(tmp_path, capsys):
    """the launcher-facing surface of test_model/DCN.py: DeepModel(parser).fit(train, val) prints the per-epoch line and
    exports; predict() reloads the export and prints the AUC; the export is self-describing"""
    import pickle
    import deep_learning_b200.models.dcn as alg_model
    from deep_learning_b200.models._base import read_export_meta
    hp = O.HParams(hidden_units=[32, 16], batch_size=128, deep_feature_size=20, l2_reg=1e-4, learning_rate=1e-2, epochs=2)
    hp.model_pb = str(tmp_path / "pb")
    pk = lambda b: pickle.dumps({"features": b["features"].numpy().tolist(), "labels": b["label"].numpy().tolist()})
    train = [pk(_batch(hp, s, 128)) for s in range(4)]
    val = [pk(_batch(hp, 50 + s, 128 if s else 60)) for s in range(2)]
    m = alg_model.DeepModel(hp)
    m.fit(train, val)
    want = m._auc(val)
    auc = m.predict(val)
    out = capsys.readouterr().out
    assert out.count("valid-auc=") == 2 and "val of auc:" in out
    assert abs(auc - want) < 1e-12 and m.engine.global_step == 8
    meta = read_export_meta(hp.model_pb)
    assert meta["alg"] == "dcn" and meta["signature"]["outputs"]["output"]["tensor"] == "Sigmoid:0"
    assert meta["hparams"]["deep_feature_size"] == 20
