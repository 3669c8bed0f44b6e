"""Diagnostic (not a test): distribution of the TF32-path gradient error against the fp64 oracle, and how
much of it is explained by ReLU-mask flips (a pre-activation within TF32 rounding of zero changes side;
the function value is continuous there, its gradient is not).  Run on a GPU box:

    python tests/diag_tf32.py
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

from oracle import ctr_oracle as O   # noqa: E402
import ctr_helpers as H              # noqa: E402


def study(hp, tag):
    from deep_learning_b200.engine import CtrEngine
    from deep_learning_b200.spec import spec_from_hparams
    p = O.init_params("deepfm", hp, seed=2019, dtype=torch.float64)
    st = O.adam_init(p)
    spec = spec_from_hparams("deepfm", hp)
    for prec in ("fp32", "tf32"):
        eng = CtrEngine(spec, hp.batch_size, learning_rate=hp.learning_rate, precision=prec, adam_mode="lazy")
        eng.set_weights({k: v.numpy() for k, v in p.items()})
        b = H.make_batch("deepfm", hp, seed=2020)
        ob = H.oracle_batch(b, torch.float64)
        import copy
        st2 = copy.deepcopy(st)
        _, p_new = O.train_step("deepfm", dict(p), st2, ob, hp, adam_mode="lazy")
        eng.train_step(b)
        got_m, _ = eng.get_optimizer_state()
        # oracle activations, canonical column order irrelevant for a flip COUNT per sample
        x = None
        with torch.no_grad():
            # replicate the tower on the oracle side to get the masks
            cate = ob["cate_feats"].long()
            emb = p["feats_emb"][cate].reshape(cate.shape[0], -1)
            x = torch.cat([ob["cont_feats"], ob["vector_feats"], emb], dim=1)
            masks = []
            for i in range(len(hp.hidden_units)):
                x = torch.relu(x @ p["deep_%d" % i] + p["deep_bias_%d" % i])
                masks.append((x > 0).numpy())
        flips = 0
        flipped_samples = np.zeros(hp.batch_size, dtype=bool)
        for i, mk in enumerate(masks):
            dm = (eng.acts[i][:, :hp.hidden_units[i]].cpu().numpy() > 0)
            diff = dm != mk
            flips += int(diff.sum())
            flipped_samples |= diff.any(axis=1)
        print("[%s/%s] relu flips: %d units, %d of %d samples" % (tag, prec, flips, int(flipped_samples.sum()), hp.batch_size))
        dev_masks = [torch.from_numpy(eng.acts[i][:, :hp.hidden_units[i]].cpu().numpy() > 0) for i in range(len(masks))]
        st3 = copy.deepcopy(st)
        with O.relu_branches(dev_masks):
            O.train_step("deepfm", dict(p), st3, ob, hp, adam_mode="lazy")
        worst = max(H.rel_err(got_m[n].reshape(-1), st3.m[n].numpy().reshape(-1)) for n in st3.m)
        print("  same-branch oracle: worst max-norm rel err of m over all params = %.3g" % worst)
        for n in sorted(st3.m):
            print("    same-branch m[%-16s] max %.3g" % (n, H.rel_err(got_m[n].reshape(-1), st3.m[n].numpy().reshape(-1))))
        for n in sorted(st2.m):
            ref = st2.m[n].numpy().reshape(-1)
            a = got_m[n].reshape(-1).astype(np.float64)
            scale = max(float(np.max(np.abs(ref))), 1e-30)
            d = np.abs(a - ref) / scale
            l2 = float(np.linalg.norm(a - ref) / max(np.linalg.norm(ref), 1e-30))
            nz = np.abs(ref) > 0
            print("  m[%-16s] max %.3g  p99.9 %.3g  p99 %.3g  median(nz) %.3g  relL2 %.3g  frac>2e-3 %.4f" % (
                n, d.max(), np.percentile(d, 99.9), np.percentile(d, 99), float(np.median(d[nz])) if nz.any() else 0.0, l2,
                float((d > 2e-3).mean())))
        del eng


if __name__ == "__main__":
    study(H.hp_c1(), "c1")
    study(H.hp_criteo(B=1024, rows=50000), "criteo1k")
