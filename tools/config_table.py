"""One timed line per BASELINE.json config (the rows of BASELINE.md section 5); run on the GPU box.

    python tools/config_table.py > gpurun_out/config_table.txt

c1 models/deepfm.py B=256 (FP32 tower, TF1 dense Adam: the reference's exact semantics; tiny, launch-bound)
c2 models/deepfm_multi.py B=16K, k=16, 13 cont + 20 single + 6 multi-valued fields (top_n 16), TF32, lazy Adam
c3 models/wdl.py Criteo shape B=64K, k=16, MLP 400x3, 1e7-row table, TF32, lazy Adam
c5 predict-only DeepFM, c4 shapes, B = 1K .. 1M (scores only)
(c4 is bench.py itself.)  CUDA events, 5 warm-up steps, batches rotate; synthetic Zipf ids."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import ctr_helpers as H                                           # noqa: E402  (batch + hyper-parameter helpers only)
from deep_learning_b200 import data as D                          # noqa: E402
from deep_learning_b200.engine import CtrEngine                   # noqa: E402
from deep_learning_b200.models._base import init_weights          # noqa: E402
from deep_learning_b200.spec import spec_from_hparams             # noqa: E402


def timed(fn, n, warm=5):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def train_line(tag, alg, hp, precision, adam_mode, steps, big_table=False):
    spec = spec_from_hparams(alg, hp)
    eng = CtrEngine(spec, hp.batch_size, learning_rate=hp.learning_rate, precision=precision, adam_mode=adam_mode)
    if big_table:
        eng.init_table_on_device(seed=2019)
        eng.set_dense_weights(init_weights(spec, 2019, with_table=False))
    else:
        eng.set_weights(init_weights(spec, 2019))
    bs = [{k: v.cuda() for k, v in H.make_batch(alg, hp, seed=2020 + i).items()} for i in range(4)]
    ms = timed(lambda i: eng.train_step(bs[i % 4], next_batch=bs[(i + 1) % 4]), steps)
    u = int(eng.n_unique.item()) if spec.ids_ld else 0
    print("%-4s %-13s B=%-6d %s / %-8s train: %8.3f ms/step  %10.2f K samples/s   U/B=%.2f  table %d rows" % (
        tag, alg, hp.batch_size, precision, adam_mode, ms, hp.batch_size / ms, u / hp.batch_size, spec.table_rows), flush=True)
    del eng
    torch.cuda.empty_cache()


def main():
    torch.cuda.set_device(0)
    train_line("c1", "deepfm", H.hp_c1(), "fp32", "tf_dense", 200)
    train_line("c1", "deepfm", H.hp_c1(), "tf32", "lazy", 200)
    train_line("c2", "deepfm_multi", H.hp_multi(B=16384, cate_feats_size=2_000_001), "tf32", "lazy", 50, big_table=True)
    train_line("c3", "wdl", H.hp_criteo(B=65536, rows=10_000_000), "tf32", "lazy", 30, big_table=True)
    # c5: predict-only sweep at the c4 shapes
    R = 100_000_000
    hp = H.hp_criteo(B=1 << 20, rows=R)
    spec = spec_from_hparams("deepfm", hp)
    eng = CtrEngine(spec, 1 << 20, precision="tf32", train=False)
    eng.init_table_on_device(seed=2019)
    eng.set_dense_weights(init_weights(spec, 2019, with_table=False))
    sizes = D.field_sizes(R, spec.Fc)
    for B in (1 << 10, 1 << 12, 1 << 14, 1 << 16, 1 << 18, 1 << 20):
        bs = [D.synthetic_batch(B, spec.C, spec.Fc, sizes, seed=3000 + i, device="cuda") for i in range(2 if B > (1 << 18) else 4)]
        ms = timed(lambda i: eng.predict(bs[i % len(bs)]), 40 if B <= (1 << 16) else 10)
        print("c5   deepfm predict B=%-8d tf32: %8.3f ms/batch  %10.2f M samples/s" % (B, ms, B / ms / 1e3), flush=True)
        if B <= (1 << 16):      # the forward pass replayed from a CUDA graph: what a serving loop at small batches would call
            msg = timed(lambda i: eng.predict_graphed(bs[i % len(bs)]), 40)
            print("c5   deepfm predict B=%-8d tf32: %8.3f ms/batch  %10.2f M samples/s   (predict_graphed)" % (B, msg, B / msg / 1e3),
                  flush=True)
        del bs


if __name__ == "__main__":
    main()
