"""Stand-alone timing of the embedding kernels at the config-4/5 shape (not a test):

    [CTR_K1_GENERAL=1] [DIST=zipf|uniform] [B=65536] [ROWS=100000000] python tests/bench_embed.py

K1 (ctr_fm_gather_fwd) alone on a 100 M-row table, rotating over 4 batches / X0 buffers (> 126 MB L2), CUDA events on the
launching stream.  Prints the duration, the algorithmic GB/s (1932 B per sample) and the fraction of
the measured HBM peak; with SWEEP=1 the predict-only batch sweep of config 5 (1K .. 1M)."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from deep_learning_b200 import _lib, data as D        # noqa: E402
from deep_learning_b200._lib import EmbedCfg, ptr     # noqa: E402


def main():
    Bmax = int(os.environ.get("B", 65536))
    R = int(os.environ.get("ROWS", 100_000_000))
    dist = os.environ.get("DIST", "zipf")
    rnd = int(os.environ.get("RND", 1))
    peak = 6539.5
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    hd = _lib.Handle(0)
    st = torch.cuda.current_stream().cuda_stream
    k, Fc, C, stride = 16, 26, 13, 24
    x0_ld, fm_ld = 432, 56
    table = torch.empty(R + C, stride, device="cuda")
    table.uniform_(-0.01, 0.01)
    sizes = D.field_sizes(R, Fc)
    NB = 4
    sweep = [1024, 4096, 16384, 65536, 262144, 1048576] if os.environ.get("SWEEP") else [Bmax]
    for B in sweep:
        nb = NB if B <= 262144 else 2
        bs = [D.synthetic_batch(B, C, Fc, sizes, seed=2020 + i, dist=dist, device="cuda") for i in range(nb)]
        ids = [b["cate_feats"].to(torch.int32).contiguous() for b in bs]
        cont = [b["cont_feats"].contiguous() for b in bs]
        x0 = [torch.empty(B, x0_ld, device="cuda") for _ in range(nb)]
        fm = [torch.empty(B, fm_ld, device="cuda") for _ in range(nb)]
        se = [torch.empty(B, k, device="cuda") for _ in range(nb)]
        cfg = EmbedCfg(B=B, id_bytes=4, ids_ld=Fc, Fc=Fc, Fm=0, M=0, C=C, V=0, k=k, has_first=1, use_fm=1, fm_cont=1, zero_row0=0,
                       count_rule=0, combiner=0, table_stride=stride, x0_ld=x0_ld, fm_ld=fm_ld, grad_stride=20, round_tf32=rnd)
        crow = table[R:]

        def run(i):
            hd.call("ctr_fm_gather_fwd", st, cfg, ptr(table), ptr(ids[i]), 0, ptr(cont[i]), None, ptr(crow), None, ptr(x0[i]),
                    ptr(fm[i]), ptr(se[i]))
        n = 40 if B <= 65536 else 12
        for i in range(5):
            run(i % nb)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(n):
            run(i % nb)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / n * 1e3
        alg = 1932.0 * B
        uniq = int(torch.unique(ids[0]).numel())
        print("K1 B=%7d %s ids (U/B=%.2f): %7.1f us  algorithmic %.0f GB/s = %.1f %% of %.0f; with the X0 tile (%d B/sample) %.0f GB/s"
              % (B, dist, uniq / B, us, alg / us / 1e3, 100 * alg / us / 1e3 / peak, peak, x0_ld * 4, (alg + x0_ld * 4.0 * B) / us / 1e3))
        del bs, ids, cont, x0, fm, se


if __name__ == "__main__":
    main()
