"""Stand-alone timing of the id plan kernels at the config-4 shape (not a test):  python tests/bench_sort.py
ctr_sort_ids (26 x 64K Zipf ids + 13 continuous rows, 27-bit keys) and ctr_unique_segments, CUDA events, alone on the GPU."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from deep_learning_b200 import _lib, data as D        # noqa: E402
from deep_learning_b200._lib import ptr               # noqa: E402


def main():
    B, R, Fc, C = int(os.environ.get("B", 65536)), 100_000_000, 26, 13
    hd = _lib.Handle(0)
    st = torch.cuda.current_stream().cuda_stream
    sizes = D.field_sizes(R, Fc)
    ids = [D.synthetic_batch(B, C, Fc, sizes, seed=2020 + i, device="cuda")["cate_feats"].to(torch.int32).contiguous() for i in range(4)]
    n = B * Fc + C
    u32 = dict(dtype=torch.int32, device="cuda")
    keys, perm, uniq, uidx = (torch.zeros(n, **u32) for _ in range(4))
    seg, nu = torch.zeros(n + 2, **u32), torch.zeros(1, **u32)
    wsn = max(hd.query("ctr_sort_ws_bytes", n), 4 * (n // 2048 + 8))
    ws = torch.zeros(wsn, dtype=torch.uint8, device="cuda")
    for world, lb in ((1, 27), (8, 24)):
        kb = lb + (world - 1).bit_length()

        def sort(i):
            hd.call("ctr_sort_ids", st, ptr(ids[i]), 4, B, Fc, 0, Fc, 0, 0, Fc, C, world, lb, kb, 0, None, ptr(keys), ptr(perm), ptr(ws), wsn)

        def uniqf(i):
            hd.call("ctr_unique_segments", st, ptr(keys), n, ptr(uniq), ptr(seg), ptr(uidx), ptr(nu), ptr(ws), wsn, None)
        for name, fn in (("ctr_sort_ids", sort), ("ctr_unique_segments", uniqf)):
            for i in range(4):
                fn(i)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(20):
                fn(i % 4)
            e1.record()
            torch.cuda.synchronize()
            print("world %d  %-20s n=%d key_bits=%d: %7.1f us" % (world, name, n, kb, e0.elapsed_time(e1) / 20 * 1e3))


if __name__ == "__main__":
    main()
