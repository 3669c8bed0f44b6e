"""Per-tile phase times of the persistent CTA-pair GEMM (block 0):  CTR_GEMM_MODE=pp CTR_GEMM_DBG=8 python tests/trace_gemm_pp.py"""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from deep_learning_b200 import _lib           # noqa: E402
from deep_learning_b200._lib import ptr       # noqa: E402

B, K, N, ldx = 65536, 400, 400, 404
LDY = int(os.environ.get("LDY", N + 4))
hd = _lib.Handle(0)
st = torch.cuda.current_stream().cuda_stream
X = torch.randn(B, ldx, device="cuda"); Y = torch.zeros(B, LDY, device="cuda")
W = torch.randn(K, N, device="cuda") * 0.05; Wt = torch.zeros(N, K, device="cuda"); bias = torch.randn(N, device="cuda")
hd.call("ctr_transpose", st, ptr(W), K, N, ptr(Wt), K, 1)
dgrad = len(sys.argv) > 1 and sys.argv[1] == "dgrad"
dZ = torch.randn(B, N, device="cuda"); dX = torch.zeros(B, ldx, device="cuda")
for _ in range(3):
    if dgrad:
        hd.call("ctr_mlp_bwd_data", st, 1, B, K, N, ptr(dZ), N, ptr(W), ptr(X), ldx, 1.0, ptr(dX), ldx)
    else:
        hd.call("ctr_mlp_fwd", st, 1, B, K, N, ptr(X), ldx, ptr(W), ptr(Wt), ptr(bias), 1, 1.0, 0, ptr(Y), LDY)
torch.cuda.synchronize()
full = np.zeros((4096 + 64, 8), dtype=np.uint64)
assert hd.lib.ctr_debug_gemm_trace(full.ctypes.data, 4096 + 64) == 0
t = full[:148].astype(np.int64)
fine = full[4096:].reshape(-1).astype(np.int64)
t0 = t[0, 0]
print("block 0: start 0, setup done %.2f, all MMAs issued %.2f, epilogues done %.2f, exit %.2f us" % tuple((t[0, i] - t0) / 1e3 for i in (1, 2, 4, 6)))
print("kernel span %.1f us" % ((t[:, 6].max() - t[:, 0].min()) / 1e3))
for tl in range(10):
    v = fine[256 + 4 * tl: 256 + 4 * tl + 4]
    if v[0] <= 0:
        break
    print("tile %d: mma start %.2f  mma issued %.2f  acc ready %.2f  epilogue done %.2f   (main loop %.2f us, epilogue %.2f us)" % (
        (tl,) + tuple((x - t0) / 1e3 for x in v) + ((v[2] - v[0]) / 1e3, (v[3] - v[2]) / 1e3)))
