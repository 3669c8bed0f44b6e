"""Per-CTA phase timeline of the TF32 forward GEMM (not a test):  CTR_GEMM_DBG=8 [CTR_GEMM_MODE=..] python tests/trace_gemm.py"""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["CTR_GEMM_DBG"] = str(int(os.environ.get("CTR_GEMM_DBG", "8")) | 8)
from deep_learning_b200 import _lib           # noqa: E402
from deep_learning_b200._lib import ptr       # noqa: E402

B, K, N, ldx = 65536, 400, 400, 404
hd = _lib.Handle(0)
st = torch.cuda.current_stream().cuda_stream
X = torch.randn(B, ldx, device="cuda"); Y = torch.zeros(B, N + 4, device="cuda")
W = torch.randn(K, N, device="cuda") * 0.05; Wt = torch.zeros(N, K, device="cuda"); bias = torch.randn(N, device="cuda")
hd.call("ctr_transpose", st, ptr(W), K, N, ptr(Wt), K, 1)
for _ in range(3):
    hd.call("ctr_mlp_fwd", st, 1, B, K, N, ptr(X), ldx, ptr(W), ptr(Wt), ptr(bias), 1, 1.0, 0, ptr(Y), N + 4)
torch.cuda.synchronize()
n = 1024 if os.environ.get("CTR_GEMM_MODE", "split2").startswith("sp") else 512
full = np.zeros((4096 + 64, 8), dtype=np.uint64)
fn = hd.lib.ctr_debug_gemm_trace
fn.argtypes = [C.c_void_p, C.c_int]
assert fn(full.ctypes.data, 4096 + 64) == 0
t = full[:n].astype(np.int64)
fine = full[4096:].reshape(-1).astype(np.int64)
t0 = t[:, 0].min()
names = ["start->setup", "setup->mma_done", "mma_done->acc_ready", "acc_ready->epi_done", "epi_done->synced", "synced->freed"]
d = np.stack([t[:, 1] - t[:, 0], t[:, 2] - t[:, 1], t[:, 3] - t[:, 2], t[:, 4] - t[:, 3], t[:, 5] - t[:, 4], t[:, 6] - t[:, 5]], 1) / 1e3
print("kernel span: %.1f us; CTA lifetime median %.2f us" % ((t[:, 6].max() - t0) / 1e3, np.median(t[:, 6] - t[:, 0]) / 1e3))
for i, nm in enumerate(names):
    print("  %-22s median %6.2f us   p10 %6.2f  p90 %6.2f" % (nm, np.median(d[:, i]), np.percentile(d[:, i], 10), np.percentile(d[:, i], 90)))
order = np.argsort(t[:, 0])
print("CTA start times (us) of the first 8 and around the first wave boundary:", np.round((t[order[:8], 0] - t0) / 1e3, 2),
      np.round((t[order[146:152], 0] - t0) / 1e3, 2))

base = t[0, 0]
prod = (fine[0:14] - base) / 1e3
mma = (fine[64:78] - base) / 1e3
print("block 0 producer got a free stage at (us):", np.round(prod, 2))
print("block 0 MMA issuer saw a full stage at (us):", np.round(mma, 2))
ep0 = fine[128:128 + 64].reshape(-1, 4)[:, :2]
ep0 = ep0[ep0[:, 0] > 0]
print("block 0 epilogue warp 0, per chunk [ldtm done, chunk done] (us):", np.round((ep0 - base) / 1e3, 2).tolist())
# global phase picture: how many CTAs are in the main loop / in the epilogue at each instant
T = (t[:, 6].max() - t0) / 1e3
grid_t = np.linspace(0, T, 41)
ml = [(int(np.sum((t[:, 1] - t0 <= x * 1e3) & (t[:, 2] - t0 > x * 1e3))), int(np.sum((t[:, 3] - t0 <= x * 1e3) & (t[:, 4] - t0 > x * 1e3)))) for x in grid_t]
print("time(us): CTAs in main loop / in epilogue")
print("  ".join("%.0f:%d/%d" % (x, a, b) for x, (a, b) in zip(grid_t, ml)))
