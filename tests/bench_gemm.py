"""Stand-alone timing of the MLP GEMM entry points at the config-4 shapes (not a test):

    CTR_GEMM_MODE=split2|pair|single python tests/bench_gemm.py

CUDA events on the launching stream, 5 warm-up + 20 timed launches per op, rotating over 3 buffer sets (> 126 MB L2)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from deep_learning_b200 import _lib           # noqa: E402
from deep_learning_b200._lib import ptr       # noqa: E402


def main():
    B = int(os.environ.get("B", 65536))
    hd = _lib.Handle(0)
    st = torch.cuda.current_stream().cuda_stream
    NSET = 3

    def timeit(fn, n=20):
        for i in range(5):
            fn(i % NSET)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(n):
            fn(i % NSET)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n * 1e3

    for (K, N, ldx) in [(429, 400, 432), (400, 400, 404)]:
        X = [torch.randn(B, ldx, device="cuda") for _ in range(NSET)]
        Y = [torch.zeros(B, N + 4, device="cuda") for _ in range(NSET)]
        dZ = [torch.randn(B, N, device="cuda") for _ in range(NSET)]
        dX = [torch.zeros(B, K if K % 4 == 0 else K + (4 - K % 4), device="cuda") for _ in range(NSET)]
        W = torch.randn(K, N, device="cuda") * 0.05
        Wt = torch.zeros(N, (K + 3) // 4 * 4, device="cuda")
        hd.call("ctr_transpose", st, ptr(W), K, N, ptr(Wt), Wt.shape[1], 1)
        bias = torch.randn(N, device="cuda")
        dW, db = torch.zeros(K, N, device="cuda"), torch.zeros(N, device="cuda")
        wsn = hd.query("ctr_mlp_bwd_weight_ws_bytes", 1, B, K, N)
        ws = torch.zeros(wsn, dtype=torch.uint8, device="cuda")
        flops = 2.0 * B * K * N
        t = timeit(lambda i: hd.call("ctr_mlp_fwd", st, 1, B, K, N, ptr(X[i]), ldx, ptr(W), ptr(Wt), ptr(bias), 1, 1.0, 0,
                                     ptr(Y[i]), N + 4))
        print("fwd   K=%d N=%d: %7.1f us  %6.1f TFLOP/s  (HBM floor %.1f us)" % (K, N, t, flops / t / 1e6, (B * (ldx + N) * 4) / 6.5395e6))
        t = timeit(lambda i: hd.call("ctr_mlp_bwd_data", st, 1, B, K, N, ptr(dZ[i]), N, ptr(W), ptr(X[i]), ldx, 1.0,
                                     ptr(dX[i]), dX[i].shape[1]))
        print("dgrad K=%d N=%d: %7.1f us  %6.1f TFLOP/s  (HBM floor %.1f us)" % (K, N, t, flops / t / 1e6, (B * (N + 2 * K) * 4) / 6.5395e6))
        t = timeit(lambda i: hd.call("ctr_mlp_bwd_weight", st, 1, B, K, N, ptr(X[i]), ldx, ptr(dZ[i]), N, ptr(dW), ptr(db),
                                     ptr(ws), wsn, 0))
        print("wgrad K=%d N=%d: %7.1f us  %6.1f TFLOP/s  (HBM floor %.1f us)" % (K, N, t, flops / t / 1e6, (B * (K + N) * 4) / 6.5395e6))
        nb = hd.query("ctr_mlp_relu_bits_bytes", B, N)
        bits = [torch.zeros(nb // 4, dtype=torch.int32, device="cuda") for _ in range(NSET)]
        t = timeit(lambda i: hd.call("ctr_mlp_fwd_bits", st, B, K, N, ptr(X[i]), ldx, ptr(Wt), ptr(bias), 1.0, 0, ptr(Y[i]), N + 4,
                                     ptr(bits[i])))
        print("fwd + sign bits      : %7.1f us" % t)
        if K == N:      # the bits of a [B, N] activation mask the data gradient whose output is [B, K = N]
            t = timeit(lambda i: hd.call("ctr_mlp_bwd_data_bits", st, B, K, N, ptr(dZ[i]), N, ptr(W), ptr(bits[i]), 1.0,
                                         ptr(dX[i]), dX[i].shape[1]))
            print("dgrad with sign bits : %7.1f us" % t)
        # the library GEMM on the same operands (cuBLAS TF32 through torch.matmul: the bare product, no bias / ReLU / mask
        # epilogue and no tf32 rounding of the output), as the yardstick for what this shape allows
        torch.backends.cuda.matmul.allow_tf32 = True
        Yc = [torch.zeros(B, N, device="cuda") for _ in range(NSET)]
        dXc = [torch.zeros(B, K, device="cuda") for _ in range(NSET)]
        dWc = torch.zeros(K, N, device="cuda")
        t = timeit(lambda i: torch.mm(X[i][:, :K], W, out=Yc[i]))
        print("  cuBLAS fwd  : %7.1f us  %6.1f TFLOP/s" % (t, flops / t / 1e6))
        t = timeit(lambda i: torch.mm(dZ[i], W.t(), out=dXc[i]))
        print("  cuBLAS dgrad: %7.1f us  %6.1f TFLOP/s" % (t, flops / t / 1e6))
        t = timeit(lambda i: torch.mm(X[i][:, :K].t(), dZ[i], out=dWc))
        print("  cuBLAS wgrad: %7.1f us  %6.1f TFLOP/s" % (t, flops / t / 1e6))
        del Yc, dXc


if __name__ == "__main__":
    main()
