"""The measured parity table (run on the GPU box; output committed under profiles/).

    python tools/parity_report.py > gpurun_out/parity_report.txt

For each of the 13 reference model files: the CUDA path (FP32 tower, TF1 dense Adam, free-running over the reference's
own `fit` schedule) against the float64 fixture of that run, next to the float32 stand-in's distance to the same
fixture (the floor float32 arithmetic sets).  Same numbers tests/test_ref_gpu.py asserts on; here they are printed."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import test_ref_gpu as T   # noqa: E402


def main():
    print("%-20s %-24s %10s %10s %7s" % ("model", "quantity (worst of kind)", "cuda err", "fp32 floor", "ratio"))
    worst_all = 0.0
    for alg in T.ALGS:
        rows = T.replay(alg)
        kinds = {}
        for what, err, floor in rows:
            kind = what.split()[0] if not what.startswith("final") else what
            if kind not in kinds or err > kinds[kind][0]:
                kinds[kind] = (err, floor)
        for kind, (err, floor) in kinds.items():
            print("%-20s %-24s %10.2e %10.2e %7.1f" % (alg, kind, err, floor, err / max(floor, 1e-30)))
            worst_all = max(worst_all, err)
    print("worst over all models and quantities: %.2e (bar %.0e)" % (worst_all, T.BAR))


if __name__ == "__main__":
    main()
