"""Prints the world-N parity runs of tests/mgpu_parity.py with the ranks as threads on ONE GPU (loopback comm): the log
committed under profiles/ as evidence beside the torchrun runs on real peers.

    python tests/loopback_parity.py [world alg mode ...] > gpurun_out/parity_loopback.txt"""
import os
import sys
import traceback

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import mgpu_parity                                        # noqa: E402
from deep_learning_b200.comm import LoopbackWorld         # noqa: E402

DEFAULT = [(2, "deepfm", "lazy"), (2, "deepfm", "tf_dense"), (2, "wdl", "lazy"), (2, "deepfm_multi", "tf_dense"),
           (2, "wdl_textline", "lazy"), (4, "deepfm", "lazy"), (8, "deepfm", "lazy"), (8, "deepfm", "tf_dense"),
           (8, "deepfm_multi", "lazy"), (8, "deepfm_multi", "tf_dense"), (4, "dnn_multi", "lazy")]

if __name__ == "__main__":
    a = sys.argv[1:]
    runs = [(int(a[i]), a[i + 1], a[i + 2]) for i in range(0, len(a), 3)] if a else DEFAULT
    bad = 0
    for world, alg, mode in runs:
        print("==== world %d %s %s" % (world, alg, mode), flush=True)
        try:
            LoopbackWorld(world).run(lambda comm: mgpu_parity.run(comm, alg, mode))
        except BaseException:        # noqa: BLE001 -- keep going: this is a report
            bad += 1
            traceback.print_exc(file=sys.stdout)
            print("FAILED world %d %s %s" % (world, alg, mode), flush=True)
    print("loopback parity: %d of %d runs failed" % (bad, len(runs)))
    sys.exit(1 if bad else 0)
