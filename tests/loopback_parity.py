"""Prints the world-N parity runs of tests/mgpu_parity.py with the ranks as threads on ONE GPU (loopback comm): the log
committed under profiles/ as evidence beside the torchrun runs on real peers.

    python tests/loopback_parity.py > gpurun_out/parity_loopback.txt"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import mgpu_parity                                        # noqa: E402
from deep_learning_b200.comm import LoopbackWorld         # noqa: E402

if __name__ == "__main__":
    for world, alg, mode in [(2, "deepfm", "lazy"), (2, "deepfm", "tf_dense"), (2, "wdl", "lazy"), (2, "deepfm_multi", "tf_dense"),
                             (2, "wdl_textline", "lazy"), (4, "deepfm", "lazy"), (8, "deepfm", "lazy"), (8, "deepfm", "tf_dense"),
                             (8, "deepfm_multi", "lazy")]:
        print("==== world %d %s %s" % (world, alg, mode), flush=True)
        LoopbackWorld(world).run(lambda comm: mgpu_parity.run(comm, alg, mode))
