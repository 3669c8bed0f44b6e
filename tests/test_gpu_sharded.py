"""-m gpu: the row-sharded multi-GPU engine against the oracle on the concatenated batch.
world=1 is a loopback through the full exchange path (NCCL with one rank) and runs on any GPU box; world=2 needs
two GPUs and is skipped otherwise (the driver's scaling run and `gpurun --gpus 2` exercise it)."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(world, alg, adam_mode):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "mgpu_parity.py"), alg, adam_mode]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0 and "MGPU PARITY OK" in r.stdout, r.stdout[-3000:] + "\n" + r.stderr[-3000:]


@pytest.mark.parametrize("alg,adam_mode", [("deepfm", "lazy"), ("deepfm", "tf_dense"), ("deepfm_multi", "lazy"),
                                           ("wdl", "lazy"), ("dnn", "lazy"), ("wdl_textline", "lazy")])
def test_sharded_loopback_world1(alg, adam_mode):
    _run(1, alg, adam_mode)


@pytest.mark.parametrize("alg,adam_mode", [("deepfm", "lazy"), ("deepfm_multi", "tf_dense"), ("wdl", "lazy")])
def test_sharded_world2(alg, adam_mode):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    _run(2, alg, adam_mode)
