"""Host-side cost of one training step (not a test): wall time the Python driver needs to ENQUEUE a step, with the time it
spends blocked on the GPU (the one host read of the id-plan split sizes, sharded engine only) taken out.  If this is
close to the measured step time the step is host-bound and kernel-side savings do not show.

    python tests/host_timing.py            |  torchrun ... tests/host_timing.py
"""
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench as Bn                                                   # noqa: E402
from deep_learning_b200 import data as D                             # noqa: E402
from deep_learning_b200.engine import CtrEngine                      # noqa: E402
from deep_learning_b200.spec import spec_from_hparams                # noqa: E402
from deep_learning_b200.models._base import init_weights             # noqa: E402


def main():
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    B, R = int(os.environ.get("B", 65536)), int(os.environ.get("ROWS", 100_000_000))
    hp = Bn.hparams(B, R)
    spec = spec_from_hparams("deepfm", hp)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        from deep_learning_b200.sharded import ShardedCtrEngine
        eng = ShardedCtrEngine(spec, B, device=dev, learning_rate=hp.learning_rate, precision="tf32", adam_mode="lazy")
    else:
        eng = CtrEngine(spec, B, device=dev, learning_rate=hp.learning_rate, precision="tf32", adam_mode="lazy")
    eng.init_table_on_device(seed=2019)
    eng.set_dense_weights(init_weights(spec, 2019, with_table=False))
    sizes = D.field_sizes(R, spec.Fc)
    NB = 4
    batches = [D.synthetic_batch(B, spec.C, spec.Fc, sizes, seed=2020 + i * world + rank, device=dev) for i in range(NB)]
    for i in range(6):
        eng.train_step(batches[i % NB], next_batch=batches[(i + 1) % NB])
    torch.cuda.synchronize()
    steps = 12
    eng.host_prof = {"sync_s": 0.0}
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    t0 = time.perf_counter()
    for i in range(steps):
        eng.train_step(batches[i % NB], next_batch=batches[(i + 1) % NB])
    t1 = time.perf_counter()
    e1.record()
    torch.cuda.synchronize()
    gpu_ms = e0.elapsed_time(e1) / steps
    host_ms = (t1 - t0) * 1e3 / steps
    blocked_ms = eng.host_prof["sync_s"] * 1e3 / steps
    if rank == 0:
        print("world %d: GPU step %.3f ms | host enqueue loop %.3f ms/step, of which blocked on the GPU %.3f ms -> host work %.3f ms/step"
              % (world, gpu_ms, host_ms, blocked_ms, host_ms - blocked_ms))
    if os.environ.get("CPROFILE") and rank == 0:
        import cProfile
        import pstats
        pr = cProfile.Profile()
        pr.enable()
        for i in range(steps):
            eng.train_step(batches[i % NB], next_batch=batches[(i + 1) % NB])
        pr.disable()
        torch.cuda.synchronize()
        pstats.Stats(pr).sort_stats("cumulative").print_stats(35)
    elif os.environ.get("CPROFILE"):
        for i in range(steps):
            eng.train_step(batches[i % NB], next_batch=batches[(i + 1) % NB])
        torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
