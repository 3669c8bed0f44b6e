"""Main-stream phase picture of the training step (not a test).  One GPU, or under torchrun for the sharded engine:

    python tests/phase_timing.py
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tests/phase_timing.py

Each phase is the time between two events on the main stream, so it includes any wait on the side stream / NCCL."""
import os
import sys
from collections import OrderedDict

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench as Bn                                                   # noqa: E402
from deep_learning_b200 import data as D                             # noqa: E402
from deep_learning_b200.engine import CtrEngine                      # noqa: E402
from deep_learning_b200.spec import spec_from_hparams                # noqa: E402
from deep_learning_b200.models._base import init_weights                  # noqa: E402


def main():
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    B, R = int(os.environ.get("B", 65536)), int(os.environ.get("ROWS", 100_000_000))
    alg = os.environ.get("ALG", "deepfm")          # ALG=wdl ROWS=10000000: BASELINE.json configs[2]
    if alg == "deepfm":
        hp = Bn.hparams(B, R)
    else:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import ctr_helpers as H
        hp = H.hp_criteo(B=B, rows=R)
    spec = spec_from_hparams(alg, hp)
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
    if alg == "deepfm":
        batches = [D.synthetic_batch(B, spec.C, spec.Fc, sizes, seed=2020 + i * world + rank, device=dev) for i in range(NB)]
    else:
        batches = [{k: v.to(dev) for k, v in H.make_batch(alg, hp, seed=2020 + i * world + rank).items()} for i in range(NB)]
    pf = not os.environ.get("NOPREFETCH")
    for i in range(6):
        eng.train_step(batches[i % NB], next_batch=batches[(i + 1) % NB] if pf else None)
    torch.cuda.synchronize()
    steps = 20
    names = os.environ.get("TIMERS", "")
    if names:     # per-entry-point durations (events on the launching stream, so side-stream kernels are seen too)
        eng.h.timers = {n: [] for n in names.split(",")}
    eng.phase_marks = []
    for i in range(steps):
        eng.train_step(batches[i % NB], next_batch=batches[(i + 1) % NB] if pf else None)
    torch.cuda.synchronize()
    marks, eng.phase_marks = eng.phase_marks, None
    acc = OrderedDict()
    for (n0, e0), (n1, e1) in zip(marks[:-1], marks[1:]):
        key = "%s -> %s" % (n0, n1)
        acc.setdefault(key, []).append(e0.elapsed_time(e1) * 1e3)
    if rank == 0:
        tot = 0.0
        for k, v in acc.items():
            avg = sum(v) / steps
            tot += avg
            print("%-32s %8.1f us" % (k, avg))
        print("%-32s %8.1f us per step" % ("total", tot))
        if names:
            for n, ev in eng.h.timers.items():
                if ev:
                    print("  %-30s %8.1f us x %.1f per step" % (n, sum(a.elapsed_time(b) for a, b in ev) * 1e3 / len(ev), len(ev) / steps))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
