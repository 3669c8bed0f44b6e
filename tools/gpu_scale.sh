#!/bin/bash
# One 8-GPU box visit: bench.py at N = 2, 4, 8 (the driver's scaling run), world-8 parity for one configuration, phase and
# host timing at N = 8.     /usr/local/graft/bin/gpurun --gpus 8 --timeout 900 -- 'bash tools/gpu_scale.sh'
set -x
mkdir -p gpurun_out
for N in 2 4 8; do
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N"
  timeout 240 $TR bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
  tail -c 600 gpurun_out/bench_n$N.json | cut -c1-10 > /dev/null
done
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520"
timeout 200 $TR tests/mgpu_parity.py deepfm lazy > gpurun_out/parity_n8.log 2>&1; echo "parity deepfm lazy rc $?" >> gpurun_out/parity_n8.log
timeout 200 $TR tests/phase_timing.py > gpurun_out/phases_n8.txt 2>&1
timeout 150 $TR tests/host_timing.py > gpurun_out/host_n8.txt 2>&1
for N in 2 4 8; do python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/bench_n$N.json").read().strip().splitlines()[-1])
    print($N, d["value"], d["ms_per_step"], d["e2e"]["value"], d["clocks"])
except Exception as e:
    print($N, "no line", e)
PY
done
tail -n 4 gpurun_out/parity_n8.log; tail -n 16 gpurun_out/phases_n8.txt
