#!/bin/bash
# One GPU-box visit (run under gpurun): GPU tests, parity table, bench, phase picture, launch list.  N=${1:-1} ranks.
#   /usr/local/graft/bin/gpurun [--gpus N] --timeout 900 -- 'bash tools/gpu_round.sh N'
set -x
mkdir -p gpurun_out
N=${1:-1}
if [ "$N" = "1" ]; then
  python -m pytest tests -m gpu -q -s > gpurun_out/pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/pytest.log
  timeout 300 python tools/parity_report.py > gpurun_out/parity_report.txt 2>&1
  timeout 600 python tests/loopback_parity.py > gpurun_out/parity_loopback.txt 2>&1
  timeout 400 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err
  timeout 200 python tests/phase_timing.py > gpurun_out/phases_n1.txt 2>&1
  timeout 300 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
      --log-file gpurun_out/launches_n1.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/ncu_launch.log 2>&1
else
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
  for cfg in "deepfm lazy" "deepfm tf_dense" "wdl lazy" "deepfm_multi lazy"; do
    timeout 300 $TR tests/mgpu_parity.py $cfg >> gpurun_out/parity_n$N.log 2>&1; echo "parity $cfg rc $?" >> gpurun_out/parity_n$N.log
  done
  timeout 300 $TR bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
  timeout 300 $TR tests/phase_timing.py > gpurun_out/phases_n$N.txt 2>&1
  timeout 200 $TR tests/host_timing.py > gpurun_out/host_n$N.txt 2>&1
fi
tail -n 3 gpurun_out/*.log | cut -c1-300
