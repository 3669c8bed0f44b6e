#!/bin/bash
# One GPU-box visit (run under gpurun): GPU tests, bench, phase picture, launch list.  N=${1:-1} ranks.
#   /usr/local/graft/bin/gpurun [--gpus N] --timeout 900 -- 'bash tools/gpu_round.sh N'
set -x
mkdir -p gpurun_out
N=${1:-1}
if [ "$N" = "1" ]; then
  python -m pytest tests -m gpu -q -s > gpurun_out/pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/pytest.log
  timeout 300 python tools/parity_report.py > gpurun_out/parity_report.txt 2>&1
  timeout 300 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err
  timeout 200 python tests/phase_timing.py > gpurun_out/phases_n1.txt 2>&1
  # A/B knobs that are in the tree but default off
  CTR_GEMM_DBG=16 timeout 300 python -m pytest tests/test_gpu_kernels.py -q -k 'gemm or mlp' > gpurun_out/pytest_gemm_direct_epilogue.log 2>&1
  CTR_GEMM_DBG=16 timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_n1_direct_epilogue.json 2> gpurun_out/bench_n1_direct_epilogue.err
  CTR_SORT_AT=head timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_n1_sort_at_head.json 2> gpurun_out/bench_n1_sort_at_head.err
  timeout 300 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
      --log-file gpurun_out/launches_n1.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
  # one-off probes for open design questions; each bounded, none touches the library
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/probe_gather4 tools/probe_gather4.cu \
      && timeout 120 /tmp/probe_gather4 > gpurun_out/probe_gather4.txt 2>&1
else
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
  timeout 300 $TR tests/mgpu_parity.py deepfm lazy > gpurun_out/parity_n$N.log 2>&1; echo "parity rc $?" >> gpurun_out/parity_n$N.log
  timeout 300 $TR bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
  timeout 300 $TR tests/phase_timing.py > gpurun_out/phases_n$N.txt 2>&1
  timeout 200 $TR tests/host_timing.py > gpurun_out/host_n$N.txt 2>&1
fi
tail -3 gpurun_out/*.log | cut -c1-300
