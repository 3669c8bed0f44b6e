set -x
mkdir -p gpurun_out
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
python -m pytest tests/test_gpu_peer_kernels.py tests/test_gpu_bwd_kernels.py tests/test_gpu_kernels.py -m gpu -x -q > gpurun_out/pytest_s.log 2>&1; echo "pytest rc $?" >> gpurun_out/pytest_s.log; tail -3 gpurun_out/pytest_s.log
for cfg in "deepfm lazy" "wdl_textline lazy"; do
  timeout 300 $TR tests/mgpu_parity.py $cfg > gpurun_out/parity_s$N.log 2>&1; echo "parity $cfg rc $?" | tee -a gpurun_out/parity_s$N.rc
  tail -2 gpurun_out/parity_s$N.log
done
timeout 300 $TR tests/host_timing.py > gpurun_out/host_s$N.log 2>&1; grep "^world" gpurun_out/host_s$N.log
timeout 300 $TR bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_s$N.json 2> gpurun_out/bench_s$N.err
head -c 330 gpurun_out/bench_s$N.json; tail -3 gpurun_out/bench_s$N.err
