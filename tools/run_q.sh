set -x
mkdir -p gpurun_out
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
for cfg in "deepfm lazy" "wdl lazy" "deepfm_multi tf_dense"; do
  timeout 300 $TR tests/mgpu_parity.py $cfg > gpurun_out/parity_q$N.log 2>&1; echo "parity $cfg rc $?" | tee -a gpurun_out/parity_q$N.rc
  tail -3 gpurun_out/parity_q$N.log
done
timeout 300 $TR tests/host_timing.py > gpurun_out/host_q$N.log 2>&1; grep "^world" gpurun_out/host_q$N.log
timeout 300 $TR bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_q$N.json 2> gpurun_out/bench_q$N.err
head -c 400 gpurun_out/bench_q$N.json; tail -3 gpurun_out/bench_q$N.err
