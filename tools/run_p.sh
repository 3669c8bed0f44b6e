set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
CPROFILE=1 timeout 200 python tests/host_timing.py > gpurun_out/host_p1.log 2>&1
CPROFILE=1 timeout 300 $TR tests/host_timing.py > gpurun_out/host_p2.log 2>&1
grep "^world" gpurun_out/host_p1.log gpurun_out/host_p2.log
