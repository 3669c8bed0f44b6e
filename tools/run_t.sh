set -x
mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $TR bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_t$N.json 2> gpurun_out/bench_t$N.err
head -c 330 gpurun_out/bench_t$N.json; tail -3 gpurun_out/bench_t$N.err
TIMERS=ctr_gather_rows_push,ctr_embed_bwd_finalize_push,ctr_segment_sum_rows,ctr_adam_rows,ctr_peer_barrier timeout 200 $TR tests/phase_timing.py > gpurun_out/ph_t$N.log 2>&1
grep -v "^\*\|OMP\|^$" gpurun_out/ph_t$N.log
