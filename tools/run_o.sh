set -x
mkdir -p gpurun_out
N=${1:-2}
python -m pytest tests/test_gpu_sharded.py tests/test_gpu_kernels.py -m gpu -x -q > gpurun_out/pytest_o$N.log 2>&1; echo "pytest rc $?" >> gpurun_out/pytest_o$N.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $TR tests/mgpu_parity.py deepfm lazy > gpurun_out/parity_o$N.log 2>&1; echo "parity rc $?" >> gpurun_out/parity_o$N.log
timeout 300 $TR bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_o$N.json 2> gpurun_out/bench_o$N.err
TIMERS=ctr_sort_ids,ctr_unique_segments,ctr_merge_sorted_segments,ctr_gather_rows_push,ctr_embed_bwd_finalize_push,ctr_segment_sum_rows,ctr_adam_rows,ctr_embed_bwd_partials,ctr_fm_gather_fwd,ctr_peer_barrier,ctr_head_fwd,ctr_dense_update timeout 300 $TR tests/phase_timing.py > gpurun_out/ph_o$N.log 2>&1
tail -3 gpurun_out/pytest_o$N.log; tail -4 gpurun_out/parity_o$N.log; cat gpurun_out/ph_o$N.log; head -c 300 gpurun_out/bench_o$N.json
