set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_k.log 2>&1; echo "pytest rc $?" >> gpurun_out/pytest_k.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
TIMERS=ctr_sort_ids,ctr_unique_segments,ctr_merge_sorted_segments,ctr_gather_rows_push,ctr_embed_bwd_finalize_push,ctr_segment_sum_rows,ctr_adam_rows,ctr_embed_bwd_partials,ctr_fm_gather_fwd,ctr_peer_barrier,ctr_shard_counts,ctr_invert_perm timeout 300 $TR tests/phase_timing.py > gpurun_out/ph2_k.log 2>&1
timeout 300 $TR bench.py --gpus 2 --steps 30 --warmup 5 > gpurun_out/bench_n2k.json 2> gpurun_out/bench_n2k.err
timeout 300 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_k.json 2> gpurun_out/bench_k.err
TIMERS=ctr_sort_ids,ctr_unique_segments,ctr_fm_gather_fwd,ctr_head_fwd,ctr_embed_bwd_partials,ctr_adam_rows timeout 200 python tests/phase_timing.py > gpurun_out/ph1_k.log 2>&1
tail -3 gpurun_out/pytest_k.log
