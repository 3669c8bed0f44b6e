set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_m.log 2>&1; echo "pytest rc $?" >> gpurun_out/pytest_m.log
timeout 300 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_m.json 2> gpurun_out/bench_m.err
TIMERS=ctr_sort_ids,ctr_unique_segments,ctr_fm_gather_fwd,ctr_head_fwd,ctr_embed_bwd_partials,ctr_adam_rows,ctr_fm_cont_bwd,ctr_dense_update,ctr_mlp_fwd,ctr_mlp_bwd_data,ctr_mlp_bwd_weight timeout 200 python tests/phase_timing.py > gpurun_out/ph1_m.log 2>&1
tail -15 gpurun_out/pytest_m.log; cat gpurun_out/ph1_m.log; head -c 400 gpurun_out/bench_m.json
