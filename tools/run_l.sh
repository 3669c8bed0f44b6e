set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_l.log 2>&1; echo "pytest rc $?" >> gpurun_out/pytest_l.log
timeout 300 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_l.json 2> gpurun_out/bench_l.err
SWEEP=1 timeout 200 python tests/bench_embed.py > gpurun_out/sweep_l.txt 2>&1
TIMERS=ctr_sort_ids,ctr_unique_segments,ctr_fm_gather_fwd,ctr_head_fwd,ctr_embed_bwd_partials,ctr_adam_rows,ctr_fm_cont_bwd,ctr_adam_dense,ctr_mlp_fwd,ctr_mlp_bwd_data,ctr_mlp_bwd_weight timeout 200 python tests/phase_timing.py > gpurun_out/ph1_l.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:"head_fwd|adam_rows_vec" -c 4 -o gpurun_out/ncu_head_adam_l python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_l.log 2>&1
tail -3 gpurun_out/pytest_l.log; cat gpurun_out/bench_l.json | head -c 600
