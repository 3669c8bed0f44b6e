set -x
mkdir -p gpurun_out
python -m pytest tests/test_gpu_kernels.py tests/test_gpu_parity.py tests/test_gpu_golden.py -m gpu -x -q > gpurun_out/pytest_n.log 2>&1; echo "pytest rc $?" >> gpurun_out/pytest_n.log
timeout 300 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_n.json 2> gpurun_out/bench_n.err
TIMERS=ctr_sort_ids,ctr_unique_segments,ctr_fm_gather_fwd,ctr_head_fwd,ctr_embed_bwd_partials,ctr_adam_rows,ctr_fm_cont_bwd,ctr_dense_update,ctr_mlp_fwd,ctr_mlp_bwd_data,ctr_mlp_bwd_weight timeout 200 python tests/phase_timing.py > gpurun_out/ph1_n.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:"head_fwd|embed_bwd_partials" -c 2 -o gpurun_out/ncu_head_part_n python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_n.log 2>&1
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_n.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launch_n.log 2>&1
tail -4 gpurun_out/pytest_n.log; cat gpurun_out/ph1_n.log; head -c 300 gpurun_out/bench_n.json
