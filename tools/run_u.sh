set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_u.log 2>&1; echo "pytest rc $?" >> gpurun_out/pytest_u.log
timeout 300 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_u.json 2> gpurun_out/bench_u.err
timeout 200 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_u.json 2> gpurun_out/bench_ref_u.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_u.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launch_u.log 2>&1
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_u.log 2>&1
tail -3 gpurun_out/pytest_u.log; head -c 300 gpurun_out/bench_u.json; echo; cat gpurun_out/bench_ref_u.json | head -c 600; echo; cat gpurun_out/smoke_u.log | tail -2
