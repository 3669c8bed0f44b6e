set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_w.log 2>&1; echo "pytest rc $?" >> gpurun_out/pytest_w.log
timeout 200 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_w.json 2> gpurun_out/bench_w.err
tail -25 gpurun_out/pytest_w.log | cut -c1-300; head -c 300 gpurun_out/bench_w.json
