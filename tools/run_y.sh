set -x
mkdir -p gpurun_out
python -m pytest tests/test_gpu_bwd_kernels.py tests/test_gpu_sharded.py -m gpu -q > gpurun_out/pytest_y.log 2>&1; echo "pytest rc $?" >> gpurun_out/pytest_y.log
tail -6 gpurun_out/pytest_y.log | cut -c1-400
