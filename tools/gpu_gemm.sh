#!/bin/bash
# One GPU-box visit for the GEMM question (run under gpurun): kernel parity tests, then the GEMM entry points beside
# cuBLAS on the same operands for several producer widths, and the per-tile phase trace of the persistent pair kernel.
#   /usr/local/graft/bin/gpurun --timeout 600 -- 'bash tools/gpu_gemm.sh'
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_bwd_kernels.py -m gpu -q -x > gpurun_out/pytest_kernels.log 2>&1
for L in ${LANES:-1 2 4 6}; do
  CTR_GEMM_TMA_LANES=$L timeout 200 python tests/bench_gemm.py > gpurun_out/bench_gemm_l$L.txt 2>&1
done
CTR_GEMM_DBG=8 timeout 100 python tests/trace_gemm_pp.py > gpurun_out/trace_pp_fwd.txt 2>&1
CTR_GEMM_DBG=8 timeout 100 python tests/trace_gemm_pp.py dgrad > gpurun_out/trace_pp_dgrad.txt 2>&1
tail -n 3 gpurun_out/pytest_kernels.log; grep -v cuBLAS gpurun_out/bench_gemm_l*.txt; cat gpurun_out/trace_pp_fwd.txt
