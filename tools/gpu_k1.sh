#!/bin/bash
# K1 variants alone on the GPU (tests/bench_embed.py), then the embed parity tests.
mkdir -p gpurun_out
python -m pytest tests/test_gpu_embed_kernels.py tests/test_gpu_faults.py tests/test_gpu_golden.py -q > gpurun_out/pytest_k1.log 2>&1; echo "rc $?" >> gpurun_out/pytest_k1.log
( for d in zipf uniform; do
  echo "== regs $d"; CTR_K1=regs DIST=$d python tests/bench_embed.py
  for ns in 4 6; do for c in 0 2 3; do
    echo "== staged NS=$ns CTAS=$c $d"; CTR_K1_NS=$ns CTR_K1_CTAS=$c DIST=$d python tests/bench_embed.py
  done; done
done
echo "== sweep staged"; SWEEP=1 python tests/bench_embed.py ) > gpurun_out/k1_variants.txt 2>&1
tail -n 5 gpurun_out/pytest_k1.log
