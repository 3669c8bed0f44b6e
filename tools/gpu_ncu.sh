#!/bin/bash
# ncu --set full captures of selected kernels of the bench step (one launch each); read here with `ncu -i ... --page raw --csv`
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras"
$CMD > gpurun_out/ncu_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/ncu_plain.log; exit 1; }
for k in "$@"; do
  timeout 300 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:$k -c 1 -f \
      -o gpurun_out/prof_$k $CMD > gpurun_out/ncu_$k.log 2>&1
  echo "ncu $k rc $?"
done
