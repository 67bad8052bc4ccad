#!/bin/bash
mkdir -p gpurun_out
for c in 0 1; do
  for cfg in "sor2sma 512 40 - f64" "jacobi 512 40 - f64" "pbicgstab 512 12 sor2sma f64"; do
    CZB_MARCH_CFG=$c timeout 300 python scratch/solver_bench.py $cfg 2>&1 | tail -1
  done
done | tee gpurun_out/f64_bench.log
CZB_MARCH_CFG=1 timeout 600 python -m pytest tests -m gpu -x -q -k "f64 or double" > gpurun_out/f64_tests.log 2>&1; tail -2 gpurun_out/f64_tests.log
