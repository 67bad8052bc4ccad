#!/bin/bash
mkdir -p gpurun_out
for cfg in "sor2sma 256 100" "sor2sma 512 60" "sor2sma 1024 20" "jacobi 512 60" "pbicgstab 512 12 sor2sma" "pbicgstab 1024 6 sor2sma" "sor2sma 512 40 - f64" "pbicgstab 512 12 sor2sma f64"; do
  timeout 300 python scratch/solver_bench.py $cfg 2>&1 | tail -1
  CZB_ROUNDS=8 timeout 300 python scratch/solver_bench.py $cfg 2>&1 | tail -1
done | tee gpurun_out/njc_bench.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu4.log 2>&1; tail -2 gpurun_out/pytest_gpu4.log
