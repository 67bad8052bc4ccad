#!/bin/bash
# GPU check of the line-SOR kernels: tests + timings
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "pcr or lsor or solve" > gpurun_out/pcr_tests.log 2>&1; echo "tests rc=$?" 
tail -3 gpurun_out/pcr_tests.log
for cfg in "pcr_rb 512 40" "pcr_rb 1024 20" "pcr_j_esa 512 40" "pcr_rb 256 60"; do
  timeout 300 python scratch/solver_bench.py $cfg 2>&1 | tail -1
  CZB_PCR_NOBULK=1 timeout 300 python scratch/solver_bench.py $cfg 2>&1 | tail -1
  CZB_PCR_NOTMEM=1 timeout 300 python scratch/solver_bench.py $cfg 2>&1 | tail -1
done | tee gpurun_out/pcr_bench.log
