#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
for ny in 128 256 1024; do python scratch/slab_bench.py 1024 $ny 1024 sor2sma 40 >> $O/exp9.log 2>&1; done
python scratch/solver_bench.py sor2sma 512 100 >> $O/exp9.log 2>&1
python scratch/solver_bench.py pbicgstab 512 6 sor2sma f32 1.1 >> $O/exp9.log 2>&1
python scratch/call_overhead.py 1024 128 1024 sor2sma >> $O/exp9.log 2>&1
python scratch/call_overhead.py 1024 1024 1024 sor2sma >> $O/exp9.log 2>&1
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_solve.py tests/test_gpu_fullsize.py -m gpu -q -x > $O/gpu_tests9.log 2>&1; echo "gpu tests rc=$?" | tee -a $O/gpu_tests9.log
cat $O/exp9.log | cut -c1-400; tail -3 $O/gpu_tests9.log
