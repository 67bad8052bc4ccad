#!/bin/bash
# Round 2, GPU call 4 (one B200): is the slab instantiation (SIG) of the two-colour kernel still slower than the plain one?
mkdir -p gpurun_out
O=gpurun_out
for ny in 128 512; do
  python scratch/slab_bench.py 1024 $ny 1024 sor2sma 40 >> $O/exp4.log 2>&1
  CZB_EXP_FORCE_SIG=1 python scratch/slab_bench.py 1024 $ny 1024 sor2sma 40 >> $O/exp4.log 2>&1
  CZB_EXP_FORCE_SIG=1 CZB_EXP_NO_REMAP=1 python scratch/slab_bench.py 1024 $ny 1024 sor2sma 40 >> $O/exp4.log 2>&1
done
timeout 900 python -m pytest tests/test_gpu_solve.py tests/test_gpu_kernels.py -m gpu -q -x > $O/gpu_tests4.log 2>&1; echo "gpu tests rc=$?" | tee -a $O/gpu_tests4.log
( cd /tmp && CZB_PROFILE=1 $OLDPWD/cubez_b200/cz_f32 512 512 512 pbicgstab 20 1.1 sor2sma > $OLDPWD/$O/cz_profile.log 2>&1; cp profiling.txt $OLDPWD/$O/profiling_pbicgstab_512.txt )
( cd /tmp && CZB_PROFILE=1 $OLDPWD/cubez_b200/cz_f32 512 512 512 pcr_rb 50 1.5 >> $OLDPWD/$O/cz_profile.log 2>&1; cp profiling.txt $OLDPWD/$O/profiling_pcr_rb_512.txt )
cat $O/exp4.log | cut -c1-200; tail -3 $O/gpu_tests4.log; cat $O/profiling_pbicgstab_512.txt
