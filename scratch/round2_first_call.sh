#!/bin/bash
# First gpurun call of round 2 (one B200):  gpurun --timeout 2400 -- 'bash scratch/round2_first_call.sh'
# Everything written at the end of round 1 without a GPU gets its first hardware run here.
mkdir -p gpurun_out
# 1. stretched-grid kernels: the 42 parity cases, directly (no child process / xfail wrapper)
timeout 1500 python -m pytest tests/gpu_maf_cases.py -m gpu -q -x > gpurun_out/maf_cases.log 2>&1; echo "maf cases rc=$?" | tee -a gpurun_out/maf_cases.log
# 2. the whole GPU suite
timeout 2400 python -m pytest tests -m gpu -x -q -rxX > gpurun_out/gpu_tests.log 2>&1; echo "gpu tests rc=$?" | tee -a gpurun_out/gpu_tests.log
# 3. bench line incl. the new time_to_convergence key
python bench.py > gpurun_out/bench_line.json 2> gpurun_out/bench.err; echo "bench rc=$?"
# 4. time the MAF solvers (device resident) and take a launch list of one of them
for s in jacobi_maf sor2sma_maf pcr_rb_maf; do
  cubez_b200/cz_f32 512 512 512 $s 40 1.0 2>&1 | grep -i "solver loop\|flop count" | sed "s/^/$s 512^3: /" >> gpurun_out/maf_timing.log
done
cubez_b200/cz_f32 512 512 512 pbicgstab_maf 6 1.1 sor2sma_maf 2>&1 | grep -i "solver loop\|flop count" | sed "s/^/pbicgstab_maf+sor2sma_maf 512^3: /" >> gpurun_out/maf_timing.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/maf_launches.csv \
  cubez_b200/cz_f32 512 512 512 sor2sma_maf 10 1.5 > gpurun_out/ncu_maf.log 2>&1
tail -3 gpurun_out/maf_cases.log gpurun_out/gpu_tests.log; cat gpurun_out/maf_timing.log; cut -c1-300 gpurun_out/bench_line.json
