#!/bin/bash
# Round 2, GPU call H (one B200): PBiCGstab + sor2sma at 1024^3 -- launch list of one run and a full capture of the operator kernel.
mkdir -p gpurun_out
O=gpurun_out
timeout 200 python scratch/solver_bench.py pbicgstab 1024 4 sor2sma f32 1.1 > $O/bicg_plain.log 2>&1 && \
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ -c 300 --csv --log-file $O/bicg_launches.csv \
  python scratch/solver_bench.py pbicgstab 1024 4 sor2sma f32 1.1 > $O/ncuH1.log 2>&1; echo "ncu list rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_lin_pk --launch-skip 3 --launch-count 1 \
  -f -o $O/prof_ax python scratch/solver_bench.py pbicgstab 1024 4 sor2sma f32 1.1 > $O/ncuH2.log 2>&1; echo "ncu ax rc=$?"
cat $O/bicg_plain.log | cut -c1-300
