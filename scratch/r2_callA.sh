#!/bin/bash
# Round 2, session 2, GPU call A (one B200):  gpurun --timeout 1700 -- 'bash scratch/r2_callA.sh'
# The records of the first session's calls were lost with its container: everything judged is re-measured here and
# copied to profiles/r02_* right after the call.
mkdir -p gpurun_out
O=gpurun_out
{ echo "== fortran probe"; for c in gfortran flang flang-new nvfortran ifort ifx f95 f77; do printf "%s: " $c; (command -v $c && $c --version 2>&1 | head -1) || echo absent; done
  echo "== host"; nproc; lscpu | grep -i "model name\|^CPU(s)\|socket\|numa node(s)"; free -g | head -2; nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv,noheader; } > $O/probe.log 2>&1
timeout 1000 python -m pytest tests -m gpu -q -rxXs --durations=12 > $O/gpu_tests.log 2>&1; echo "gpu tests rc=$?" | tee -a $O/gpu_tests.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/smoke.log
timeout 700 python bench.py > $O/r02_bench_line.json 2> $O/bench.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > $O/r02_reference_line.json 2> $O/reference.err; echo "reference rc=$?"
# launch list of the bench command (after it has exited 0 without ncu), then one full capture of the two dominant kernels
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-also --no-ttc > $O/b2.log 2>&1 && \
timeout 500 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv \
  python bench.py --steps 5 --warmup 3 --no-cpu --no-also --no-ttc > $O/ncu1.log 2>&1; echo "ncu list rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_rb2_pk --launch-skip 8 --launch-count 1 \
  -f -o $O/prof_rb2pk python scratch/solver_bench.py sor2sma 1024 6 > $O/ncu2.log 2>&1; echo "ncu rb2 rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_pcr_warp_tm --launch-skip 8 --launch-count 1 \
  -f -o $O/prof_pcr python scratch/solver_bench.py pcr_rb 512 6 > $O/ncu3.log 2>&1; echo "ncu pcr rc=$?"
# panel: every solver family at the BASELINE sizes, device resident
for c in "jacobi 64 2000" "sor2sma 256 400" "sor2sma 512 100" "jacobi 512 100" "pcr_rb 256 200" "pcr_rb 512 100" "pcr_rb 1024 20" \
         "pcr_rb_esa 512 50" "pcr_rb_esa 1024 10" "pcr_j_esa 512 50" "pbicgstab 512 6 sor2sma f32 1.1" "pbicgstab 1024 4 sor2sma f32 1.1" \
         "sor2sma 512 40 - f64" "jacobi 512 40 - f64" "pcr_rb 512 20 - f64" "pbicgstab 512 6 sor2sma f64 1.1" \
         "jacobi_maf 512 20" "sor2sma_maf 512 20" "pcr_rb_maf 512 10" "pbicgstab_maf 512 4 sor2sma_maf f32 1.1"; do
  timeout 120 python scratch/solver_bench.py $c >> $O/panel.log 2>&1
done
for ny in 128 256 512; do timeout 120 python scratch/slab_bench.py 1024 $ny 1024 sor2sma 40 >> $O/panel.log 2>&1; done
tail -4 $O/gpu_tests.log; head -10 $O/probe.log; cut -c1-1200 $O/r02_bench_line.json; cut -c1-500 $O/r02_reference_line.json; cat $O/panel.log | cut -c1-250
