#!/bin/bash
# Round 2, GPU call 1 (one B200):  gpurun --timeout 2400 -- 'bash scratch/r2_call1.sh'
# Measurement and parity first: Fortran probe, whole GPU suite (incl. the new direct-oracle
# full-size tests and the MAF cases collected directly), smoke, the bench line with every new key, the reference arm,
# first MAF timings + launch list, then the experiments that steer this round's kernel work.
mkdir -p gpurun_out
O=gpurun_out
{ echo "== fortran probe"; for c in gfortran flang flang-new nvfortran ifort ifx f95 f77; do printf "%s: " $c; (command -v $c && $c --version 2>&1 | head -1) || echo absent; done
  echo "== host"; nproc; lscpu | grep -i "model name\|^CPU(s)\|socket\|numa node(s)"; free -g | head -2; nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv,noheader; } > $O/probe.log 2>&1
timeout 2000 python -m pytest tests -m gpu -q -rxXs --durations=15 > $O/gpu_tests.log 2>&1; echo "gpu tests rc=$?" | tee -a $O/gpu_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/smoke.log
timeout 900 python bench.py > $O/r02_bench_line.json 2> $O/bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > $O/r02_reference_line.json 2> $O/reference.err; echo "reference rc=$?"
# MAF: first timings (device resident, fixed iteration window) and a launch list
for s in jacobi_maf sor2sma_maf pcr_rb_maf pcr_rb_esa_maf; do
  python scratch/solver_bench.py $s 512 20 >> $O/maf_timing.log 2>&1
done
python scratch/solver_bench.py pbicgstab_maf 512 5 sor2sma_maf f32 1.1 >> $O/maf_timing.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/maf_launches.csv \
  python scratch/solver_bench.py sor2sma_maf 512 6 > $O/ncu_maf.log 2>&1
# --- experiments -------------------------------------------------------------------------------------------
# (a) per-chunk overhead of the two-colour sweep on the slabs of 2 / 4 / 8 GPUs
for ny in 512 256 128; do
  for njc in 0 1 2 3 4 5 6 8; do
    if [ $njc = 0 ]; then python scratch/slab_bench.py 1024 $ny 1024 sor2sma 40 >> $O/slab_njc.log 2>&1
    else CZB_NJC=$njc python scratch/slab_bench.py 1024 $ny 1024 sor2sma 40 >> $O/slab_njc.log 2>&1; fi
  done
done
# (b) small / medium grids: where the launch latency and the tile shape bite
for c in "jacobi 64 2000" "sor2sma 64 2000" "sor2sma 128 1000" "sor2sma 256 400" "sor2sma 512 100" "jacobi 256 400" "pcr_rb 256 200" "pcr_rb 512 100" "pcr_rb 1024 20" "pcr_rb_esa 1024 10" "pcr_j_esa 512 50"; do
  python scratch/solver_bench.py $c >> $O/small_grids.log 2>&1
done
for njc in 2 4 8 16 32; do CZB_NJC=$njc python scratch/solver_bench.py sor2sma 256 400 >> $O/small_grids.log 2>&1; done
# (c) double precision
for c in "sor2sma 512 40 - f64" "jacobi 512 40 - f64" "pcr_rb 512 20 - f64" "pbicgstab 512 6 sor2sma f64 1.1" "sor2sma 1024 10 - f64"; do
  python scratch/solver_bench.py $c >> $O/f64.log 2>&1
done
tail -4 $O/gpu_tests.log; cat $O/probe.log | head -12; cat $O/maf_timing.log; cut -c1-1500 $O/r02_bench_line.json; cat $O/r02_reference_line.json | cut -c1-600
