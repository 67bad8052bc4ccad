#!/bin/bash
# Round 2, GPU call 6 (one B200): chunk-count model, shell copy, whole GPU suite after the multi-rank rewrite.
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --durations=8 > $O/gpu_tests6.log 2>&1; echo "gpu tests rc=$?" | tee -a $O/gpu_tests6.log
for c in "sor2sma 512 100" "sor2sma 256 400" "sor2sma 128 1000" "sor2sma 64 2000" "jacobi 512 100" "jacobi 256 400" "jacobi 64 2000" "sor2sma 1024 30" "jacobi 1024 30"; do
  python scratch/solver_bench.py $c >> $O/exp6.log 2>&1
done
for njc in 2 3 4 6 8 13; do CZB_NJC=$njc python scratch/solver_bench.py sor2sma 512 100 >> $O/exp6.log 2>&1; done
for ny in 128 256; do python scratch/slab_bench.py 1024 $ny 1024 sor2sma 40 >> $O/exp6.log 2>&1; done
python bench.py --no-ttc > $O/bench6.json 2> $O/bench6.err; echo "bench rc=$?"
tail -12 $O/gpu_tests6.log; cat $O/exp6.log | cut -c1-250; cut -c1-700 $O/bench6.json
