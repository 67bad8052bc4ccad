#!/bin/bash
# Round 2, session 2, GPU call C (one B200): persistent small-grid loop, new shell copy, launch list of OUR kernels.
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --durations=8 > $O/gpu_testsC.log 2>&1; echo "gpu tests rc=$?" | tee -a $O/gpu_testsC.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > $O/smokeC.log 2>&1; echo "smoke rc=$?" | tee -a $O/smokeC.log
for c in "jacobi 64 2000" "sor2sma 64 2000" "jacobi 128 500" "sor2sma 128 500" "sor2sma 160 300" "jacobi 160 300" "sor2sma 200 200"; do
  timeout 120 python scratch/solver_bench.py $c >> $O/panelC.log 2>&1
  CZB_PERSIST_MB=0 timeout 120 python scratch/solver_bench.py $c >> $O/panelC.log 2>&1
  CZB_PERSIST_MB=200 timeout 120 python scratch/solver_bench.py $c >> $O/panelC.log 2>&1
done
timeout 120 python scratch/call_overhead.py 1024 1024 1024 sor2sma >> $O/panelC.log 2>&1
timeout 120 python scratch/call_overhead.py 64 64 64 jacobi >> $O/panelC.log 2>&1
timeout 700 python bench.py > $O/r02c_bench_line.json 2> $O/benchC.err; echo "bench rc=$?"
timeout 500 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ -c 400 --csv --log-file $O/launchesC.csv \
  python bench.py --steps 5 --warmup 3 --no-cpu --no-also --no-ttc > $O/ncu1C.log 2>&1; echo "ncu list rc=$?"
tail -5 $O/gpu_testsC.log; cat $O/smokeC.log; cat $O/panelC.log | cut -c1-330; cut -c1-400 $O/r02c_bench_line.json
python - <<'PY'
import json
d=[json.loads(l) for l in open("gpurun_out/r02c_bench_line.json") if l.startswith("{")][-1]
print({k:(v["iterations"], round(v["solver_loop_ms"],2)) for k,v in d.get("time_to_convergence",{}).items()})
print(d["windows_ms"], d["roofline"]["ms_per_launch"], d["e2e"]["value"])
PY
