#!/bin/bash
# Round 2, GPU call G (one B200): final state -- no-alloc test alone, whole GPU suite, smoke, the default bench line,
# one full ncu capture of the Jacobi marching kernel.
mkdir -p gpurun_out
O=gpurun_out
timeout 200 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/noalloc_alone.log 2>&1; echo "noalloc alone rc=$?" | tee -a $O/noalloc_alone.log
timeout 900 python -m pytest tests -m gpu -q -x > $O/gpu_testsG.log 2>&1; echo "gpu tests rc=$?" | tee -a $O/gpu_testsG.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > $O/smokeG.log 2>&1; echo "smoke rc=$?" | tee -a $O/smokeG.log
timeout 800 python bench.py > $O/r02_final_bench_line.json 2> $O/benchG.err; echo "bench rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_lin_pk --launch-skip 6 --launch-count 1 \
  -f -o $O/prof_jac python scratch/solver_bench.py jacobi 1024 6 > $O/ncuG.log 2>&1; echo "ncu jac rc=$?"
tail -3 $O/noalloc_alone.log; tail -3 $O/gpu_testsG.log; cat $O/smokeG.log
python - <<'PY'
import json
d=[json.loads(l) for l in open("gpurun_out/r02_final_bench_line.json") if l.startswith("{")][-1]
print("value %.1f ms/step %.4f loop %.4f frac %.3f e2e %.1f clocks %s" % (d["value"], d["ms_per_step"], d["roofline"]["ms_per_launch"], d["roofline"]["frac"], d["e2e"]["value"], d["clocks"]))
print({k: round(v["value"],1) for k,v in d.get("also",{}).items()}, d.get("cpu_baseline",{}).get("value"))
print({k:(v["iterations"], v.get("oracle_iterations"), round(v["solver_loop_ms"],1)) for k,v in d.get("time_to_convergence",{}).items()})
print(json.dumps(d["e2e"].get("to_convergence"))[:600])
PY
