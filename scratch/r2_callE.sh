#!/bin/bash
# Round 2, GPU call E (one B200): the allocation-free entry point alone in a fresh process, then the whole GPU suite
# (which collects it last), smoke, a short bench line.
mkdir -p gpurun_out
O=gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/noalloc_alone.log 2>&1; echo "noalloc alone rc=$?" | tee -a $O/noalloc_alone.log
timeout 900 python -m pytest tests -m gpu -q -x --durations=6 > $O/gpu_testsE.log 2>&1; echo "gpu tests rc=$?" | tee -a $O/gpu_testsE.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > $O/smokeE.log 2>&1; echo "smoke rc=$?" | tee -a $O/smokeE.log
timeout 300 python bench.py --no-cpu --no-ttc > $O/benchE.json 2> $O/benchE.err; echo "bench rc=$?"
tail -15 $O/noalloc_alone.log | cut -c1-250; tail -6 $O/gpu_testsE.log | cut -c1-250; cat $O/smokeE.log; cut -c1-300 $O/benchE.json
