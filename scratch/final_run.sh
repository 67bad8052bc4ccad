#!/bin/bash
# round-end validation on one B200: full GPU test suite, smoke, default bench, launch list
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu5.log 2>&1; echo "pytest rc=$?"; tail -1 gpurun_out/pytest_gpu5.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
timeout 900 python bench.py > gpurun_out/bench_final.log 2>gpurun_out/bench_final.err; echo "bench rc=$?"; tail -1 gpurun_out/bench_final.log | cut -c1-200
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --no-also > gpurun_out/b2.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv python bench.py --steps 5 --warmup 3 --no-cpu --no-also > gpurun_out/ncu1.log 2>&1; echo "ncu rc=$?"
