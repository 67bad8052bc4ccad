#!/bin/bash
# Round 2, GPU call 3 (TWO B200): where do the 13-19 % go that a slab loses in a multi-rank run?
mkdir -p gpurun_out
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
port=29600
run2() {  # label, env...
  label=$1; shift
  port=$((port+1))
  env "$@" timeout 300 $TR --master-port $port bench.py --gpus 2 --steps 20 --warmup 3 --no-ttc --no-also --no-cpu > $O/x.json 2>> $O/exp.err
  python - "$label" <<'PY'
import json,sys
try:
    d=[json.loads(l) for l in open("gpurun_out/x.json") if l.startswith("{")][-1]
    r=d["roofline"]; print("%-44s value %7.1f ms/step %.4f loop max %.4f min %.4f parity %s" % (sys.argv[1], d["value"], d["ms_per_step"], r["ms_per_launch"], r["ms_per_launch_fastest_rank"], d["parity_vs_n1"].get("ok")))
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
}
for ny in 256 1024; do
  echo "== NY=$ny (2 ranks)" >> $O/exp.log
  run2 "p2p ny=$ny"                       CZB_BENCH_NY=$ny >> $O/exp.log
  run2 "p2p noallreduce ny=$ny"           CZB_BENCH_NY=$ny CZB_EXP_NOALLREDUCE=1 >> $O/exp.log
  run2 "p2p nocomm ny=$ny"                CZB_BENCH_NY=$ny CZB_EXP_NOCOMM=1 >> $O/exp.log
  run2 "p2p nocomm nopeerstore ny=$ny"    CZB_BENCH_NY=$ny CZB_EXP_NOCOMM=1 CZB_EXP_NOPEERSTORE=1 >> $O/exp.log
  run2 "nccl ny=$ny"                      CZB_BENCH_NY=$ny CZB_NO_P2P=1 >> $O/exp.log
  run2 "nccl nocomm ny=$ny"               CZB_BENCH_NY=$ny CZB_NO_P2P=1 CZB_EXP_NOCOMM=1 >> $O/exp.log
  run2 "plain (no overlap) ny=$ny"        CZB_BENCH_NY=$ny CZB_NO_OVERLAP=1 >> $O/exp.log
done
echo "== single rank, slab instantiation forced" >> $O/exp.log
for ny in 128 512; do
  python scratch/slab_bench.py 1024 $ny 1024 sor2sma 40 >> $O/exp.log 2>&1
  CZB_EXP_FORCE_SIG=1 python scratch/slab_bench.py 1024 $ny 1024 sor2sma 40 >> $O/exp.log 2>&1
done
echo "== graphs + folded convergence test (single rank)" >> $O/exp.log
for c in "jacobi 64 2000" "sor2sma 64 2000" "sor2sma 128 1000" "sor2sma 256 400" "pcr_rb 256 200" "psor 64 100"; do
  python scratch/solver_bench.py $c >> $O/exp.log 2>&1
  CZB_NO_GRAPH=1 python scratch/solver_bench.py $c >> $O/exp.log 2>&1
done
echo "== f64 line SOR on the warp kernel" >> $O/exp.log
for c in "pcr_rb 512 20 - f64" "pcr_rb 256 40 - f64" "pcr_j_esa 512 20 - f64"; do python scratch/solver_bench.py $c >> $O/exp.log 2>&1; done
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_solve.py tests/test_gpu_oracle_fullsize.py -m gpu -q -x > $O/gpu_tests3.log 2>&1; echo "gpu tests rc=$?" | tee -a $O/gpu_tests3.log
timeout 600 $TR --master-port 29533 tests/mgpu_worker.py > $O/mgpu3.log 2>&1; echo "mgpu rc=$?" | tee -a $O/mgpu3.log
cat $O/exp.log | cut -c1-260; tail -3 $O/gpu_tests3.log; grep -h "FAIL\|rc=" $O/mgpu3.log | tail
