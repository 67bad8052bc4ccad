#!/bin/bash
# Round 2, GPU call 5 (TWO B200): split launches (edge chunks on a high-priority stream + DMA halo push, interior on the
# solve stream) -- parity on 2 slabs, then timing against the NCCL exchange and the plain loop.
mkdir -p gpurun_out
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29533 tests/mgpu_worker.py > $O/mgpu5.log 2>&1; echo "mgpu p2p rc=$?" | tee -a $O/mgpu5.log
CZB_NO_P2P=1 timeout 600 $TR --master-port 29534 tests/mgpu_worker.py > $O/mgpu5_nccl.log 2>&1; echo "mgpu nccl rc=$?" | tee -a $O/mgpu5_nccl.log
timeout 300 python -m pytest tests/test_gpu_multi.py -m gpu -q -x > $O/gpu_multi5.log 2>&1; echo "gpu multi rc=$?" | tee -a $O/gpu_multi5.log
port=29600
run2() {
  label=$1; shift
  port=$((port+1))
  env "$@" timeout 300 $TR --master-port $port bench.py --gpus 2 --steps 20 --warmup 3 --no-ttc --no-cpu > $O/x.json 2>> $O/exp5.err
  python - "$label" <<'PY'
import json,sys
try:
    d=[json.loads(l) for l in open("gpurun_out/x.json") if l.startswith("{")][-1]
    r=d["roofline"]; print("%-36s value %7.1f ms/step %.4f loop max %.4f min %.4f parity %s also %s" % (sys.argv[1], d["value"], d["ms_per_step"], r["ms_per_launch"], r["ms_per_launch_fastest_rank"], d["parity_vs_n1"].get("ok"), {k: round(v["value"],1) for k,v in d.get("also",{}).items()}))
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
}
for ny in 256 1024; do
  echo "== NY=$ny (2 ranks)" >> $O/exp5.log
  run2 "split+dma ny=$ny"        CZB_BENCH_NY=$ny >> $O/exp5.log
  run2 "split+nccl ny=$ny"       CZB_BENCH_NY=$ny CZB_NO_P2P=1 >> $O/exp5.log
  run2 "plain ny=$ny"            CZB_BENCH_NY=$ny CZB_NO_OVERLAP=1 >> $O/exp5.log
done
for njc in 3 4 6 8; do run2 "split+dma ny=256 njc=$njc" CZB_BENCH_NY=256 CZB_NJC=$njc >> $O/exp5.log; done
grep -h "mgpu\|rc=" $O/mgpu5.log $O/mgpu5_nccl.log | grep -v " OK" | tail -20; tail -3 $O/gpu_multi5.log; cat $O/exp5.log | cut -c1-300; tail -5 $O/exp5.err
