#!/bin/bash
# Round 2, GPU call F (TWO B200): slabs + bricks parity after the BiCGstab-on-bricks fix, the FULL 2-GPU bench line
# (time_to_convergence on the ranks), then on one GPU the N = 1 time-to-convergence child with e2e.to_convergence.
mkdir -p gpurun_out
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 500 python -m pytest tests/test_gpu_multi.py -m gpu -q -x > $O/gpu_multiF.log 2>&1; echo "gpu multi rc=$?" | tee -a $O/gpu_multiF.log
timeout 500 $TR --master-port 29602 bench.py --gpus 2 --steps 20 --warmup 3 > $O/b2_full.json 2>> $O/expF.err; echo "bench full rc=$?"
CUDA_VISIBLE_DEVICES=0 timeout 280 python bench.py --ttc-worker > $O/ttc_worker.json 2>> $O/expF.err; echo "ttc worker rc=$?"
grep -h "bricks\|FAIL\|passed\|failed" $O/gpu_multiF.log | cut -c1-200 | tail -16
python - <<'PY'
import json
try:
    d=[json.loads(l) for l in open("gpurun_out/b2_full.json") if l.startswith("{")][-1]
    print("value %.1f parity %s cpu %s" % (d["value"], d["parity_vs_n1"].get("ok"), d.get("cpu_baseline")))
    print({k:(v["iterations"], v.get("oracle_iterations"), round(v["solver_loop_ms"],1)) for k,v in d.get("time_to_convergence",{}).items()})
except Exception as e: print("b2_full FAILED", e)
try:
    d=[json.loads(l) for l in open("gpurun_out/ttc_worker.json") if l.startswith("{")][-1]
    print(json.dumps(d.get("e2e_to_convergence")))
except Exception as e: print("ttc worker FAILED", e)
PY
tail -5 $O/expF.err | cut -c1-300
