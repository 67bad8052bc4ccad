#!/bin/bash
# Round 2, GPU call 7 (TWO B200): re-check after the chunk model / shell copy; cz --gpus; 8-GPU-sized slabs.
mkdir -p gpurun_out
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29533 tests/mgpu_worker.py > $O/mgpu7.log 2>&1; echo "mgpu rc=$?" | tee -a $O/mgpu7.log
timeout 300 python -m pytest tests/test_gpu_multi.py -m gpu -q -x > $O/gpu_multi7.log 2>&1; echo "gpu multi rc=$?" | tee -a $O/gpu_multi7.log
CZB_BENCH_NY=256 timeout 300 $TR --master-port 29601 bench.py --gpus 2 --steps 20 --warmup 3 --no-ttc --no-cpu > $O/b7_ny256.json 2>> $O/exp7.err
timeout 600 $TR --master-port 29602 bench.py --gpus 2 --steps 20 --warmup 3 > $O/b7_full.json 2>> $O/exp7.err; echo "bench full rc=$?"
timeout 300 $TR --master-port 29603 bench.py --gpus 2 --steps 20 --warmup 3 --weak --no-ttc --no-cpu --no-also > $O/b7_weak.json 2>> $O/exp7.err
grep -h "FAIL\|rc=" $O/mgpu7.log | tail -5; tail -3 $O/gpu_multi7.log
python - <<'PY'
import json
for f in ("b7_ny256","b7_full","b7_weak"):
    try:
        d=[json.loads(l) for l in open("gpurun_out/%s.json"%f) if l.startswith("{")][-1]
        r=d["roofline"]; print(f, "value %.1f ms/step %.4f loop %.4f parity %s also %s" % (d["value"], d["ms_per_step"], r["ms_per_launch"], d["parity_vs_n1"].get("ok"), {k: round(v["value"],1) for k,v in d.get("also",{}).items()}))
        if "time_to_convergence" in d: print({k:(v["iterations"], round(v["solver_loop_ms"],1)) for k,v in d["time_to_convergence"].items()})
    except Exception as e: print(f, "FAILED", e)
PY
tail -3 $O/exp7.err
