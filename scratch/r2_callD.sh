#!/bin/bash
# Round 2, session 2, GPU call D (TWO B200): slabs + bricks parity, cz driver with divisions, 2-GPU bench line.
mkdir -p gpurun_out
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 700 python -m pytest tests/test_gpu_multi.py -m gpu -q -x > $O/gpu_multiD.log 2>&1; echo "gpu multi rc=$?" | tee -a $O/gpu_multiD.log
timeout 400 $TR --master-port 29602 bench.py --gpus 2 --steps 20 --warmup 3 --no-cpu --no-ttc > $O/b2_strong.json 2>> $O/expD.err; echo "bench strong rc=$?"
timeout 300 $TR --master-port 29603 bench.py --gpus 2 --steps 20 --warmup 3 --weak --no-ttc --no-cpu --no-also > $O/b2_weak.json 2>> $O/expD.err; echo "bench weak rc=$?"
tail -30 $O/gpu_multiD.log | cut -c1-300
python - <<'PY'
import json
for f in ("b2_strong","b2_weak"):
    try:
        d=[json.loads(l) for l in open("gpurun_out/%s.json"%f) if l.startswith("{")][-1]
        r=d.get("roofline",{}); print(f, "value %.1f ms/step %.4f loop %s fastest %s parity %s also %s" % (d["value"], d["ms_per_step"], r.get("ms_per_launch"), r.get("ms_per_launch_fastest_rank"), d.get("parity_vs_n1",{}).get("ok"), {k: round(v["value"],1) for k,v in d.get("also",{}).items()}), d.get("windows_ms"))
    except Exception as e: print(f, "FAILED", e)
PY
tail -3 $O/expD.err
