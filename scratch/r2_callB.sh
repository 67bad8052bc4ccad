#!/bin/bash
# Round 2, session 2, GPU call B (EIGHT B200): parity on 8 slabs, strong and weak scaling lines, NCCL transport A/B.
mkdir -p gpurun_out
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29533 tests/mgpu_worker.py > $O/mgpu8.log 2>&1; echo "mgpu8 rc=$?" | tee -a $O/mgpu8.log
timeout 400 $TR --master-port 29602 bench.py --gpus 8 --steps 20 --warmup 3 --no-cpu --no-ttc > $O/b8_strong.json 2>> $O/exp8.err; echo "bench strong rc=$?"
timeout 300 $TR --master-port 29603 bench.py --gpus 8 --steps 20 --warmup 3 --weak --no-ttc --no-cpu --no-also > $O/b8_weak.json 2>> $O/exp8.err; echo "bench weak rc=$?"
CZB_NO_P2P=1 timeout 300 $TR --master-port 29604 bench.py --gpus 8 --steps 20 --warmup 3 --no-ttc --no-cpu --no-also > $O/b8_strong_nccl.json 2>> $O/exp8.err
( cd /tmp && timeout 120 $OLDPWD/cubez_b200/cz_f32 512 512 512 pbicgstab 400 1.1 sor2sma --gpus 8 > $OLDPWD/$O/cz8.log 2>&1 )
grep -h "FAIL\|rc=" $O/mgpu8.log | tail -5
python - <<'PY'
import json
for f in ("b8_strong","b8_weak","b8_strong_nccl"):
    try:
        d=[json.loads(l) for l in open("gpurun_out/%s.json"%f) if l.startswith("{")][-1]
        r=d.get("roofline",{}); print(f, "value %.1f ms/step %.4f loop %s fastest %s parity %s also %s" % (d["value"], d["ms_per_step"], r.get("ms_per_launch"), r.get("ms_per_launch_fastest_rank"), d.get("parity_vs_n1",{}).get("ok"), {k: round(v["value"],1) for k,v in d.get("also",{}).items()}), d.get("windows_ms"))
    except Exception as e: print(f, "FAILED", e)
PY
tail -12 $O/cz8.log; tail -3 $O/exp8.err
