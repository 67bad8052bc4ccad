#!/bin/bash
# Round 2, GPU call I (FOUR B200): the N = 4 strong-scaling line of sor2sma (headline only, to keep the 4x charge small).
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
timeout 150 $TR --master-port 29602 bench.py --gpus 4 --steps 20 --warmup 3 --no-cpu --no-ttc --no-also > gpurun_out/b4_strong.json 2> gpurun_out/expI.err; echo "bench rc=$?"
python - <<'PY'
import json
d=[json.loads(l) for l in open("gpurun_out/b4_strong.json") if l.startswith("{")][-1]
print("value %.1f ms/step %.4f loop %.4f parity %s clocks %s" % (d["value"], d["ms_per_step"], d["roofline"]["ms_per_launch"], d["parity_vs_n1"].get("ok"), d["clocks"]), d["windows_ms"])
PY
