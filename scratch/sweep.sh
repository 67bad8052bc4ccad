#!/bin/bash
# usage: scratch/sweep.sh "VAR=val VAR2=val" ...   -> one bench line summary per configuration
for cfgs in "$@"; do
  env $cfgs python bench.py --steps 20 --warmup 3 --no-cpu > gpurun_out/sweep_tmp.log 2>&1
  python - "$cfgs" <<'PY'
import json,sys
l=open("gpurun_out/sweep_tmp.log").read().strip().splitlines()
try:
    d=json.loads(l[-1]); print(sys.argv[1], "GLUPS %.1f ms %.3f res %.15g" % (d["value"], d["roofline"]["ms_per_launch"], d["config"]["residual_last"]))
except Exception as e:
    print(sys.argv[1], "FAILED", l[-3:])
PY
done
