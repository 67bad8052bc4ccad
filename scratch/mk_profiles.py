#!/usr/bin/env python
"""Turn gpurun_out/launches.csv and a --set full .ncu-rep into the tracked summaries under profiles/."""
import csv, re, subprocess, sys, os
from collections import defaultdict
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r01"
rep = sys.argv[2] if len(sys.argv) > 2 else "gpurun_out/prof_rb2pk.ncu-rep"
lc = sys.argv[3] if len(sys.argv) > 3 else "gpurun_out/launches.csv"
cmd = sys.argv[4] if len(sys.argv) > 4 else "python bench.py --steps 5 --warmup 3 --no-cpu --no-also"

rows = list(csv.reader(open(os.path.join(ROOT, lc))))
hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
h = rows[hi]; ik, iv = h.index("Kernel Name"), h.index("Metric Value")
tot, cnt = defaultdict(float), defaultdict(int)
for r in rows[hi + 1:]:
    if len(r) > iv:
        n = re.sub(r"\(.*", "", r[ik]).replace("czb::<unnamed>::", "")[:70]
        tot[n] += float(r[iv].replace(",", "")); cnt[n] += 1
T = sum(tot.values())
with open(os.path.join(ROOT, "profiles", tag + "_launches.csv"), "w") as f:
    f.writelines(open(os.path.join(ROOT, lc)).readlines())
with open(os.path.join(ROOT, "profiles", tag + "_launches_summary.md"), "w") as f:
    f.write("# %s - ncu launch list of `%s` (1x B200)\n\n" % (tag, cmd))
    f.write("`ncu --metrics gpu__time_duration.sum --clock-control none -c 200` (cold-cache, serialised: compare SHARES). "
            "Raw list: profiles/%s_launches.csv\n\n| kernel | launches | total us | share |\n|---|---|---|---|\n" % tag)
    for k, v in sorted(tot.items(), key=lambda x: -x[1]):
        f.write("| `%s` | %d | %.1f | %.1f%% |\n" % (k, cnt[k], v / 1000, 100 * v / T))

out = subprocess.run(["ncu", "-i", os.path.join(ROOT, rep), "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(out.splitlines()))
H, U, V = r[0], r[1], r[2]
m = {k: (V[i], U[i]) for i, k in enumerate(H)}
keys = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__cycles_active.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__m_xbar2l1tex_read_bytes.sum", "smsp__inst_executed.sum", "sm__inst_executed.avg.per_cycle_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread", "launch__block_size", "launch__grid_size",
        "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed_op_tma_ld.sum", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"]
keys += sorted(k for k in m if re.match(r"smsp__average_warps_issue_stalled_.*_per_issue_active.ratio", k))
with open(os.path.join(ROOT, "profiles", tag + "_rb2_metrics.csv"), "w") as f:
    f.write("metric,unit,value\n")
    for k in keys:
        if k in m:
            f.write("%s,%s,%s\n" % (k, m[k][1], m[k][0]))
print(open(os.path.join(ROOT, "profiles", tag + "_rb2_metrics.csv")).read())
