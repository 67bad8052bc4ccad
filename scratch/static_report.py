#!/usr/bin/env python
"""Static figures of the shipped libraries (no GPU needed): per-kernel registers / stack (= spills) / shared memory
from `cuobjdump -res-usage`, and the SASS mnemonics that prove which hardware paths the code uses
(`cuobjdump -sass`: UTMALDG / UBLKCP = TMA, LDTM / STTM = tensor memory, FFMA2 / FADD2 / FMUL2 = packed FP32,
UTC*MMA = tensor-core MMA, which this path must NOT have).  Writes profiles/r02_static.md."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
    return dict(zip(names, out))


def short(name):
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"\(anonymous namespace\)::", "", name)
    depth, cut = 0, len(name)                          # drop the argument list: first '(' outside template brackets
    for i, ch in enumerate(name):
        if ch == "<":
            depth += 1
        elif ch == ">":
            depth -= 1
        elif ch == "(" and depth == 0:
            cut = i
            break
    name = name[:cut]
    name = name.replace("czb::", "")
    return name


def res_usage(so):
    txt = subprocess.run(["cuobjdump", "-res-usage", so], capture_output=True, text=True).stdout
    rows, src = [], None
    lines = txt.split("\n")
    for i, ln in enumerate(lines):
        m = re.match(r"identifier = (.*)", ln)
        if m:
            src = os.path.basename(m.group(1))
        m = re.match(r"\s*Function (\S+):", ln)
        if m:
            u = dict(kv.split(":") for kv in re.findall(r"[A-Z]+(?:\[\d+\])?:\d+", lines[i + 1]))
            cmem = sum(int(v) for k, v in u.items() if k.startswith("CONSTANT"))
            rows.append((src, m.group(1), int(u.get("REG", 0)), int(u.get("STACK", 0)), int(u.get("SHARED", 0)), cmem))
    return rows


def sass_counts(so):
    p = subprocess.Popen(["cuobjdump", "-sass", so], stdout=subprocess.PIPE, text=True)
    per, tot, fn = collections.defaultdict(collections.Counter), collections.Counter(), None
    pat = re.compile(r"\b(UTMALDG|UTMASTG|UBLKCP|UTMAPF|LDTM|STTM|UTCATOMSWS|UTC[A-Z]*MMA|FFMA2|FADD2|FMUL2|SHFL|LDG|STG|LDS|STS|"
                     r"BAR|SYNCS|DADD|DMUL|DFMA|F2F|MUFU|FFMA|FADD|FMUL|RED|ATOMG|ATOM)\b")
    for ln in p.stdout:
        m = re.match(r"\s*Function : (\S+)", ln)
        if m:
            fn = m.group(1)
            continue
        if "/*" not in ln:
            continue
        m = pat.search(ln)
        if m:
            per[fn][m.group(1)] += 1
            tot[m.group(1)] += 1
        per[fn]["_all"] += 1 if re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln) else 0
    p.wait()
    return per, tot


def main():
    out = ["# Static figures of the shipped libraries (round 2; made without a GPU by `scratch/static_report.py`)", "",
           "Build: `nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -fmad=false` (`cubez_b200/csrc/Makefile`).",
           "`STACK` is the per-thread local-memory frame (0 = nothing spilled); `SHARED` is static shared memory only (the",
           "marching and line kernels take their tiles as dynamic shared memory at launch).", ""]
    for real in ("f32", "f64"):
        so = os.path.join(ROOT, "cubez_b200", "libczb200_%s.so" % real)
        rows = res_usage(so)
        per, tot = sass_counts(so)
        dm = demangle([r[1] for r in rows])
        out += ["## libczb200_%s.so" % real, "",
                "SASS mnemonic totals: " + ", ".join("`%s` %d" % (k, tot[k]) for k in
                                                     ("UTMALDG", "UBLKCP", "LDTM", "STTM", "UTCATOMSWS", "FFMA2", "FADD2", "FMUL2",
                                                      "SHFL", "DADD") if tot[k]) +
                "; tensor-core MMA (`UTC*MMA`): %d." % sum(v for k, v in tot.items() if k.endswith("MMA")), "",
                "| source | kernel | regs | stack B | static smem B | SASS instr | TMA | TMEM ld/st | packed f32 | SHFL |",
                "|---|---|---|---|---|---|---|---|---|---|"]
        for src, fn, reg, stack, shared, cmem in sorted(rows, key=lambda r: (r[0], short(dm[r[1]]))):
            c = per.get(fn, {})
            out.append("| %s | `%s` | %d | %d | %d | %d | %d | %d | %d | %d |" % (
                src, short(dm[fn]).replace("|", "\\|"), reg, stack, shared, c.get("_all", 0),
                c.get("UTMALDG", 0) + c.get("UBLKCP", 0) + c.get("UTMASTG", 0), c.get("LDTM", 0) + c.get("STTM", 0),
                c.get("FFMA2", 0) + c.get("FADD2", 0) + c.get("FMUL2", 0), c.get("SHFL", 0)))
        spilled = [short(dm[r[1]]) for r in rows if r[3] > 0]
        out += ["", "%d kernels; with a stack frame: %s." % (len(rows), ", ".join("`%s`" % s for s in spilled) or "none"), ""]
    path = os.path.join(ROOT, "profiles", "r02_static.md")
    open(path, "w").write("\n".join(out))
    print("wrote", path)


if __name__ == "__main__":
    sys.exit(main())
