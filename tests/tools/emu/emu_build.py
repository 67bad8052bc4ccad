"""TEST INFRASTRUCTURE: builds tests/tools/emu/libemu_maf_{f32,f64}.so -- the device code of cubez_b200/csrc/maf.cu
compiled for the CPU (see emu_maf.cpp).  The only source transformation is the launch syntax:
    k<<<grid, block, smem, stream>>>(args);   ->   emu_launch(grid, block, [&]{ k(args); });
applied to maf.cu and to the headers it includes from cubez_b200/csrc (cells.cuh); everything else is compiled as is."""
import os
import re
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(HERE)))
CSRC = os.path.join(ROOT, "cubez_b200", "csrc")
LAUNCH = re.compile(r"(\b\w+(?:<[^<>;(){}]*>)?)\s*<<<\s*([^;]*?)>>>\s*\((.*?)\);", re.S)


def _rewrite(text):
    def sub(m):
        cfg = [c.strip() for c in m.group(2).split(",")]
        return "emu_launch(dim3(%s), dim3(%s), [&] { %s(%s); });" % (cfg[0], cfg[1], m.group(1), m.group(3))
    return LAUNCH.sub(sub, text)


def generate(outdir):
    """maf.cu with cells.cuh inlined (both rewritten) -> maf_emu_gen.inc; kernels.cuh / common.cuh are included as is."""
    maf = open(os.path.join(CSRC, "maf.cu")).read()
    cells = open(os.path.join(CSRC, "cells.cuh")).read().replace("#pragma once", "")
    cells = cells.replace('#include "kernels.cuh"', "")
    maf = maf.replace('#include "cells.cuh"', _rewrite(cells))
    gen = _rewrite(maf)
    assert "<<<" not in gen
    path = os.path.join(outdir, "maf_emu_gen.inc")
    with open(path, "w") as f:
        f.write(gen)
    return path


def build(outdir=None):
    outdir = outdir or os.path.join(HERE, "_build")
    os.makedirs(outdir, exist_ok=True)
    generate(outdir)
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    cuda_inc = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "include")
    libs = {}
    for real, flag in (("f32", []), ("f64", ["-D_REAL_IS_DOUBLE_"])):
        so = os.path.join(outdir, "libemu_maf_%s.so" % real)
        cmd = [cxx, "-O2", "-std=c++17", "-ffp-contract=off", "-fPIC", "-shared", "-w", "-I" + outdir, "-I" + CSRC,
               "-I" + cuda_inc] + flag + ["-o", so, os.path.join(HERE, "emu_maf.cpp")]
        subprocess.check_call(cmd)
        libs[real] = so
    return libs


if __name__ == "__main__":
    print(build())
