"""TEST INFRASTRUCTURE: builds tests/tools/emu/_build/libczb200_emu_{f32,f64}.so -- the library compiled for the CPU:
host code and every kernel that is plain CUDA C++ (kernels.cu, maf.cu, solver.cu's loop-control kernels) over an
emulation of the CUDA built-ins (emu_prelude.h: one thread per block, launches executed sequentially) and of the
runtime calls (emu_cudart.cpp: device memory = malloc).  With -ffp-contract=off the REAL arithmetic is the IEEE
arithmetic nvcc emits under -fmad=false, so fields must equal the oracle bit for bit.  What this cannot show is anything
that depends on real concurrency (barriers, races) or on fused.cu / pcr.cu (TMA, tensor memory, inline PTX: replaced by
emu_stubs.cpp) -- that stays with the -m gpu tests.  The only source transformations are the launch syntax:
    k<<<grid, block, smem, stream>>>(args);   ->   emu_launch(grid, block, [&]{ k(args); });
and `extern __shared__ T name[];` -> `static T name[1 << 16];`, applied to the .cu sources and to cells.cuh; everything
else is compiled as is."""
import os
import re
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(HERE)))
CSRC = os.path.join(ROOT, "cubez_b200", "csrc")
DYN_SHARED = re.compile(r"extern\s+__shared__\s+(\w+)\s+(\w+)\[\];")
LAUNCH = re.compile(r"(\b\w+(?:<[^<>;(){}]*>)?)\s*<<<\s*([^;]*?)>>>\s*\((.*?)\);", re.S)


def _rewrite(text):
    def sub(m):
        cfg = [c.strip() for c in m.group(2).split(",")]
        return "emu_launch(dim3(%s), dim3(%s), [&] { %s(%s); });" % (cfg[0], cfg[1], m.group(1), m.group(3))
    text = DYN_SHARED.sub(r"static \1 \2[1 << 16];", text)   # dynamic shared memory -> a static pool (512 KB)
    return LAUNCH.sub(sub, text)


LIB_SOURCES = ("kernels.cu", "maf.cu", "pcr.cu", "noalloc.cu", "runtime.cu", "abi.cu", "solver.cu", "comm.cpp")


def _pcr_plain(text):
    """pcr.cu without its warp-per-line kernels (warp shuffles, packed f32x2 PTX, tensor memory, TMA bulk copies: device
    only): keeps k_pcr_tables, the block-per-line kernel k_pcr_lines, the table management and the tail of
    launch_pcr_lines that launches k_pcr_lines -- the path the GPU takes for double precision and for lines the warp
    kernels do not cover.  Marker based; fails loudly if pcr.cu is reorganised."""
    a = text.index("// One WARP per line, the right-hand side d in registers.")
    a = text.rindex("\n// ----", 0, a) + 1
    b = text.index("}  // namespace\n\nvoid pcr_build_tables")
    text = text[:a] + text[b:]
    a = text.index("#ifndef _REAL_IS_DOUBLE_\n  static const bool no_warp")
    b = text.index("#endif\n  if (T.last4) {", a)
    return text[:a] + text[b + len("#endif\n"):]


def build_lib(outdir=None, sanitize=None):
    """libczb200_emu_{f32,f64}.so: the library's host code + every kernel that is plain CUDA C++ (kernels.cu, maf.cu,
    the loop-control kernels of solver.cu, the block-per-line part of pcr.cu) compiled for the CPU over emu_prelude.h /
    emu_cudart.cpp; fused.cu (TMA, inline PTX) is replaced by emu_stubs.cpp, the warp-per-line kernels of pcr.cu are cut.  sanitize="address": AddressSanitizer build
    (every "device" allocation is a malloc block, so an out-of-bounds access of a kernel is reported)."""
    outdir = outdir or os.path.join(HERE, "_build" + ("_" + sanitize if sanitize else ""))
    opt = ["-O1", "-g", "-fsanitize=" + sanitize, "-fno-omit-frame-pointer"] if sanitize else ["-O2"]
    gen = os.path.join(outdir, "lib")
    os.makedirs(gen, exist_ok=True)
    hdr = os.path.join(ROOT, "include", "czb200.h")
    done = {r: os.path.join(outdir, "libczb200_emu_%s.so" % r) for r in ("f32", "f64")}
    deps = [hdr] + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh", ".h", ".cpp"))] + \
           [os.path.join(HERE, f) for f in os.listdir(HERE) if f.endswith((".cpp", ".h", ".py"))]
    if all(os.path.exists(so) and os.path.getmtime(so) >= max(os.path.getmtime(d) for d in deps) for so in done.values()):
        return done
    with open(os.path.join(gen, "cells.cuh"), "w") as f:      # found first by the generated sources' #include "cells.cuh"
        f.write(_rewrite(open(os.path.join(CSRC, "cells.cuh")).read()))
    srcs = []
    for name in LIB_SOURCES:
        text = open(os.path.join(CSRC, name)).read()
        if name == "pcr.cu":
            text = _pcr_plain(text)
        text = _rewrite(text)
        assert "<<<" not in text and "asm(" not in text and "asm volatile" not in text, name
        text = text.replace('"../../include/czb200.h"', '"%s"' % hdr)
        out = os.path.join(gen, name.replace(".cu", "_emu.cpp").replace("comm.cpp", "comm_emu.cpp"))
        with open(out, "w") as f:
            f.write('#include "emu_prelude.h"\n' + text)
        srcs.append(out)
    srcs += [os.path.join(HERE, "emu_cudart.cpp"), os.path.join(HERE, "emu_stubs.cpp")]
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    cuda_inc = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "include")
    libs = {}
    for real, flag in (("f32", []), ("f64", ["-D_REAL_IS_DOUBLE_"])):
        so = os.path.join(outdir, "libczb200_emu_%s.so" % real)
        cmd = [cxx] + opt + ["-std=c++17", "-ffp-contract=off", "-fPIC", "-shared", "-w", "-I" + HERE, "-I" + CSRC,
               "-I" + cuda_inc] + flag + ["-Wl,-Bsymbolic", "-o", so] + srcs + ["-ldl"]
        subprocess.check_call(cmd)
        libs[real] = so
    return libs


def build_nccl(outdir=None):
    """libnccl_emu.so: file-based stand-in for the NCCL calls of comm.cpp (nccl_emu.cpp), for 2-process CPU runs."""
    outdir = outdir or os.path.join(HERE, "_build")
    os.makedirs(outdir, exist_ok=True)
    so = os.path.join(outdir, "libnccl_emu.so")
    src = os.path.join(HERE, "nccl_emu.cpp")
    if not (os.path.exists(so) and os.path.getmtime(so) >= os.path.getmtime(src)):
        cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
        cuda_inc = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "include")
        subprocess.check_call([cxx, "-O2", "-std=c++17", "-fPIC", "-shared", "-w", "-I" + cuda_inc, "-o", so, src])
    return so


def build_cz(outdir=None):
    """cz_emu_f32: the product's `cz` driver (cubez_b200/csrc/cz_main.cpp, unchanged) linked to the emulated library."""
    outdir = outdir or os.path.join(HERE, "_build")
    libs = build_lib(outdir)
    exe = os.path.join(outdir, "cz_emu_f32")
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    subprocess.check_call([cxx, "-O2", "-std=c++17", "-I" + os.path.join(ROOT, "include"), "-o", exe,
                           os.path.join(CSRC, "cz_main.cpp"), libs["f32"], "-Wl,-rpath," + outdir])
    return exe


def build_simt(outdir=None):
    """simt_line_{f32,f64}: maf.cu's line kernel with every CUDA thread of a block as a std::thread (simt_line.cpp),
    built with ThreadSanitizer as a stand-alone program."""
    outdir = outdir or os.path.join(HERE, "_build")
    gen = os.path.join(outdir, "simt")
    os.makedirs(gen, exist_ok=True)
    maf = open(os.path.join(CSRC, "maf.cu")).read()
    cells = open(os.path.join(CSRC, "cells.cuh")).read().replace("#pragma once", "").replace('#include "kernels.cuh"', "")
    text = _rewrite(maf.replace('#include "cells.cuh"', cells))
    assert "<<<" not in text
    with open(os.path.join(gen, "maf_simt_gen.inc"), "w") as f:
        f.write(text)
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    cuda_inc = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "include")
    libs = {}
    for real, flag in (("f32", []), ("f64", ["-D_REAL_IS_DOUBLE_"])):
        exe = os.path.join(outdir, "simt_line_%s" % real)
        subprocess.check_call([cxx, "-O1", "-g", "-std=c++20", "-ffp-contract=off", "-fsanitize=thread", "-w",
                               "-I" + gen, "-I" + CSRC, "-I" + cuda_inc] + flag +
                              ["-o", exe, os.path.join(HERE, "simt_line.cpp"), "-lpthread"])
        libs[real] = exe
    return libs


if __name__ == "__main__":
    print(build_lib())
