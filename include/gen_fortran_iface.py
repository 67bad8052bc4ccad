#!/usr/bin/env python
"""Generates include/czb200_iface.F90 from include/czb200.h: a Fortran module of `bind(C)` interface blocks for
every entry point of libczb200 (north_star: "callable from Fortran via ISO_C_BINDING").

    python include/gen_fortran_iface.py            # rewrites include/czb200_iface.F90
    python include/gen_fortran_iface.py --check    # exit 1 if the file on disk is stale

PART 1 (the reference's kernel layer, src/cz_cpp/cz_Ffunc.h:16-578): the reference's Fortran routines take every
argument by reference and are linked by the compiler's lower-case + underscore mangling; the same exported symbols
are reached here through bind(C, name='jacobi_') with NON-value dummies, so a Fortran caller's source
(`call jacobi(p, sz, idx, g, cf, omg, b, res, wk2, flop)`) is unchanged.  Whether a dummy is an array or a scalar
cannot be read off the C prototype (all are pointers); it follows the reference's declarations
(src/cz_f90/cz_solver.f90, cz_blas.f90, cz_utility.f90, cz_maf.f90) through the tables below.
PART 2 (czb_*): C-style entry points, scalars by value.
The pattern is the reference's own bind(C) module, src/cz_f90/nvtx.f90:1-86.
"""
import os
import re
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
HDR = os.path.join(HERE, "czb200.h")
OUT = os.path.join(HERE, "czb200_iface.F90")

# PART 1 dummies that are SCALARS in the reference's Fortran (everything else is an array)
SCALAR = {"g", "omg", "res", "flop", "ip", "color", "pn", "ofst", "s", "beta", "d", "dh"}
SCALAR_IN = {"blas_triad_": {"a"}, "blas_bicg_2_": {"a", "b"}, "blas_dot1_": {"r"}, "blas_dot2_": {"r"}}
ARRAY_IN = {"fileout_t_": {"s"}}          # the field to write is called s there (cz_utility.f90:17)
# PART 2 pointer arguments that are scalars passed by reference
P2_SCALAR = {"res", "flop", "dot_w", "dot_self", "h2d", "d2h"}


def prototypes(text):
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    body = text[text.index('extern "C" {'):]
    for m in re.finditer(r"(?m)^([A-Za-z_][\w \*]*?)\b(\w+)\s*\(([^;{}]*?)\)\s*;", body):
        ret, name, args = m.group(1).strip(), m.group(2), " ".join(m.group(3).split())
        if name in ("struct",) or ret.startswith("typedef"):
            continue
        yield ret, name, [a.strip() for a in args.split(",")] if args not in ("void", "") else []


def split_arg(a):
    m = re.match(r"^(.*?)(\w+)\s*(\[\d*\])?$", a)
    ctype, name, arr = m.group(1).strip(), m.group(2), m.group(3)
    return ctype, name, arr


def ftype(ctype):
    c = ctype.replace("const", "").replace(" ", "")
    base = c.rstrip("*")
    return {"int": "integer(c_int)", "double": "real(c_double)", "CZB_REAL": "real(czb_real)", "float": "real(c_float)",
            "size_t": "integer(c_size_t)", "longlong": "integer(c_long_long)", "char": "character(kind=c_char)",
            "void": "type(c_ptr)", "czb_ctx": "type(c_ptr)"}[base], c.endswith("*")


def emit(ret, name, args):
    part1 = name.endswith("_") and not name.startswith("czb_")
    fname = name[:-1] if part1 else name
    names, decls = [], []
    for a in args:
        ctype, an, arr = split_arg(a)
        ft, is_ptr = ftype(ctype)
        an_f = an
        names.append(an_f)
        if part1:
            scalar = (an in SCALAR or an in SCALAR_IN.get(name, ())) and an not in ARRAY_IN.get(name, ())
            if ft.startswith("character"):
                decls.append("character(kind=c_char) :: %s(*)" % an_f)
            else:
                decls.append("%s :: %s%s" % (ft, an_f, "" if scalar else "(*)"))
        else:
            if ft.startswith("character"):
                decls.append("character(kind=c_char), intent(in) :: %s(*)" % an_f)
            elif ft == "type(c_ptr)":
                decls.append("type(c_ptr), value :: %s" % an_f)           # opaque handles and raw device pointers
            elif arr or is_ptr:
                decls.append("%s :: %s%s" % (ft, an_f, "" if an in P2_SCALAR else "(*)"))
            else:
                decls.append("%s, value :: %s" % (ft, an_f))
    # Fortran is case-insensitive: X and x of the MAF routines would clash -> rename the coordinate vectors
    seen, final = set(), []
    for n in names:
        f = n
        if f.lower() in seen:
            f = n + "c"
        seen.add(f.lower())
        final.append(f)
    decls = [re.sub(r":: %s(\(\*\))?$" % re.escape(n), ":: %s\\1" % f, d) for n, f, d in zip(names, final, decls)]
    rt = None if ret == "void" else ret
    lines = []
    kind = "subroutine" if rt is None else "function"
    bind = " bind(C, name='%s')" % name
    head = "%s %s(%s)" % (kind, fname, ", ".join(final))
    if len(head) + len(bind) + 4 <= 118:
        lines.append("    " + head + bind)
    else:       # wrap inside the dummy list only (never inside the bind clause)
        cur = "%s %s(" % (kind, fname)
        for i, f in enumerate(final):
            piece = f + (", " if i + 1 < len(final) else ")")
            if len(cur) + len(piece) > 110:
                lines.append("    " + cur.rstrip() + " &")
                cur = "      "
            cur += piece
        lines.append("    " + cur + " &")
        lines.append("     " + bind)
    lines.append("      import :: c_int, c_double, c_float, c_size_t, c_long_long, c_char, c_ptr, czb_real")
    for d in decls:
        lines.append("      " + d)
    if rt is not None:
        r = rt.replace("const", "").strip()
        lines.append("      %s :: %s" % ("type(c_ptr)" if r.endswith("*") else ftype(r)[0], fname))
        lines.append("    end function %s" % fname)
    else:
        lines.append("    end subroutine %s" % fname)
    return lines


def generate():
    text = open(HDR).read()
    out = ["! include/czb200_iface.F90 -- GENERATED by include/gen_fortran_iface.py from include/czb200.h; do not edit.",
           "!",
           "! ISO_C_BINDING interfaces of libczb200_f32.so / libczb200_f64.so (compile with -D_REAL_IS_DOUBLE_ for the",
           "! double-precision library, exactly like the reference's real_type option).  PART 1 are the reference's own",
           "! Fortran kernel names (src/cz_f90/*.f90 behind src/cz_cpp/cz_Ffunc.h:16-578), every argument by reference:",
           "!     use czb200",
           "!     call jacobi(p, sz, idx, g, cf, omg, b, res, wk2, flop)      ! runs on the GPU",
           "! Arrays may be host arrays (staged per call) or device memory obtained from czb_malloc and associated with",
           "! c_f_pointer (zero copy).  PART 2 are the czb_* extensions (device memory, solver context, fused solve loop).",
           "! Pattern: the reference's bind(C) module src/cz_f90/nvtx.f90.",
           "module czb200",
           "  use iso_c_binding",
           "  implicit none",
           "#ifdef _REAL_IS_DOUBLE_",
           "  integer, parameter :: czb_real = c_double",
           "#else",
           "  integer, parameter :: czb_real = c_float",
           "#endif",
           "  interface"]
    n1 = n2 = 0
    for ret, name, args in prototypes(text):
        out += emit(ret, name, args)
        if name.startswith("czb_"):
            n2 += 1
        else:
            n1 += 1
    out += ["  end interface", "end module czb200", ""]
    return "\n".join(out), n1, n2


def main():
    src, n1, n2 = generate()
    if "--check" in sys.argv:
        ok = os.path.exists(OUT) and open(OUT).read() == src
        print("czb200_iface.F90 %s (%d reference symbols, %d czb_* entry points)" % ("up to date" if ok else "STALE", n1, n2))
        sys.exit(0 if ok else 1)
    open(OUT, "w").write(src)
    print("wrote %s: %d reference symbols, %d czb_* entry points" % (OUT, n1, n2))


if __name__ == "__main__":
    main()
