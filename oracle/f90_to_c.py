#!/usr/bin/env python3
"""oracle/f90_to_c.py -- TEST INFRASTRUCTURE ONLY (pins the parity oracle; never shipped, never timed).

Mechanical source-to-source translation of the reference's Fortran 90 kernels into C, so that the
reference's OWN statements -- read from the unmodified files where they lie under
/root/reference/src/cz_f90 -- can be executed in an image that has no Fortran front-end
(gfortran / flang / nvfortran are absent here and on the GPU box).

    python oracle/f90_to_c.py --real f32|f64 [--openmp] --out DIR  FILE.f90 [FILE.f90 ...]

writes one DIR/<name>.c per input file; every Fortran subroutine `foo` becomes `void foo_(...)` with
gfortran's calling convention (lower case, trailing underscore, every argument by reference), i.e. the
symbols `src/cz_cpp/cz_Ffunc.h` declares, so the reference's unmodified C++ driver links against the result.

What "mechanical" means here -- the translator knows Fortran, not CubeZ:
  * a C-preprocessor pass with no macro defined (`_OPENACC`, `__NEC__`, `_SVR`, `_aurora_` undefined: the
    plain-CPU build of `cmake/CompileOptionSelector.cmake`);
  * `!...` comments dropped -- that includes every `!$OMP` / `!$acc` / `!dir$` directive, exactly what a Fortran
    compiler does without `-fopenmp`: the statements run in the reference's 1-thread order;
  * free-form continuation lines joined, names lower-cased (Fortran is case-insensitive);
  * declarations -> C declarations; explicit-shape arrays keep their declared lower bounds and column-major
    order through per-array (lower bound, extent) descriptors evaluated at entry, as Fortran does;
  * expressions parsed with Fortran's grammar (`**` right-associative above `* /` above unary/binary `+ -`,
    so `-e*ap*a(kl)` is `-((e*ap)*a(kl))`; same-level operators left to right) and emitted FULLY
    parenthesised, so the C compiler has no freedom the Fortran one does not have;
  * Fortran's type rules: default-REAL literals (`18.0`) are REAL, `d`-exponent literals double,
    integer/REAL/double promotion per operation, integer division truncates, `2**n` is integer;
    intrinsics by argument type (`abs`, `max`, `min`, `mod`, `sqrt`, `sin`, `sinh`, `asin`, `dble`, `real`);
  * `--real f64` = gfortran's `-fdefault-real-8 -fdefault-double-8` (CompileOptionSelector.cmake:193).
Compile the output with `-O2 -ffp-contract=off` (baseline x86-64 gfortran emits no FMA either).
`-DF2C_BOUNDS` adds an index check on every array reference (records the Fortran file:line, then performs the access
exactly as the unchecked build does).
`--openmp` keeps the `!$OMP` sentinel lines and maps them to C pragmas with the same meaning (PARALLEL, DO, PARALLEL DO;
SCHEDULE(static), COLLAPSE, REDUCTION(+), PRIVATE, FIRSTPRIVATE, END DO NOWAIT): the associated loops in OpenMP's
canonical form, sequential loop variables privatised as Fortran does implicitly, privatised arrays as per-thread
buffers, reductions on dummy arguments through a local accumulator.  Compile that output with the reference's GNU flags
(-O3 -fopenmp): it is the CPU arm bench.py times and the kernel library of oracle/_ref/cz_f90c_omp_*, the reference
program.  A subroutine whose directives the mapping does not cover is translated serially and listed in
openmp_report.txt.
Anything outside the subset raises an error naming the file and line; nothing is skipped silently.
"""
import argparse
import os
import re
import sys

# ----------------------------------------------------------------------------------------------
# 1. source -> logical statements
# ----------------------------------------------------------------------------------------------

def preprocess(path, defines):
    """cpp-like pass: #ifdef/#ifndef/#if NAME==V/#else/#endif with `defines`; returns [(file, line, text)]."""
    out = []
    stack = []          # (parent_active, this_branch_taken)
    active = True
    with open(path, encoding="utf-8", errors="replace") as f:
        lines = f.read().split("\n")
    for n, raw in enumerate(lines, 1):
        s = raw.strip()
        if s.startswith("#"):
            d = s[1:].strip()
            m = re.match(r"(ifdef|ifndef|if|else|endif|elif)\b\s*(.*)", d)
            if not m:
                raise SyntaxError("%s:%d: unsupported directive %r" % (path, n, s))
            kw, rest = m.group(1), m.group(2).strip()
            if kw in ("ifdef", "ifndef", "if"):
                if kw == "ifdef":
                    c = rest in defines
                elif kw == "ifndef":
                    c = rest not in defines
                else:
                    mm = re.match(r"(\w+)\s*==\s*(\w+)$", rest)
                    if not mm:
                        raise SyntaxError("%s:%d: unsupported #if %r" % (path, n, rest))
                    c = str(defines.get(mm.group(1), 0)) == mm.group(2)
                stack.append((active, c))
                active = active and c
            elif kw == "else":
                parent, taken = stack.pop()
                stack.append((parent, True))
                active = parent and not taken
            elif kw == "endif":
                parent, _ = stack.pop()
                active = parent
            else:
                raise SyntaxError("%s:%d: #elif not supported" % (path, n))
            continue
        if active:
            out.append((path, n, raw))
    if stack:
        raise SyntaxError("%s: unterminated #if" % path)
    return out


def strip_comment(text):
    """Remove a trailing `! ...` comment (string literals respected)."""
    q = None
    for i, ch in enumerate(text):
        if q:
            if ch == q:
                q = None
        elif ch in "'\"":
            q = ch
        elif ch == "!":
            return text[:i]
    return text


def logical_statements(path, defines, openmp=False):
    """Join continuation lines, lower-case outside strings, split at `;`, expand `include`.  openmp: `!$omp` sentinel
    lines (with their own `&` continuation) are kept as statements "!$omp <directive>" instead of being comments."""
    stmts = []
    cur, cur_loc = "", None
    omp, omp_loc = "", None
    for (f, n, raw) in preprocess(path, defines):
        if openmp and raw.strip().lower().startswith("!$omp"):
            body = strip_comment(raw.strip()[5:]).strip()
            if omp and body.startswith("&"):
                body = body[1:].lstrip()
            if omp_loc is None:
                omp_loc = (f, n)
            if body.endswith("&"):
                omp += body[:-1] + " "
                continue
            stmts.append((omp_loc[0], omp_loc[1], "!$omp " + " ".join((omp + body).lower().split())))
            omp, omp_loc = "", None
            continue
        t = strip_comment(raw).strip()
        if not t:
            continue
        if cur and t.startswith("&"):
            t = t[1:].lstrip()
        if cur_loc is None:
            cur_loc = (f, n)
        if t.endswith("&"):
            cur += t[:-1] + " "
            continue
        cur += t
        for piece in cur.split(";"):
            piece = piece.strip()
            if piece:
                stmts.append((cur_loc[0], cur_loc[1], lower_outside_strings(piece)))
        cur, cur_loc = "", None
    if cur:
        raise SyntaxError("%s: dangling continuation" % path)
    # includes
    out = []
    for (f, n, s) in stmts:
        m = re.match(r"include\s+['\"](.+)['\"]$", s)
        if m:
            inc = os.path.join(os.path.dirname(f), m.group(1))
            out.extend(logical_statements(inc, defines, openmp))
        else:
            out.append((f, n, s))
    return out


def lower_outside_strings(s):
    out, q = [], None
    for ch in s:
        if q:
            out.append(ch)
            if ch == q:
                q = None
        else:
            if ch in "'\"":
                q = ch
                out.append(ch)
            else:
                out.append(ch.lower())
    return "".join(out)

# ----------------------------------------------------------------------------------------------
# 2. expressions
# ----------------------------------------------------------------------------------------------

TOK = re.compile(r"""\s*(?:
    (?P<num>(?:\d+\.\d*|\.\d+|\d+)(?:[ed][+-]?\d+)?(?:_\w+)?)
  | (?P<dot>\.(?:lt|le|gt|ge|eq|ne|and|or|not|true|false)\.)
  | (?P<name>[a-z_]\w*)
  | (?P<op>\*\*|==|/=|<=|>=|=>|[-+*/(),:=<>])
)""", re.X)

DOTMAP = {".lt.": "<", ".le.": "<=", ".gt.": ">", ".ge.": ">=", ".eq.": "==", ".ne.": "/="}
RANK = {"int": 0, "real": 1, "double": 2}


def tokenize(s, where):
    toks, pos = [], 0
    s = s.rstrip()
    while pos < len(s):
        m = TOK.match(s, pos)
        if not m or m.end() == pos:
            raise SyntaxError("%s: cannot tokenise %r at %r" % (where, s, s[pos:]))
        pos = m.end()
        if m.group("num") is not None:
            # "1.and." style ambiguities do not occur in the subset; "1.e0" handled by the regex
            toks.append(("num", m.group("num")))
        elif m.group("dot") is not None:
            d = m.group("dot")
            toks.append(("op", DOTMAP.get(d, d)))
        elif m.group("name") is not None:
            toks.append(("name", m.group("name")))
        else:
            toks.append(("op", m.group("op")))
    return toks


class Node:
    __slots__ = ("kind", "type", "a", "b", "val", "args")

    def __init__(self, kind, type_, a=None, b=None, val=None, args=None):
        self.kind, self.type, self.a, self.b, self.val, self.args = kind, type_, a, b, val, args


def promote(t1, t2, where):
    if t1 not in RANK or t2 not in RANK:
        raise TypeError("%s: arithmetic on %s/%s" % (where, t1, t2))
    return t1 if RANK[t1] >= RANK[t2] else t2


class Parser:
    """Fortran expression grammar (F2008 R7xx): level-5 .or. > .and. > .not. > relational > additive (with the
    leading sign) > multiplicative > power (right assoc.) > primary."""

    def __init__(self, toks, syms, where):
        self.t, self.i, self.syms, self.where = toks, 0, syms, where

    def peek(self):
        return self.t[self.i] if self.i < len(self.t) else (None, None)

    def next(self):
        tok = self.peek()
        self.i += 1
        return tok

    def accept(self, val):
        if self.peek() == ("op", val):
            self.i += 1
            return True
        return False

    def expect(self, val):
        if not self.accept(val):
            raise SyntaxError("%s: expected %r, found %r" % (self.where, val, self.peek()))

    def done(self):
        return self.i >= len(self.t)

    def expr(self):
        n = self.and_()
        while self.accept(".or."):
            n = Node("bin", "logical", n, self.and_(), "||")
        return n

    def and_(self):
        n = self.not_()
        while self.accept(".and."):
            n = Node("bin", "logical", n, self.not_(), "&&")
        return n

    def not_(self):
        if self.accept(".not."):
            return Node("un", "logical", self.not_(), val="!")
        return self.rel()

    def rel(self):
        n = self.add()
        k, v = self.peek()
        if k == "op" and v in ("<", "<=", ">", ">=", "==", "/="):
            self.next()
            r = self.add()
            promote(n.type, r.type, self.where)
            return Node("bin", "logical", n, r, "!=" if v == "/=" else v)
        return n

    def add(self):
        k, v = self.peek()
        if k == "op" and v in "+-":
            self.next()
            n = self.mul()
            if v == "-":
                n = Node("un", n.type, n, val="-")
        else:
            n = self.mul()
        while True:
            k, v = self.peek()
            if k == "op" and v in ("+", "-"):
                self.next()
                r = self.mul()
                n = Node("bin", promote(n.type, r.type, self.where), n, r, v)
            else:
                return n

    def mul(self):
        n = self.power()
        while True:
            k, v = self.peek()
            if k == "op" and v in ("*", "/"):
                self.next()
                r = self.power()
                n = Node("bin", promote(n.type, r.type, self.where), n, r, v)
            else:
                return n

    def power(self):
        n = self.primary()
        if self.accept("**"):
            r = self.power()            # right associative
            if n.type != "int" or r.type != "int":
                raise TypeError("%s: only integer ** integer is in the subset" % self.where)
            return Node("pow", "int", n, r)
        return n

    def arglist(self):
        args = []
        if self.accept(")"):
            return args
        while True:
            k, v = self.peek()
            # keyword argument (kind=8)
            if k == "name" and self.i + 1 < len(self.t) and self.t[self.i + 1] == ("op", "="):
                self.next(); self.next()
                args.append(("kw", v, self.expr()))
            else:
                args.append(self.expr())
            if self.accept(")"):
                return args
            self.expect(",")

    def primary(self):
        k, v = self.next()
        if k == "num":
            return number(v, self.where)
        if k == "op" and v == "(":
            n = self.expr()
            self.expect(")")
            return Node("paren", n.type, n)
        if k == "op" and v in (".true.", ".false."):
            return Node("lit", "logical", val="1" if v == ".true." else "0")
        if k == "name":
            sym = self.syms.get(v)
            if self.accept("("):
                args = self.arglist()
                if sym is not None and sym.dims is not None:
                    for a in args:
                        if isinstance(a, tuple) or a.type != "int":
                            raise TypeError("%s: non-integer subscript of %s" % (self.where, v))
                    if len(args) != len(sym.dims):
                        raise TypeError("%s: rank mismatch on %s" % (self.where, v))
                    return Node("aref", sym.type, val=v, args=args)
                if sym is not None:
                    raise TypeError("%s: %s is a scalar but is subscripted" % (self.where, v))
                return intrinsic(v, args, self.where)
            if sym is None:
                raise NameError("%s: undeclared name %r (implicit none)" % (self.where, v))
            if sym.param is not None:
                return Node("lit", sym.type, val=sym.param)
            return Node("var", sym.type, val=v)
        raise SyntaxError("%s: unexpected token %r" % (self.where, (k, v)))


def number(text, where):
    m = re.match(r"^(\d+\.\d*|\.\d+|\d+)(?:([ed])([+-]?\d+))?(?:_(\w+))?$", text)
    mant, ech, exp, kind = m.groups()
    if kind is not None:
        raise SyntaxError("%s: kind suffix on literal %r not in the subset" % (where, text))
    if "." not in mant and ech is None:
        return Node("lit", "int", val=mant)
    c = mant
    if c.endswith("."):
        c += "0"
    if c.startswith("."):
        c = "0" + c
    if exp is not None:
        c += "e" + exp
    return Node("reallit", "double" if ech == "d" else "real", val=c)


FUN1 = {"sqrt": "sqrt", "sin": "sin", "cos": "cos", "sinh": "sinh", "cosh": "cosh", "asin": "asin",
        "acos": "acos", "exp": "exp", "log": "log", "tanh": "tanh", "atan": "atan"}


def intrinsic(name, args, where):
    pos = [a for a in args if not isinstance(a, tuple)]
    kw = {a[1]: a[2] for a in args if isinstance(a, tuple)}
    if name == "dble" and len(pos) == 1 and not kw:
        return Node("cast", "double", pos[0])
    if name == "real" and len(pos) in (1, 2):
        kind = kw.get("kind", pos[1] if len(pos) == 2 else None)
        if kind is None:
            return Node("cast", "real", pos[0])
        if kind.kind == "lit" and kind.val == "8":
            return Node("cast", "double", pos[0])
        if kind.kind == "lit" and kind.val == "4":
            return Node("cast", "real4", pos[0])
        raise TypeError("%s: real(kind=%r)" % (where, kind.val))
    if name == "int" and len(pos) == 1 and not kw:
        return Node("cast", "int", pos[0])
    if name == "float" and len(pos) == 1 and not kw:
        return Node("cast", "real", pos[0])
    if name == "mod" and len(pos) == 2 and not kw:
        t = promote(pos[0].type, pos[1].type, where)
        return Node("call", t, val="mod", args=pos)
    if name == "abs" and len(pos) == 1 and not kw:
        return Node("call", pos[0].type, val="abs", args=pos)
    if name in ("max", "min") and len(pos) >= 2 and not kw:
        t = pos[0].type
        for a in pos[1:]:
            if a.type != t:
                raise TypeError("%s: %s() with mixed argument types" % (where, name))
        return Node("call", t, val=name, args=pos)
    if name in FUN1 and len(pos) == 1 and not kw:
        if pos[0].type == "int":
            raise TypeError("%s: %s(integer)" % (where, name))
        return Node("call", pos[0].type, val=name, args=pos)
    raise NameError("%s: unknown function / undeclared array %r" % (where, name))

# ----------------------------------------------------------------------------------------------
# 3. C emission
# ----------------------------------------------------------------------------------------------


class Sym:
    def __init__(self, name, type_, dims=None, is_arg=False, allocatable=False, param=None, char=False):
        self.name, self.type, self.dims = name, type_, dims
        self.is_arg, self.allocatable, self.param, self.char = is_arg, allocatable, param, char


class Emitter:
    def __init__(self, real):
        self.real = real            # "f32" | "f64"

    def ctype(self, t):
        if t == "int":
            return "int"
        if t == "double":
            return "double"
        if t == "real":
            return "float" if self.real == "f32" else "double"
        if t == "real4":
            return "float"
        if t == "logical":
            return "int"
        raise TypeError(t)

    def isfloat32(self, t):
        return (t == "real" and self.real == "f32") or t == "real4"

    def gen(self, n, syms, loc):
        k = n.kind
        if k == "lit":
            return n.val
        if k == "reallit":
            return n.val + ("f" if self.isfloat32(n.type) else "")
        if k == "paren":
            return "(" + self.gen(n.a, syms, loc) + ")"
        if k == "var":
            s = syms[n.val]
            if s.dims is not None:
                raise TypeError("%s: whole-array reference to %s" % (loc, n.val))
            return "(*v_%s)" % n.val if s.is_arg else "v_%s" % n.val
        if k == "aref":
            return "v_%s[%s]" % (n.val, self.index(n, syms, loc))
        if k == "un":
            return "(%s%s)" % (n.val, self.gen(n.a, syms, loc))
        if k == "bin":
            return "(%s %s %s)" % (self.gen(n.a, syms, loc), n.val, self.gen(n.b, syms, loc))
        if k == "pow":
            return "f2c_ipow(%s, %s)" % (self.gen(n.a, syms, loc), self.gen(n.b, syms, loc))
        if k == "cast":
            return "((%s)%s)" % (self.ctype(n.type), self.gen(n.a, syms, loc))
        if k == "call":
            a = [self.gen(x, syms, loc) for x in n.args]
            f = n.val
            t = n.type
            if f == "mod":
                if t == "int":
                    return "(%s %% %s)" % (a[0], a[1])
                return "%s(%s, %s)" % ("fmodf" if self.isfloat32(t) else "fmod", a[0], a[1])
            if f == "abs":
                if t == "int":
                    return "abs(%s)" % a[0]
                return "%s(%s)" % ("fabsf" if self.isfloat32(t) else "fabs", a[0])
            if f in ("max", "min"):
                sfx = "i" if t == "int" else ("f" if self.isfloat32(t) else "d")
                e = a[0]
                for x in a[1:]:
                    e = "f2c_%s_%s(%s, %s)" % (f, sfx, e, x)
                return e
            return "%s%s(%s)" % (FUN1[f], "f" if self.isfloat32(t) else "", a[0])
        raise TypeError("%s: node %s" % (loc, k))

    def index(self, n, syms, loc):
        """column-major offset with the declared lower bounds: sum_d (i_d - lb_d) * prod_{e<d} n_e"""
        name = n.val
        r = len(n.args)
        e = None
        for d in range(r, 0, -1):
            sub = self.gen(n.args[d - 1], syms, loc)
            term = "F2C_SUB(%s, %s_lb%d, %s_n%d, \"%s\", %d, \"%s\")" % (sub, name, d, name, d, name, d, loc)
            e = term if e is None else "(%s + %s_n%d * %s)" % (term, name, d, e)
        return e


PRELUDE = r"""/* GENERATED by oracle/f90_to_c.py from %(src)s (--real %(real)s) -- do not edit, do not commit.
 * TEST INFRASTRUCTURE: the reference's own Fortran statements, mechanically translated, for pinning the oracle. */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef F2C_BOUNDS
long f2c_sub(long i, long lb, long n, const char* a, int d, const char* loc);   /* f2c_runtime.c */
#define F2C_SUB(i, lb, n, a, d, loc) f2c_sub((long)(i), lb, n, a, d, loc)
#else
#define F2C_SUB(i, lb, n, a, d, loc) ((long)(i) - (lb))
#endif
static inline int f2c_ipow(int b, int e) { int r = 1; if (e < 0) return (b == 1) ? 1 : ((b == -1) ? ((e & 1) ? -1 : 1) : 0); while (e-- > 0) r *= b; return r; }
static inline int f2c_max_i(int a, int b) { return a > b ? a : b; }
static inline int f2c_min_i(int a, int b) { return a < b ? a : b; }
static inline float f2c_max_f(float a, float b) { return a > b ? a : b; }
static inline float f2c_min_f(float a, float b) { return a < b ? a : b; }
static inline double f2c_max_d(double a, double b) { return a > b ? a : b; }
static inline double f2c_min_d(double a, double b) { return a < b ? a : b; }
/* automatic / allocated / per-thread arrays: zero-filled, with slack behind the declared extent -- the reference's padded
 * line routines index up to 3*2^(pn-2) elements past it on line lengths other than 2^m - 2 (F2C_BOUNDS reports them) */
static void* f2c_alloc(long n, size_t sz) { void* p = calloc((size_t)(n > 0 ? n : 1) + 4096, sz); if (!p) { fprintf(stderr, "f2c: out of memory\n"); abort(); } return p; }
"""

DECL_RE = re.compile(r"^(integer|real|double\s+precision|character\s*\*\s*\d+|logical)\b(.*)$")


def split_top(s, sep=","):
    out, depth, cur = [], 0, ""
    for ch in s:
        if ch == "(":
            depth += 1
        elif ch == ")":
            depth -= 1
        if ch == sep and depth == 0:
            out.append(cur.strip())
            cur = ""
        else:
            cur += ch
    if cur.strip():
        out.append(cur.strip())
    return out


def matching_paren(s, i):
    depth = 0
    for j in range(i, len(s)):
        if s[j] == "(":
            depth += 1
        elif s[j] == ")":
            depth -= 1
            if depth == 0:
                return j
    raise SyntaxError("unbalanced parentheses in %r" % s)


class Unit:
    """one subroutine"""

    def __init__(self, name, args, loc):
        self.name, self.args, self.loc = name, args, loc
        self.syms = {}
        self.order = []
        self.body = []          # (file, line, text)


def parse_units(stmts):
    units, cur = [], None
    for (f, n, s) in stmts:
        loc = "%s:%d" % (os.path.basename(f), n)
        m = re.match(r"^subroutine\s+(\w+)\s*\((.*)\)$", s)
        if m:
            if cur is not None:
                raise SyntaxError("%s: nested subroutine" % loc)
            cur = Unit(m.group(1), [a.strip() for a in m.group(2).split(",") if a.strip()], loc)
            continue
        if re.match(r"^end(\s+subroutine(\s+\w+)?)?$", s):
            if cur is None:
                raise SyntaxError("%s: end without subroutine" % loc)
            units.append(cur)
            cur = None
            continue
        if cur is None:
            raise SyntaxError("%s: statement outside a subroutine: %r" % (loc, s))
        cur.body.append((f, n, s))
    if cur is not None:
        raise SyntaxError("unterminated subroutine %s" % cur.name)
    return units


def parse_decl(u, s, loc):
    """returns True if `s` is a specification statement"""
    if s == "implicit none":
        return True
    m = re.match(r"^parameter\s*\((.*)\)$", s)
    if m:
        for item in split_top(m.group(1)):
            nm, val = [x.strip() for x in item.split("=", 1)]
            if nm not in u.syms:
                raise NameError("%s: parameter %s has no type" % (loc, nm))
            if not re.match(r"^-?\d+$", val):
                raise SyntaxError("%s: only integer literal parameters are in the subset" % loc)
            u.syms[nm].param = val
        return True
    m = DECL_RE.match(s)
    if not m:
        return False
    tword, rest = m.group(1), m.group(2).strip()
    # `real(...)  = ` would be an assignment to an array called real: not in the subset
    char = tword.startswith("character")
    type_ = {"integer": "int", "real": "real", "logical": "logical"}.get(tword)
    if tword.startswith("double"):
        type_ = "double"
    if char:
        type_ = "char"
    attrs_dims, allocatable = None, False
    if "::" in rest:
        attrs, ents = rest.split("::", 1)
        for a in split_top(attrs):
            if not a:
                continue
            mm = re.match(r"^dimension\s*\((.*)\)$", a)
            if mm:
                attrs_dims = split_top(mm.group(1))
            elif a == "allocatable":
                allocatable = True
            elif re.match(r"^intent\s*\(", a):
                pass
            else:
                raise SyntaxError("%s: attribute %r not in the subset" % (loc, a))
    else:
        if rest.startswith(","):
            raise SyntaxError("%s: attributes without `::`" % loc)
        ents = rest
    for e in split_top(ents):
        mm = re.match(r"^(\w+)\s*(?:\((.*)\))?$", e)
        if not mm:
            raise SyntaxError("%s: entity %r" % (loc, e))
        nm = mm.group(1)
        dims = split_top(mm.group(2)) if mm.group(2) is not None else attrs_dims
        if nm in u.syms:
            raise NameError("%s: %s declared twice" % (loc, nm))
        u.syms[nm] = Sym(nm, type_, dims=dims, is_arg=nm in u.args, allocatable=allocatable, char=char)
        u.order.append(nm)
    return True


class OmpUnsupported(Exception):
    pass


def omp_clauses(text, u, em, where, extra_private=()):
    """Fortran OpenMP clause list -> dict(loop=[C clauses of the worksharing loop], data=[C data-sharing clauses],
    pre=[C lines before the construct], open=[C lines at the top of the construct's block, run by every thread],
    close=[... at its bottom], post=[... after it]).
    Local scalars map to the C clause of the same name.  What C cannot express directly is restated with the same
    meaning: a privatised ARRAY (dummy or allocatable: a pointer in C) becomes a per-thread buffer declared inside the
    block under the same name (copied from the original for firstprivate); a reduction on a dummy-argument scalar
    (a pointer in C) goes through a local accumulator that is added to the argument after the construct."""
    cl = {"loop": [], "data": [], "pre": [], "open": [], "close": [], "post": []}
    listed = set()
    rest = text.strip()
    while rest:
        m = re.match(r"^(schedule|collapse|reduction|private|firstprivate)\s*\(([^)]*)\)\s*", rest)
        if not m:
            raise OmpUnsupported("%s: clause %r" % (where, rest))
        kw, arg = m.group(1), m.group(2).strip()
        rest = rest[m.end():]
        if kw == "schedule":
            if arg != "static":
                raise OmpUnsupported("%s: schedule(%s)" % (where, arg))
            cl["loop"].append("schedule(static)")
            continue
        if kw == "collapse":
            cl["loop"].append("collapse(%d)" % int(arg))
            continue
        if kw == "reduction":
            op, arg = [x.strip() for x in arg.split(":", 1)]
            if op != "+":
                raise OmpUnsupported("%s: reduction(%s)" % (where, op))
        scal = []
        for nm in [x.strip() for x in arg.split(",") if x.strip()]:
            sy = u.syms.get(nm)
            if sy is None or sy.param is not None or sy.char:
                raise OmpUnsupported("%s: %s(%s)" % (where, kw, nm))
            listed.add(nm)
            ct = em.ctype(sy.type)
            if sy.dims is None and not sy.is_arg:
                scal.append("v_" + nm)
            elif sy.dims is None and kw == "reduction":
                cl["pre"].append("%s f2c_red_%s = 0;" % (ct, nm))
                cl["data"].append("reduction(+:f2c_red_%s)" % nm)
                cl["open"].append("%s* v_%s = &f2c_red_%s;   /* reduction(+:%s) on a dummy argument */" % (ct, nm, nm, nm))
                cl["post"].append("(*v_%s) += f2c_red_%s;" % (nm, nm))
            elif sy.dims is not None and kw in ("private", "firstprivate"):
                total = " * ".join("%s_n%d" % (nm, d) for d in range(1, len(sy.dims) + 1))
                if kw == "firstprivate":
                    cl["pre"].append("%s* const f2c_orig_%s = v_%s;" % (ct, nm, nm))
                cl["open"].append("%s* v_%s = (%s*)f2c_alloc(%s, sizeof(%s));   /* %s(%s): per-thread copy */"
                                  % (ct, nm, ct, total, ct, kw, nm))
                if kw == "firstprivate":
                    cl["open"].append("memcpy(v_%s, f2c_orig_%s, (size_t)(%s) * sizeof(%s));" % (nm, nm, total, ct))
                cl["close"].append("free(v_%s);" % nm)
            else:
                raise OmpUnsupported("%s: %s(%s)" % (where, kw, nm))
        if scal:
            cl["data"].append("%s(%s%s)" % (kw, "+:" if kw == "reduction" else "", ", ".join(scal)))
    more = [nm for nm in extra_private if nm not in listed]
    if more:
        # Fortran: the iteration variable of a sequential loop inside a parallel construct is private (OpenMP 2.9.1.1);
        # in C it is shared unless listed
        cl["data"].append("private(%s)" % ", ".join("v_" + nm for nm in more))
    return cl


def translate_unit(u, em, openmp=False):
    C = []
    # ---- specification part
    k = 0
    while k < len(u.body):
        f, n, s = u.body[k]
        if not parse_decl(u, s, "%s:%d" % (os.path.basename(f), n)):
            break
        k += 1
    exe = u.body[k:]
    for a in u.args:
        if a not in u.syms:
            raise NameError("%s: dummy argument %s of %s has no declaration" % (u.loc, a, u.name))
    # ---- signature
    sig = []
    for a in u.args:
        s = u.syms[a]
        sig.append("char* v_%s" % a if s.char else "%s* v_%s" % (em.ctype(s.type), a))
    C.append("/* %s: subroutine %s */" % (u.loc, u.name))
    C.append("void %s_(%s) {" % (u.name, ", ".join(sig)))
    # ---- locals
    for nm in u.order:
        s = u.syms[nm]
        if s.is_arg or s.param is not None:
            continue
        if s.char:
            raise TypeError("%s: local character variable" % u.loc)
        if s.dims is None:
            C.append("  %s v_%s = 0;" % (em.ctype(s.type), nm))
        else:
            C.append("  %s* v_%s = NULL;" % (em.ctype(s.type), nm))
    # ---- array descriptors: constant-bound arrays first (sz, idx), then those that depend on them
    def bounds(dim, where):
        if dim == ":":
            return None, None
        parts = split_top(dim, ":")
        if len(parts) == 1:
            lo, hi = "1", parts[0]
        elif len(parts) == 2:
            lo, hi = parts
        else:
            raise SyntaxError("%s: dimension %r" % (where, dim))
        out = []
        for x in (lo, hi):
            p = Parser(tokenize(x, where), u.syms, where)
            node = p.expr()
            if not p.done() or node.type != "int":
                raise SyntaxError("%s: bound %r" % (where, x))
            out.append(em.gen(node, u.syms, where))
        return out[0], out[1]

    arrays = [nm for nm in u.order if u.syms[nm].dims is not None]
    const = [nm for nm in arrays if all(re.match(r"^[\d:\s+-]*$", d) for d in u.syms[nm].dims)]
    for nm in const + [a for a in arrays if a not in const]:
        s = u.syms[nm]
        total = []
        for d, dim in enumerate(s.dims, 1):
            lo, hi = bounds(dim, u.loc)
            if lo is None:
                if not s.allocatable:
                    raise SyntaxError("%s: assumed shape %s" % (u.loc, nm))
                C.append("  long %s_lb%d = 1, %s_n%d = 0;" % (nm, d, nm, d))
            else:
                C.append("  long %s_lb%d = %s, %s_n%d = (long)(%s) - %s_lb%d + 1; if (%s_n%d < 0) %s_n%d = 0;"
                         % (nm, d, lo, nm, d, hi, nm, d, nm, d, nm, d))
                total.append("%s_n%d" % (nm, d))
            C.append("  (void)%s_lb%d; (void)%s_n%d;" % (nm, d, nm, d))
        if not s.is_arg and not s.allocatable:
            C.append("  v_%s = (%s*)f2c_alloc(%s, sizeof(%s));   /* automatic array */"
                     % (nm, em.ctype(s.type), " * ".join(total), em.ctype(s.type)))
    # ---- executable part
    ind = 1
    blocks = []        # "do" | "if" | "omp" | "do_omp_outer" | "do_omp_inner"
    omp_skip = [0]
    tmp = [0]

    def expr(text, where, want=None):
        p = Parser(tokenize(text, where), u.syms, where)
        node = p.expr()
        if not p.done():
            raise SyntaxError("%s: trailing tokens in %r" % (where, text))
        if want == "logical" and node.type != "logical":
            raise TypeError("%s: condition %r is not logical" % (where, text))
        if want == "int" and node.type != "int":
            raise TypeError("%s: %r is not integer" % (where, text))
        return em.gen(node, u.syms, where)

    def simple(s, where):
        """one action statement -> C text (no trailing newline)"""
        if s == "cycle":
            if not any(b.startswith("do") for b in blocks):
                raise SyntaxError("%s: cycle outside do" % where)
            return "continue;"
        if s == "exit":
            if not any(b.startswith("do") for b in blocks):
                raise SyntaxError("%s: exit outside do" % where)
            if blocks and blocks[-1].startswith("do_omp"):
                raise OmpUnsupported("%s: exit from a worksharing loop" % where)
            return "break;"
        if s == "return":
            return "goto f2c_return;"
        if s == "continue":
            return ";"
        m = re.match(r"^allocate\s*\((.*)\)$", s)
        if m:
            out = []
            for item in split_top(m.group(1)):
                mm = re.match(r"^(\w+)\s*\((.*)\)$", item)
                nm = mm.group(1)
                sy = u.syms.get(nm)
                if sy is None or not sy.allocatable:
                    raise TypeError("%s: allocate of %s" % (where, nm))
                dims = split_top(mm.group(2))
                if len(dims) != len(sy.dims):
                    raise TypeError("%s: allocate rank of %s" % (where, nm))
                tot = []
                for d, dim in enumerate(dims, 1):
                    lo, hi = bounds(dim, where)
                    out.append("%s_lb%d = %s; %s_n%d = (long)(%s) - %s_lb%d + 1; if (%s_n%d < 0) %s_n%d = 0;"
                               % (nm, d, lo, nm, d, hi, nm, d, nm, d, nm, d))
                    tot.append("%s_n%d" % (nm, d))
                out.append("v_%s = (%s*)f2c_alloc(%s, sizeof(%s));" % (nm, em.ctype(sy.type), " * ".join(tot), em.ctype(sy.type)))
            return " ".join(out)
        m = re.match(r"^deallocate\s*\((.*)\)$", s)
        if m:
            out = []
            for nm in split_top(m.group(1)):
                sy = u.syms.get(nm)
                if sy is None or not sy.allocatable:
                    raise TypeError("%s: deallocate of %s" % (where, nm))
                out.append("free(v_%s); v_%s = NULL;" % (nm, nm))
            return " ".join(out)
        # assignment: lhs = rhs
        toks = tokenize(s, where)
        p = Parser(toks, u.syms, where)
        lhs = p.primary()
        if lhs.kind not in ("var", "aref"):
            raise SyntaxError("%s: not an assignment: %r" % (where, s))
        p.expect("=")
        rhs = p.expr()
        if not p.done():
            raise SyntaxError("%s: trailing tokens in %r" % (where, s))
        if lhs.type == "logical" or rhs.type == "logical":
            raise TypeError("%s: logical assignment" % where)
        # Fortran converts the value to the type of the variable on assignment; so does C
        return "%s = %s;" % (em.gen(lhs, u.syms, where), em.gen(rhs, u.syms, where))

    DO_RE = re.compile(r"^do\s+(\w+)\s*=\s*(.*)$")

    def region_do_vars(start, end_marker):
        """iteration variables of every do statement between exe[start] and the matching `!$omp end ...`"""
        names = []
        for (_, _, t) in exe[start:]:
            if t.startswith("!$omp " + end_marker):
                return names
            mm = DO_RE.match(t)
            if mm and mm.group(1) not in names:
                names.append(mm.group(1))
        raise OmpUnsupported("%s: no `!$omp %s`" % (u.loc, end_marker))

    omp_for = None      # pending (pragma text, where) for the next do statement(s)
    omp_open = []       # constructs whose block is open: dict(kind, close, post)

    def open_block(cl, where, pragma):
        """pre lines; [pragma]; `{`; per-thread open lines"""
        nonlocal ind
        pad0 = "  " * ind
        C.append("%s/* %s */ {" % (pad0, where))
        for ln in cl["pre"]:
            C.append("%s  %s" % (pad0, ln))
        if pragma:
            C.append(pragma)
        C.append("%s  {" % pad0)
        for ln in cl["open"]:
            C.append("%s    %s" % (pad0, ln))
        ind += 2

    def close_block(entry, where):
        nonlocal ind
        ind -= 2
        pad0 = "  " * ind
        for ln in entry["close"]:
            C.append("%s    %s" % (pad0, ln))
        C.append("%s  }   /* %s */" % (pad0, where))
        for ln in entry["post"]:
            C.append("%s  %s" % (pad0, ln))
        C.append("%s}" % pad0)

    for pos, (f, n, s) in enumerate(exe):
        where = "%s:%d" % (os.path.basename(f), n)
        pad = "  " * ind
        if s.startswith("!$omp "):
            d = s[6:].strip()
            if d.startswith("end "):
                kind = "parallel do" if d.startswith("end parallel do") else ("parallel" if d.startswith("end parallel") else "do")
                if not d.startswith("end " + kind) or not omp_open or omp_open[-1]["kind"] != kind or blocks[-1:] != ["omp"]:
                    raise OmpUnsupported("%s: unmatched `%s`" % (where, d))
                blocks.pop()
                close_block(omp_open.pop(), where + ": " + d)
                continue
            if omp_for is not None:
                raise OmpUnsupported("%s: directive between `do` directive and loop" % where)
            if d.startswith("parallel do"):
                cl = omp_clauses(d[len("parallel do"):], u, em, where, region_do_vars(pos + 1, "end parallel do"))
                open_block(cl, where, "#pragma omp parallel " + " ".join(cl["data"]))
                omp_for = ("#pragma omp for " + " ".join(cl["loop"]), where)
                kind = "parallel do"
            elif d.startswith("parallel"):
                cl = omp_clauses(d[len("parallel"):], u, em, where, region_do_vars(pos + 1, "end parallel"))
                if cl["loop"]:
                    raise OmpUnsupported("%s: loop clause on `parallel`" % where)
                open_block(cl, where, "#pragma omp parallel " + " ".join(cl["data"]))
                kind = "parallel"
            elif d.startswith("do"):
                nowait = ""
                for (_, _, t) in exe[pos + 1:]:
                    if t.startswith("!$omp end do"):
                        nowait = " nowait" if "nowait" in t else ""
                        break
                cl = omp_clauses(d[2:], u, em, where)
                if cl["post"] or cl["pre"] and not cl["open"]:
                    raise OmpUnsupported("%s: reduction on a dummy argument at a `do` directive" % where)
                open_block(cl, where, None)       # a plain block every thread of the team runs: per-thread copies
                omp_for = ("#pragma omp for " + " ".join(cl["loop"] + cl["data"]) + nowait, where)
                kind = "do"
            else:
                raise OmpUnsupported("%s: directive %r" % (where, d))
            omp_open.append({"kind": kind, "close": cl["close"], "post": cl["post"]})
            blocks.append("omp")
            continue
        if DECL_RE.match(s) or s.startswith("parameter"):
            raise SyntaxError("%s: declaration after the first executable statement" % where)
        m = DO_RE.match(s)
        if omp_for is not None:
            # the loops associated with the worksharing construct, in OpenMP's canonical form: upper bounds hoisted
            # (Fortran evaluates them once anyway), unit stride, nothing between the collapsed loops
            pragma, pwhere = omp_for
            omp_for = None
            mc = re.search(r"collapse\((\d+)\)", pragma)
            ncol = int(mc.group(1)) if mc else 1
            loops = []
            for (ff, nn, t) in exe[pos:pos + ncol]:
                mm = DO_RE.match(t)
                if not mm or len(split_top(mm.group(2))) != 2:
                    raise OmpUnsupported("%s: collapse(%d) needs %d unit-stride do statements" % (pwhere, ncol, ncol))
                w2 = "%s:%d" % (os.path.basename(ff), nn)
                lo, hi = split_top(mm.group(2))
                sy = u.syms.get(mm.group(1))
                if sy is None or sy.type != "int" or sy.dims is not None or sy.is_arg:
                    raise TypeError("%s: do variable %s" % (w2, mm.group(1)))
                tmp[0] += 1
                loops.append((mm.group(1), expr(lo, w2, "int"), expr(hi, w2, "int"), tmp[0], w2))
            C.append("%s/* %s */ { const int %s;" % (pad, pwhere, ", ".join("f2c_hi%d = %s" % (t, hi) for (_, _, hi, t, _) in loops)))
            C.append(pragma)
            for k2, (var, lo, hi, t, w2) in enumerate(loops):
                C.append("%sfor (v_%s = %s; v_%s <= f2c_hi%d; v_%s += 1) {   /* %s */" % (pad, var, lo, var, t, var, w2))
                blocks.append("do_omp_outer" if k2 == 0 else "do_omp_inner")
                ind += 1
            skip_do = ncol - 1
            omp_skip[0] = skip_do
            continue
        if omp_skip[0] > 0:
            if not m:
                raise OmpUnsupported("%s: statement between collapsed loops" % where)
            omp_skip[0] -= 1
            continue
        if m and u.syms.get(m.group(1)) is not None and len(split_top(m.group(2))) in (2, 3):
            var = m.group(1)
            sy = u.syms[var]
            if sy.type != "int" or sy.dims is not None or sy.is_arg:
                raise TypeError("%s: do variable %s" % (where, var))
            parts = split_top(m.group(2))
            e1 = expr(parts[0], where, "int")
            e2 = expr(parts[1], where, "int")
            e3 = expr(parts[2], where, "int") if len(parts) == 3 else "1"
            tmp[0] += 1
            t = tmp[0]
            # bounds and stride evaluated once, iteration count fixed at entry (F2008 8.1.6.6)
            C.append("%s/* %s */ { const int f2c_hi%d = %s, f2c_st%d = %s;" % (pad, where, t, e2, t, e3))
            C.append("%sfor (v_%s = %s; (f2c_st%d > 0) ? (v_%s <= f2c_hi%d) : (v_%s >= f2c_hi%d); v_%s += f2c_st%d) {"
                     % (pad, var, e1, t, var, t, var, t, var, t))
            blocks.append("do")
            ind += 1
            continue
        if re.match(r"^end\s*do$", s):
            kind = blocks.pop() if blocks else None
            if kind not in ("do", "do_omp_outer", "do_omp_inner"):
                raise SyntaxError("%s: end do without do" % where)
            ind -= 1
            C.append("%s%s" % ("  " * ind, "}" if kind == "do_omp_inner" else "} }"))
            continue
        if s.startswith("if") and re.match(r"^if\s*\(", s):
            i0 = s.index("(")
            i1 = matching_paren(s, i0)
            cond, tail = s[i0 + 1:i1], s[i1 + 1:].strip()
            if tail == "then":
                C.append("%s/* %s */ if (%s) {" % (pad, where, expr(cond, where, "logical")))
                blocks.append("if")
                ind += 1
            else:
                C.append("%s/* %s */ if (%s) { %s }" % (pad, where, expr(cond, where, "logical"), simple(tail, where)))
            continue
        m = re.match(r"^else\s*if\s*\((.*)\)\s*then$", s)
        if m:
            if not blocks or blocks[-1] != "if":
                raise SyntaxError("%s: else if without if" % where)
            C.append("%s} else if (%s) { /* %s */" % ("  " * (ind - 1), expr(m.group(1), where, "logical"), where))
            continue
        if s == "else":
            if not blocks or blocks[-1] != "if":
                raise SyntaxError("%s: else without if" % where)
            C.append("%s} else {" % ("  " * (ind - 1)))
            continue
        if re.match(r"^end\s*if$", s):
            if not blocks or blocks.pop() != "if":
                raise SyntaxError("%s: end if without if" % where)
            ind -= 1
            C.append("%s}" % ("  " * ind))
            continue
        C.append("%s/* %s */ %s" % (pad, where, simple(s, where)))
    if blocks:
        raise SyntaxError("%s: unterminated block in %s" % (u.loc, u.name))
    C.append("  goto f2c_return;")
    C.append("f2c_return:")
    for nm in u.order:
        s = u.syms[nm]
        if s.dims is not None and not s.is_arg:
            C.append("  free(v_%s);" % nm)
    C.append("  return;")
    C.append("}")
    C.append("")
    return "\n".join(C)


RUNTIME = r"""/* GENERATED by oracle/f90_to_c.py -- bounds-check runtime of the F2C_BOUNDS build (test infrastructure).
 * A subscript outside the DECLARED bounds is recorded (first occurrence per source line / array / dimension) and the
 * access then proceeds exactly as in the unchecked build, i.e. as compiled Fortran without -fcheck=bounds does.
 * F2C_BOUNDS_ABORT=1 in the environment aborts at the first violation instead. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#define F2C_MAXV 256
static struct { const char* loc; const char* a; int d; long i, lo, hi; long count; } f2c_v[F2C_MAXV];
static int f2c_nv = 0;
long f2c_sub(long i, long lb, long n, const char* a, int d, const char* loc) {
  if (i < lb || i >= lb + n) {
    int k;
    if (getenv("F2C_BOUNDS_ABORT")) {
      fprintf(stderr, "F2C_BOUNDS: %s: subscript %d of %s = %ld outside [%ld, %ld]\n", loc, d, a, i, lb, lb + n - 1);
      abort();
    }
    for (k = 0; k < f2c_nv; ++k)
      if (f2c_v[k].loc == loc && f2c_v[k].a == a && f2c_v[k].d == d) { f2c_v[k].count++; break; }
    if (k == f2c_nv && f2c_nv < F2C_MAXV) {
      f2c_v[k].loc = loc; f2c_v[k].a = a; f2c_v[k].d = d; f2c_v[k].i = i; f2c_v[k].lo = lb; f2c_v[k].hi = lb + n - 1;
      f2c_v[k].count = 1; f2c_nv++;
    }
  }
  return i - lb;
}
int f2c_bounds_count(void) { return f2c_nv; }
void f2c_bounds_reset(void) { f2c_nv = 0; }
/* one violation per line: "file:line array dim first_index lo hi count" */
int f2c_bounds_report(char* buf, int len) {
  int k, pos = 0;
  if (len > 0) buf[0] = 0;
  for (k = 0; k < f2c_nv && pos < len - 160; ++k)
    pos += snprintf(buf + pos, (size_t)(len - pos), "%s %s %d %ld %ld %ld %ld\n", f2c_v[k].loc, f2c_v[k].a, f2c_v[k].d,
                    f2c_v[k].i, f2c_v[k].lo, f2c_v[k].hi, f2c_v[k].count);
  return f2c_nv;
}
"""


def translate_file(path, real, defines=None, openmp=False, report=None):
    """openmp: map the reference's `!$OMP` directives to C pragmas, subroutine by subroutine; a subroutine whose
    directives need what the mapping does not cover (dummy arguments or arrays in private / reduction clauses) is
    translated without them, i.e. serially.  report: dict name -> "openmp" | reason it stayed serial."""
    em = Emitter(real)
    out = [PRELUDE % {"src": path, "real": real}]
    serial = {u.name: u for u in parse_units(logical_statements(path, defines or {}))}
    units = parse_units(logical_statements(path, defines or {}, openmp)) if openmp else list(serial.values())
    for u in units:
        if not openmp:
            out.append(translate_unit(u, em))
            continue
        try:
            text = translate_unit(u, em, openmp=True)
            how = "openmp" if "#pragma omp" in text else "no directives"
        except OmpUnsupported as ex:
            text = translate_unit(serial[u.name], em)
            how = "serial: " + str(ex)
        if report is not None:
            report[u.name] = how
        out.append(text)
    return "\n".join(out), [u.name for u in units]


def signatures(path, defines=None):
    """[{name, args: [{name, type: int|real|double|char, dims: [str, ...] | None}]}] read from the declarations."""
    out = []
    for u in parse_units(logical_statements(path, defines or {})):
        for (f, n, s) in u.body:
            if not parse_decl(u, s, "%s:%d" % (os.path.basename(f), n)):
                break
        out.append({"name": u.name, "file": os.path.basename(path),
                    "args": [{"name": a, "type": u.syms[a].type, "dims": u.syms[a].dims} for a in u.args]})
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--real", choices=("f32", "f64"), required=True)
    ap.add_argument("--out", required=True)
    ap.add_argument("--openmp", action="store_true", help="map the !$OMP directives to C pragmas where possible")
    ap.add_argument("files", nargs="+")
    a = ap.parse_args()
    os.makedirs(a.out, exist_ok=True)
    report = {}
    for f in a.files:
        text, names = translate_file(f, a.real, openmp=a.openmp, report=report)
        dst = os.path.join(a.out, os.path.splitext(os.path.basename(f))[0] + ".c")
        with open(dst, "w") as fh:
            fh.write(text)
        print("%s -> %s: %s" % (f, dst, " ".join(names)), file=sys.stderr)
    with open(os.path.join(a.out, "f2c_runtime.c"), "w") as fh:
        fh.write(RUNTIME)
    if a.openmp:
        with open(os.path.join(a.out, "openmp_report.txt"), "w") as fh:
            for k, v in report.items():
                fh.write("%s: %s\n" % (k, v))


if __name__ == "__main__":
    main()
