"""include/czb200_iface.F90 (bind(C) interface module, north_star "callable from Fortran via ISO_C_BINDING") is generated
from include/czb200.h: it must be up to date, name every exported symbol, and stay within free-form Fortran's limits.
No Fortran compiler exists in this image, so this is a structural check, not a compile."""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cubez_b200 import ffi  # noqa: E402

F90 = os.path.join(ROOT, "include", "czb200_iface.F90")


def test_interface_module_is_up_to_date():
    p = subprocess.run([sys.executable, os.path.join(ROOT, "include", "gen_fortran_iface.py"), "--check"],
                       capture_output=True, text=True)
    assert p.returncode == 0, p.stdout + p.stderr


def test_every_exported_symbol_has_an_interface():
    src = open(F90).read()
    bound = set(re.findall(r"bind\(C, name='(\w+)'\)", src))
    missing = [s for s in list(ffi.REFERENCE_SYMBOLS) + list(ffi.CZB_SYMBOLS) if s not in bound]
    assert not missing, missing
    for name, nargs in ffi.REFERENCE_SYMBOLS.items():      # argument counts of the reference ABI (cz_Ffunc.h)
        m = re.search(r"subroutine %s\(([^)]*)\)\s*(?:&\s*)?bind\(C, name='%s'\)" % (name[:-1], name), src.replace("&\n", ""))
        assert m, name
        assert len([a for a in m.group(1).split(",") if a.strip()]) == nargs, name


def test_free_form_limits_and_balance():
    lines = open(F90).read().splitlines()
    assert max(len(l) for l in lines) <= 132
    assert sum(1 for l in lines if re.match(r"\s*(subroutine|function) ", l)) == \
        sum(1 for l in lines if re.match(r"\s*end (subroutine|function) ", l))
    # Fortran is case-insensitive: no dummy may appear twice in one argument list
    for m in re.finditer(r"(?:subroutine|function) \w+\(([^)]*)\)", "\n".join(lines).replace("&\n", "")):
        names = [a.strip().lower() for a in m.group(1).split(",") if a.strip()]
        assert len(names) == len(set(names)), m.group(0)
