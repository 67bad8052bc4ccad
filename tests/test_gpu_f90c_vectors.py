"""GPU parity against OUTPUTS OF THE REFERENCE'S OWN FORTRAN: every entry point of the reference's kernel ABI, called on
libczb200 with host pointers, must reproduce the committed golden vectors written by the reference's Fortran text executed
through oracle/f90_to_c.py (tests/golden/f90c_digests.json; generator tests/golden/make_f90c_golden.py) -- every array
argument BIT FOR BIT (non-unit coefficients, stretched mesh, line lengths 21 / 29 / 48 / 62, four consecutive calls),
residual sums and dot products within the accumulation error (the library accumulates REAL products in double).
Needs neither /root/reference nor oracle/: only the committed JSON."""
import json
import os
import sys

import pytest

import f90c_harness as H

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))


@pytest.mark.parametrize("real", ["f32", "f64"])
def test_gpu_library_reproduces_the_reference_fortran_vectors(real):
    import cubez_b200
    import make_f90c_golden as MG
    digests = json.load(open(os.path.join(HERE, "golden", "f90c_digests.json")))
    sigs = {s["name"]: s for s in json.load(open(H.SIG_JSON))}
    lib = cubez_b200.lib.load(real)
    assert lib.czb_init(0) == 0
    n = H.check_library_against_fortran_vectors(lib, real, MG.SIZES[:4], digests, sigs, MG.run_routine, MG.record)
    assert n >= 130
    assert lib.czb_launch_count() > 0
