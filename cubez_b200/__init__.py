"""cubez_b200 -- B200 (sm_100a) implementation of the CubeZ iterative-solver hot path
(7-point Poisson/Laplace sweeps: Jacobi, red-black SOR, PBiCGstab, line-SOR with PCR) behind the
reference's Fortran-kernel C ABI.  See DESIGN.md / INTEGRATION.md."""
from . import decomp, ffi, lib  # noqa: F401
from .solver import CZ, SOLVERS, GUIDE  # noqa: F401

__all__ = ["CZ", "SOLVERS", "GUIDE", "decomp", "ffi", "lib"]
