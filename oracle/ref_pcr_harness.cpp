// oracle/ref_pcr_harness.cpp -- TEST INFRASTRUCTURE.  Runs the REFERENCE's own C++ PCR, CZ::pcr / CZ::pcr_kernel of
// src/cz_cpp/cz_pcr.cpp (compiled unmodified, where it lies under /root/reference, by oracle/Makefile), on a
// tridiagonal system read from stdin and prints the coefficients and right-hand side after the requested number of
// stages as C99 hex floats, i.e. exactly.  The reference's example programs example/tri/test3.cpp / test5.cpp call the
// same routine; this file only replaces their PMlib timing scaffold by exact input / output.  The outputs pin the PCR
// stage recurrences of the oracle port (oracle_pcr_stages_ = cz_solver.f90:584-608 restated): bit for bit in the
// double build; in the float build the reference's C++ evaluates `1.0 / (1.0 - ap*c[iL] - cp*a[iR])` with double
// literals (one rounding to float) where the Fortran stays in REAL, so float agrees to a few ulp only.
//
//   stdin :  n stages            then 3 lines of n hex floats: a[1..n], c[1..n], d[1..n]   (a[1] and c[n] should be 0)
//   stdout:  3 lines of n hex floats: a, c, d after `stages` stages
#include "cz.h"
#include <cstdio>
#include <cstdlib>
#include <vector>

int main() {
  int n = 0, stages = 0;
  if (scanf("%d %d", &n, &stages) != 2 || n < 1 || stages < 0) { fprintf(stderr, "usage: n stages, then a c d\n"); return 2; }
  // arrays of n + 2 entries, cells 1..n, both end entries zero (cz_pcr.cpp clamps neighbour indices to 0 and n+1)
  std::vector<REAL_TYPE> a(n + 2, 0), c(n + 2, 0), d(n + 2, 0), a1(n + 2, 0), c1(n + 2, 0), d1(n + 2, 0);
  std::vector<REAL_TYPE>* in[3] = {&a, &c, &d};
  for (int q = 0; q < 3; ++q)
    for (int i = 1; i <= n; ++i) {
      double v;
      if (scanf("%la", &v) != 1) { fprintf(stderr, "short input\n"); return 2; }
      (*in[q])[i] = (REAL_TYPE)v;
    }
  CZ z;
  double flop = 0.0;
  z.pcr(n, stages, d.data(), a.data(), c.data(), d1.data(), a1.data(), c1.data(), flop);   // the reference, unmodified
  // CZ::pcr swaps its LOCAL pointers after every stage: after an odd number of stages the newest values are in the
  // second set of arrays
  const bool odd = (stages & 1) != 0;
  std::vector<REAL_TYPE>* out[3] = {odd ? &a1 : &a, odd ? &c1 : &c, odd ? &d1 : &d};
  for (int q = 0; q < 3; ++q) {
    for (int i = 1; i <= n; ++i) printf("%a ", (double)(*out[q])[i]);
    printf("\n");
  }
  return 0;
}
