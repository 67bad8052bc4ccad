/* Checks the division used by the fused kernels (cubez_b200/csrc/fused.cu: div_fast):
 *   a/6  ==  fma(fma(-6, as*y, as), y, as*y) * 2^-64,  as = a*2^64,  y = RN(1/6)
 * against IEEE division for every float with biased exponent 0, 1 and 4..183 (-0 excluded).
 *   ./div6_check            sampled (every 61st bit pattern)
 *   ./div6_check full       exhaustive (all bit patterns of the set; ~20 s on 8 cores)
 * Exit code 0 iff there is no mismatch.  Needs a CPU with FMA. */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
int main(int argc, char** argv) {
  const long long step = (argc > 1 && !strcmp(argv[1], "full")) ? 1 : 61;
  const float b = 6.0f, y = 1.0f / 6.0f;
  long long bad = 0, n = 0;
#pragma omp parallel for reduction(+ : bad, n) schedule(static)
  for (long long u = 0; u < (1LL << 32); u += step) {
    uint32_t bits = (uint32_t)u; float a; memcpy(&a, &bits, 4);
    const uint32_t e = (bits >> 23) & 0xff;
    if (e > 183 || e == 2 || e == 3 || bits == 0x80000000u) continue;   /* the `rare` set of div_guard_fail */
    const float ref = a / b;
    const float as = a * 0x1p64f;
    const float q = as * y;
    const float r = fmaf(-b, q, as);
    const float q2 = fmaf(r, y, q) * 0x1p-64f;
    uint32_t rb, qb; memcpy(&rb, &ref, 4); memcpy(&qb, &q2, 4);
    n++;
    if (rb != qb) bad++;
  }
  printf("checked %lld values, mismatches %lld\n", n, bad);
  return bad != 0;
}
