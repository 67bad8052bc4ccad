/* Checks the division used by the fused kernels (cubez_b200/csrc/fused.cu: div_dd<true>):
 *   a/6  ==  fma(fma(-6, a*y, a), y, a*y),  y = RN(1/6)
 * against IEEE division for float operands with biased exponent in [27,247].
 *   ./div6_check            sampled (every 61st bit pattern, ~70 M values)
 *   ./div6_check full       exhaustive (all 2^32 bit patterns; ~20 s on 8 cores)
 * Exit code 0 iff there is no mismatch inside the guarded range.  Needs a CPU with FMA. */
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
    if (e < 27 || e > 247) continue;
    const float ref = a / b;
    const float q = a * y;
    const float r = fmaf(-b, q, a);
    const float q2 = fmaf(r, y, q);
    uint32_t rb, qb; memcpy(&rb, &ref, 4); memcpy(&qb, &q2, 4);
    n++;
    if (rb != qb) bad++;
  }
  printf("checked %lld values, mismatches %lld\n", n, bad);
  return bad != 0;
}
