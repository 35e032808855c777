/* roundcheck.c -- exhaustive proof of the rounding identity the GPU activation quantizers rely on.
 *
 * quantize_row_q8_0/q8_1_reference (llama.cpp-b2430/ggml-quants.c:1194-1203, 1409-1421) compute
 * q = roundf(x * id), round-half-AWAY-from-zero, with |x * id| <= 127 (+ rounding).  CUDA has no single
 * instruction for that, so lamm_quant_dev.cuh uses
 *        q = __float2int_rz(__fadd_rn(t, copysignf(0.49999997f, t)))
 * (0.49999997f = the largest float below 0.5).  This program checks the identity for EVERY f32 with
 * |t| <= 256 (2.26e9 bit patterns, both signs, subnormals included):
 *        gcc -O2 -o /tmp/roundcheck tools/roundcheck.c -lm && /tmp/roundcheck
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

int main(void) {
  const float half_below = 0.49999997f;
  uint64_t checked = 0, bad = 0;
  for (uint32_t sign = 0; sign < 2; sign++) {
    for (uint32_t bits = 0; bits <= 0x43800000u; bits++) {          /* 0 .. 256.0f */
      const uint32_t b = bits | (sign << 31);
      float t;
      memcpy(&t, &b, 4);
      const int want = (int)roundf(t);
      volatile float s = t + copysignf(half_below, t);              /* f32 round-to-nearest-even add */
      const int got = (int)s;                                        /* truncation toward zero */
      checked++;
      if (want != got) {
        if (bad < 10) printf("MISMATCH t=%a roundf=%d trick=%d\n", t, want, got);
        bad++;
      }
    }
  }
  printf("%llu values checked, %llu mismatches\n", (unsigned long long)checked, (unsigned long long)bad);
  return bad != 0;
}
