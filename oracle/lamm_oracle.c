/*
 * lamm_oracle.c -- CPU oracle (see lamm_oracle.h).  TEST INFRASTRUCTURE ONLY.
 *
 * Build: gcc -std=c11 -O2 -ffp-contract=off -fPIC -shared (oracle/Makefile).
 * -ffp-contract=off matters: the reference compiles ggml-quants.c with
 * -std=c11, for which GCC never fuses a*b+c, so x*id and iscale*x are rounded
 * to f32 before the next operation.
 */
#define _GNU_SOURCE
#include "lamm_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ */
/* type table (llama.cpp-b2430/ggml.c:458-775, rows used by the path)  */
/* ------------------------------------------------------------------ */
int orc_blck_size(int type) {
  switch (type) {
  case ORC_TYPE_F32: case ORC_TYPE_F16: return 1;
  case ORC_TYPE_Q4_0: case ORC_TYPE_Q4_1: case ORC_TYPE_Q5_0:
  case ORC_TYPE_Q5_1: case ORC_TYPE_Q8_0: case ORC_TYPE_Q8_1: return 32;
  case ORC_TYPE_Q2_K: case ORC_TYPE_Q8_K: return 256;
  case ORC_TYPE_Q3_K: case ORC_TYPE_Q4_K: case ORC_TYPE_Q5_K: case ORC_TYPE_Q6_K: return 256;
  default: return 0;
  }
}

size_t orc_type_size(int type) {
  switch (type) {
  case ORC_TYPE_F32: return 4;
  case ORC_TYPE_F16: return 2;
  case ORC_TYPE_Q4_0: return sizeof(orc_q4_0);
  case ORC_TYPE_Q4_1: return sizeof(orc_q4_1);
  case ORC_TYPE_Q5_0: return sizeof(orc_q5_0);
  case ORC_TYPE_Q5_1: return sizeof(orc_q5_1);
  case ORC_TYPE_Q8_0: return sizeof(orc_q8_0);
  case ORC_TYPE_Q8_1: return sizeof(orc_q8_1);
  case ORC_TYPE_Q2_K: return sizeof(orc_q2_K);
  case ORC_TYPE_Q8_K: return sizeof(orc_q8_K);
  case ORC_TYPE_Q3_K: return sizeof(orc_q3_K);
  case ORC_TYPE_Q4_K: return sizeof(orc_q4_K);
  case ORC_TYPE_Q5_K: return sizeof(orc_q5_K);
  case ORC_TYPE_Q6_K: return sizeof(orc_q6_K);
  default: return 0;
  }
}

/* src/loongarch_matmul.cpp:37-42 pairs */
int orc_vec_dot_type(int type) {
  switch (type) {
  case ORC_TYPE_F32: return ORC_TYPE_F32;
  case ORC_TYPE_F16: return ORC_TYPE_F16;            /* ggml.c:486-496 */
  case ORC_TYPE_Q4_0: case ORC_TYPE_Q5_0: case ORC_TYPE_Q8_0: return ORC_TYPE_Q8_0;
  case ORC_TYPE_Q4_1: case ORC_TYPE_Q5_1: case ORC_TYPE_Q8_1: return ORC_TYPE_Q8_1;
  case ORC_TYPE_Q2_K: return ORC_TYPE_Q8_K;
  case ORC_TYPE_Q3_K: case ORC_TYPE_Q4_K: case ORC_TYPE_Q5_K: case ORC_TYPE_Q6_K: return ORC_TYPE_Q8_K;
  default: return -1;
  }
}

size_t orc_row_size(int type, int64_t k) {
  int bs = orc_blck_size(type);
  if (bs == 0) return 0;
  return (size_t)(k / bs) * orc_type_size(type);
}

_Static_assert(sizeof(orc_q4_0) == 18, "q4_0");
_Static_assert(sizeof(orc_q4_1) == 20, "q4_1");
_Static_assert(sizeof(orc_q5_0) == 22, "q5_0");
_Static_assert(sizeof(orc_q5_1) == 24, "q5_1");
_Static_assert(sizeof(orc_q8_0) == 34, "q8_0");
_Static_assert(sizeof(orc_q8_1) == 36, "q8_1");
_Static_assert(sizeof(orc_q2_K) == 84, "q2_K");
_Static_assert(sizeof(orc_q8_K) == 292, "q8_K");

/* ------------------------------------------------------------------ */
/* fp16 (IEEE binary16, round-to-nearest-even on narrowing)            */
/* ------------------------------------------------------------------ */
static inline uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

float orc_fp16_to_fp32(uint16_t h) {
  uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
  uint32_t exp = (h >> 10) & 0x1Fu;
  uint32_t man = h & 0x3FFu;
  if (exp == 0x1F) return u2f(sign | 0x7F800000u | (man << 13)); /* inf/nan */
  if (exp == 0) {
    if (man == 0) return u2f(sign);
    /* subnormal: man * 2^-24, exact in f32 */
    float v = (float)man * 0x1.0p-24f;
    return u2f(f2u(v) | sign);
  }
  return u2f(sign | ((exp + 112u) << 23) | (man << 13));
}

uint16_t orc_fp32_to_fp16(float f) {
  uint32_t u = f2u(f);
  uint16_t sign = (uint16_t)((u >> 16) & 0x8000u);
  uint32_t a = u & 0x7FFFFFFFu;
  if (a > 0x7F800000u) return (uint16_t)(sign | 0x7E00u);       /* nan */
  if (a >= 0x477FF000u) return (uint16_t)(sign | 0x7C00u);      /* >= 65520 -> inf */
  if (a < 0x33000000u) return sign;                             /* < 2^-25 -> 0 */
  int32_t e = (int32_t)(a >> 23) - 127;
  uint32_t m = (a & 0x7FFFFFu) | 0x800000u;                     /* 24-bit */
  int shift;                                                    /* bits dropped */
  uint32_t base;
  if (e < -14) { shift = 13 + (-14 - e); base = 0; }            /* subnormal */
  else { shift = 13; base = (uint32_t)(e + 15) << 10; m &= 0x7FFFFFu; }
  uint32_t q = m >> shift;
  uint32_t rem = m & ((1u << shift) - 1u);
  uint32_t half = 1u << (shift - 1);
  if (rem > half || (rem == half && (q & 1u))) q++;
  /* mantissa carry rolls into the exponent field, which is what IEEE wants */
  return (uint16_t)(sign | (base + q));
}

/* ------------------------------------------------------------------ */
/* activation quantizers                                               */
/* ------------------------------------------------------------------ */
/* ggml-quants.c:1182-1205 */
void orc_quantize_row_q8_0(const float *x, void *vy, int64_t k) {
  orc_q8_0 *y = (orc_q8_0 *)vy;
  const int64_t nb = k / 32;
  for (int64_t i = 0; i < nb; i++) {
    const float *xb = x + i * 32;
    float amax = 0.0f;
    for (int j = 0; j < 32; j++) {
      float a = fabsf(xb[j]);
      amax = amax > a ? amax : a;          /* MAX(amax, |v|): NaN never wins */
    }
    const float d = amax / 127.0f;
    const float id = d ? 1.0f / d : 0.0f;
    y[i].d = orc_fp32_to_fp16(d);
    for (int j = 0; j < 32; j++) {
      const float v = xb[j] * id;
      y[i].qs[j] = (int8_t)roundf(v);      /* half away from zero */
    }
  }
}

/* ggml-quants.c:1396-1429 */
void orc_quantize_row_q8_1(const float *x, void *vy, int64_t k) {
  orc_q8_1 *y = (orc_q8_1 *)vy;
  const int64_t nb = k / 32;
  for (int64_t i = 0; i < nb; i++) {
    const float *xb = x + i * 32;
    float amax = 0.0f;
    for (int j = 0; j < 32; j++) {
      float a = fabsf(xb[j]);
      amax = amax > a ? amax : a;
    }
    const float d = amax / 127.0f;
    const float id = d ? 1.0f / d : 0.0f;
    y[i].d = orc_fp32_to_fp16(d);
    int sum = 0;
    for (int j = 0; j < 32; j++) {
      const float v = xb[j] * id;
      y[i].qs[j] = (int8_t)roundf(v);
      sum += y[i].qs[j];
    }
    const float s = (float)sum * d;        /* int -> f32, one f32 multiply */
    y[i].s = orc_fp32_to_fp16(s);
  }
}

/* ggml-quants.c:1767-1772: round-to-nearest-even through the 1.5*2^23 magic */
static inline int orc_nearest_int(float v) {
  float t = v + 12582912.0f;
  return (int)(f2u(t) & 0x007FFFFFu) - 0x00400000;
}

/* ggml-quants.c:3981-4018 */
void orc_quantize_row_q8_K(const float *x, void *vy, int64_t k) {
  orc_q8_K *y = (orc_q8_K *)vy;
  const int64_t nb = k / 256;
  for (int64_t i = 0; i < nb; i++) {
    const float *xb = x + i * 256;
    float max = 0.0f, amax = 0.0f;
    for (int j = 0; j < 256; j++) {
      float ax = fabsf(xb[j]);
      if (ax > amax) { amax = ax; max = xb[j]; }   /* first occurrence wins */
    }
    if (!amax) {
      y[i].d = 0.0f;
      memset(y[i].qs, 0, 256);
      continue;                                    /* bsums untouched (9-4) */
    }
    const float iscale = -127.0f / max;
    for (int j = 0; j < 256; j++) {
      const float p = iscale * xb[j];
      int v = orc_nearest_int(p);
      y[i].qs[j] = (int8_t)(v < 127 ? v : 127);
    }
    for (int g = 0; g < 16; g++) {
      int sum = 0;
      for (int l = 0; l < 16; l++) sum += y[i].qs[g * 16 + l];
      y[i].bsums[g] = (int16_t)sum;
    }
    y[i].d = 1.0f / iscale;
  }
}

int orc_quantize_act(int vec_dot_type, const float *x, int64_t ldx, void *y,
                     int64_t n_rows, int64_t k) {
  size_t rs = orc_row_size(vec_dot_type, k);
  for (int64_t r = 0; r < n_rows; r++) {
    const float *xr = x + r * ldx;
    char *yr = (char *)y + r * rs;
    switch (vec_dot_type) {
    case ORC_TYPE_Q8_0: orc_quantize_row_q8_0(xr, yr, k); break;
    case ORC_TYPE_Q8_1: orc_quantize_row_q8_1(xr, yr, k); break;
    case ORC_TYPE_Q8_K: orc_quantize_row_q8_K(xr, yr, k); break;
    case ORC_TYPE_F32: memcpy(yr, xr, (size_t)k * 4); break;
    case ORC_TYPE_F16:                               /* ggml_fp32_to_fp16_row, ggml.c:354-372 */
      for (int64_t i = 0; i < k; i++) ((uint16_t *)yr)[i] = orc_fp32_to_fp16(xr[i]);
      break;
    default: return -1;
    }
  }
  return 0;
}

/* ------------------------------------------------------------------ */
/* weight quantizers -- synthetic tensor generation only               */
/* ------------------------------------------------------------------ */
static inline uint8_t clip_u8(int v, int hi) { return (uint8_t)(v < hi ? v : hi); }

/* symmetric 4/5-bit: ggml-quants.c:1002-1037 (q4_0), :1085-1127 (q5_0).
 * The signed extreme maps to -2^(bits-1); the (int8_t) cast truncates.    */
static void quant_sym(const float *x, int bits, float *d_out, uint8_t q[32]) {
  float amax = 0.0f, max = 0.0f;
  for (int j = 0; j < 32; j++) {
    float a = fabsf(x[j]);
    if (amax < a) { amax = a; max = x[j]; }
  }
  const int half = 1 << (bits - 1);
  const float d = max / (float)(-half);
  const float id = d ? 1.0f / d : 0.0f;
  *d_out = d;
  for (int j = 0; j < 32; j++) {
    const float v = x[j] * id;
    q[j] = clip_u8((int8_t)(v + ((float)half + 0.5f)), 2 * half - 1); /* x + 8.5f / 16.5f */
  }
}

/* asymmetric 4/5-bit: ggml-quants.c:1044-1078 (q4_1), :1133-1175 (q5_1) */
static void quant_asym(const float *x, int bits, float *d_out, float *m_out,
                       uint8_t q[32]) {
  float mn = 3.402823466e+38f, mx = -3.402823466e+38f;
  for (int j = 0; j < 32; j++) {
    if (x[j] < mn) mn = x[j];
    if (x[j] > mx) mx = x[j];
  }
  const int levels = (1 << bits) - 1;
  const float d = (mx - mn) / (float)levels;
  const float id = d ? 1.0f / d : 0.0f;
  *d_out = d; *m_out = mn;
  for (int j = 0; j < 32; j++) {
    const float v = (x[j] - mn) * id;
    if (bits == 4) q[j] = clip_u8((int8_t)(v + 0.5f), 15);
    else q[j] = (uint8_t)(v + 0.5f);
  }
}

static void pack_nibbles(const uint8_t q[32], uint8_t qs[16]) {
  for (int j = 0; j < 16; j++) qs[j] = (uint8_t)((q[j] & 0x0F) | ((q[j + 16] & 0x0F) << 4));
}
static void pack_high_bits(const uint8_t q[32], uint8_t qh[4]) {
  uint32_t h = 0;
  for (int j = 0; j < 32; j++) h |= (uint32_t)((q[j] >> 4) & 1u) << j;
  memcpy(qh, &h, 4);
}

int orc_quantize_weights(int type, const float *x, void *vy, int64_t n_rows,
                         int64_t k) {
  if (type == ORC_TYPE_F32) { memcpy(vy, x, (size_t)(n_rows * k) * 4); return 0; }
  if (type == ORC_TYPE_F16) {
    for (int64_t i = 0; i < n_rows * k; i++) ((uint16_t *)vy)[i] = orc_fp32_to_fp16(x[i]);
    return 0;
  }
  if (orc_blck_size(type) != 32 || k % 32) return -1;
  const int64_t nb = n_rows * (k / 32);
  uint8_t q[32];
  float d, m;
  for (int64_t i = 0; i < nb; i++) {
    const float *xb = x + i * 32;
    switch (type) {
    case ORC_TYPE_Q4_0: {
      orc_q4_0 *y = (orc_q4_0 *)vy + i;
      quant_sym(xb, 4, &d, q);
      y->d = orc_fp32_to_fp16(d); pack_nibbles(q, y->qs);
    } break;
    case ORC_TYPE_Q4_1: {
      orc_q4_1 *y = (orc_q4_1 *)vy + i;
      quant_asym(xb, 4, &d, &m, q);
      y->d = orc_fp32_to_fp16(d); y->m = orc_fp32_to_fp16(m); pack_nibbles(q, y->qs);
    } break;
    case ORC_TYPE_Q5_0: {
      orc_q5_0 *y = (orc_q5_0 *)vy + i;
      quant_sym(xb, 5, &d, q);
      y->d = orc_fp32_to_fp16(d); pack_nibbles(q, y->qs); pack_high_bits(q, y->qh);
    } break;
    case ORC_TYPE_Q5_1: {
      orc_q5_1 *y = (orc_q5_1 *)vy + i;
      quant_asym(xb, 5, &d, &m, q);
      y->d = orc_fp32_to_fp16(d); y->m = orc_fp32_to_fp16(m);
      pack_nibbles(q, y->qs); pack_high_bits(q, y->qh);
    } break;
    case ORC_TYPE_Q8_0: orc_quantize_row_q8_0(xb, (orc_q8_0 *)vy + i, 32); break;
    default: return -1;
    }
  }
  return 0;
}

/* splitmix64: tiny seeded generator for synthetic Q2_K super-blocks */
static inline uint64_t sm64(uint64_t *s) {
  uint64_t z = (*s += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

/* Uniform random bit patterns for scales/qs; d, dmin ~ fp16(U(0.001, 0.1)):
 * finite, never NaN/Inf (SURVEY 8d "alternative fast path").               */
void orc_random_q2_K(void *vy, int64_t n_blocks, uint64_t seed) {
  orc_q2_K *y = (orc_q2_K *)vy;
  uint64_t s = seed;
  for (int64_t i = 0; i < n_blocks; i++) {
    uint64_t r;
    for (int j = 0; j < 16; j += 8) { r = sm64(&s); memcpy(y[i].scales + j, &r, 8); }
    for (int j = 0; j < 64; j += 8) { r = sm64(&s); memcpy(y[i].qs + j, &r, 8); }
    r = sm64(&s);
    float u0 = (float)(r & 0xFFFFFF) / 16777216.0f;
    float u1 = (float)((r >> 24) & 0xFFFFFF) / 16777216.0f;
    y[i].d = orc_fp32_to_fp16(0.001f + 0.099f * u0);
    y[i].dmin = orc_fp32_to_fp16(0.001f + 0.099f * u1);
  }
}

int orc_random_kquant(int type, void *vy, int64_t n_blocks, uint64_t seed) {
  if (type == ORC_TYPE_Q2_K) { orc_random_q2_K(vy, n_blocks, seed); return 0; }
  const size_t bs = orc_type_size(type);
  size_t off_d, off_dmin = 0;                 /* byte offsets of the fp16 super-block scales */
  switch (type) {
  case ORC_TYPE_Q3_K: off_d = offsetof(orc_q3_K, d); break;
  case ORC_TYPE_Q4_K: off_d = 0; off_dmin = 2; break;
  case ORC_TYPE_Q5_K: off_d = 0; off_dmin = 2; break;
  case ORC_TYPE_Q6_K: off_d = offsetof(orc_q6_K, d); break;
  default: return -1;
  }
  uint8_t *y = (uint8_t *)vy;
  uint64_t s = seed;
  for (int64_t i = 0; i < n_blocks; i++, y += bs) {
    for (size_t j = 0; j < bs; j += 8) {
      uint64_t r = sm64(&s);
      memcpy(y + j, &r, bs - j < 8 ? bs - j : 8);
    }
    uint64_t r = sm64(&s);
    uint16_t h = orc_fp32_to_fp16(0.001f + 0.099f * ((float)(r & 0xFFFFFF) / 16777216.0f));
    memcpy(y + off_d, &h, 2);
    if (off_dmin) {
      h = orc_fp32_to_fp16(0.001f + 0.099f * ((float)((r >> 24) & 0xFFFFFF) / 16777216.0f));
      memcpy(y + off_dmin, &h, 2);
    }
  }
  return 0;
}

/* ------------------------------------------------------------------ */
/* unpack helpers: one block -> 32 (or 256) integer weights            */
/* ------------------------------------------------------------------ */
static inline void unpack_q4(const uint8_t qs[16], int q[32]) {
  for (int j = 0; j < 16; j++) { q[j] = qs[j] & 0x0F; q[j + 16] = qs[j] >> 4; }
}
static inline void unpack_q5(const uint8_t qs[16], const uint8_t qh[4], int q[32]) {
  uint32_t h; memcpy(&h, qh, 4);
  unpack_q4(qs, q);
  for (int j = 0; j < 32; j++) q[j] |= (int)((h >> j) & 1u) << 4;
}
/* Q2_K element order: ggml-quants.c:2116-2143 / src/lamm_kernel_q2_k.hpp:49-71.
 * sub[e] = index of the 16-element sub-block that element e belongs to.   */
static inline void unpack_q2K(const orc_q2_K *b, int q[256]) {
  int e = 0;
  for (int n = 0; n < 2; n++)
    for (int sh = 0; sh < 8; sh += 2)
      for (int l = 0; l < 32; l++) q[e++] = (b->qs[n * 32 + l] >> sh) & 3;
}

/* ---- K-quants of SURVEY 8f-1 -----------------------------------------
 * One super-block -> 256 signed integer weights q[e] in element order, the
 * integer scale sc[g] of every 16-element sub-block g = e/16 and, for the
 * formats with a minimum, the integer min mn[g]:
 *     w[e] = f(d) * sc[e/16] * q[e] - f(dmin) * mn[e/16]
 * get_scale_min_k4: ggml-quants.c:2027-2034.                              */
static inline void scale_min_k4(int j, const uint8_t *q, int *d, int *m) {
  if (j < 4) { *d = q[j] & 63; *m = q[j + 4] & 63; }
  else { *d = (q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4); *m = (q[j + 4] >> 4) | ((q[j] >> 6) << 4); }
}
typedef struct { int q[256]; int sc[16]; int mn[16]; float d, dmin; } orc_kq;

static void unpack_kq(int type, const void *vb, orc_kq *o) {
  memset(o->mn, 0, sizeof o->mn);
  o->dmin = 0.0f;
  switch (type) {
  case ORC_TYPE_Q4_K: {          /* ggml-quants.c:2851-2874 / 7301-7358 */
    const orc_q4_K *b = (const orc_q4_K *)vb;
    o->d = orc_fp16_to_fp32(b->d); o->dmin = orc_fp16_to_fp32(b->dmin);
    for (int g = 0; g < 8; g++) {              /* 32-element groups */
      int sc, m; scale_min_k4(g, b->scales, &sc, &m);
      o->sc[2 * g] = o->sc[2 * g + 1] = sc;
      o->mn[2 * g] = o->mn[2 * g + 1] = m;
      for (int l = 0; l < 32; l++) {
        const uint8_t by = b->qs[32 * (g >> 1) + l];
        o->q[32 * g + l] = (g & 1) ? (by >> 4) : (by & 0xF);
      }
    }
  } break;
  case ORC_TYPE_Q5_K: {          /* ggml-quants.c:3134-3160 / 7968-8029 */
    const orc_q5_K *b = (const orc_q5_K *)vb;
    o->d = orc_fp16_to_fp32(b->d); o->dmin = orc_fp16_to_fp32(b->dmin);
    for (int g = 0; g < 8; g++) {
      int sc, m; scale_min_k4(g, b->scales, &sc, &m);
      o->sc[2 * g] = o->sc[2 * g + 1] = sc;
      o->mn[2 * g] = o->mn[2 * g + 1] = m;
      for (int l = 0; l < 32; l++) {
        const uint8_t by = b->qs[32 * (g >> 1) + l];
        o->q[32 * g + l] = ((g & 1) ? (by >> 4) : (by & 0xF)) + (((b->qh[l] >> g) & 1) ? 16 : 0);
      }
    }
  } break;
  case ORC_TYPE_Q6_K: {          /* ggml-quants.c:3383-3412 / 8695-8737 */
    const orc_q6_K *b = (const orc_q6_K *)vb;
    o->d = orc_fp16_to_fp32(b->d);
    for (int n = 0; n < 2; n++)
      for (int l = 0; l < 32; l++) {
        const uint8_t la = b->ql[64 * n + l], lb = b->ql[64 * n + 32 + l], h = b->qh[32 * n + l];
        int *q = o->q + 128 * n;
        q[l]      = (int)((la & 0xF) | (((h >> 0) & 3) << 4)) - 32;
        q[l + 32] = (int)((lb & 0xF) | (((h >> 2) & 3) << 4)) - 32;
        q[l + 64] = (int)((la >> 4)  | (((h >> 4) & 3) << 4)) - 32;
        q[l + 96] = (int)((lb >> 4)  | (((h >> 6) & 3) << 4)) - 32;
      }
    for (int g = 0; g < 16; g++) o->sc[g] = b->scales[g];
  } break;
  default: {                     /* Q3_K: ggml-quants.c:2550-2598 / 6583-6646 */
    const orc_q3_K *b = (const orc_q3_K *)vb;
    o->d = orc_fp16_to_fp32(b->d);
    for (int g = 0; g < 16; g++) {
      const int lo = (g < 8) ? (b->scales[g] & 0xF) : (b->scales[g - 8] >> 4);
      const int hi = (b->scales[8 + (g & 3)] >> (2 * (g >> 2))) & 3;
      o->sc[g] = (lo | (hi << 4)) - 32;
    }
    int e = 0;
    for (int n = 0; n < 2; n++)
      for (int j = 0; j < 4; j++)
        for (int l = 0; l < 32; l++)
          o->q[e++] = (int)((b->qs[32 * n + l] >> (2 * j)) & 3) - (((b->hmask[l] >> (4 * n + j)) & 1) ? 0 : 4);
  } break;
  }
}
static int is_kq_ext(int t) {
  return t == ORC_TYPE_Q3_K || t == ORC_TYPE_Q4_K || t == ORC_TYPE_Q5_K || t == ORC_TYPE_Q6_K;
}

int orc_dequantize_row(int type, const void *vx, float *y, int64_t k) {
  if (is_kq_ext(type)) {
    orc_kq u;
    const size_t bs = orc_type_size(type);
    for (int64_t i = 0; i < k / 256; i++) {
      unpack_kq(type, (const char *)vx + (size_t)i * bs, &u);
      for (int e = 0; e < 256; e++) {
        /* dequantize_row_q{4,5}_K: d1 = d*sc, m1 = min*m, y = d1*q - m1; q{3,6}_K: (d*sc)*q */
        const float dl = u.d * (float)u.sc[e / 16];
        y[i * 256 + e] = (type == ORC_TYPE_Q4_K || type == ORC_TYPE_Q5_K)
                             ? dl * (float)u.q[e] - u.dmin * (float)u.mn[e / 16]
                             : dl * (float)u.q[e];
      }
    }
    return 0;
  }
  int q[256];
  switch (type) {
  case ORC_TYPE_F32: memcpy(y, vx, (size_t)k * 4); return 0;
  case ORC_TYPE_F16:
    for (int64_t i = 0; i < k; i++) y[i] = orc_fp16_to_fp32(((const uint16_t *)vx)[i]);
    return 0;
  case ORC_TYPE_Q4_0: {
    const orc_q4_0 *x = (const orc_q4_0 *)vx;
    for (int64_t i = 0; i < k / 32; i++) {
      float d = orc_fp16_to_fp32(x[i].d); unpack_q4(x[i].qs, q);
      for (int j = 0; j < 32; j++) y[i * 32 + j] = (float)(q[j] - 8) * d;
    }
  } return 0;
  case ORC_TYPE_Q4_1: {
    const orc_q4_1 *x = (const orc_q4_1 *)vx;
    for (int64_t i = 0; i < k / 32; i++) {
      float d = orc_fp16_to_fp32(x[i].d), m = orc_fp16_to_fp32(x[i].m);
      unpack_q4(x[i].qs, q);
      for (int j = 0; j < 32; j++) y[i * 32 + j] = (float)q[j] * d + m;
    }
  } return 0;
  case ORC_TYPE_Q5_0: {
    const orc_q5_0 *x = (const orc_q5_0 *)vx;
    for (int64_t i = 0; i < k / 32; i++) {
      float d = orc_fp16_to_fp32(x[i].d); unpack_q5(x[i].qs, x[i].qh, q);
      for (int j = 0; j < 32; j++) y[i * 32 + j] = (float)(q[j] - 16) * d;
    }
  } return 0;
  case ORC_TYPE_Q5_1: {
    const orc_q5_1 *x = (const orc_q5_1 *)vx;
    for (int64_t i = 0; i < k / 32; i++) {
      float d = orc_fp16_to_fp32(x[i].d), m = orc_fp16_to_fp32(x[i].m);
      unpack_q5(x[i].qs, x[i].qh, q);
      for (int j = 0; j < 32; j++) y[i * 32 + j] = (float)q[j] * d + m;
    }
  } return 0;
  case ORC_TYPE_Q8_0: {
    const orc_q8_0 *x = (const orc_q8_0 *)vx;
    for (int64_t i = 0; i < k / 32; i++) {
      float d = orc_fp16_to_fp32(x[i].d);
      for (int j = 0; j < 32; j++) y[i * 32 + j] = (float)x[i].qs[j] * d;
    }
  } return 0;
  case ORC_TYPE_Q8_1: {
    const orc_q8_1 *x = (const orc_q8_1 *)vx;
    for (int64_t i = 0; i < k / 32; i++) {
      float d = orc_fp16_to_fp32(x[i].d);
      for (int j = 0; j < 32; j++) y[i * 32 + j] = (float)x[i].qs[j] * d;
    }
  } return 0;
  case ORC_TYPE_Q2_K: {
    const orc_q2_K *x = (const orc_q2_K *)vx;
    for (int64_t i = 0; i < k / 256; i++) {
      float d = orc_fp16_to_fp32(x[i].d), mn = orc_fp16_to_fp32(x[i].dmin);
      unpack_q2K(&x[i], q);
      for (int e = 0; e < 256; e++) {
        uint8_t sc = x[i].scales[e / 16];
        float dl = d * (float)(sc & 0xF), ml = mn * (float)(sc >> 4);
        y[i * 256 + e] = dl * (float)q[e] - ml;
      }
    }
  } return 0;
  case ORC_TYPE_Q8_K: {
    const orc_q8_K *x = (const orc_q8_K *)vx;
    for (int64_t i = 0; i < k / 256; i++)
      for (int e = 0; e < 256; e++) y[i * 256 + e] = x[i].d * (float)x[i].qs[e];
  } return 0;
  default: return -1;
  }
}

/* ------------------------------------------------------------------ */
/* per-format block terms (the kernel formulas of SURVEY 8a)           */
/* ------------------------------------------------------------------ */
/* integer dot of 32 weights against 32 int8 activations */
static inline int idot32(const int q[32], const int8_t *x) {
  int s = 0;
  for (int j = 0; j < 32; j++) s += q[j] * (int)x[j];
  return s;
}

/* One (W block, Xq block) term, returned as the exact f32 factors so that the
 * caller can combine them in double (oracle of record) or in f32 (naive
 * order).  term = scale * isum + extra.                                     */
typedef struct { float scale; int isum; float extra_a, extra_b; int extra_i; } orc_term;
/* value = (double)scale*isum + (double)extra_a*extra_b            (32-blocks)
 * Q2_K:  = (double)scale*isum - (double)extra_a*extra_i  with scale=x.d*f(w.d),
 *          extra_a = x.d*f(w.dmin), extra_i = sum bsums*min                   */

/* src/lamm_kernel_q4_0.hpp:21-37 */
static inline orc_term term_q4_0(const orc_q4_0 *w, const orc_q8_0 *x) {
  int q[32]; unpack_q4(w->qs, q);
  for (int j = 0; j < 32; j++) q[j] -= 8;
  orc_term t = { orc_fp16_to_fp32(w->d) * orc_fp16_to_fp32(x->d), idot32(q, x->qs), 0, 0, 0 };
  return t;
}
/* src/lamm_kernel_q4_1.hpp:7-24 */
static inline orc_term term_q4_1(const orc_q4_1 *w, const orc_q8_1 *x) {
  int q[32]; unpack_q4(w->qs, q);
  orc_term t = { orc_fp16_to_fp32(w->d) * orc_fp16_to_fp32(x->d), idot32(q, x->qs),
                 orc_fp16_to_fp32(w->m), orc_fp16_to_fp32(x->s), 0 };
  return t;
}
/* src/lamm_kernel_q5_0.hpp:49-67 (simd form; the naive one is defect 9-1) */
static inline orc_term term_q5_0(const orc_q5_0 *w, const orc_q8_0 *x) {
  int q[32]; unpack_q5(w->qs, w->qh, q);
  for (int j = 0; j < 32; j++) q[j] -= 16;
  orc_term t = { orc_fp16_to_fp32(w->d) * orc_fp16_to_fp32(x->d), idot32(q, x->qs), 0, 0, 0 };
  return t;
}
/* src/lamm_kernel_q5_1.hpp:34-56 */
static inline orc_term term_q5_1(const orc_q5_1 *w, const orc_q8_1 *x) {
  int q[32]; unpack_q5(w->qs, w->qh, q);
  orc_term t = { orc_fp16_to_fp32(w->d) * orc_fp16_to_fp32(x->d), idot32(q, x->qs),
                 orc_fp16_to_fp32(w->m), orc_fp16_to_fp32(x->s), 0 };
  return t;
}
/* src/lamm_kernel_q8_0.hpp:15-30 (signed x signed; simd form is defect 9-2) */
static inline orc_term term_q8_0(const orc_q8_0 *w, const orc_q8_0 *x) {
  int q[32];
  for (int j = 0; j < 32; j++) q[j] = w->qs[j];
  orc_term t = { orc_fp16_to_fp32(w->d) * orc_fp16_to_fp32(x->d), idot32(q, x->qs), 0, 0, 0 };
  return t;
}
/* src/lamm_kernel_q2_k.hpp:28-75 */
static inline orc_term term_q2_K(const orc_q2_K *w, const orc_q8_K *x) {
  int q[256]; unpack_q2K(w, q);
  int isum = 0, summs = 0;
  for (int g = 0; g < 16; g++) {
    int sub = 0;
    for (int l = 0; l < 16; l++) sub += q[g * 16 + l] * (int)x->qs[g * 16 + l];
    isum += (w->scales[g] & 0xF) * sub;
    summs += (int)x->bsums[g] * (w->scales[g] >> 4);
  }
  orc_term t = { x->d * orc_fp16_to_fp32(w->d), isum,
                 x->d * orc_fp16_to_fp32(w->dmin), 0, summs };
  return t;
}

/* ggml_vec_dot_q{3,4,5,6}_K_q8_K, scalar forms: integer sub-block dots times the
 * integer scales, the min term from bsums; term = x.d*f(w.d)*isum - x.d*f(w.dmin)*summs */
static inline orc_term term_kq(int type, const void *w, const orc_q8_K *x) {
  orc_kq u; unpack_kq(type, w, &u);
  int isum = 0, summs = 0;
  for (int g = 0; g < 16; g++) {
    int sub = 0;
    for (int l = 0; l < 16; l++) sub += u.q[g * 16 + l] * (int)x->qs[g * 16 + l];
    isum += u.sc[g] * sub;
    summs += (int)x->bsums[g] * u.mn[g];
  }
  orc_term t = { x->d * u.d, isum, x->d * u.dmin, 0, summs };
  return t;
}

static inline orc_term block_term(int w_type, const char *w, const char *x, int64_t kb) {
  if (is_kq_ext(w_type))
    return term_kq(w_type, w + (size_t)kb * orc_type_size(w_type), (const orc_q8_K *)x + kb);
  switch (w_type) {
  case ORC_TYPE_Q4_0: return term_q4_0((const orc_q4_0 *)w + kb, (const orc_q8_0 *)x + kb);
  case ORC_TYPE_Q4_1: return term_q4_1((const orc_q4_1 *)w + kb, (const orc_q8_1 *)x + kb);
  case ORC_TYPE_Q5_0: return term_q5_0((const orc_q5_0 *)w + kb, (const orc_q8_0 *)x + kb);
  case ORC_TYPE_Q5_1: return term_q5_1((const orc_q5_1 *)w + kb, (const orc_q8_1 *)x + kb);
  case ORC_TYPE_Q8_0: return term_q8_0((const orc_q8_0 *)w + kb, (const orc_q8_0 *)x + kb);
  default:            return term_q2_K((const orc_q2_K *)w + kb, (const orc_q8_K *)x + kb);
  }
}

static int supported_w(int t) {
  return t == ORC_TYPE_F32 || t == ORC_TYPE_F16 || t == ORC_TYPE_Q4_0 || t == ORC_TYPE_Q4_1 ||
         t == ORC_TYPE_Q5_0 || t == ORC_TYPE_Q5_1 || t == ORC_TYPE_Q8_0 ||
         t == ORC_TYPE_Q2_K || is_kq_ext(t);
}

/* one output element, double accumulation */
static double dot_f64(int w_type, const char *wrow, const char *xrow, int64_t K) {
  double acc = 0.0;
  if (w_type == ORC_TYPE_F32) {
    const float *a = (const float *)wrow, *b = (const float *)xrow;
    for (int64_t k = 0; k < K; k++) acc += (double)a[k] * (double)b[k];
    return acc;
  }
  if (w_type == ORC_TYPE_F16) {                      /* ggml_vec_dot_f16: products of the two halves */
    const uint16_t *a = (const uint16_t *)wrow, *b = (const uint16_t *)xrow;
    for (int64_t k = 0; k < K; k++)
      acc += (double)orc_fp16_to_fp32(a[k]) * (double)orc_fp16_to_fp32(b[k]);
    return acc;
  }
  const int64_t nb = K / orc_blck_size(w_type);
  for (int64_t kb = 0; kb < nb; kb++) {
    orc_term t = block_term(w_type, wrow, xrow, kb);
    if (w_type == ORC_TYPE_Q2_K || is_kq_ext(w_type))
      acc += (double)t.scale * (double)t.isum - (double)t.extra_a * (double)t.extra_i;
    else
      acc += (double)t.scale * (double)t.isum + (double)t.extra_a * (double)t.extra_b;
  }
  return acc;
}

/* one output element, f32 running sum in the naive kernels' order */
static float dot_f32(int w_type, const char *wrow, const char *xrow, int64_t K) {
  float acc = 0.0f;
  if (w_type == ORC_TYPE_F32) {
    const float *a = (const float *)wrow, *b = (const float *)xrow;
    for (int64_t k = 0; k < K; k++) acc += a[k] * b[k];
    return acc;
  }
  if (w_type == ORC_TYPE_F16) {
    const uint16_t *a = (const uint16_t *)wrow, *b = (const uint16_t *)xrow;
    for (int64_t k = 0; k < K; k++) acc += orc_fp16_to_fp32(a[k]) * orc_fp16_to_fp32(b[k]);
    return acc;
  }
  const int64_t nb = K / orc_blck_size(w_type);
  for (int64_t kb = 0; kb < nb; kb++) {
    orc_term t = block_term(w_type, wrow, xrow, kb);
    if (w_type == ORC_TYPE_Q2_K || is_kq_ext(w_type)) {
      const float a = t.scale * (float)t.isum;
      const float b = t.extra_a * (float)t.extra_i;
      acc += a - b;
    } else if (w_type == ORC_TYPE_Q4_1 || w_type == ORC_TYPE_Q5_1) {
      const float a = t.scale * (float)t.isum;
      const float b = t.extra_a * t.extra_b;
      acc += a + b;
    } else {
      acc += t.scale * (float)t.isum;
    }
  }
  return acc;
}

static int check_shape(int w_type, int64_t K) {
  if (!supported_w(w_type)) return -1;
  if (K <= 0 || K % orc_blck_size(w_type)) return -2;
  return 0;
}

int orc_mul_mat_q(int w_type, const void *w, int64_t w_stride, const void *xq,
                  int64_t xq_stride, float *c, int64_t ldc, int64_t M,
                  int64_t N, int64_t K) {
  int rc = check_shape(w_type, K);
  if (rc) return rc;
  for (int64_t j = 0; j < N; j++)
    for (int64_t i = 0; i < M; i++)
      c[j * ldc + i] = (float)dot_f64(w_type, (const char *)w + i * w_stride,
                                      (const char *)xq + j * xq_stride, K);
  return 0;
}

int orc_mul_mat_q_f32(int w_type, const void *w, int64_t w_stride,
                      const void *xq, int64_t xq_stride, float *c, int64_t ldc,
                      int64_t M, int64_t N, int64_t K) {
  int rc = check_shape(w_type, K);
  if (rc) return rc;
  for (int64_t j = 0; j < N; j++)
    for (int64_t i = 0; i < M; i++)
      c[j * ldc + i] = dot_f32(w_type, (const char *)w + i * w_stride,
                               (const char *)xq + j * xq_stride, K);
  return 0;
}

/* src/loongarch_matmul.cpp:64-143 + ggml.c:10865-10887 for one batch plane */
int orc_mul_mat(int w_type, const void *w, int64_t w_stride, const float *x,
                int64_t ldx, float *c, int64_t ldc, int64_t M, int64_t N,
                int64_t K, void *xq_out) {
  int rc = check_shape(w_type, K);
  if (rc) return rc;
  const int vt = orc_vec_dot_type(w_type);
  const size_t rs = orc_row_size(vt, K);
  void *xq = xq_out ? xq_out : malloc(rs * (size_t)N);
  if (!xq) return -3;
  /* calloc-like for Q8_K bsums of all-zero blocks when we own the buffer */
  if (!xq_out) memset(xq, 0, rs * (size_t)N);
  orc_quantize_act(vt, x, ldx, xq, N, K);
  rc = orc_mul_mat_q(w_type, w, w_stride, xq, (int64_t)rs, c, ldc, M, N, K);
  if (!xq_out) free(xq);
  return rc;
}

/* ggml_mul_mat_id (ggml.c:10992-11170): row j of X against expert ids[j][id] */
int orc_mul_mat_id(int w_type, const void *const *ws, int n_as, int64_t w_stride,
                   const int32_t *ids, int64_t ids_stride, int id, const float *x,
                   int64_t ldx, float *c, int64_t ldc, int64_t M, int64_t N, int64_t K) {
  int rc = check_shape(w_type, K);
  if (rc) return rc;
  for (int64_t j = 0; j < N; j++) {
    const int32_t e = ids[j * ids_stride + id];
    if (e < 0 || e >= n_as) return -4;
    rc = orc_mul_mat(w_type, ws[e], w_stride, x + j * ldx, ldx, c + j * ldc, ldc, M, 1, K, NULL);
    if (rc) return rc;
  }
  return 0;
}

/* ---- threaded f32 variant (timed CPU "port" baseline) -------------- */
typedef struct {
  int w_type; const char *w; int64_t w_stride; const char *xq; int64_t xq_stride;
  float *c; int64_t ldc, N, K, i0, i1;
} mt_job;

static void *mt_worker(void *p) {
  mt_job *j = (mt_job *)p;
  for (int64_t n = 0; n < j->N; n++)
    for (int64_t i = j->i0; i < j->i1; i++)
      j->c[n * j->ldc + i] = dot_f32(j->w_type, j->w + i * j->w_stride,
                                     j->xq + n * j->xq_stride, j->K);
  return NULL;
}

/* ------------------------------------------------------------------ */
/* the ops either side of the path                        */
/* ------------------------------------------------------------------ */
/* ggml_compute_forward_rms_norm_f32 (ggml.c:10355-10405): the sum of the f32 squares is kept in a double
 * (ggml_float) in element order, mean = (float)(sum / K), scale = 1.0f / sqrtf(mean + eps), y = x * scale
 * (ggml_vec_scale_f32); then llm_build_norm's ggml_mul by the norm weight (llama.cpp:5171-5201), one f32
 * product per element.  norm_w may be NULL (no weight).                                                       */
int orc_rms_norm_mul(const float *x, int64_t ldx, const float *norm_w, float eps, float *y, int64_t ldy,
                     int64_t n_rows, int64_t k) {
  if (!x || !y || n_rows < 0 || k <= 0 || !(eps > 0.0f)) return -1;
  for (int64_t r = 0; r < n_rows; r++) {
    const float *xr = x + r * ldx;
    float *yr = y + r * ldy;
    double sum = 0.0;
    for (int64_t i = 0; i < k; i++) sum += (double)(xr[i] * xr[i]);
    const float mean = (float)(sum / (double)k);
    const float scale = 1.0f / sqrtf(mean + eps);
    for (int64_t i = 0; i < k; i++) {
      float v = xr[i] * scale;
      if (norm_w) v = v * norm_w[i];
      yr[i] = v;
    }
  }
  return 0;
}

/* ggml_vec_silu_f32 with GGML_SILU_FP16 (ggml.c:157, 1894-1902): the argument is rounded to fp16 and looked up in a
 * 65 536-entry table that ggml_init filled with fp16(silu(f32(h))), silu(f) = f / (1 + expf(-f)) (ggml.c:1883-1885,
 * 2711-2719).  The value is therefore a function of the fp16 pattern only; expf is the host libm's, as in ggml.    */
float orc_silu(float x) {
  const float f = orc_fp16_to_fp32(orc_fp32_to_fp16(x));
  return orc_fp16_to_fp32(orc_fp32_to_fp16(f / (1.0f + expf(-f))));
}

/* llm_build_ffn, LLM_FFN_SILU + LLM_FFN_PAR (llama.cpp:5203-5290): y = silu(gate) * up, one f32 product            */
int orc_silu_mul(const float *gate, int64_t ldg, const float *up, int64_t ldu, float *y, int64_t ldy, int64_t n_rows,
                 int64_t k) {
  if (!gate || !up || !y || n_rows < 0 || k <= 0) return -1;
  for (int64_t r = 0; r < n_rows; r++)
    for (int64_t i = 0; i < k; i++) y[r * ldy + i] = orc_silu(gate[r * ldg + i]) * up[r * ldu + i];
  return 0;
}

int orc_mul_mat_mt(int w_type, const void *w, int64_t w_stride, const float *x,
                   int64_t ldx, float *c, int64_t ldc, int64_t M, int64_t N,
                   int64_t K, int nth, void *xq_scratch) {
  int rc = check_shape(w_type, K);
  if (rc) return rc;
  if (nth < 1) nth = 1;
  if (nth > 256) nth = 256;
  const int vt = orc_vec_dot_type(w_type);
  const size_t rs = orc_row_size(vt, K);
  void *xq = xq_scratch ? xq_scratch : calloc((size_t)N, rs);
  if (!xq) return -3;
  orc_quantize_act(vt, x, ldx, xq, N, K);   /* single-threaded, as ggml INIT */
  pthread_t th[256]; mt_job jobs[256];
  const int64_t per = M / nth;
  for (int t = 0; t < nth; t++) {
    mt_job jb = { w_type, (const char *)w, w_stride, (const char *)xq, (int64_t)rs,
                  c, ldc, N, K, t * per, (t == nth - 1) ? M : (t + 1) * per };
    jobs[t] = jb;
    if (t > 0) pthread_create(&th[t], NULL, mt_worker, &jobs[t]);
  }
  mt_worker(&jobs[0]);
  for (int t = 1; t < nth; t++) pthread_join(th[t], NULL);
  if (!xq_scratch) free(xq);
  return 0;
}
