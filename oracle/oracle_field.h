/* ORACLE — TEST INFRASTRUCTURE ONLY. Never linked into or called by the product path.
 *
 * CPU restatement of the reference's field arithmetic (reference = /root/reference,
 * NethermindEth/starkpack-winterfell, pure Rust; no Rust toolchain on this image so the
 * reference itself cannot be built: see DESIGN.md "oracle").
 *
 *   f64  : Goldilocks p = 2^64 - 2^32 + 1, elements are Montgomery words x*2^64 mod p in [0,p)
 *          math/src/field/f64/mod.rs:37-40 (M, R2), :57-59 (new), :119-164 (exp, inv),
 *          :275-282 (as_int), :314-366 (add/sub/mul), :651-661 (mont_red_cst)
 *   f128 : q = 2^128 - 45*2^40 + 1, canonical u128
 *          math/src/field/f128/mod.rs:35-38 (M, G), :57-59 (new), :418-474 (add/sub/mul)
 *
 * Parity pinning: tests/test_oracle_field.py checks every function here against Python
 * big-int arithmetic and the constants / identities of the reference's own tests
 * (f64/tests.rs:150-160, f128/tests.rs:129-141).
 */
#ifndef ORACLE_FIELD_H
#define ORACLE_FIELD_H
#include <stdint.h>

typedef unsigned __int128 u128;

/* ---------------------------------------------------------------- f64 (Montgomery) */
#define F64_M 0xFFFFFFFF00000001ULL
#define F64_R2 0xFFFFFFFE00000001ULL
#define F64_ROOT_CANON 7277203076849721926ULL /* f64/mod.rs:264 */
#define F64_TWO_ADICITY 32
#define F64_GENERATOR_CANON 7ULL

/* f64/mod.rs:651-661 */
static inline uint64_t f64_mont_red(u128 x) {
  uint64_t xl = (uint64_t)x, xh = (uint64_t)(x >> 64);
  uint64_t a = xl + (xl << 32);
  uint64_t e = a < xl;
  uint64_t b = a - (a >> 32) - e;
  uint64_t r = xh - b;
  uint64_t c = xh < b;
  return r - (uint64_t)(uint32_t)(0u - (uint32_t)c);
}
/* f64/mod.rs:57-59 */
static inline uint64_t f64_new(uint64_t v) { return f64_mont_red((u128)v * (u128)F64_R2); }
/* f64/mod.rs:275-282 */
static inline uint64_t f64_as_int(uint64_t x) { return f64_mont_red((u128)x); }
/* f64/mod.rs:314-324 */
static inline uint64_t f64_add(uint64_t a, uint64_t b) {
  uint64_t t = F64_M - b;
  uint64_t x1 = a - t;
  uint64_t c = a < t;
  return x1 - (uint64_t)(uint32_t)(0u - (uint32_t)c);
}
/* f64/mod.rs:336-344 */
static inline uint64_t f64_sub(uint64_t a, uint64_t b) {
  uint64_t x1 = a - b;
  uint64_t c = a < b;
  return x1 - (uint64_t)(uint32_t)(0u - (uint32_t)c);
}
/* f64/mod.rs:356-361 */
static inline uint64_t f64_mul(uint64_t a, uint64_t b) { return f64_mont_red((u128)a * (u128)b); }

static inline uint64_t f64_exp(uint64_t b, uint64_t e) {
  uint64_t r = f64_new(1);
  while (e) {
    if (e & 1) r = f64_mul(r, b);
    b = f64_mul(b, b);
    e >>= 1;
  }
  return r;
}
static inline uint64_t f64_inv(uint64_t x) { return f64_exp(x, F64_M - 2); }

/* ---------------------------------------------------------------- f128 (canonical) */
#define F128_M ((((u128)0xFFFFFFFFFFFFFFFFULL) << 64) + (u128)0xFFFFD30000000001ULL)
#define F128_TWO_ADICITY 40
static inline u128 f128_root(void) { /* 23953097886125630542083529559205016746, f128/mod.rs:38 */
  return (((u128)0x120532e7b364080aULL) << 64) | (u128)0x86b8723e1920f4aaULL;
}

static inline u128 f128_new(u128 v) { return v < F128_M ? v : v - F128_M; }
/* f128/mod.rs:418-426 */
static inline u128 f128_add(u128 a, u128 b) {
  u128 z = F128_M - b;
  return a < z ? F128_M - z + a : a - z;
}
/* f128/mod.rs:429-435 */
static inline u128 f128_sub(u128 a, u128 b) { return a < b ? F128_M - b + a : a - b; }

/* (a*b) mod q. The reference (f128/mod.rs:437-474) does two 128x64 schoolbook products with
 * folding steps; the result is the unique canonical residue, so any exact reduction agrees.
 * Here: full 256-bit product, then fold with 2^128 == 45*2^40 - 1 (mod q). */
static inline u128 f128_mul(u128 a, u128 b) {
  uint64_t a0 = (uint64_t)a, a1 = (uint64_t)(a >> 64), b0 = (uint64_t)b, b1 = (uint64_t)(b >> 64);
  u128 p00 = (u128)a0 * b0, p01 = (u128)a0 * b1, p10 = (u128)a1 * b0, p11 = (u128)a1 * b1;
  u128 mid = (p00 >> 64) + (uint64_t)p01 + (uint64_t)p10;
  u128 lo = ((u128)(uint64_t)mid << 64) | (uint64_t)p00;
  u128 hi = p11 + (p01 >> 64) + (p10 >> 64) + (mid >> 64);
  /* value = hi*2^128 + lo == hi*(45*2^40 - 1) + lo */
  const u128 K = ((u128)45 << 40) - 1; /* < 2^46 */
  while (hi) {
    uint64_t h0 = (uint64_t)hi, h1 = (uint64_t)(hi >> 64);
    u128 t0 = (u128)h0 * (uint64_t)K;          /* < 2^110 */
    u128 t1 = (u128)h1 * (uint64_t)K;          /* < 2^110, weight 2^64 */
    u128 add = t0 + (t1 << 64);                /* low 128 bits of hi*K */
    u128 carry = (t1 >> 64) + (add < t0);
    u128 s = lo + add;
    carry += (s < lo);
    lo = s;
    hi = carry;
  }
  while (lo >= F128_M) lo -= F128_M;
  return lo;
}
static inline u128 f128_exp(u128 b, u128 e) {
  u128 r = 1;
  while (e) {
    if (e & 1) r = f128_mul(r, b);
    b = f128_mul(b, b);
    e >>= 1;
  }
  return r;
}
static inline u128 f128_inv(u128 x) { return x ? f128_exp(x, F128_M - 2) : 0; }
#endif
