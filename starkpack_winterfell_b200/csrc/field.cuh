// Field arithmetic for the trace-commitment path, host + device.
//
//   Gl64 : Goldilocks p = 2^64 - 2^32 + 1. The reference stores Montgomery words x*2^64 mod p
//          (math/src/field/f64/mod.rs:48-53). The NTT is linear, so the kernels multiply those words by
//          PLAIN-form twiddles with plain modular multiplication and the data stays in Montgomery form
//          (SURVEY F5). All functions here are therefore plain arithmetic mod p on canonical inputs in
//          [0,p), results canonical. `mont_red` is the reference's reduction (f64/mod.rs:651-661) used
//          for as_int (hashing) and new.
//   F128 : q = 2^128 - 45*2^40 + 1, canonical (math/src/field/f128/mod.rs:35,418-474).
//
// Everything is __host__ __device__ so the host plan builder (twiddle tables) and the CPU unit tests use
// exactly the code the kernels run.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define WG_HD __host__ __device__ __forceinline__
#else
#define WG_HD inline
#endif

namespace wg {

WG_HD void mul64wide(uint64_t a, uint64_t b, uint64_t& hi, uint64_t& lo) {
#if defined(__CUDA_ARCH__)
  lo = a * b;
  hi = __umul64hi(a, b);
#else
  unsigned __int128 p = (unsigned __int128)a * b;
  lo = (uint64_t)p;
  hi = (uint64_t)(p >> 64);
#endif
}

// ------------------------------------------------------------------------------------------------ Gl64
// Elements are MONTGOMERY WORDS x*2^64 mod p in [0,p), exactly the reference's BaseElement(u64); mul is the
// reference's Montgomery multiplication (f64/mod.rs:356-361 with mont_red_cst :651-661), so twiddle tables are
// Montgomery words too. On the device add/sub/mont_red are written as 32-bit carry chains in PTX: the
// straightforward C forms compile to 10 / 10 / 33 SASS instructions, these to 8 / 5 / ~17, and the product's
// four IMAD.WIDE run on the FMA pipe while the carry chains run on the ALU pipe.
#if defined(__CUDACC__)
// run-time 1 and -1: opaque to ptxas, so multiply-adds by them stay IMAD (FMA pipe) instead of being folded to IADD3
static __constant__ uint32_t K_ONE = 1u;
static __constant__ uint32_t K_MONE = 0xFFFFFFFFu;
#endif

struct Gl64 {
  typedef uint64_t T;
  static constexpr int kFieldId = 0;
  static constexpr int kElemBytes = 8;
  static constexpr int kTwoAdicity = 32;                         // f64/mod.rs:252
  static constexpr uint64_t P = 0xFFFFFFFF00000001ULL;            // f64/mod.rs:37
  static constexpr uint64_t EPS = 0xFFFFFFFFULL;                  // 2^64 mod p (= Montgomery form of 1)
  static constexpr uint64_t ROOT_CANON = 7277203076849721926ULL;  // f64/mod.rs:264 (order 2^32)
  static constexpr uint64_t GENERATOR_CANON = 7;                  // f64/mod.rs:248

  static WG_HD T zero() { return 0; }
  static WG_HD T one() { return EPS; }
  static WG_HD bool is_zero(T a) { return a == 0; }

  // Device add/sub/mul below are written so that every carry-CONSUMING step that does not have to produce a carry is a
  // multiply-add (mad/madc with the run-time constants K_ONE / K_MONE, which ptxas cannot fold): those issue on the
  // FMA pipe as IMAD / IMAD.X, while carry-producing steps must be IADD3 on the ALU pipe. The NTT kernels are bound
  // by integer issue with the ALU pipe the busy one, so this balances the two pipes (about 17 ALU : 16 FMA per
  // butterfly instead of 22 : 7).
  static WG_HD T add(T a, T b) {
#if defined(__CUDA_ARCH__)
    // a - (p - b), then + p on borrow (the reference's formulation, f64/mod.rs:314-324)
    uint32_t a0 = (uint32_t)a, a1 = (uint32_t)(a >> 32), b0 = (uint32_t)b, b1 = (uint32_t)(b >> 32), d0, d1;
    asm("{\n\t.reg .u32 m, bit, t0, t1, n0, n1;\n\t"
        "mad.lo.u32 n0, %4, %7, %7;\n\t"    // ~b0
        "add.cc.u32 n0, n0, 2;\n\t"          // 1 - b0; carry = no borrow
        "madc.lo.u32 n1, %5, %7, 0xFFFFFFFE;\n\t"  // ~b1 - 1 + carry = 0xFFFFFFFF - b1 - borrow
        "sub.cc.u32 t0, %2, n0;\n\tsubc.cc.u32 t1, %3, n1;\n\tsubc.u32 m, 0, 0;\n\t"
        "mul.lo.u32 bit, m, %7;\n\t"         // borrow as 0/1
        "add.cc.u32 %0, t0, bit;\n\t"        // + p = (+1, +0xFFFFFFFF + carry)
        "madc.lo.u32 %1, t1, %6, m;\n\t}"
        : "=&r"(d0), "=&r"(d1)
        : "r"(a0), "r"(a1), "r"(b0), "r"(b1), "r"(K_ONE), "r"(K_MONE));
    return ((uint64_t)d1 << 32) | d0;
#else
    uint64_t s = a + b;
    if (s < a || s >= P) s += EPS;
    return s;
#endif
  }
  static WG_HD T sub(T a, T b) {
#if defined(__CUDA_ARCH__)
    uint32_t a0 = (uint32_t)a, a1 = (uint32_t)(a >> 32), b0 = (uint32_t)b, b1 = (uint32_t)(b >> 32), d0, d1;
    asm("{\n\t.reg .u32 m, bit, t0, t1;\n\t"
        "sub.cc.u32 t0, %2, %4;\n\tsubc.cc.u32 t1, %3, %5;\n\tsubc.u32 m, 0, 0;\n\t"
        "mul.lo.u32 bit, m, %7;\n\t"
        "add.cc.u32 %0, t0, bit;\n\t"
        "madc.lo.u32 %1, t1, %6, m;\n\t}"
        : "=&r"(d0), "=&r"(d1)
        : "r"(a0), "r"(a1), "r"(b0), "r"(b1), "r"(K_ONE), "r"(K_MONE));
    return ((uint64_t)d1 << 32) | d0;
#else
    uint64_t d = a - b;
    if (a < b) d -= EPS;  // + p
    return d;
#endif
  }
  static WG_HD T neg(T a) { return a ? P - a : 0; }

  // reference Montgomery reduction x*2^-64 mod p (f64/mod.rs:651-661), x = xh*2^64 + xl < p*2^64; result in [0,p).
  // Regrouped as r = (xh + a1 + e) - (a1:l0) with a1 = (l1 + l0) mod 2^32, e its carry (xh + 2^32 cannot overflow
  // for xh < p - 2^32, which holds for products of reduced elements).
  static WG_HD uint64_t mont_red(uint64_t xh, uint64_t xl) {
#if defined(__CUDA_ARCH__)
    uint32_t l0 = (uint32_t)xl, l1 = (uint32_t)(xl >> 32), h0 = (uint32_t)xh, h1 = (uint32_t)(xh >> 32), r0, r1;
    asm("{\n\t.reg .u32 a1, s0, s1, m;\n\t"
        "add.cc.u32 a1, %3, %2;\n\t"
        "addc.cc.u32 s0, %4, a1;\n\taddc.u32 s1, %5, 0;\n\t"
        "sub.cc.u32 s0, s0, %2;\n\tsubc.cc.u32 s1, s1, a1;\n\tsubc.u32 m, 0, 0;\n\t"
        "mul.lo.u32 a1, m, %7;\n\t"
        "add.cc.u32 %0, s0, a1;\n\t"
        "madc.lo.u32 %1, s1, %6, m;\n\t}"
        : "=&r"(r0), "=&r"(r1)
        : "r"(l0), "r"(l1), "r"(h0), "r"(h1), "r"(K_ONE), "r"(K_MONE));
    return ((uint64_t)r1 << 32) | r0;
#else
    uint64_t a = xl + (xl << 32);
    uint64_t e = a < xl ? 1 : 0;
    uint64_t b = a - (a >> 32) - e;
    uint64_t r = xh - b;
    if (xh < b) r -= EPS;
    return r;
#endif
  }
  // Montgomery product a*b*2^-64 mod p of two words in [0,p)
  static WG_HD T mul(T a, T b) {
#if defined(__CUDA_ARCH__) && defined(WG_MUL_I128)
    // 128-bit product left to the compiler: mul.lo.u64 / mul.hi.u64 share their partial products and ptxas chains the
    // four IMAD.WIDE through their carry predicates (IMAD.WIDE.X), i.e. the accumulation runs on the FMA pipe:
    // 6 ALU + 10 FMA instructions with the reduction, against 10 + 8 for the hand-written carry chains below
    const unsigned __int128 x = (unsigned __int128)a * b;
    return mont_red((uint64_t)(x >> 64), (uint64_t)x);
#elif defined(__CUDA_ARCH__)
    // four 32x32->64 products (IMAD.WIDE, FMA pipe), two carry chains to assemble (h1:h0:l1:t0), then the
    // Montgomery reduction of mont_red() inline; 10 ALU-pipe + 8 FMA-pipe SASS instructions
    uint32_t a0 = (uint32_t)a, a1 = (uint32_t)(a >> 32), b0 = (uint32_t)b, b1 = (uint32_t)(b >> 32), r0, r1;
    asm("{\n\t.reg .u64 t, u, v, w;\n\t.reg .u32 t0, t1, u0, u1, v0, v1, w0, w1, l1, h0, h1, a1, s0, s1, m;\n\t"
        "mul.wide.u32 t, %2, %4;\n\tmul.wide.u32 v, %2, %5;\n\tmul.wide.u32 w, %3, %4;\n\tmul.wide.u32 u, %3, %5;\n\t"
        "mov.b64 {t0, t1}, t;\n\tmov.b64 {v0, v1}, v;\n\tmov.b64 {w0, w1}, w;\n\tmov.b64 {u0, u1}, u;\n\t"
        "add.cc.u32 l1, t1, v0;\n\taddc.cc.u32 h0, u0, v1;\n\taddc.u32 h1, u1, 0;\n\t"
        "add.cc.u32 l1, l1, w0;\n\taddc.cc.u32 h0, h0, w1;\n\taddc.u32 h1, h1, 0;\n\t"
        "add.cc.u32 a1, l1, t0;\n\t"
        "addc.cc.u32 s0, h0, a1;\n\taddc.u32 s1, h1, 0;\n\t"
        "sub.cc.u32 s0, s0, t0;\n\tsubc.cc.u32 s1, s1, a1;\n\tsubc.u32 m, 0, 0;\n\t"
        "mul.lo.u32 u0, m, %7;\n\t"
        "add.cc.u32 %0, s0, u0;\n\t"
        "madc.lo.u32 %1, s1, %6, m;\n\t}"
        : "=&r"(r0), "=&r"(r1)
        : "r"(a0), "r"(a1), "r"(b0), "r"(b1), "r"(K_ONE), "r"(K_MONE));
    return ((uint64_t)r1 << 32) | r0;
#else
    uint64_t hi, lo;
    mul64wide(a, b, hi, lo);
    return mont_red(hi, lo);
#endif
  }

  // (hi*2^64 + lo) mod p, plain reduction (2^64 = 2^32 - 1, 2^96 = -1); host-side table building only
  static WG_HD T reduce128(uint64_t hi, uint64_t lo) {
    uint64_t hh = hi >> 32, hl = hi & EPS;
    uint64_t t0 = lo - hh;
    if (lo < hh) t0 -= EPS;
    uint64_t t1 = hl * EPS;
    uint64_t r = t0 + t1;
    if (r < t0) r += EPS;
    if (r >= P) r -= P;
    return r;
  }
  // internal word -> canonical integer (f64/mod.rs:275-282). Input any u64, output in [0,p).
  static WG_HD uint64_t as_int(uint64_t w) { return mont_red(0, w); }
  // canonical integer -> internal Montgomery word (f64/mod.rs:57-59): v * 2^64 mod p
  static WG_HD uint64_t to_mont(uint64_t v) { return reduce128(v >= P ? v - P : v, 0); }
  static WG_HD T from_canon(uint64_t v) { return to_mont(v); }
  static WG_HD T from_u64(uint64_t v) { return to_mont(v); }

  // 2^e mod p for 0 <= e < 192 (2^96 = -1), canonical
  static WG_HD constexpr uint64_t pow2_mod(int e) {
    // not recursive on purpose: a recursive constexpr is not inlined in device code and ends up as a run-time CALL
    const bool negate = e >= 96;
    const int f = negate ? e - 96 : e;
    const uint64_t v = f < 64 ? (1ULL << (f & 63)) : ((1ULL << ((f - 32) & 63)) - (1ULL << ((f - 64) & 63)));
    return negate ? P - v : v;
  }
  // Montgomery reduction of the 128-bit integer x << S (0 <= S < 64): x * 2^S * 2^-64 mod p, canonical. This is how the
  // kernels multiply by the power-of-two roots of unity: the high half of x << S is tiny, so no product is needed and,
  // for S >= 32, the low word is zero and half of the reduction disappears (12 / 10 SASS instructions instead of the 19
  // of a product with a constant).
  template <int S>
  static WG_HD T mont_shift(T x) {
    static_assert(S >= 0 && S < 64, "shift out of range");
#if defined(__CUDA_ARCH__)
    uint32_t x0 = (uint32_t)x, x1 = (uint32_t)(x >> 32), r0, r1;
    if (S == 0) return mont_red(0, x);
    if (S < 32) {
      const uint32_t l0 = x0 << (S & 31), l1 = __funnelshift_l(x0, x1, S & 31), h0 = x1 >> ((32 - S) & 31);
      asm("{\n\t.reg .u32 a1, s0, s1, m;\n\t"
          "add.cc.u32 a1, %3, %2;\n\t"
          "addc.cc.u32 s0, %4, a1;\n\taddc.u32 s1, 0, 0;\n\t"
          "sub.cc.u32 s0, s0, %2;\n\tsubc.cc.u32 s1, s1, a1;\n\tsubc.u32 m, 0, 0;\n\t"
          "mul.lo.u32 a1, m, %6;\n\t"
          "add.cc.u32 %0, s0, a1;\n\t"
          "madc.lo.u32 %1, s1, %5, m;\n\t}"
          : "=&r"(r0), "=&r"(r1)
          : "r"(l0), "r"(l1), "r"(h0), "r"(K_ONE), "r"(K_MONE));
    } else {
      constexpr int Q = S - 32;
      const uint32_t l1 = Q ? x0 << (Q & 31) : x0, h0 = Q ? __funnelshift_l(x0, x1, Q & 31) : x1,
                     h1 = Q ? x1 >> ((32 - Q) & 31) : 0u;
      asm("{\n\t.reg .u32 s0, s1, m, bit;\n\t"
          "add.cc.u32 s0, %3, %2;\n\taddc.u32 s1, %4, 0;\n\t"
          "sub.cc.u32 s1, s1, %2;\n\tsubc.u32 m, 0, 0;\n\t"
          "mul.lo.u32 bit, m, %6;\n\t"
          "add.cc.u32 %0, s0, bit;\n\t"
          "madc.lo.u32 %1, s1, %5, m;\n\t}"
          : "=&r"(r0), "=&r"(r1)
          : "r"(l1), "r"(h0), "r"(h1), "r"(K_ONE), "r"(K_MONE));
    }
    return ((uint64_t)r1 << 32) | r0;
#else
    return S == 0 ? mont_red(0, x) : mont_red(x >> ((64 - S) & 63), x << (S & 63));
#endif
  }
  // The order-16 root is a power of two: w16 = 2^12 (ROOT^(2^26) = 8), and 2^96 = -1. Multiplying a
  // Montgomery word by the PLAIN constant 2^k is the Montgomery product with 2^(k+64); with e = (k + 64) mod 192 that
  // constant is +2^e for e < 64 and -2^(e-96) for 96 <= e < 160 (a shift, mont_shift above), and a two-bit value
  // otherwise (e in [64,96) or [160,192): idx 1, 2 forward, 6, 7 inverse), which stays a product.
  template <bool INV>
  static WG_HD constexpr int root16_e(int idx) { return ((INV ? 192 - 12 * (idx & 15) : 12 * (idx & 15)) + 64) % 192; }
  template <bool INV>
  static WG_HD constexpr bool root16_neg(int idx) {  // the shift form yields -x*w
    return root16_e<INV>(idx) >= 96 && root16_e<INV>(idx) < 160;
  }
  // |x * w16^(+-idx)| up to the sign root16_neg reports
  template <bool INV>
  static WG_HD T mul_root16_abs(T x, int idx) {
    const int e = root16_e<INV>(idx);
    switch (e >= 96 && e < 160 ? e - 96 : e) {
      case 4: return mont_shift<4>(x);
      case 16: return mont_shift<16>(x);
      case 28: return mont_shift<28>(x);
      case 40: return mont_shift<40>(x);
      case 52: return mont_shift<52>(x);
      default: return mul(x, pow2_mod(e));
    }
  }
  // x * w16^(+-idx)
  template <bool INV>
  static WG_HD T mul_root16(T x, int idx) {
    idx &= 15;
    if (idx == 0) return x;
    if (idx == 8) return neg(x);
    const T r = mul_root16_abs<INV>(x, idx);
    return root16_neg<INV>(idx) ? neg(r) : r;
  }
  // (x - y) * w16^(+-idx): the sign of the shift form is absorbed by swapping the operands of the subtraction
  template <bool INV>
  static WG_HD T sub_mul_root16(T x, T y, int idx) {
    idx &= 15;
    if (idx == 0) return sub(x, y);
    if (idx == 8) return sub(y, x);
    return root16_neg<INV>(idx) ? mul_root16_abs<INV>(sub(y, x), idx) : mul_root16_abs<INV>(sub(x, y), idx);
  }
};

// ------------------------------------------------------------------------------------------------ F128
struct u128 {
  uint64_t lo, hi;
};

WG_HD bool u128_lt(u128 a, u128 b) { return a.hi < b.hi || (a.hi == b.hi && a.lo < b.lo); }
WG_HD u128 u128_add(u128 a, u128 b, uint64_t& carry) {
  u128 r;
  r.lo = a.lo + b.lo;
  uint64_t c0 = r.lo < a.lo ? 1 : 0;
  uint64_t t = a.hi + b.hi;
  uint64_t c1 = t < a.hi ? 1 : 0;
  r.hi = t + c0;
  uint64_t c2 = r.hi < t ? 1 : 0;
  carry = c1 | c2;
  return r;
}
WG_HD u128 u128_sub(u128 a, u128 b) {  // wrapping
  u128 r;
  r.lo = a.lo - b.lo;
  r.hi = a.hi - b.hi - (a.lo < b.lo ? 1 : 0);
  return r;
}

struct F128 {
  typedef u128 T;
  static constexpr int kFieldId = 1;
  static constexpr int kElemBytes = 16;
  static constexpr int kTwoAdicity = 40;                      // f128/mod.rs:169
  static constexpr uint64_t M_LO = 0xFFFFD30000000001ULL;      // q = 2^128 - 45*2^40 + 1 (f128/mod.rs:35)
  static constexpr uint64_t M_HI = 0xFFFFFFFFFFFFFFFFULL;
  static constexpr uint64_t K = (45ULL << 40) - 1;             // 2^128 mod q
  static constexpr uint64_t ROOT_LO = 0x86b8723e1920f4aaULL;   // 23953097886125630542083529559205016746
  static constexpr uint64_t ROOT_HI = 0x120532e7b364080aULL;   // (f128/mod.rs:38), order 2^40
  static constexpr uint64_t GENERATOR_CANON = 3;               // f128/mod.rs:165

  static WG_HD T modulus() { return T{M_LO, M_HI}; }
  static WG_HD T zero() { return T{0, 0}; }
  static WG_HD T one() { return T{1, 0}; }
  static WG_HD T from_u64(uint64_t v) { return T{v, 0}; }
  static WG_HD T from_canon(uint64_t v) { return T{v, 0}; }
  static WG_HD bool is_zero(T a) { return (a.lo | a.hi) == 0; }
  static WG_HD T canon(T a) { return u128_lt(a, modulus()) ? a : u128_sub(a, modulus()); }

  // Device add / sub / mul are 32-bit carry chains in PTX (the portable C forms compile to 34 / 27 / 145 SASS
  // instructions, these to 15 / 11 / ~110). -q mod 2^128 = K = 45*2^40 - 1 = (0, 0, 0x2CFF, 0xFFFFFFFF).
  static WG_HD T add(T a, T b) {
#if defined(__CUDA_ARCH__)
    // s = a + b (129 bits), t = s + K = s - q mod 2^128; a + b >= q exactly when one of the two additions carries out
    T r;
    asm("{\n\t.reg .u32 a0, a1, a2, a3, b0, b1, b2, b3, s0, s1, s2, s3, t0, t1, t2, t3, c;\n\t.reg .pred p;\n\t"
        "mov.b64 {a0, a1}, %2;\n\tmov.b64 {a2, a3}, %3;\n\tmov.b64 {b0, b1}, %4;\n\tmov.b64 {b2, b3}, %5;\n\t"
        "add.cc.u32 s0, a0, b0;\n\taddc.cc.u32 s1, a1, b1;\n\taddc.cc.u32 s2, a2, b2;\n\taddc.cc.u32 s3, a3, b3;\n\taddc.u32 c, 0, 0;\n\t"
        "add.cc.u32 t0, s0, 0xFFFFFFFF;\n\taddc.cc.u32 t1, s1, 0x2CFF;\n\taddc.cc.u32 t2, s2, 0;\n\taddc.cc.u32 t3, s3, 0;\n\taddc.u32 c, c, 0;\n\t"
        "setp.ne.u32 p, c, 0;\n\t"
        "selp.u32 s0, t0, s0, p;\n\tselp.u32 s1, t1, s1, p;\n\tselp.u32 s2, t2, s2, p;\n\tselp.u32 s3, t3, s3, p;\n\t"
        "mov.b64 %0, {s0, s1};\n\tmov.b64 %1, {s2, s3};\n\t}"
        : "=l"(r.lo), "=l"(r.hi)
        : "l"(a.lo), "l"(a.hi), "l"(b.lo), "l"(b.hi));
    return r;
#else
    uint64_t c;
    T s = u128_add(a, b, c);
    // a,b < q: true sum < 2q. Subtract q when the sum overflowed 2^128 or is >= q.
    if (c || !u128_lt(s, modulus())) s = u128_sub(s, modulus());
    return s;
#endif
  }
  static WG_HD T sub(T a, T b) {
#if defined(__CUDA_ARCH__)
    // d = a - b mod 2^128, then + q under the borrow mask; q = (0xFFFFFFFF, 0xFFFFFFFF, 0xFFFFD300, 1)
    T r;
    asm("{\n\t.reg .u32 a0, a1, a2, a3, b0, b1, b2, b3, d0, d1, d2, d3, m, q0, q1;\n\t"
        "mov.b64 {a0, a1}, %2;\n\tmov.b64 {a2, a3}, %3;\n\tmov.b64 {b0, b1}, %4;\n\tmov.b64 {b2, b3}, %5;\n\t"
        "sub.cc.u32 d0, a0, b0;\n\tsubc.cc.u32 d1, a1, b1;\n\tsubc.cc.u32 d2, a2, b2;\n\tsubc.cc.u32 d3, a3, b3;\n\tsubc.u32 m, 0, 0;\n\t"
        "and.b32 q0, m, 1;\n\tand.b32 q1, m, 0xFFFFD300;\n\t"
        "add.cc.u32 d0, d0, q0;\n\taddc.cc.u32 d1, d1, q1;\n\taddc.cc.u32 d2, d2, m;\n\taddc.u32 d3, d3, m;\n\t"
        "mov.b64 %0, {d0, d1};\n\tmov.b64 %1, {d2, d3};\n\t}"
        : "=l"(r.lo), "=l"(r.hi)
        : "l"(a.lo), "l"(a.hi), "l"(b.lo), "l"(b.hi));
    return r;
#else
    T d = u128_sub(a, b);
    if (u128_lt(a, b)) {
      uint64_t c;
      d = u128_add(d, modulus(), c);
    }
    return d;
#endif
  }
  static WG_HD T neg(T a) { return is_zero(a) ? a : u128_sub(modulus(), a); }

  // fold v = h*2^128 + l  ->  l + h*K (h < 2^64), returns 128-bit result plus carry word
  static WG_HD T fold64(uint64_t h, T l, uint64_t& carry) {
    uint64_t phi, plo;
    mul64wide(h, K, phi, plo);
    return u128_add(l, T{plo, phi}, carry);
  }
  static WG_HD T mul(T a, T b) {
#if defined(__CUDA_ARCH__)
    // 256-bit product from sixteen 32x32->64 products: those with i+j even sit on 64-bit boundaries, those with i+j
    // odd are offset by 32 bits, so each class is summed with 64-bit carry chains and the two sums are merged once.
    // Reduction (2^128 = K = 45*2^40 - 1 mod q): T = H*2^128 + L  ->  U = L + (H*45 << 40) - H  (< 2^174), its top
    // 46 bits H2 folded the same way -> V < 2^128 + 2^92, the single carry folded as + K, then one conditional - q.
    T r;
    asm("{\n\t.reg .u32 a0, a1, a2, a3, b0, b1, b2, b3, t0, t1, t2, t3, t4, t5, t6, t7;\n\t"
        ".reg .u32 e0l, e0h, e1l, e1h, e2l, e2h, e3l, e3h, o0l, o0h, o1l, o1h, o2l, o2h, o3l, o3h;\n\t"
        ".reg .u32 g0, g1, g2, g3, g4, wl, wh, m1, m2, m3, m4, m5, u0, u1, u2, u3, u4, u5, c3, msk, k1, s0, s1, s2, s3;\n\t"
        ".reg .u64 p00, p01, p02, p03, p10, p11, p12, p13, p20, p21, p22, p23, p30, p31, p32, p33;\n\t"
        ".reg .u64 E1, E2, E3, O0, O1, O2, O3, w;\n\t.reg .pred p;\n\t"
        "mov.b64 {a0, a1}, %2;\n\tmov.b64 {a2, a3}, %3;\n\tmov.b64 {b0, b1}, %4;\n\tmov.b64 {b2, b3}, %5;\n\t"
        "mul.wide.u32 p00, a0, b0;\n\tmul.wide.u32 p01, a0, b1;\n\tmul.wide.u32 p02, a0, b2;\n\tmul.wide.u32 p03, a0, b3;\n\t"
        "mul.wide.u32 p10, a1, b0;\n\tmul.wide.u32 p11, a1, b1;\n\tmul.wide.u32 p12, a1, b2;\n\tmul.wide.u32 p13, a1, b3;\n\t"
        "mul.wide.u32 p20, a2, b0;\n\tmul.wide.u32 p21, a2, b1;\n\tmul.wide.u32 p22, a2, b2;\n\tmul.wide.u32 p23, a2, b3;\n\t"
        "mul.wide.u32 p30, a3, b0;\n\tmul.wide.u32 p31, a3, b1;\n\tmul.wide.u32 p32, a3, b2;\n\tmul.wide.u32 p33, a3, b3;\n\t"
        // even class: (E3:E2:E1:p00)
        "add.cc.u64 E1, p02, p11;\n\taddc.cc.u64 E2, p13, p22;\n\taddc.u64 E3, p33, 0;\n\t"
        "add.cc.u64 E1, E1, p20;\n\taddc.cc.u64 E2, E2, p31;\n\taddc.u64 E3, E3, 0;\n\t"
        // odd class, 32 bits up: (O3:O2:O1:O0)
        "add.cc.u64 O0, p01, p10;\n\taddc.cc.u64 O1, p03, p12;\n\taddc.cc.u64 O2, p23, p32;\n\taddc.u64 O3, 0, 0;\n\t"
        "add.cc.u64 O1, O1, p21;\n\taddc.cc.u64 O2, O2, 0;\n\taddc.u64 O3, O3, 0;\n\t"
        "add.cc.u64 O1, O1, p30;\n\taddc.cc.u64 O2, O2, 0;\n\taddc.u64 O3, O3, 0;\n\t"
        "mov.b64 {t0, e0h}, p00;\n\tmov.b64 {e1l, e1h}, E1;\n\tmov.b64 {e2l, e2h}, E2;\n\tmov.b64 {e3l, e3h}, E3;\n\t"
        "mov.b64 {o0l, o0h}, O0;\n\tmov.b64 {o1l, o1h}, O1;\n\tmov.b64 {o2l, o2h}, O2;\n\tmov.b64 {o3l, o3h}, O3;\n\t"
        "add.cc.u32 t1, e0h, o0l;\n\taddc.cc.u32 t2, e1l, o0h;\n\taddc.cc.u32 t3, e1h, o1l;\n\taddc.cc.u32 t4, e2l, o1h;\n\t"
        "addc.cc.u32 t5, e2h, o2l;\n\taddc.cc.u32 t6, e3l, o2h;\n\taddc.u32 t7, e3h, o3l;\n\t"
        // G = H * 45 (5 limbs)
        "mul.wide.u32 w, t4, 45;\n\tmov.b64 {g0, wh}, w;\n\t"
        "mul.wide.u32 w, t5, 45;\n\tmov.b64 {wl, g2}, w;\n\tadd.cc.u32 g1, wl, wh;\n\t"
        "mul.wide.u32 w, t6, 45;\n\tmov.b64 {wl, g3}, w;\n\taddc.cc.u32 g2, g2, wl;\n\t"
        "mul.wide.u32 w, t7, 45;\n\tmov.b64 {wl, g4}, w;\n\taddc.cc.u32 g3, g3, wl;\n\taddc.u32 g4, g4, 0;\n\t"
        // M = G << 40: limbs 1..5
        "shl.b32 m1, g0, 8;\n\tshf.l.wrap.b32 m2, g0, g1, 8;\n\tshf.l.wrap.b32 m3, g1, g2, 8;\n\tshf.l.wrap.b32 m4, g2, g3, 8;\n\t"
        "shf.l.wrap.b32 m5, g3, g4, 8;\n\t"
        // U = L + M - H (6 limbs)
        "add.cc.u32 u1, t1, m1;\n\taddc.cc.u32 u2, t2, m2;\n\taddc.cc.u32 u3, t3, m3;\n\taddc.cc.u32 u4, m4, 0;\n\taddc.u32 u5, m5, 0;\n\t"
        "sub.cc.u32 u0, t0, t4;\n\tsubc.cc.u32 u1, u1, t5;\n\tsubc.cc.u32 u2, u2, t6;\n\tsubc.cc.u32 u3, u3, t7;\n\t"
        "subc.cc.u32 u4, u4, 0;\n\tsubc.u32 u5, u5, 0;\n\t"
        // second fold: H2 = (u5:u4) < 2^46; G2 = H2 * 45 < 2^52 = (g1:g0); N = G2 << 40 = limbs 1, 2
        "mul.wide.u32 w, u4, 45;\n\tmov.b64 {g0, wh}, w;\n\tmad.lo.u32 g1, u5, 45, wh;\n\t"
        "shl.b32 m1, g0, 8;\n\tshf.l.wrap.b32 m2, g0, g1, 8;\n\t"
        "add.cc.u32 u1, u1, m1;\n\taddc.cc.u32 u2, u2, m2;\n\taddc.cc.u32 u3, u3, 0;\n\taddc.u32 c3, 0, 0;\n\t"
        "sub.cc.u32 u0, u0, u4;\n\tsubc.cc.u32 u1, u1, u5;\n\tsubc.cc.u32 u2, u2, 0;\n\tsubc.cc.u32 u3, u3, 0;\n\tsubc.u32 c3, c3, 0;\n\t"
        // third fold: + K under the carry
        "neg.s32 msk, c3;\n\tand.b32 k1, msk, 0x2CFF;\n\t"
        "add.cc.u32 u0, u0, msk;\n\taddc.cc.u32 u1, u1, k1;\n\taddc.cc.u32 u2, u2, 0;\n\taddc.u32 u3, u3, 0;\n\t"
        // canonical: - q when >= q
        "add.cc.u32 s0, u0, 0xFFFFFFFF;\n\taddc.cc.u32 s1, u1, 0x2CFF;\n\taddc.cc.u32 s2, u2, 0;\n\taddc.cc.u32 s3, u3, 0;\n\taddc.u32 c3, 0, 0;\n\t"
        "setp.ne.u32 p, c3, 0;\n\t"
        "selp.u32 u0, s0, u0, p;\n\tselp.u32 u1, s1, u1, p;\n\tselp.u32 u2, s2, u2, p;\n\tselp.u32 u3, s3, u3, p;\n\t"
        "mov.b64 %0, {u0, u1};\n\tmov.b64 %1, {u2, u3};\n\t}"
        : "=l"(r.lo), "=l"(r.hi)
        : "l"(a.lo), "l"(a.hi), "l"(b.lo), "l"(b.hi));
    return r;
#else
    // 256-bit schoolbook product w3..w0 (64-bit words)
    uint64_t h00, l00, h01, l01, h10, l10, h11, l11;
    mul64wide(a.lo, b.lo, h00, l00);
    mul64wide(a.lo, b.hi, h01, l01);
    mul64wide(a.hi, b.lo, h10, l10);
    mul64wide(a.hi, b.hi, h11, l11);
    uint64_t w0 = l00;
    uint64_t w1 = h00 + l01;
    uint64_t c1 = w1 < h00 ? 1 : 0;
    w1 += l10;
    c1 += w1 < l10 ? 1 : 0;
    uint64_t w2 = h01 + h10;
    uint64_t c2 = w2 < h01 ? 1 : 0;
    w2 += l11;
    c2 += w2 < l11 ? 1 : 0;
    w2 += c1;
    c2 += w2 < c1 ? 1 : 0;
    uint64_t w3 = h11 + c2;
    // value = (w3:w2)*2^128 + (w1:w0) == (w3:w2)*K + (w1:w0); K < 2^46
    // (w3:w2)*K = w2*K + (w3*K << 64)
    uint64_t p2h, p2l, p3h, p3l;
    mul64wide(w2, K, p2h, p2l);
    mul64wide(w3, K, p3h, p3l);
    // t = (w1:w0) + (p2h:p2l) + (p3l << 64), overflow word o collects p3h and carries
    uint64_t c;
    T t = u128_add(T{w0, w1}, T{p2l, p2h}, c);
    uint64_t o = p3h + c;  // p3h < 2^46
    uint64_t th = t.hi + p3l;
    o += th < t.hi ? 1 : 0;
    t.hi = th;
    // second fold: o < 2^47, o*K < 2^93
    t = fold64(o, t, c);
    // third fold: c in {0,1}; if it wrapped the remainder is tiny, so adding K cannot wrap again
    if (c) {
      uint64_t c2b;
      t = u128_add(t, T{K, 0}, c2b);
    }
    return canon(t);
#endif
  }
  // w16^idx, w16 = ROOT^(2^36) (order 16); literals so the compiler folds them at unrolled call sites
  static WG_HD T root16(int idx) {
    switch (idx & 15) {
      case 1: return T{0xab1ac6d8062c4665ULL, 0x9c511064995809f9ULL};
      case 2: return T{0x21273cb06325d9a9ULL, 0xa2184c20306200b8ULL};
      case 3: return T{0xb0d002c22bc23d17ULL, 0x2033e32141f648cdULL};
      case 4: return T{0xbcb1099aefe38f3eULL, 0xab3b34a97d5ad72bULL};
      case 5: return T{0x5b00e519a0a78877ULL, 0x39e874e797e3f285ULL};
      case 6: return T{0x615afcf9833aa8e8ULL, 0xf6fdec9e92e80782ULL};
      case 7: return T{0x5490ae4d6cef124cULL, 0x984cbf8ca1f67c30ULL};
      case 8: return T{0xffffd30000000000ULL, 0xffffffffffffffffULL};
      case 9: return T{0x54e50c27f9d3b99cULL, 0x63aeef9b66a7f606ULL};
      case 10: return T{0xded8964f9cda2658ULL, 0x5de7b3dfcf9dff47ULL};
      case 11: return T{0x4f2fd03dd43dc2eaULL, 0xdfcc1cdebe09b732ULL};
      case 12: return T{0x434ec965101c70c3ULL, 0x54c4cb5682a528d4ULL};
      case 13: return T{0xa4feede65f58778aULL, 0xc6178b18681c0d7aULL};
      case 14: return T{0x9ea4d6067cc55719ULL, 0x090213616d17f87dULL};
      case 15: return T{0xab6f24b29310edb5ULL, 0x67b340735e0983cfULL};
      default: return T{1, 0};
    }
  }
  template <bool INV>
  static WG_HD T mul_root16(T x, int idx) {
    idx &= 15;
    if (idx == 0) return x;
    if (idx == 8) return neg(x);
    return mul(x, root16(INV ? 16 - idx : idx));
  }
  template <bool INV>
  static WG_HD T sub_mul_root16(T x, T y, int idx) {
    idx &= 15;
    if (idx == 8) return sub(y, x);
    return mul_root16<INV>(sub(x, y), idx);
  }
};

// ------------------------------------------------------------------------------------------------ generic helpers
template <class F>
WG_HD typename F::T f_pow(typename F::T b, uint64_t e) {
  typename F::T r = F::one();
  while (e) {
    if (e & 1) r = F::mul(r, b);
    b = F::mul(b, b);
    e >>= 1;
  }
  return r;
}
// a^(p-2); only used on the host for n^-1 and inverse roots
inline uint64_t gl64_inv(uint64_t a) { return f_pow<Gl64>(a, Gl64::P - 2); }
inline u128 f128_inv(u128 a) {
  // exponent q-2 = (M_HI:M_LO) - 2, square-and-multiply over 128 bits
  u128 e = u128_sub(F128::modulus(), u128{2, 0});
  u128 r = F128::one(), b = a;
  for (int i = 0; i < 128; i++) {
    uint64_t bit = i < 64 ? (e.lo >> i) & 1 : (e.hi >> (i - 64)) & 1;
    if (bit) r = F128::mul(r, b);
    b = F128::mul(b, b);
  }
  return r;
}
template <class F>
inline typename F::T f_inv(typename F::T a);
template <>
inline uint64_t f_inv<Gl64>(uint64_t a) { return gl64_inv(a); }
template <>
inline u128 f_inv<F128>(u128 a) { return f128_inv(a); }

// root of unity of order 2^log_n as a PLAIN field value (math/src/field/traits.rs:254-263)
template <class F>
inline typename F::T f_root_of_unity(int log_n);
template <>
inline uint64_t f_root_of_unity<Gl64>(int log_n) {
  uint64_t r = Gl64::to_mont(Gl64::ROOT_CANON);
  for (int i = 0; i < Gl64::kTwoAdicity - log_n; i++) r = Gl64::mul(r, r);
  return r;
}
template <>
inline u128 f_root_of_unity<F128>(int log_n) {
  u128 r{F128::ROOT_LO, F128::ROOT_HI};
  for (int i = 0; i < F128::kTwoAdicity - log_n; i++) r = F128::mul(r, r);
  return r;
}

}  // namespace wg
