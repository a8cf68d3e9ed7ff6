// Lazily reduced Goldilocks arithmetic for the register butterflies (device only).
//
// Gl64 (field.cuh) keeps every value canonical in [0,p): a modular add is 9 SASS instructions, a sub 6. Inside a
// radix-16 butterfly most of those corrections are unnecessary: 2^64 = eps (mod p) with eps = 2^32 - 1, so a 64-bit
// wrap-around is repaired by adding eps, and a value may stay anywhere in [0, 2^64) ("lazy") as long as the operation
// that consumes it tolerates that. The rules used here (proved in the comments of each function):
//     add(a, b), sub(a, b): a ANY u64, b <= p            -> result any u64, congruent mod p
//     mul(a, w), mont_shift(a): a ANY u64, w < p         -> result canonical (< p)
//     canon(a): any u64 -> [0,p)
// A decimation-in-time butterfly (t = w*y; x + t, x - t) meets them by construction: the second operand of every
// add/sub is a fresh product. Only butterflies with a trivial twiddle take their y as it is, and those canonicalise it
// first (7 of the 32 butterflies of a radix-16 network; the 8 of the first level read canonical inputs).
// The reference's arithmetic (math/src/field/f64/mod.rs:314-366) is exact, so the lazily reduced network yields the
// same words after the final canonicalisation.
#pragma once
#include "field.cuh"

#if defined(__CUDACC__)
namespace wg {

static __constant__ uint32_t K_ZERO = 0u;

struct GlLazy {
  typedef uint64_t T;
  static constexpr uint64_t P = Gl64::P;

  // a + b. If the 64-bit sum wraps, the true sum is s + 2^64 = s + eps (mod p); s + eps cannot wrap again because
  // s = a + b - 2^64 <= a - 2^32 + 1 - ... : with b <= p = 2^64 - eps, s + eps <= a < 2^64.
  static __device__ __forceinline__ T add(T a, T b) {
    uint32_t a0 = (uint32_t)a, a1 = (uint32_t)(a >> 32), b0 = (uint32_t)b, b1 = (uint32_t)(b >> 32);
    uint64_t r;
    asm("{\n\t.reg .u32 s0, s1, c;\n\t.reg .u64 s;\n\t"
        "add.cc.u32 s0, %1, %3;\n\taddc.cc.u32 s1, %2, %4;\n\taddc.u32 c, 0, 0;\n\t"
        "mov.b64 s, {s0, s1};\n\t"
        "mad.wide.u32 %0, c, %5, s;\n\t}"
        : "=l"(r)
        : "r"(a0), "r"(a1), "r"(b0), "r"(b1), "r"(K_MONE));
    return r;
  }
  // the same with the carry read on the FMA pipe (IMAD.X) instead of the ALU pipe
  static __device__ __forceinline__ T add_b(T a, T b) {
    uint32_t a0 = (uint32_t)a, a1 = (uint32_t)(a >> 32), b0 = (uint32_t)b, b1 = (uint32_t)(b >> 32);
    uint64_t r;
    asm("{\n\t.reg .u32 s0, s1, c;\n\t.reg .u64 s;\n\t"
        "add.cc.u32 s0, %1, %3;\n\taddc.cc.u32 s1, %2, %4;\n\tmadc.lo.u32 c, %5, %5, 0;\n\t"
        "mov.b64 s, {s0, s1};\n\t"
        "mad.wide.u32 %0, c, %6, s;\n\t}"
        : "=l"(r)
        : "r"(a0), "r"(a1), "r"(b0), "r"(b1), "r"(K_ZERO), "r"(K_MONE));
    return r;
  }
  // a - b. On a borrow the true difference is d - 2^64 = d - eps (mod p), i.e. d + 0xFFFFFFFF00000001 (mod 2^64); with the
  // borrow mask m = 0xFFFFFFFF that is d + m*0xFFFFFFFF (= 0xFFFFFFFE00000001) + 2^32. d - eps cannot wrap below zero:
  // d = a - b + 2^64 >= 2^64 - p = eps for b <= p.
  static __device__ __forceinline__ T sub(T a, T b) {
    uint32_t a0 = (uint32_t)a, a1 = (uint32_t)(a >> 32), b0 = (uint32_t)b, b1 = (uint32_t)(b >> 32);
    uint64_t r;
    asm("{\n\t.reg .u32 d0, d1, m;\n\t.reg .u64 d;\n\t"
        "sub.cc.u32 d0, %1, %3;\n\tsubc.cc.u32 d1, %2, %4;\n\tsubc.u32 m, 0, 0;\n\t"
        "sub.u32 d1, d1, m;\n\t"
        "mov.b64 d, {d0, d1};\n\t"
        "mad.wide.u32 %0, m, %5, d;\n\t}"
        : "=l"(r)
        : "r"(a0), "r"(a1), "r"(b0), "r"(b1), "r"(K_MONE));
    return r;
  }
  static __device__ __forceinline__ T sub_b(T a, T b) {
    uint32_t a0 = (uint32_t)a, a1 = (uint32_t)(a >> 32), b0 = (uint32_t)b, b1 = (uint32_t)(b >> 32);
    uint64_t r;
    asm("{\n\t.reg .u32 d0, d1, m;\n\t.reg .u64 d;\n\t"
        "sub.cc.u32 d0, %1, %3;\n\tsubc.cc.u32 d1, %2, %4;\n\tsubc.u32 m, 0, 0;\n\t"
        "mad.lo.u32 d1, m, %5, d1;\n\t"
        "mov.b64 d, {d0, d1};\n\t"
        "mad.wide.u32 %0, m, %6, d;\n\t}"
        : "=l"(r)
        : "r"(a0), "r"(a1), "r"(b0), "r"(b1), "r"(K_MONE), "r"(K_MONE));
    return r;
  }
  // any u64 -> [0,p): a >= p exactly when a + eps wraps
  static __device__ __forceinline__ T canon(T a) {
    uint32_t a0 = (uint32_t)a, a1 = (uint32_t)(a >> 32);
    uint64_t r;
    asm("{\n\t.reg .u32 t, c;\n\t"
        "add.cc.u32 t, %1, 0xFFFFFFFF;\n\taddc.cc.u32 t, %2, 0;\n\taddc.u32 c, 0, 0;\n\t"
        "mad.wide.u32 %0, c, %4, %3;\n\t}"
        : "=l"(r)
        : "r"(a0), "r"(a1), "l"(a), "r"(K_MONE));
    return r;
  }

  // Montgomery product a*w*2^-64 mod p, a ANY u64, w < p; result in [0,p). The 128-bit product is accumulated through
  // IMAD.WIDE's 64-bit addend (no carry chain: each partial sum provably fits 64 bits), then the reference's reduction
  // (f64/mod.rs:651-661) regrouped as in Gl64::mont_red; xh <= p - 2 for a < 2^64, so xh + a1 + e < 2^64 still holds.
  static __device__ __forceinline__ T mul_w(T a, T w) {
    uint32_t a0 = (uint32_t)a, a1 = (uint32_t)(a >> 32), b0 = (uint32_t)w, b1 = (uint32_t)(w >> 32);
    uint64_t r;
    asm("{\n\t.reg .u64 t, v, x, u, s;\n\t.reg .u32 t0, t1, v0, v1, x0, x1, u0, u1, q, s0, s1, m;\n\t"
        "mul.wide.u32 t, %1, %3;\n\t"              // a0*b0
        "mov.b64 {t0, t1}, t;\n\t"
        "mul.wide.u32 v, %1, %4;\n\t"              // a0*b1
        "mad.wide.u32 v, t1, 1, v;\n\t"            // + t1            (< 2^64)
        "mov.b64 {v0, v1}, v;\n\t"
        "mul.wide.u32 x, %2, %3;\n\t"              // a1*b0
        "mad.wide.u32 x, v0, 1, x;\n\t"            // + v0            (< 2^64): x0 = limb 1
        "mov.b64 {x0, x1}, x;\n\t"
        "mul.wide.u32 u, %2, %4;\n\t"              // a1*b1
        "mad.wide.u32 u, v1, 1, u;\n\t"
        "mad.wide.u32 u, x1, 1, u;\n\t"            // (u1:u0) = limbs 3, 2
        "mov.b64 {u0, u1}, u;\n\t"
        // reduction: q = (x0 + t0) mod 2^32 with carry e; s = (u1:u0) + q + e - (q:t0); + p on borrow
        "add.cc.u32 q, x0, t0;\n\t"
        "addc.cc.u32 s0, u0, q;\n\taddc.u32 s1, u1, 0;\n\t"
        "sub.cc.u32 s0, s0, t0;\n\tsubc.cc.u32 s1, s1, q;\n\tsubc.u32 m, 0, 0;\n\t"
        "sub.u32 s1, s1, m;\n\t"
        "mov.b64 s, {s0, s1};\n\t"
        "mad.wide.u32 %0, m, %5, s;\n\t}"
        : "=l"(r)
        : "r"(a0), "r"(a1), "r"(b0), "r"(b1), "r"(K_MONE));
    return r;
  }
  // same product with the two middle accumulations on the ALU pipe (pipe balance)
  static __device__ __forceinline__ T mul_w2(T a, T w) {
    uint32_t a0 = (uint32_t)a, a1 = (uint32_t)(a >> 32), b0 = (uint32_t)w, b1 = (uint32_t)(w >> 32);
    uint64_t r;
    asm("{\n\t.reg .u64 t, v, x, u, s;\n\t.reg .u32 t0, t1, v0, v1, x0, x1, u0, u1, l1, h0, h1, q, s0, s1, m;\n\t"
        "mul.wide.u32 t, %1, %3;\n\tmul.wide.u32 v, %1, %4;\n\tmul.wide.u32 x, %2, %3;\n\tmul.wide.u32 u, %2, %4;\n\t"
        "mov.b64 {t0, t1}, t;\n\tmov.b64 {v0, v1}, v;\n\tmov.b64 {x0, x1}, x;\n\t"
        "mad.wide.u32 u, v1, 1, u;\n\t"
        "mad.wide.u32 u, x1, 1, u;\n\t"
        "mov.b64 {u0, u1}, u;\n\t"
        "add.cc.u32 l1, t1, v0;\n\taddc.cc.u32 h0, u0, 0;\n\taddc.u32 h1, u1, 0;\n\t"
        "add.cc.u32 l1, l1, x0;\n\taddc.cc.u32 h0, h0, 0;\n\taddc.u32 h1, h1, 0;\n\t"
        "add.cc.u32 q, l1, t0;\n\t"
        "addc.cc.u32 s0, h0, q;\n\taddc.u32 s1, h1, 0;\n\t"
        "sub.cc.u32 s0, s0, t0;\n\tsubc.cc.u32 s1, s1, q;\n\tsubc.u32 m, 0, 0;\n\t"
        "sub.u32 s1, s1, m;\n\t"
        "mov.b64 s, {s0, s1};\n\t"
        "mad.wide.u32 %0, m, %5, s;\n\t}"
        : "=l"(r)
        : "r"(a0), "r"(a1), "r"(b0), "r"(b1), "r"(K_MONE));
    return r;
  }

  // Montgomery reduction of x << S (Gl64::mont_shift) for ANY u64 x, canonical result; the final + p on borrow through
  // IMAD.WIDE as above.
  template <int S>
  static __device__ __forceinline__ T mont_shift(T x) {
    static_assert(S > 0 && S < 64, "shift out of range");
    uint32_t x0 = (uint32_t)x, x1 = (uint32_t)(x >> 32);
    uint64_t r;
    if (S < 32) {
      const uint32_t l0 = x0 << (S & 31), l1 = __funnelshift_l(x0, x1, S & 31), h0 = x1 >> ((32 - S) & 31);
      asm("{\n\t.reg .u32 q, s0, s1, m;\n\t.reg .u64 s;\n\t"
          "add.cc.u32 q, %2, %1;\n\t"
          "addc.cc.u32 s0, %3, q;\n\taddc.u32 s1, 0, 0;\n\t"
          "sub.cc.u32 s0, s0, %1;\n\tsubc.cc.u32 s1, s1, q;\n\tsubc.u32 m, 0, 0;\n\t"
          "sub.u32 s1, s1, m;\n\t"
          "mov.b64 s, {s0, s1};\n\t"
          "mad.wide.u32 %0, m, %4, s;\n\t}"
          : "=l"(r)
          : "r"(l0), "r"(l1), "r"(h0), "r"(K_MONE));
    } else {
      constexpr int Q = S - 32;
      const uint32_t l1 = Q ? x0 << (Q & 31) : x0, h0 = Q ? __funnelshift_l(x0, x1, Q & 31) : x1,
                     h1 = Q ? x1 >> ((32 - Q) & 31) : 0u;
      asm("{\n\t.reg .u32 s0, s1, m;\n\t.reg .u64 s;\n\t"
          "add.cc.u32 s0, %2, %1;\n\taddc.u32 s1, %3, 0;\n\t"
          "sub.cc.u32 s1, s1, %1;\n\tsubc.u32 m, 0, 0;\n\t"
          "sub.u32 s1, s1, m;\n\t"
          "mov.b64 s, {s0, s1};\n\t"
          "mad.wide.u32 %0, m, %4, s;\n\t}"
          : "=l"(r)
          : "r"(l1), "r"(h0), "r"(h1), "r"(K_MONE));
    }
    return r;
  }

  // |x * w16^(+-idx)| up to the sign Gl64::root16_neg reports (x any u64, result canonical)
  template <bool INV, int IDX>
  static __device__ __forceinline__ T mul_root16_abs(T x) {
    constexpr int e = Gl64::root16_e<INV>(IDX);
    constexpr int sh = (e >= 96 && e < 160) ? e - 96 : e;
    if constexpr (sh > 0 && sh < 64) {
      return mont_shift<sh>(x);
    } else {
      return mul_w(x, Gl64::pow2_mod(e));
    }
  }
};

// One decimation-in-time butterfly on registers: (x, y) -> (x + w*y, x - w*y), w = w16^(+-IDX).
template <bool INV, int IDX>
__device__ __forceinline__ void dit_bfly(uint64_t& x, uint64_t& y) {
  if constexpr ((IDX & 15) == 0) {
    const uint64_t t = y;  // caller guarantees t <= p
    y = GlLazy::sub(x, t);
    x = GlLazy::add(x, t);
  } else {
    const uint64_t t = GlLazy::mul_root16_abs<INV, (IDX & 15)>(y);
    if constexpr (Gl64::root16_neg<INV>(IDX & 15)) {  // t = -(w*y)
      y = GlLazy::add(x, t);
      x = GlLazy::sub(x, t);
    } else {
      y = GlLazy::sub(x, t);
      x = GlLazy::add(x, t);
    }
  }
}

template <int LR>
__host__ __device__ constexpr int bitrev_k(int x) {
  int r = 0;
  for (int i = 0; i < LR; i++) r |= ((x >> i) & 1) << (LR - 1 - i);
  return r;
}

// Natural-order in, natural-order out length-2^LR transform (LR <= 4) of registers a[]: X[k] = sum_n a[n] w^(+-nk),
// w of order 2^LR. Inputs must be <= p (canonical); outputs are lazy (any u64, congruent mod p).
template <int LR, bool INV, int HALF, int G, int K>
struct DitStage {
  static __device__ __forceinline__ void run(uint64_t (&b)[1 << LR]) {
    constexpr int R = 1 << LR;
    if constexpr (G < R) {
      if constexpr (K < HALF) {
        constexpr int IDX = K * (8 / HALF);  // w_{2*HALF}^K as a power of w16
        if constexpr (IDX == 0 && HALF > 1) b[G + K + HALF] = GlLazy::canon(b[G + K + HALF]);
        dit_bfly<INV, IDX>(b[G + K], b[G + K + HALF]);
        DitStage<LR, INV, HALF, G, K + 1>::run(b);
      } else {
        DitStage<LR, INV, HALF, G + 2 * HALF, 0>::run(b);
      }
    } else if constexpr (2 * HALF < R) {
      DitStage<LR, INV, 2 * HALF, 0, 0>::run(b);
    }
  }
};

template <int LR, bool INV>
__device__ __forceinline__ void radix_dit_lazy(uint64_t (&a)[1 << LR]) {
  constexpr int R = 1 << LR;
  if constexpr (LR == 0) return;
  uint64_t b[R];
#pragma unroll
  for (int i = 0; i < R; i++) b[bitrev_k<LR>(i)] = a[i];
  DitStage<LR, INV, 1, 0, 0>::run(b);
#pragma unroll
  for (int i = 0; i < R; i++) a[i] = b[i];
}

}  // namespace wg
#endif
