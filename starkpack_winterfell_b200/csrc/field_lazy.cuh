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

  // a + b, a ANY u64, b <= p. If the 64-bit sum wraps, the true sum is s + 2^64 = s + eps (mod p), and s + eps cannot wrap
  // again: s = a + b - 2^64 <= a - eps, so s + eps <= a. (PTX's CC.CF is the hardware carry: after an add it is the
  // carry-out, and subc subtracts 1 - CF, so the carry word is read with addc / madc, never with subc.)
  static __device__ __forceinline__ T add(T a, T b) { return add_b(a, b); }
  // the same with the carry word and its mask on the FMA pipe: 3 ALU + 3 FMA
  static __device__ __forceinline__ T add_b(T a, T b) {
    uint32_t a0 = (uint32_t)a, a1 = (uint32_t)(a >> 32), b0 = (uint32_t)b, b1 = (uint32_t)(b >> 32), r0, r1;
    asm("{\n\t.reg .u32 c, m;\n\t"
        "add.cc.u32 %0, %2, %4;\n\taddc.cc.u32 %1, %3, %5;\n\tmadc.lo.u32 c, %7, %7, 0;\n\t"
        "mul.lo.u32 m, c, %8;\n\t"
        "add.cc.u32 %0, %0, m;\n\t"
        "madc.lo.u32 %1, %1, %6, 0;\n\t}"
        : "=&r"(r0), "=&r"(r1)
        : "r"(a0), "r"(a1), "r"(b0), "r"(b1), "r"(K_ONE), "r"(K_ZERO), "r"(K_MONE));
    return ((uint64_t)r1 << 32) | r0;
  }
  // a - b, a ANY u64, b <= p: Gl64::sub's sequence is already the lazy form (borrow -> + p, which cannot wrap for b <= p)
  static __device__ __forceinline__ T sub(T a, T b) { return Gl64::sub(a, b); }
  static __device__ __forceinline__ T sub_b(T a, T b) { return Gl64::sub(a, b); }
  // any u64 -> [0,p): a >= p exactly when t = a + eps wraps, and then a - p = t (mod 2^64)
  static __device__ __forceinline__ T canon(T a) {
    uint32_t a0 = (uint32_t)a, a1 = (uint32_t)(a >> 32), r0, r1;
    asm("{\n\t.reg .u32 t0, t1, c;\n\t.reg .pred p;\n\t"
        "add.cc.u32 t0, %2, 0xFFFFFFFF;\n\taddc.cc.u32 t1, %3, 0;\n\taddc.u32 c, 0, 0;\n\t"
        "setp.ne.u32 p, c, 0;\n\t"
        "selp.u32 %0, t0, %2, p;\n\tselp.u32 %1, t1, %3, p;\n\t}"
        : "=r"(r0), "=r"(r1)
        : "r"(a0), "r"(a1));
    return ((uint64_t)r1 << 32) | r0;
  }
  // products and power-of-two shifts accept ANY u64 first operand (x*w < 2^64 * p keeps the Montgomery reduction exact,
  // its high word xh <= p - 2 so xh + a1 + e cannot overflow) and return canonical values
  static __device__ __forceinline__ T mul_w(T a, T w) { return Gl64::mul(a, w); }
  static __device__ __forceinline__ T mul_w2(T a, T w) {
    // 64x64 product by the compiler (mul.lo.u64 / mul.hi.u64 share their partial products: 4 IMAD.WIDE + carries)
    const unsigned __int128 x = (unsigned __int128)a * w;
    return Gl64::mont_red((uint64_t)(x >> 64), (uint64_t)x);
  }
  template <int S>
  static __device__ __forceinline__ T mont_shift(T x) { return Gl64::mont_shift<S>(x); }

  // |x * w16^(+-idx)| up to the sign Gl64::root16_neg reports (x any u64, result canonical)
  template <bool INV, int IDX>
  static __device__ __forceinline__ T mul_root16_abs(T x) {
    constexpr int e = Gl64::root16_e<INV>(IDX);
    constexpr int sh = (e >= 96 && e < 160) ? e - 96 : e;
    if constexpr (sh > 0 && sh < 64) {
      return Gl64::mont_shift<sh>(x);
    } else {
      return Gl64::mul(x, Gl64::pow2_mod(e));
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
