// Host-side test shim (NOT part of the product path): compiles field.cuh / hash.cuh with g++ so the CPU
// test-suite can check the exact arithmetic the kernels run against Python big ints / hashlib / blake3.
#include <string.h>
#include "field.cuh"
#include "hash.cuh"
using namespace wg;
extern "C" {
uint64_t wgt_gl(int op, uint64_t a, uint64_t b) {
  switch (op) {
    case 0: return Gl64::add(a, b);
    case 1: return Gl64::sub(a, b);
    case 2: return Gl64::mul(a, b);
    case 3: return Gl64::pow2_mod((int)b);
    case 4: return Gl64::as_int(a);
    case 5: return Gl64::to_mont(a);
    case 6: return Gl64::neg(a);
    case 7: return Gl64::mul_root16<false>(a, (int)b);
    case 8: return Gl64::mul_root16<true>(a, (int)b);
    case 9: return f_root_of_unity<Gl64>((int)a);
    default: return gl64_inv(a);
  }
}
void wgt_f128(int op, const uint64_t* a, const uint64_t* b, uint64_t* out) {
  u128 x{a[0], a[1]}, y{b[0], b[1]}, r;
  switch (op) {
    case 0: r = F128::add(x, y); break;
    case 1: r = F128::sub(x, y); break;
    case 2: r = F128::mul(x, y); break;
    case 3: r = F128::neg(x); break;
    case 4: r = F128::mul_root16<false>(x, (int)b[0]); break;
    case 5: r = F128::mul_root16<true>(x, (int)b[0]); break;
    case 6: r = f_root_of_unity<F128>((int)a[0]); break;
    default: r = f128_inv(x); break;
  }
  out[0] = r.lo;
  out[1] = r.hi;
}
void wgt_hash(int hash, const uint8_t* data, uint32_t nbytes, uint8_t* out32) {
  uint32_t o[8];
  hash_words(hash, [&](uint32_t i) { uint32_t w; memcpy(&w, data + 4 * (size_t)i, 4); return w; }, nbytes, o);
  memcpy(out32, o, 32);
}
void wgt_b3_merge(const uint8_t* in64, uint8_t* out32) {
  uint32_t m[16], o[8];
  memcpy(m, in64, 64);
  b3_merge(m, o);
  memcpy(out32, o, 32);
}
}
