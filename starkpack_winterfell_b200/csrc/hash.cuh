// BLAKE3 (default mode, 32-byte output) and SHA3-256 as host+device inline functions.
//
// The reference takes both from crates that are not in its tree (blake3 ^1.3, sha3 ^0.10;
// crypto/Cargo.toml:33-35) and calls them from crypto/src/hash/blake/mod.rs:27-59 (hash / merge /
// hash_elements) and crypto/src/hash/sha/mod.rs:24-56. These are written from the public specifications.
// Input is presented through a word loader so the same code hashes a staged shared-memory row, a global
// row or a 64-byte pair of digests. All inputs on this path are multiples of 8 bytes long.
#pragma once
#include <stdint.h>
#include "field.cuh"

namespace wg {

// Blake3_192 (crypto/src/hash/blake/mod.rs:68-114) is BLAKE3 truncated to its first 24 bytes; merge hashes the 48 bytes of
// two truncated digests.
enum { HASH_BLAKE3_256 = 0, HASH_SHA3_256 = 1, HASH_BLAKE3_192 = 2 };
WG_HD constexpr int hash_digest_bytes(int hash) { return hash == HASH_BLAKE3_192 ? 24 : 32; }

// ------------------------------------------------------------------------------------------------ BLAKE3
enum { B3_CHUNK_START = 1, B3_CHUNK_END = 2, B3_PARENT = 4, B3_ROOT = 8 };

WG_HD uint32_t rotr32(uint32_t x, int n) {
#if defined(__CUDA_ARCH__)
  return __funnelshift_r(x, x, n);
#else
  return (x >> n) | (x << (32 - n));
#endif
}

// 32-bit add routed to the FMA pipe on the device: `one` is an opaque run-time 1 (a kernel argument), so ptxas keeps
// a*one+b as IMAD instead of folding it back into IADD3. BLAKE3's G is 12 ALU-pipe instructions per call when written
// with plain adds (the ALU pipe is the measured bottleneck of the hash kernels, FMA pipe ~15 % busy); with the six
// adds as IMADs it is 8 ALU + 6 FMA.
WG_HD uint32_t b3_add(uint32_t a, uint32_t b, uint32_t one) {
#if defined(__CUDA_ARCH__)
  uint32_t r;
  asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(one), "r"(b));
  return r;
#else
  (void)one;
  return a + b;
#endif
}

#define WG_B3_G(a, b, c, d, mx, my)               \
  do {                                            \
    a = b3_add(b3_add(a, b, one), (mx), one);     \
    d = rotr32(d ^ a, 16);                        \
    c = b3_add(c, d, one);                        \
    b = rotr32(b ^ c, 12);                        \
    a = b3_add(b3_add(a, b, one), (my), one);     \
    d = rotr32(d ^ a, 8);                         \
    c = b3_add(c, d, one);                        \
    b = rotr32(b ^ c, 7);                         \
  } while (0)

#define WG_B3_ROUND(m0, m1, m2, m3, m4, m5, m6, m7, m8, m9, m10, m11, m12, m13, m14, m15) \
  do {                                                                                    \
    WG_B3_G(v0, v4, v8, v12, m0, m1);                                                     \
    WG_B3_G(v1, v5, v9, v13, m2, m3);                                                     \
    WG_B3_G(v2, v6, v10, v14, m4, m5);                                                    \
    WG_B3_G(v3, v7, v11, v15, m6, m7);                                                    \
    WG_B3_G(v0, v5, v10, v15, m8, m9);                                                    \
    WG_B3_G(v1, v6, v11, v12, m10, m11);                                                  \
    WG_B3_G(v2, v7, v8, v13, m12, m13);                                                   \
    WG_B3_G(v3, v4, v9, v14, m14, m15);                                                   \
  } while (0)

// cv <- first 8 words of compress(cv, m, counter, block_len, flags). The seven message schedules are the
// BLAKE3 permutation [2,6,3,10,7,0,4,13,1,11,12,5,9,14,15,8] applied repeatedly, written out so that the
// message words stay in fixed registers.
WG_HD void b3_compress(uint32_t cv[8], const uint32_t m[16], uint64_t counter, uint32_t block_len, uint32_t flags,
                        uint32_t one = 1) {
  uint32_t v0 = cv[0], v1 = cv[1], v2 = cv[2], v3 = cv[3], v4 = cv[4], v5 = cv[5], v6 = cv[6], v7 = cv[7];
  uint32_t v8 = 0x6A09E667u, v9 = 0xBB67AE85u, v10 = 0x3C6EF372u, v11 = 0xA54FF53Au;
  uint32_t v12 = (uint32_t)counter, v13 = (uint32_t)(counter >> 32), v14 = block_len, v15 = flags;
  WG_B3_ROUND(m[0], m[1], m[2], m[3], m[4], m[5], m[6], m[7], m[8], m[9], m[10], m[11], m[12], m[13], m[14], m[15]);
  WG_B3_ROUND(m[2], m[6], m[3], m[10], m[7], m[0], m[4], m[13], m[1], m[11], m[12], m[5], m[9], m[14], m[15], m[8]);
  WG_B3_ROUND(m[3], m[4], m[10], m[12], m[13], m[2], m[7], m[14], m[6], m[5], m[9], m[0], m[11], m[15], m[8], m[1]);
  WG_B3_ROUND(m[10], m[7], m[12], m[9], m[14], m[3], m[13], m[15], m[4], m[0], m[11], m[2], m[5], m[8], m[1], m[6]);
  WG_B3_ROUND(m[12], m[13], m[9], m[11], m[15], m[10], m[14], m[8], m[7], m[2], m[5], m[3], m[0], m[1], m[6], m[4]);
  WG_B3_ROUND(m[9], m[14], m[11], m[5], m[8], m[12], m[15], m[1], m[13], m[3], m[0], m[10], m[2], m[6], m[4], m[7]);
  WG_B3_ROUND(m[11], m[15], m[5], m[0], m[1], m[9], m[8], m[6], m[14], m[10], m[2], m[12], m[3], m[4], m[7], m[13]);
  cv[0] = v0 ^ v8;
  cv[1] = v1 ^ v9;
  cv[2] = v2 ^ v10;
  cv[3] = v3 ^ v11;
  cv[4] = v4 ^ v12;
  cv[5] = v5 ^ v13;
  cv[6] = v6 ^ v14;
  cv[7] = v7 ^ v15;
}

WG_HD void b3_iv(uint32_t cv[8]) {
  cv[0] = 0x6A09E667u;
  cv[1] = 0xBB67AE85u;
  cv[2] = 0x3C6EF372u;
  cv[3] = 0xA54FF53Au;
  cv[4] = 0x510E527Fu;
  cv[5] = 0x9B05688Cu;
  cv[6] = 0x1F83D9ABu;
  cv[7] = 0x5BE0CD19u;
}

// parent node: compress(IV, left || right, 0, 64, PARENT | extra)
WG_HD void b3_parent(const uint32_t left[8], uint32_t right_inout[8], uint32_t extra_flags, uint32_t one = 1) {
  uint32_t m[16];
#pragma unroll
  for (int i = 0; i < 8; i++) {
    m[i] = left[i];
    m[8 + i] = right_inout[i];
  }
  b3_iv(right_inout);
  b3_compress(right_inout, m, 0, 64, B3_PARENT | extra_flags, one);
}

// Hash merge: hash(left32 || right32), a single 64-byte block (Hasher::merge, blake/mod.rs:31-33).
WG_HD void b3_merge(const uint32_t in16[16], uint32_t out[8], uint32_t one = 1) {
  b3_iv(out);
  b3_compress(out, in16, 0, 64, B3_CHUNK_START | B3_CHUNK_END | B3_ROOT, one);
}

// Blake3_192 merge: hash(left24 || right24), one 48-byte block; the caller keeps the first 6 output words
WG_HD void b3_merge48(const uint32_t in12[12], uint32_t out[8], uint32_t one = 1) {
  uint32_t m[16];
#pragma unroll
  for (int i = 0; i < 12; i++) m[i] = in12[i];
#pragma unroll
  for (int i = 12; i < 16; i++) m[i] = 0;
  b3_iv(out);
  b3_compress(out, m, 0, 48, B3_CHUNK_START | B3_CHUNK_END | B3_ROOT, one);
}

// Streaming BLAKE3 over an input of known length: feed() takes consecutive 64-byte blocks (zero padded).
// Single-chunk inputs (<= 1024 B: every config except 2^24 x 256) never touch the cv stack.
constexpr int B3_MAX_STACK = 12;  // rows up to 4 MiB
template <bool MULTI>
struct B3StreamT {
  uint32_t cv[8];
  uint32_t stack[MULTI ? B3_MAX_STACK : 1][8];  // single-chunk inputs never touch it
  int sp;
  uint32_t nbytes, pos, nchunks, one;
  WG_HD void init(uint32_t nbytes_, uint32_t one_ = 1) {
    one = one_;
    nbytes = nbytes_;
    pos = 0;
    sp = 0;
    nchunks = nbytes <= 1024 ? 1 : (nbytes + 1023) / 1024;
  }
  WG_HD bool done() const { return pos > 0 && pos >= nbytes; }
  WG_HD void feed(const uint32_t m[16]) {
    const uint32_t c = pos >> 10, off = pos & 1023;
    const uint32_t chunk_bytes = (nbytes - c * 1024) < 1024 ? (nbytes - c * 1024) : 1024;
    const bool last_in_chunk = off + 64 >= chunk_bytes;
    const uint32_t blen = last_in_chunk ? chunk_bytes - off : 64;
    uint32_t flags = (off == 0 ? B3_CHUNK_START : 0) | (last_in_chunk ? B3_CHUNK_END : 0);
    if (nchunks == 1 && last_in_chunk) flags |= B3_ROOT;
    if (off == 0) b3_iv(cv);
    b3_compress(cv, m, c, blen, flags, one);
    pos += 64;
    if (MULTI && last_in_chunk && nchunks > 1) {
      if (c + 1 < nchunks) {
        uint32_t total = c + 1;
        while ((total & 1) == 0) {
          sp--;
          b3_parent(stack[sp], cv, 0, one);
          total >>= 1;
        }
        for (int i = 0; i < 8; i++) stack[sp][i] = cv[i];
        sp++;
        pos = (c + 1) << 10;
      } else {
        while (sp > 0) {
          sp--;
          b3_parent(stack[sp], cv, sp == 0 ? B3_ROOT : 0, one);
        }
        pos = nbytes;
      }
    }
  }
};

typedef B3StreamT<true> B3Stream;

// Digest of `nbytes` (multiple of 4) delivered by ld(word_index) -> uint32 (little-endian words).
template <class LD>
WG_HD void b3_hash_words(LD ld, uint32_t nbytes, uint32_t out[8]) {
  B3StreamT<true> st;
  st.init(nbytes);
  const uint32_t nwords = nbytes / 4;
  uint32_t w0 = 0;
  do {
    uint32_t m[16];
#pragma unroll
    for (int i = 0; i < 16; i++) m[i] = (w0 + i) < nwords ? ld(w0 + i) : 0u;
    st.feed(m);
    w0 += 16;
  } while (w0 < nwords);
#pragma unroll
  for (int i = 0; i < 8; i++) out[i] = st.cv[i];
}

// ------------------------------------------------------------------------------------------------ SHA3-256
WG_HD uint64_t rotl64(uint64_t x, int n) { return n == 0 ? x : (x << n) | (x >> (64 - n)); }

WG_HD void keccak_f1600(uint64_t s[25]) {
  const uint64_t RC[24] = {0x0000000000000001ULL, 0x0000000000008082ULL, 0x800000000000808aULL, 0x8000000080008000ULL,
                           0x000000000000808bULL, 0x0000000080000001ULL, 0x8000000080008081ULL, 0x8000000000008009ULL,
                           0x000000000000008aULL, 0x0000000000000088ULL, 0x0000000080008009ULL, 0x000000008000000aULL,
                           0x000000008000808bULL, 0x800000000000008bULL, 0x8000000000008089ULL, 0x8000000000008003ULL,
                           0x8000000000008002ULL, 0x8000000000000080ULL, 0x000000000000800aULL, 0x800000008000000aULL,
                           0x8000000080008081ULL, 0x8000000000008080ULL, 0x0000000080000001ULL, 0x8000000080008008ULL};
#pragma unroll
  for (int round = 0; round < 24; round++) {
    uint64_t c0 = s[0] ^ s[5] ^ s[10] ^ s[15] ^ s[20];
    uint64_t c1 = s[1] ^ s[6] ^ s[11] ^ s[16] ^ s[21];
    uint64_t c2 = s[2] ^ s[7] ^ s[12] ^ s[17] ^ s[22];
    uint64_t c3 = s[3] ^ s[8] ^ s[13] ^ s[18] ^ s[23];
    uint64_t c4 = s[4] ^ s[9] ^ s[14] ^ s[19] ^ s[24];
    uint64_t d0 = c4 ^ rotl64(c1, 1), d1 = c0 ^ rotl64(c2, 1), d2 = c1 ^ rotl64(c3, 1), d3 = c2 ^ rotl64(c4, 1),
             d4 = c3 ^ rotl64(c0, 1);
    // theta + rho + pi: b[y][2x+3y] = rot(a[x][y] ^ d[x], r[x][y]); index = x + 5y
    uint64_t b0 = s[0] ^ d0;
    uint64_t b10 = rotl64(s[1] ^ d1, 1);
    uint64_t b20 = rotl64(s[2] ^ d2, 62);
    uint64_t b5 = rotl64(s[3] ^ d3, 28);
    uint64_t b15 = rotl64(s[4] ^ d4, 27);
    uint64_t b16 = rotl64(s[5] ^ d0, 36);
    uint64_t b1 = rotl64(s[6] ^ d1, 44);
    uint64_t b11 = rotl64(s[7] ^ d2, 6);
    uint64_t b21 = rotl64(s[8] ^ d3, 55);
    uint64_t b6 = rotl64(s[9] ^ d4, 20);
    uint64_t b7 = rotl64(s[10] ^ d0, 3);
    uint64_t b17 = rotl64(s[11] ^ d1, 10);
    uint64_t b2 = rotl64(s[12] ^ d2, 43);
    uint64_t b12 = rotl64(s[13] ^ d3, 25);
    uint64_t b22 = rotl64(s[14] ^ d4, 39);
    uint64_t b23 = rotl64(s[15] ^ d0, 41);
    uint64_t b8 = rotl64(s[16] ^ d1, 45);
    uint64_t b18 = rotl64(s[17] ^ d2, 15);
    uint64_t b3 = rotl64(s[18] ^ d3, 21);
    uint64_t b13 = rotl64(s[19] ^ d4, 8);
    uint64_t b14 = rotl64(s[20] ^ d0, 18);
    uint64_t b24 = rotl64(s[21] ^ d1, 2);
    uint64_t b9 = rotl64(s[22] ^ d2, 61);
    uint64_t b19 = rotl64(s[23] ^ d3, 56);
    uint64_t b4 = rotl64(s[24] ^ d4, 14);
    // chi
    s[0] = b0 ^ (~b1 & b2);
    s[1] = b1 ^ (~b2 & b3);
    s[2] = b2 ^ (~b3 & b4);
    s[3] = b3 ^ (~b4 & b0);
    s[4] = b4 ^ (~b0 & b1);
    s[5] = b5 ^ (~b6 & b7);
    s[6] = b6 ^ (~b7 & b8);
    s[7] = b7 ^ (~b8 & b9);
    s[8] = b8 ^ (~b9 & b5);
    s[9] = b9 ^ (~b5 & b6);
    s[10] = b10 ^ (~b11 & b12);
    s[11] = b11 ^ (~b12 & b13);
    s[12] = b12 ^ (~b13 & b14);
    s[13] = b13 ^ (~b14 & b10);
    s[14] = b14 ^ (~b10 & b11);
    s[15] = b15 ^ (~b16 & b17);
    s[16] = b16 ^ (~b17 & b18);
    s[17] = b17 ^ (~b18 & b19);
    s[18] = b18 ^ (~b19 & b15);
    s[19] = b19 ^ (~b15 & b16);
    s[20] = b20 ^ (~b21 & b22);
    s[21] = b21 ^ (~b22 & b23);
    s[22] = b22 ^ (~b23 & b24);
    s[23] = b23 ^ (~b24 & b20);
    s[24] = b24 ^ (~b20 & b21);
    s[0] ^= RC[round];
  }
}

// SHA3-256 of `nbytes` (multiple of 8) delivered as little-endian 64-bit lanes by ld64(lane_index).
template <class LD64>
WG_HD void sha3_256_lanes(LD64 ld64, uint32_t nbytes, uint32_t out[8]) {
  uint64_t s[25];
#pragma unroll
  for (int i = 0; i < 25; i++) s[i] = 0;
  const uint32_t nl = nbytes / 8;
  uint32_t pos = 0;
  while (nl - pos >= 17) {
#pragma unroll
    for (int i = 0; i < 17; i++) s[i] ^= ld64(pos + i);
    keccak_f1600(s);
    pos += 17;
  }
  const uint32_t rem = nl - pos;  // 0..16 lanes in the final block
#pragma unroll
  for (int i = 0; i < 17; i++) {
    uint64_t v = (uint32_t)i < rem ? ld64(pos + i) : 0ULL;
    if ((uint32_t)i == rem) v ^= 0x06ULL;
    if (i == 16) v ^= 0x8000000000000000ULL;
    s[i] ^= v;
  }
  keccak_f1600(s);
#pragma unroll
  for (int i = 0; i < 4; i++) {
    out[2 * i] = (uint32_t)s[i];
    out[2 * i + 1] = (uint32_t)(s[i] >> 32);
  }
}

// Generic front-end: digest of a word stream with either hash.
template <class LD>
WG_HD void hash_words(int hash, LD ld, uint32_t nbytes, uint32_t out[8]) {
  if (hash == HASH_BLAKE3_256 || hash == HASH_BLAKE3_192) {
    b3_hash_words(ld, nbytes, out);
  } else {
    sha3_256_lanes([&](uint32_t i) { return (uint64_t)ld(2 * i) | ((uint64_t)ld(2 * i + 1) << 32); }, nbytes, out);
  }
}

}  // namespace wg
