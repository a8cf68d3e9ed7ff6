/* ORACLE — TEST INFRASTRUCTURE ONLY. Never linked into or called by the product path.
 *
 * Portable BLAKE3 (default hash mode, 32-byte output) and SHA3-256 (FIPS 202).
 * The reference takes both from third-party crates that are NOT under /root/reference:
 *   blake3 ^1.3 (default-features = false)  crypto/Cargo.toml:33   call sites crypto/src/hash/blake/mod.rs:27-59
 *   sha3   ^0.10                            crypto/Cargo.toml:35   call sites crypto/src/hash/sha/mod.rs:24-56
 * (no Cargo.lock in the reference => versions unpinned). Restated here from the published
 * specifications; pinned in tests/test_oracle_hash.py against the independent Python `blake3`
 * module and hashlib.sha3_256 on lengths 0..4200 (covers empty, partial block, exact block, exact
 * chunk, 2-chunk, 3-chunk, 4-chunk+ tree shapes) and against the spec's empty-input digests.
 */
#ifndef ORACLE_HASH_H
#define ORACLE_HASH_H
#include <stdint.h>
#include <string.h>
#include <stddef.h>

/* ------------------------------------------------------------------ BLAKE3 */
static const uint32_t B3_IV[8] = {0x6A09E667u, 0xBB67AE85u, 0x3C6EF372u, 0xA54FF53Au,
                                  0x510E527Fu, 0x9B05688Cu, 0x1F83D9ABu, 0x5BE0CD19u};
static const uint8_t B3_PERM[16] = {2, 6, 3, 10, 7, 0, 4, 13, 1, 11, 12, 5, 9, 14, 15, 8};
enum { B3_CHUNK_START = 1, B3_CHUNK_END = 2, B3_PARENT = 4, B3_ROOT = 8 };

static inline uint32_t b3_rotr(uint32_t x, int n) { return (x >> n) | (x << (32 - n)); }
#define B3_G(a, b, c, d, mx, my) \
  do {                           \
    a = a + b + (mx);            \
    d = b3_rotr(d ^ a, 16);      \
    c = c + d;                   \
    b = b3_rotr(b ^ c, 12);      \
    a = a + b + (my);            \
    d = b3_rotr(d ^ a, 8);       \
    c = c + d;                   \
    b = b3_rotr(b ^ c, 7);       \
  } while (0)

/* out[0..8) = first 8 words of the compression output (chaining value / root bytes) */
static inline void b3_compress(const uint32_t cv[8], const uint32_t block[16], uint64_t counter,
                               uint32_t block_len, uint32_t flags, uint32_t out[8]) {
  uint32_t v[16], m[16], t[16];
  for (int i = 0; i < 8; i++) v[i] = cv[i];
  for (int i = 0; i < 4; i++) v[8 + i] = B3_IV[i];
  v[12] = (uint32_t)counter;
  v[13] = (uint32_t)(counter >> 32);
  v[14] = block_len;
  v[15] = flags;
  for (int i = 0; i < 16; i++) m[i] = block[i];
  for (int r = 0; r < 7; r++) {
    B3_G(v[0], v[4], v[8], v[12], m[0], m[1]);
    B3_G(v[1], v[5], v[9], v[13], m[2], m[3]);
    B3_G(v[2], v[6], v[10], v[14], m[4], m[5]);
    B3_G(v[3], v[7], v[11], v[15], m[6], m[7]);
    B3_G(v[0], v[5], v[10], v[15], m[8], m[9]);
    B3_G(v[1], v[6], v[11], v[12], m[10], m[11]);
    B3_G(v[2], v[7], v[8], v[13], m[12], m[13]);
    B3_G(v[3], v[4], v[9], v[14], m[14], m[15]);
    for (int i = 0; i < 16; i++) t[i] = m[B3_PERM[i]];
    for (int i = 0; i < 16; i++) m[i] = t[i];
  }
  for (int i = 0; i < 8; i++) out[i] = v[i] ^ v[i + 8];
}

/* chaining value of one chunk (<= 1024 bytes); root_flag = B3_ROOT when it is the only chunk */
static inline void b3_chunk_cv(const uint8_t* in, size_t len, uint64_t chunk_idx, uint32_t root_flag,
                               uint32_t cv_out[8]) {
  uint32_t cv[8];
  memcpy(cv, B3_IV, 32);
  size_t nblocks = len == 0 ? 1 : (len + 63) / 64;
  for (size_t b = 0; b < nblocks; b++) {
    uint8_t buf[64];
    size_t off = b * 64, bl = len - off < 64 ? len - off : 64;
    memset(buf, 0, 64);
    memcpy(buf, in + off, bl);
    uint32_t w[16];
    memcpy(w, buf, 64); /* little-endian host */
    uint32_t fl = 0;
    if (b == 0) fl |= B3_CHUNK_START;
    if (b == nblocks - 1) fl |= B3_CHUNK_END | root_flag;
    b3_compress(cv, w, chunk_idx, (uint32_t)bl, fl, cv);
  }
  memcpy(cv_out, cv, 32);
}

static inline void b3_parent_cv(const uint32_t l[8], const uint32_t r[8], uint32_t root_flag, uint32_t out[8]) {
  uint32_t w[16];
  memcpy(w, l, 32);
  memcpy(w + 8, r, 32);
  b3_compress(B3_IV, w, 0, 64, B3_PARENT | root_flag, out);
}

/* recursive tree hash over whole chunks: left subtree = largest power of two of chunks < total */
static void b3_subtree(const uint8_t* in, size_t len, uint64_t chunk0, uint32_t root_flag, uint32_t out[8]) {
  if (len <= 1024) {
    b3_chunk_cv(in, len, chunk0, root_flag, out);
    return;
  }
  size_t nchunks = (len + 1023) / 1024;
  size_t left = 1;
  while (left * 2 < nchunks) left *= 2;
  uint32_t l[8], r[8];
  b3_subtree(in, left * 1024, chunk0, 0, l);
  b3_subtree(in + left * 1024, len - left * 1024, chunk0 + left, 0, r);
  b3_parent_cv(l, r, root_flag, out);
}

static inline void oracle_blake3_256(const uint8_t* in, size_t len, uint8_t out[32]) {
  uint32_t cv[8];
  b3_subtree(in, len, 0, B3_ROOT, cv);
  memcpy(out, cv, 32);
}

/* ------------------------------------------------------------------ SHA3-256 */
static const uint64_t KECCAK_RC[24] = {
    0x0000000000000001ULL, 0x0000000000008082ULL, 0x800000000000808aULL, 0x8000000080008000ULL,
    0x000000000000808bULL, 0x0000000080000001ULL, 0x8000000080008081ULL, 0x8000000000008009ULL,
    0x000000000000008aULL, 0x0000000000000088ULL, 0x0000000080008009ULL, 0x000000008000000aULL,
    0x000000008000808bULL, 0x800000000000008bULL, 0x8000000000008089ULL, 0x8000000000008003ULL,
    0x8000000000008002ULL, 0x8000000000000080ULL, 0x000000000000800aULL, 0x800000008000000aULL,
    0x8000000080008081ULL, 0x8000000000008080ULL, 0x0000000080000001ULL, 0x8000000080008008ULL};
static const int KECCAK_ROT[25] = {0,  1,  62, 28, 27, 36, 44, 6,  55, 20, 3,  10, 43,
                                   25, 39, 41, 45, 15, 21, 8,  18, 2,  61, 56, 14};
static inline uint64_t k_rotl(uint64_t x, int n) { return n ? (x << n) | (x >> (64 - n)) : x; }

static inline void keccak_f1600(uint64_t a[25]) {
  for (int rnd = 0; rnd < 24; rnd++) {
    uint64_t c[5], d[5], b[25];
    for (int x = 0; x < 5; x++) c[x] = a[x] ^ a[x + 5] ^ a[x + 10] ^ a[x + 15] ^ a[x + 20];
    for (int x = 0; x < 5; x++) d[x] = c[(x + 4) % 5] ^ k_rotl(c[(x + 1) % 5], 1);
    for (int i = 0; i < 25; i++) a[i] ^= d[i % 5];
    for (int x = 0; x < 5; x++)
      for (int y = 0; y < 5; y++) b[y + 5 * ((2 * x + 3 * y) % 5)] = k_rotl(a[x + 5 * y], KECCAK_ROT[x + 5 * y]);
    for (int y = 0; y < 5; y++)
      for (int x = 0; x < 5; x++) a[x + 5 * y] = b[x + 5 * y] ^ (~b[(x + 1) % 5 + 5 * y] & b[(x + 2) % 5 + 5 * y]);
    a[0] ^= KECCAK_RC[rnd];
  }
}

static inline void oracle_sha3_256(const uint8_t* in, size_t len, uint8_t out[32]) {
  uint64_t st[25];
  memset(st, 0, sizeof st);
  const size_t rate = 136;
  while (len >= rate) {
    for (size_t i = 0; i < rate / 8; i++) {
      uint64_t w;
      memcpy(&w, in + 8 * i, 8);
      st[i] ^= w;
    }
    keccak_f1600(st);
    in += rate;
    len -= rate;
  }
  uint8_t last[136];
  memset(last, 0, rate);
  memcpy(last, in, len);
  last[len] ^= 0x06;
  last[rate - 1] ^= 0x80;
  for (size_t i = 0; i < rate / 8; i++) {
    uint64_t w;
    memcpy(&w, last + 8 * i, 8);
    st[i] ^= w;
  }
  keccak_f1600(st);
  memcpy(out, st, 32);
}

/* hash ids shared with include/winter_gpu.h */
enum { OR_HASH_BLAKE3_256 = 0, OR_HASH_SHA3_256 = 1, OR_HASH_BLAKE3_192 = 2 };
/* digest length: Blake3_192 keeps the first 24 bytes of the BLAKE3 output (crypto/src/hash/blake/mod.rs:71-83) */
static inline size_t oracle_digest_len(int hash) { return hash == OR_HASH_BLAKE3_192 ? 24 : 32; }
/* writes oracle_digest_len(hash) bytes */
static inline void oracle_hash(int hash, const uint8_t* in, size_t len, uint8_t* out) {
  if (hash == OR_HASH_SHA3_256) {
    oracle_sha3_256(in, len, out);
  } else if (hash == OR_HASH_BLAKE3_192) {
    uint8_t full[32];
    oracle_blake3_256(in, len, full);
    memcpy(out, full, 24);
  } else {
    oracle_blake3_256(in, len, out);
  }
}
#endif
