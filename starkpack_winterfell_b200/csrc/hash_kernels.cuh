// Leaf hashing and Merkle reduction kernels for sm_100a.
//
//   hash_rows_kernel  : RowMatrix::commit_to_rows / commit_to_comb_rows leaves
//                       (prover/src/matrix/row_matrix.rs:183-238): leaf_k = H(row_k(t0)[0..C0] || row_k(t1)[0..C1] ..)
//                       with every f64 element serialised canonically (as_int, math/src/field/f64/mod.rs:605-610)
//                       and f128 as raw little-endian memory (f128/mod.rs:107-112); padding columns are skipped.
//   merkle_level_kernel / merkle_top_kernel : MerkleTree::new -> build_merkle_nodes
//                       (crypto/src/merkle/mod.rs:117-136,350-374): heap layout, nodes[0] = 0, root = nodes[1].
//
// Row hashing is integer-pipe bound (about 900 instructions per 64-byte BLAKE3 block), so one lane owns one
// row and keeps the whole compression state in registers; a warp stages 128-byte (BLAKE3) / 136-byte (SHA3
// rate) slabs of its 32 rows through shared memory so global loads stay coalesced and the Montgomery ->
// canonical conversion happens once per element on the way in.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "field.cuh"
#include "hash.cuh"

namespace wg {

constexpr int HASH_MAX_TRACES = 64;  // aggregated traces, or ranks x chunks of a sharded commit (WG_SHARD_MAX_BLOCKS)

struct RowHashArgs {
  const void* lde[HASH_MAX_TRACES];      // row-major matrices
  uint32_t row_width[HASH_MAX_TRACES];   // elements per stored row (padded)
  uint32_t unit_off[HASH_MAX_TRACES + 1];  // prefix sums of hashed 8-byte units per row
  uint32_t n_traces;
  uint64_t rows;
  uint8_t* leaves;  // rows x 32 bytes
  uint32_t one;     // run-time 1 (see b3_add)
};

constexpr int HASH_WARPS = 8;

// MULTI = rows longer than one 1024-byte BLAKE3 chunk (needs the chaining-value stack)
// ONE   = a single matrix (the plain commit_to_rows): no matrix lookup at all, pointer and row pitch are loop invariants
template <class F, int HASH, bool MULTI, bool ONE>
__global__ void __launch_bounds__(HASH_WARPS * 32, 4) hash_rows_kernel(const __grid_constant__ RowHashArgs A) {
  // __grid_constant__: the per-matrix tables are indexed with a run-time matrix number; without it the 1 KB argument
  // struct would be copied to every thread's local memory
  constexpr bool IS_B3 = HASH != HASH_SHA3_256;  // Blake3_256 and its 24-byte truncation Blake3_192
  constexpr int SU = IS_B3 ? 16 : 17;             // 8-byte units per slab
  constexpr int PITCH = 2 * SU + 1;                      // words; odd => conflict-free row reads
  __shared__ uint32_t slab_all[HASH_WARPS][32 * PITCH];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  uint32_t* slab = slab_all[wid];
  const uint64_t row0 = ((uint64_t)blockIdx.x * HASH_WARPS + wid) * 32;
  if (row0 >= A.rows) return;
  const uint32_t units_total = A.unit_off[A.n_traces];
  const uint32_t nbytes = units_total * 8;

  B3StreamT<MULTI> b3;
  uint64_t ks[25];
  if (IS_B3) {
    b3.init(nbytes, A.one);
  } else {
#pragma unroll
    for (int i = 0; i < 25; i++) ks[i] = 0;
  }
  bool sha_done = false;

  // the matrix that holds unit u0 lives in registers and advances monotonically with the slabs; a slab takes the
  // indexed-constant path only when it straddles a matrix boundary
  uint32_t t_base = 0;
  const uint64_t* cur_p = (const uint64_t*)A.lde[0];
  uint64_t cur_pitch = (uint64_t)A.row_width[0] * (F::kElemBytes / 8);
  uint32_t cur_lo = 0, cur_hi = ONE ? units_total : A.unit_off[1];
  // an empty row still takes one (padding-only) block: the digest of the empty input
  for (uint32_t u0 = 0; u0 < units_total || (HASH == HASH_SHA3_256 && !sha_done) || (IS_B3 && nbytes == 0 && u0 == 0);
       u0 += SU) {
    // ---- stage SU units of each of the warp's 32 rows: issue all SU loads first (one long-scoreboard wait per
    // slab instead of one per element), then convert and store
    if (!ONE && u0 >= cur_hi && u0 < units_total) {
      while (t_base + 1 < A.n_traces && u0 >= A.unit_off[t_base + 1]) t_base++;
      cur_p = (const uint64_t*)A.lde[t_base];
      cur_pitch = (uint64_t)A.row_width[t_base] * (F::kElemBytes / 8);
      cur_lo = A.unit_off[t_base];
      cur_hi = A.unit_off[t_base + 1];
    }
    uint64_t vals[SU];
    const bool straddle = !ONE && u0 + SU > cur_hi && cur_hi < units_total;  // warp-uniform
    if (IS_B3) {
      // SU = 16 divides the warp: lane + 32*it is always unit (lane & 15) of row (lane >> 4) + 2*it, so a lane looks its
      // unit's matrix up ONCE per slab, also when the slab straddles matrices (8-column chunks of a sharded commit
      // straddle in every slab)
      const uint32_t gu = u0 + (lane & 15);
      const uint64_t* p = cur_p;
      uint64_t pitch = cur_pitch;
      uint32_t lo = cur_lo;
      if (straddle && gu >= cur_hi && gu < units_total) {
        uint32_t t = t_base + 1;
        while (t + 1 < A.n_traces && gu >= A.unit_off[t + 1]) t++;
        p = (const uint64_t*)A.lde[t];
        pitch = (uint64_t)A.row_width[t] * (F::kElemBytes / 8);
        lo = A.unit_off[t];
      }
      p += gu - lo;
      const uint64_t rbase = row0 + (lane >> 4);
      if (row0 + 32 <= A.rows) {
        // 32 valid rows (every warp but possibly the last): no row predicates, and the 16 addresses are one
        // 32 x 32 -> 64 multiply-add each off one base (FMA pipe, which has slack here; the kernel is ALU-pipe bound):
        // 110 of the slab's 1,160 ALU-pipe instructions less, hash 17.0 -> 15.6 ms (profiles/r2_ab_hashfast.jsonl)
        if (gu < units_total) {
          const uint64_t* q = p + rbase * pitch;
          const uint32_t step = 2u * (uint32_t)pitch;
#pragma unroll
          for (int it = 0; it < SU; it++) vals[it] = __ldg(q + (uint64_t)((uint32_t)it * step));
        } else {
#pragma unroll
          for (int it = 0; it < SU; it++) vals[it] = 0;
        }
      } else {
#pragma unroll
        for (int it = 0; it < SU; it++) {
          const uint64_t row = rbase + 2 * it;
          vals[it] = (gu < units_total && row < A.rows) ? __ldg(p + row * pitch) : 0;
        }
      }
    } else if (!straddle) {
#pragma unroll
      for (int it = 0; it < SU; it++) {
        const int idx = lane + 32 * it;
        const int r = idx / SU, u = idx - r * SU;
        const uint32_t gu = u0 + u;
        const uint64_t row = row0 + r;
        uint64_t v = 0;
        if (gu < units_total && row < A.rows) v = __ldg(cur_p + row * cur_pitch + (gu - cur_lo));
        vals[it] = v;
      }
    } else {
#pragma unroll
      for (int it = 0; it < SU; it++) {
        const int idx = lane + 32 * it;
        const int r = idx / SU, u = idx - r * SU;
        const uint32_t gu = u0 + u;
        const uint64_t row = row0 + r;
        uint64_t v = 0;
        if (gu < units_total && row < A.rows) {
          uint32_t t = t_base;
          while (t + 1 < A.n_traces && gu >= A.unit_off[t + 1]) t++;
          const uint64_t* p = (const uint64_t*)A.lde[t];
          v = __ldg(p + row * ((uint64_t)A.row_width[t] * (F::kElemBytes / 8)) + (gu - A.unit_off[t]));
        }
        vals[it] = v;
      }
    }
#pragma unroll
    for (int it = 0; it < SU; it++) {
      const int idx = lane + 32 * it;
      const int r = idx / SU, u = idx - r * SU;
      const uint64_t v = F::kFieldId == 0 ? Gl64::as_int(vals[it]) : vals[it];
      slab[r * PITCH + 2 * u] = (uint32_t)v;
      slab[r * PITCH + 2 * u + 1] = (uint32_t)(v >> 32);
    }
    __syncwarp();
    // ---- consume: lane owns row `lane`
    const uint32_t* my = slab + lane * PITCH;
    if (IS_B3) {
#pragma unroll 1
      for (int b = 0; b < 2; b++) {
        if ((u0 + 8 * b) * 8 < nbytes || (nbytes == 0 && b == 0)) {
          uint32_t m[16];
#pragma unroll
          for (int i = 0; i < 16; i++) m[i] = my[16 * b + i];
          b3.feed(m);
        }
      }
    } else {
      const uint32_t rem = units_total - u0 < 17u ? units_total - u0 : 17u;  // units_total >= u0 here
#pragma unroll
      for (int i = 0; i < 17; i++) {
        uint64_t v = (uint64_t)my[2 * i] | ((uint64_t)my[2 * i + 1] << 32);
        if (rem < 17u) {  // final (padded) block; staged units beyond `rem` are zero
          if ((uint32_t)i == rem) v ^= 0x06ULL;
          if (i == 16) v ^= 0x8000000000000000ULL;
        }
        ks[i] ^= v;
      }
      keccak_f1600(ks);
      if (rem < 17u) sha_done = true;
    }
    __syncwarp();
  }
  const uint64_t row = row0 + lane;
  if (row < A.rows) {
    uint4* dst = reinterpret_cast<uint4*>(A.leaves + row * 32);
    if (HASH == HASH_BLAKE3_192) {
      uint2* d24 = reinterpret_cast<uint2*>(A.leaves + row * 24);  // first 24 bytes of the BLAKE3 output
      d24[0] = make_uint2(b3.cv[0], b3.cv[1]);
      d24[1] = make_uint2(b3.cv[2], b3.cv[3]);
      d24[2] = make_uint2(b3.cv[4], b3.cv[5]);
    } else if (HASH == HASH_BLAKE3_256) {
      dst[0] = make_uint4(b3.cv[0], b3.cv[1], b3.cv[2], b3.cv[3]);
      dst[1] = make_uint4(b3.cv[4], b3.cv[5], b3.cv[6], b3.cv[7]);
    } else {
      dst[0] = make_uint4((uint32_t)ks[0], (uint32_t)(ks[0] >> 32), (uint32_t)ks[1], (uint32_t)(ks[1] >> 32));
      dst[1] = make_uint4((uint32_t)ks[2], (uint32_t)(ks[2] >> 32), (uint32_t)ks[3], (uint32_t)(ks[3] >> 32));
    }
  }
}

// ---------------------------------------------------------------------------------------------- Merkle
// Digests are DW 32-bit words (8 for the 32-byte hashes, 6 for Blake3_192); nodes are addressed in 8-byte units so both
// sizes share the kernels: digest i starts at unit i * DW / 2.
template <int HASH>
struct DigestCfg {
  static constexpr int DW = hash_digest_bytes(HASH) / 4;
  static constexpr int U = DW / 2;  // 8-byte units per digest
};

// out = H(left || right) for the two consecutive digests at `pair`
template <int HASH>
__device__ __forceinline__ void merge_pair(const uint2* __restrict__ pair, uint32_t out[8], uint32_t one) {
  constexpr int U = DigestCfg<HASH>::U;
  uint32_t m[16];
#pragma unroll
  for (int k = 0; k < 2 * U; k++) {
    const uint2 v = pair[k];
    m[2 * k] = v.x;
    m[2 * k + 1] = v.y;
  }
  if (HASH == HASH_BLAKE3_256) {
    b3_merge(m, out, one);
  } else if (HASH == HASH_BLAKE3_192) {
    b3_merge48(m, out, one);
  } else {
    sha3_256_lanes([&](uint32_t i) { return (uint64_t)m[2 * i] | ((uint64_t)m[2 * i + 1] << 32); }, 64, out);
  }
}
template <int HASH>
__device__ __forceinline__ void store_digest(uint2* dst, const uint32_t o[8]) {
#pragma unroll
  for (int k = 0; k < DigestCfg<HASH>::U; k++) dst[k] = make_uint2(o[2 * k], o[2 * k + 1]);
}

// parents[i] = H(children[2i] || children[2i+1]), i < n_parents
template <int HASH>
__global__ void __launch_bounds__(256) merkle_level_kernel(const uint2* __restrict__ children, uint2* __restrict__ parents,
                                                           uint64_t n_parents, uint32_t one) {
  constexpr int U = DigestCfg<HASH>::U;
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_parents) return;
  uint32_t o[8];
  merge_pair<HASH>(children + 2 * U * i, o, one);
  store_digest<HASH>(parents + U * i, o);
}

// Finishes the tree in one block: nodes[n .. 2n) hold level n (n <= 1024 a power of two); computes
// nodes[n/2 .. n), ..., nodes[1], and zeroes nodes[0].
template <int HASH>
__global__ void __launch_bounds__(512) merkle_top_kernel(uint2* nodes, uint32_t n, uint32_t one) {
  constexpr int U = DigestCfg<HASH>::U;
  for (uint32_t lvl = n / 2; lvl >= 1; lvl >>= 1) {
    for (uint32_t i = threadIdx.x; i < lvl; i += blockDim.x) {
      const uint32_t node = lvl + i;
      uint32_t o[8];
      merge_pair<HASH>(nodes + (size_t)U * (2 * node), o, one);
      store_digest<HASH>(nodes + (size_t)U * node, o);
    }
    __syncthreads();
  }
  if (threadIdx.x < U) nodes[threadIdx.x] = make_uint2(0, 0);
}

}  // namespace wg
