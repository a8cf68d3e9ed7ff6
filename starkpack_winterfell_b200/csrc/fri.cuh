// FRI commit-phase kernels for sm_100a (the next caller of the hash / Merkle / NTT kernels).
//
//   fri_transpose_kernel : utils::transpose_slice::<_, N> (utils/core/src/lib.rs:206-225) as used by
//                          FriProver::build_layer (fri/src/prover/mod.rs:191-196): row i of the layer holds the N
//                          evaluations i, i + rows, ..., i + (N-1)*rows, so one Merkle leaf de-commits a whole coset.
//   fri_fold_kernel      : folding::apply_drp (fri/src/folding/mod.rs:115-146): per row an inverse length-N transform,
//                          coefficient k scaled by N^-1 * s_i^k with s_i = offset^-1 * g^-i (get_inv_offsets :186-193),
//                          then evaluation at alpha. Here the scaling is folded into the evaluation point:
//                          result_i = N^-1 * sum_k c_k * (alpha * s_i)^k  (exact field arithmetic, same element).
// Leaves are hashed by hash_rows_kernel on the transposed matrix (row_width = elements = N) and the tree is built by
// the Merkle kernels, exactly as for trace rows. Base field only (E = B), as in BASELINE configs 1, 3, 4
// (FieldExtension::None).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "field.cuh"
#include "ntt.cuh"

namespace wg {

template <class T>
__global__ void __launch_bounds__(256) fri_transpose_kernel(const T* __restrict__ in, T* __restrict__ out, uint64_t rows,
                                                            uint32_t log_n) {
  // one thread per output element, consecutive threads write consecutive addresses (full 32-byte sectors); the reads of
  // a warp fall on N runs of 32/N consecutive elements, every sector of which is consumed by this or a neighbouring
  // warp straight from L1/L2
  const uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t total = rows << log_n;
  if (idx >= total) return;
  const uint64_t i = idx >> log_n, j = idx & ((1u << log_n) - 1u);
  out[idx] = in[i + j * rows];
}

constexpr int FRI_MAX_LOG_DOMAIN = 40;

template <class F>
struct FriFoldArgs {
  typedef typename F::T T;
  const T* rows_in;  // [rows][N] transposed evaluations
  T* out;            // [rows]
  uint64_t rows;
  T alpha_sinv0;     // alpha * offset^-1
  T n_inv;           // N^-1
  T ginv_pow[FRI_MAX_LOG_DOMAIN];  // g^-(2^b), g = root of unity of order rows*N
  uint32_t log_rows;
};

template <class F, int LOG_N>
__global__ void __launch_bounds__(128) fri_fold_kernel(const FriFoldArgs<F> A) {
  typedef typename F::T T;
  constexpr int N = 1 << LOG_N;
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= A.rows) return;
  T a[N];
  const T* src = A.rows_in + i * N;
#pragma unroll
  for (int j = 0; j < N; j++) a[j] = src[j];
  radix_ntt<F, LOG_N, true>(a);  // c_k = sum_j v_j w_N^(-jk), natural order
  // x = alpha * offset^-1 * g^-i
  T x = A.alpha_sinv0;
#pragma unroll 1
  for (uint32_t b = 0; b < A.log_rows; b++)
    if ((i >> b) & 1) x = F::mul(x, A.ginv_pow[b]);
  T r = a[N - 1];
#pragma unroll
  for (int k = N - 2; k >= 0; k--) r = F::add(F::mul(r, x), a[k]);
  A.out[i] = F::mul(r, A.n_inv);
}

}  // namespace wg
