// Transforms longer than the two-pass plan's 2^24 points (device kernels; orchestration in api.cu).
//
// The reference's FFT accepts any power of two up to the field's two-adicity (math/src/fft/mod.rs:291-295: 2^32 for f64,
// 2^40 for f128). The on-chip passes (ntt.cuh) cover N = K1*K2 with K1, K2 <= 2^12. Longer transforms take one more
// decimation step by R = 2^LR <= 16, written as the polynomial identity it is:
//
//     P(x) = sum_{r < R} x^r Q_r(x^R),   Q_r(y) = sum_m a[m*R + r] y^m          (deg Q_r < M = n / R)
//
// * LDE / evaluation over the coset {offset * w^k}: the R sub-polynomials of every column become R*C columns of
//   length M (split_deinterleave_kernel), the existing LDE evaluates them over the domain {offset^R * w_{MB}^k'}
//   (k' < M*B, the R-th powers of the big domain's points), and split_lde_post_kernel combines, for every k' and
//   column, the R values Q_r(x^R) into the R evaluations at the points x*w_R^j that share that R-th power:
//   X[k' + j*M*B] = sum_r (x_k' w_R^j)^r Q_r(...) = R-point transform of (x_k'^r Q_r), x_k' = offset * w_{nB}^k'.
// * interpolation is the same identity read backwards: split_interp_pre_kernel takes the R evaluations
//   v[j2 + M*j], inverts the R-point transform and removes w_n^(j2 r), leaving Q_r(w_M^j2); the existing interpolation
//   recovers every Q_r and split_interleave_kernel stores a[m*R + r].
// Longer than 2^28: the inner call splits again (api.cu recursion), so the limit is the two-adicity.
#pragma once
#include "ntt.cuh"

namespace wg {

// two-level power table (api.cu Tables::fill_two): value(k) = lo[k & mask] * hi[k >> lb]
template <class F>
struct PowTab {
  const typename F::T* lo;
  const typename F::T* hi;
  uint32_t lb;
};
template <class F>
__device__ __forceinline__ typename F::T powtab_at(const PowTab<F>& t, uint64_t k) {
  return F::mul(t.lo[k & ((1u << t.lb) - 1)], t.hi[k >> t.lb]);
}

// out[(c*R + r)*M + m] = in[c*n + m*R + r]    (grid: x over m, y over columns)
template <class F, int LR>
__global__ void __launch_bounds__(256) split_deinterleave_kernel(const typename F::T* __restrict__ in, typename F::T* __restrict__ out,
                                                                 uint64_t M) {
  typedef typename F::T T;
  constexpr int R = 1 << LR;
  const uint64_t m = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= M) return;
  const uint64_t c = blockIdx.y;
  const T* src = in + c * (M * R) + m * R;
  T* dst = out + c * (M * R) + m;
#pragma unroll
  for (int r = 0; r < R; r++) dst[(uint64_t)r * M] = src[r];
}

// out[c*n + m*R + r] = in[(c*R + r)*M + m]
template <class F, int LR>
__global__ void __launch_bounds__(256) split_interleave_kernel(const typename F::T* __restrict__ in, typename F::T* __restrict__ out,
                                                               uint64_t M) {
  typedef typename F::T T;
  constexpr int R = 1 << LR;
  const uint64_t m = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= M) return;
  const uint64_t c = blockIdx.y;
  const T* src = in + c * (M * R) + m;
  T* dst = out + c * (M * R) + m * R;
#pragma unroll
  for (int r = 0; r < R; r++) dst[r] = src[(uint64_t)r * M];
}

// Interpolation, first step. v: C columns of n = R*M evaluations; e: R*C columns of M values,
// e[(c*R + r)*M + j2] = w_n^(-j2 r) * R^-1 * sum_j w_R^(-j r) v[c*n + j2 + M*j]      (= Q_r(w_M^j2))
// tab: powers of w_n^-1 below M.
template <class F, int LR>
__global__ void __launch_bounds__(256) split_interp_pre_kernel(const typename F::T* __restrict__ v, typename F::T* __restrict__ e,
                                                               uint64_t M, PowTab<F> tab, typename F::T r_inv) {
  typedef typename F::T T;
  constexpr int R = 1 << LR;
  typedef RadixNet<F, LR, true> Net;
  const uint64_t j2 = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j2 >= M) return;
  const uint64_t c = blockIdx.y;
  const T* src = v + c * (M * R) + j2;
  T a[R];
#pragma unroll
  for (int j = 0; j < R; j++) a[j] = src[(uint64_t)j * M];
  Net::run(a);
  const T y = powtab_at<F>(tab, j2);
  T p = r_inv;
  T* dst = e + c * (M * R) + j2;
#pragma unroll
  for (int r = 0; r < R; r++) {
    dst[(uint64_t)r * M] = F::mul(a[r], p);  // the product canonicalises a lazily reduced a[r]
    p = F::mul(p, y);
  }
}

// LDE, last step. e: row-major [MB][rw_e], row k' holds Q_r(x_k'^R) of column c at e[k'*rw_e + c*R + r] (c < nc_in).
// out: row-major [R*MB][rw_out]; columns c0 + c, c < nc_out: out[(k' + j*MB)*rw_out + c0 + c] = sum_r (x_k' w_R^j)^r e_r
// for c < nc_in and 0 for the pad columns nc_in <= c < nc_out. tab: x_k' = offset * w_{nB}^k'.
// block = (bx columns) x (256 / bx rows), bx chosen by the host from the sweep width: stores of a row are contiguous
// across threadIdx.x.
template <class F, int LR>
__global__ void __launch_bounds__(256) split_lde_post_kernel(const typename F::T* __restrict__ e, uint64_t rw_e,
                                                             typename F::T* __restrict__ out, uint64_t rw_out, uint32_t nc_in,
                                                             uint32_t nc_out, uint64_t MB, PowTab<F> tab) {
  typedef typename F::T T;
  constexpr int R = 1 << LR;
  typedef RadixNet<F, LR, false> Net;
  const uint64_t k = (uint64_t)blockIdx.x * blockDim.y + threadIdx.y;
  const uint32_t c = blockIdx.y * blockDim.x + threadIdx.x;
  if (k >= MB || c >= nc_out) return;
  T a[R];
  if (c < nc_in) {
    const T* src = e + k * rw_e + (uint64_t)c * R;
    const T x = powtab_at<F>(tab, k);
    a[0] = src[0];
    T p = x;
#pragma unroll
    for (int r = 1; r < R; r++) {
      a[r] = F::mul(src[r], p);
      if (r + 1 < R) p = F::mul(p, x);
    }
    Net::run(a);
#pragma unroll
    for (int j = 0; j < R; j++) a[j] = Net::canon(a[j]);
  } else {
#pragma unroll
    for (int j = 0; j < R; j++) a[j] = F::zero();
  }
  T* dst = out + k * rw_out + c;
#pragma unroll
  for (int j = 0; j < R; j++) dst[(uint64_t)j * MB * rw_out] = a[j];
}

}  // namespace wg
