// Launch dispatch for ntt_pass_kernel instantiations. Each field gets its own translation unit
// (ntt_gl64.cu / ntt_f128.cu) so the two compile in parallel; api.cu only sees the declaration below.
#pragma once
#include <atomic>
#include "ntt.cuh"

namespace wg {

constexpr int MAX_LOG_S = 12;       // largest on-chip transform (4096 points)
constexpr int MIN_LOG_S_STRIDED = 6;  // strided passes are only used for N >= 2^13, i.e. K1 >= 2^6

// grid.x is computed by the caller from pass_log_tw(); returns cudaGetLastError() of the launch
template <class F>
cudaError_t launch_ntt_pass(int layout, bool inv, int log_s, const PassArgs<F>& a, dim3 grid, cudaStream_t stream);

// dynamic shared memory of the (layout, log_s) instantiation without the optional q table
template <class F>
size_t ntt_pass_smem(int layout, int log_s);
constexpr size_t SMEM_PER_SM = 227 * 1024;

template <class F>
inline int pass_log_tw(int log_s) {
  int l = FieldCfg<F>::log_tile(log_s) - log_s;
  return l < FieldCfg<F>::MAX_LOG_TW ? l : FieldCfg<F>::MAX_LOG_TW;
}

#define WG_DEFINE_LAUNCH_NTT_PASS(F)                                                                            \
  template <int LOG_S, int LAYOUT, bool INV>                                                                    \
  static cudaError_t launch_one_##F(const PassArgs<F>& a, dim3 grid, cudaStream_t stream) {                     \
    typedef PassCfg<F, LOG_S, LAYOUT> C;                                                                        \
    static std::atomic<bool> attr_set{false};                                                                   \
    constexpr size_t QB = LAYOUT == 1 ? sizeof(F::T) << LOG_S : 0;                                              \
    constexpr size_t MAXB = C::SMEM_BYTES + QB > SMEM_PER_SM ? C::SMEM_BYTES : C::SMEM_BYTES + QB;              \
    if (!attr_set && MAXB > 48 * 1024) {                                                                        \
      cudaError_t e = cudaFuncSetAttribute(ntt_pass_kernel<F, LOG_S, LAYOUT, INV>,                              \
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MAXB);             \
      if (e != cudaSuccess) return e;                                                                           \
      attr_set = true;                                                                                          \
    }                                                                                                           \
    const size_t smem = C::SMEM_BYTES + (a.has_q ? QB : 0);                                                     \
    if (smem > MAXB) return cudaErrorInvalidValue;                                                              \
    ntt_pass_kernel<F, LOG_S, LAYOUT, INV><<<grid, C::NT, smem, stream>>>(a);                                   \
    return cudaGetLastError();                                                                                  \
  }                                                                                                             \
  template <int LAYOUT>                                                                                         \
  static size_t smem_sz_##F(int log_s) {                                                                        \
    switch (log_s) {                                                                                            \
      case 1: return PassCfg<F, 1, LAYOUT>::SMEM_BYTES;                                                         \
      case 2: return PassCfg<F, 2, LAYOUT>::SMEM_BYTES;                                                         \
      case 3: return PassCfg<F, 3, LAYOUT>::SMEM_BYTES;                                                         \
      case 4: return PassCfg<F, 4, LAYOUT>::SMEM_BYTES;                                                         \
      case 5: return PassCfg<F, 5, LAYOUT>::SMEM_BYTES;                                                         \
      case 6: return PassCfg<F, 6, LAYOUT>::SMEM_BYTES;                                                         \
      case 7: return PassCfg<F, 7, LAYOUT>::SMEM_BYTES;                                                         \
      case 8: return PassCfg<F, 8, LAYOUT>::SMEM_BYTES;                                                         \
      case 9: return PassCfg<F, 9, LAYOUT>::SMEM_BYTES;                                                         \
      case 10: return PassCfg<F, 10, LAYOUT>::SMEM_BYTES;                                                       \
      case 11: return PassCfg<F, 11, LAYOUT>::SMEM_BYTES;                                                       \
      case 12: return PassCfg<F, 12, LAYOUT>::SMEM_BYTES;                                                       \
      default: return 0;                                                                                        \
    }                                                                                                           \
  }                                                                                                             \
  template <>                                                                                                   \
  size_t ntt_pass_smem<F>(int layout, int log_s) {                                                              \
    return layout == 0 ? smem_sz_##F<0>(log_s) : smem_sz_##F<1>(log_s);                                         \
  }                                                                                                             \
  template <int LAYOUT, bool INV>                                                                               \
  static cudaError_t launch_sz_##F(int log_s, const PassArgs<F>& a, dim3 grid, cudaStream_t stream) {           \
    switch (log_s) {                                                                                            \
      case 1: if (LAYOUT == 1) return launch_one_##F<1, 1, INV>(a, grid, stream); break;                        \
      case 2: if (LAYOUT == 1) return launch_one_##F<2, 1, INV>(a, grid, stream); break;                        \
      case 3: if (LAYOUT == 1) return launch_one_##F<3, 1, INV>(a, grid, stream); break;                        \
      case 4: if (LAYOUT == 1) return launch_one_##F<4, 1, INV>(a, grid, stream); break;                        \
      case 5: if (LAYOUT == 1) return launch_one_##F<5, 1, INV>(a, grid, stream); break;                        \
      case 6: return launch_one_##F<6, LAYOUT, INV>(a, grid, stream);                                           \
      case 7: return launch_one_##F<7, LAYOUT, INV>(a, grid, stream);                                           \
      case 8: return launch_one_##F<8, LAYOUT, INV>(a, grid, stream);                                           \
      case 9: return launch_one_##F<9, LAYOUT, INV>(a, grid, stream);                                           \
      case 10: return launch_one_##F<10, LAYOUT, INV>(a, grid, stream);                                         \
      case 11: return launch_one_##F<11, LAYOUT, INV>(a, grid, stream);                                         \
      case 12: return launch_one_##F<12, LAYOUT, INV>(a, grid, stream);                                         \
      default: break;                                                                                           \
    }                                                                                                           \
    return cudaErrorInvalidValue;                                                                               \
  }                                                                                                             \
  template <>                                                                                                   \
  cudaError_t launch_ntt_pass<F>(int layout, bool inv, int log_s, const PassArgs<F>& a, dim3 grid,              \
                                 cudaStream_t stream) {                                                         \
    if (layout == 0) {                                                                                          \
      return inv ? launch_sz_##F<0, true>(log_s, a, grid, stream) : launch_sz_##F<0, false>(log_s, a, grid, stream); \
    }                                                                                                           \
    return inv ? launch_sz_##F<1, true>(log_s, a, grid, stream) : launch_sz_##F<1, false>(log_s, a, grid, stream);   \
  }

}  // namespace wg
