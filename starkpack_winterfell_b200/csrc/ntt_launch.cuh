// Launch dispatch for ntt_pass_kernel instantiations. Each field gets its own translation unit
// (ntt_gl64.cu / ntt_f128.cu) so the two compile in parallel; api.cu only sees the declaration below.
#pragma once
#include "ntt.cuh"

namespace wg {

constexpr int MAX_LOG_S = 12;       // largest on-chip transform (4096 points)
constexpr int MIN_LOG_S_STRIDED = 6;  // strided passes are only used for N >= 2^13, i.e. K1 >= 2^6

// grid.x is computed by the caller from pass_log_tw(); returns cudaGetLastError() of the launch
template <class F>
cudaError_t launch_ntt_pass(int layout, bool inv, int log_s, const PassArgs<F>& a, dim3 grid, cudaStream_t stream);

template <class F>
inline int pass_log_tw(int log_s) {
  int l = FieldCfg<F>::LOG_TILE - log_s;
  return l < FieldCfg<F>::MAX_LOG_TW ? l : FieldCfg<F>::MAX_LOG_TW;
}

#define WG_DEFINE_LAUNCH_NTT_PASS(F)                                                                            \
  template <int LOG_S, int LAYOUT, bool INV>                                                                    \
  static cudaError_t launch_one_##F(const PassArgs<F>& a, dim3 grid, cudaStream_t stream) {                     \
    typedef PassCfg<F, LOG_S, LAYOUT> C;                                                                        \
    static bool attr_set = false;                                                                               \
    if (!attr_set && C::SMEM_BYTES > 48 * 1024) {                                                               \
      cudaError_t e = cudaFuncSetAttribute(ntt_pass_kernel<F, LOG_S, LAYOUT, INV>,                              \
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM_BYTES);    \
      if (e != cudaSuccess) return e;                                                                           \
      attr_set = true;                                                                                          \
    }                                                                                                           \
    ntt_pass_kernel<F, LOG_S, LAYOUT, INV><<<grid, C::NT, C::SMEM_BYTES, stream>>>(a);                          \
    return cudaGetLastError();                                                                                  \
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
