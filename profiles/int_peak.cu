// Integer-issue micro-benchmark for sm_100a (VERDICT r1 item 3 / SURVEY appendix B item 3).
//
// Measures, on the device, the warp-instruction throughput of the integer pipes the Goldilocks NTT and the
// BLAKE3 leaf kernels live on (ALU pipe: IADD3 / LOP3 / SHF / PRMT; FMA pipe: IMAD / IMAD.WIDE / IMAD.X), alone
// and mixed at the kernels' ratio, and the throughput of candidate formulations of the field operations.
// Build:  nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -o int_peak int_peak.cu
// Run  :  ./int_peak  > int_peak.json       (one JSON object per line)
// Every test is `ITERS` rounds over `ILP` independent chains per thread, 148*k blocks x 256 threads, timed with CUDA
// events; the static SASS instruction count per round is taken from cuobjdump (scripts in profiles/README.md) and
// the reported figure is operations / s (an "operation" = one round of one chain).
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include "../starkpack_winterfell_b200/csrc/field.cuh"

using namespace wg;

#define CK(x)                                                                      \
  do {                                                                             \
    cudaError_t e_ = (x);                                                          \
    if (e_ != cudaSuccess) {                                                       \
      fprintf(stderr, "CUDA %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
      exit(1);                                                                     \
    }                                                                              \
  } while (0)

constexpr int ILP = 8;
constexpr int ITERS = 4096;

// ------------------------------------------------------------------------------------------- raw pipe ops
struct OpIadd3 {  // ALU pipe
  static __device__ __forceinline__ uint32_t f(uint32_t x, uint32_t y) {
    uint32_t r;
    asm volatile("add.u32 %0, %1, %2;" : "=r"(r) : "r"(x), "r"(y));
    return r;
  }
};
struct OpLop3 {
  static __device__ __forceinline__ uint32_t f(uint32_t x, uint32_t y) {
    uint32_t r;
    asm volatile("xor.b32 %0, %1, %2;" : "=r"(r) : "r"(x), "r"(y));
    return r;
  }
};
struct OpShf {
  static __device__ __forceinline__ uint32_t f(uint32_t x, uint32_t y) {
    uint32_t r;
    asm volatile("shf.r.wrap.b32 %0, %1, %2, 7;" : "=r"(r) : "r"(x), "r"(y));
    return r;
  }
};
struct OpPrmt {
  static __device__ __forceinline__ uint32_t f(uint32_t x, uint32_t y) {
    uint32_t r;
    asm volatile("prmt.b32 %0, %1, %2, 0x1032;" : "=r"(r) : "r"(x), "r"(y));
    return r;
  }
};
struct OpImad {  // FMA pipe
  static __device__ __forceinline__ uint32_t f(uint32_t x, uint32_t y) {
    uint32_t r;
    asm volatile("mad.lo.u32 %0, %1, %2, %1;" : "=r"(r) : "r"(x), "r"(y));
    return r;
  }
};
struct OpImadHi {
  static __device__ __forceinline__ uint32_t f(uint32_t x, uint32_t y) {
    uint32_t r;
    asm volatile("mad.hi.u32 %0, %1, %2, %1;" : "=r"(r) : "r"(x), "r"(y));
    return r;
  }
};

template <class OP>
__global__ void __launch_bounds__(256) k_raw32(uint32_t* out, uint32_t seed) {
  uint32_t x[ILP], y = seed ^ threadIdx.x;
#pragma unroll
  for (int i = 0; i < ILP; i++) x[i] = seed * (i + 1) + threadIdx.x;
#pragma unroll 1
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int u = 0; u < 4; u++) {
#pragma unroll
      for (int i = 0; i < ILP; i++) x[i] = OP::f(x[i], y);
    }
  }
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < ILP; i++) s ^= x[i];
  if (s == 0x12345678u) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// IMAD.WIDE with a 64-bit addend (the accumulate form of a 32x32 product)
__global__ void __launch_bounds__(256) k_imad_wide(uint64_t* out, uint32_t seed) {
  uint64_t x[ILP];
  uint32_t y = seed ^ threadIdx.x;
#pragma unroll
  for (int i = 0; i < ILP; i++) x[i] = (uint64_t)seed * (i + 1) + threadIdx.x;
#pragma unroll 1
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int u = 0; u < 4; u++) {
#pragma unroll
      for (int i = 0; i < ILP; i++) {
        uint32_t lo = (uint32_t)x[i];
        asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(x[i]) : "r"(lo), "r"(y));
      }
    }
  }
  uint64_t s = 0;
#pragma unroll
  for (int i = 0; i < ILP; i++) s ^= x[i];
  if (s == 0x12345678u) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// carry chain: IADD3 (carry out) + IADD3.X (carry in), a 64-bit add
__global__ void __launch_bounds__(256) k_add64(uint64_t* out, uint32_t seed) {
  uint32_t x0[ILP], x1[ILP], y0 = seed ^ threadIdx.x, y1 = seed + threadIdx.x;
#pragma unroll
  for (int i = 0; i < ILP; i++) x0[i] = seed * (i + 1) + threadIdx.x, x1[i] = seed * (i + 3);
#pragma unroll 1
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int u = 0; u < 2; u++) {
#pragma unroll
      for (int i = 0; i < ILP; i++)
        asm volatile("add.cc.u32 %0, %0, %2;\n\taddc.u32 %1, %1, %3;" : "+r"(x0[i]), "+r"(x1[i]) : "r"(y0), "r"(y1));
    }
  }
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < ILP; i++) s ^= x0[i] ^ x1[i];
  if (s == 0x12345678u) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// ALU : FMA mix, independent streams: NA IADD3 and NF IMAD per round per chain
template <int NA, int NF>
__global__ void __launch_bounds__(256) k_mix(uint32_t* out, uint32_t seed) {
  uint32_t x[ILP], z[ILP], y = seed ^ threadIdx.x;
#pragma unroll
  for (int i = 0; i < ILP; i++) x[i] = seed * (i + 1) + threadIdx.x, z[i] = seed * (i + 7) ^ threadIdx.x;
#pragma unroll 1
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < ILP; i++) {
#pragma unroll
      for (int a = 0; a < NA; a++) x[i] = OpIadd3::f(x[i], y);
#pragma unroll
      for (int a = 0; a < NF; a++) z[i] = OpImad::f(z[i], y);
    }
  }
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < ILP; i++) s ^= x[i] ^ z[i];
  if (s == 0x12345678u) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// ------------------------------------------------------------------------------------------- field-op candidates
#include "../starkpack_winterfell_b200/csrc/field_lazy.cuh"
#include "../starkpack_winterfell_b200/csrc/ntt.cuh"

// one register step as the pass kernels run it: 16-point transform, then the inter-step twiddles on outputs 1..15
template <bool INV>
__device__ __forceinline__ void radix16_dif_cur(uint64_t (&a)[16], uint64_t w) {
  radix_ntt<Gl64, 4, INV>(a);
#pragma unroll
  for (int j = 1; j < 16; j++) a[j] = Gl64::mul(a[j], w);
}
template <bool INV>
__device__ __forceinline__ void radix16_dit_lazy(uint64_t (&a)[16], uint64_t w) {
  radix_dit_lazy<4, INV>(a);
  a[0] = GlLazy::canon(a[0]);
#pragma unroll
  for (int j = 1; j < 16; j++) a[j] = GlLazy::mul_w(a[j], w);
}

struct FAddCur { static __device__ __forceinline__ uint64_t f(uint64_t a, uint64_t b) { return Gl64::add(a, b); } };
struct FSubCur { static __device__ __forceinline__ uint64_t f(uint64_t a, uint64_t b) { return Gl64::sub(a, b); } };
struct FMulCur { static __device__ __forceinline__ uint64_t f(uint64_t a, uint64_t b) { return Gl64::mul(a, b); } };
struct FSh28Cur { static __device__ __forceinline__ uint64_t f(uint64_t a, uint64_t b) { return Gl64::mont_shift<28>(a) ^ b; } };
struct FSh52Cur { static __device__ __forceinline__ uint64_t f(uint64_t a, uint64_t b) { return Gl64::mont_shift<52>(a) ^ b; } };
struct FAddLazy { static __device__ __forceinline__ uint64_t f(uint64_t a, uint64_t b) { return GlLazy::add(a, b); } };
struct FSubLazy { static __device__ __forceinline__ uint64_t f(uint64_t a, uint64_t b) { return GlLazy::sub(a, b); } };
struct FAddLazyB { static __device__ __forceinline__ uint64_t f(uint64_t a, uint64_t b) { return GlLazy::add_b(a, b); } };
struct FSubLazyB { static __device__ __forceinline__ uint64_t f(uint64_t a, uint64_t b) { return GlLazy::sub_b(a, b); } };
struct FCanon { static __device__ __forceinline__ uint64_t f(uint64_t a, uint64_t b) { return GlLazy::canon(a) ^ b; } };
struct FMulW { static __device__ __forceinline__ uint64_t f(uint64_t a, uint64_t b) { return GlLazy::mul_w(a, b); } };
struct FMulW2 { static __device__ __forceinline__ uint64_t f(uint64_t a, uint64_t b) { return GlLazy::mul_w2(a, b); } };
struct FSh28L { static __device__ __forceinline__ uint64_t f(uint64_t a, uint64_t b) { return GlLazy::mont_shift<28>(a) ^ b; } };
struct FSh52L { static __device__ __forceinline__ uint64_t f(uint64_t a, uint64_t b) { return GlLazy::mont_shift<52>(a) ^ b; } };

template <class OP>
__global__ void __launch_bounds__(256) k_field(uint64_t* out, uint64_t seed) {
  uint64_t x[ILP], y = (seed ^ threadIdx.x) | 1;
  if (y >= Gl64::P) y -= Gl64::P;
#pragma unroll
  for (int i = 0; i < ILP; i++) {
    x[i] = seed * (2 * i + 1) + threadIdx.x;
    if (x[i] >= Gl64::P) x[i] -= Gl64::P;
  }
#pragma unroll 1
  for (int it = 0; it < ITERS; it++) {
    // the second operand is another chain's previous value, so nothing about it can be hoisted out of the loop
    uint64_t o[ILP];
#pragma unroll
    for (int i = 0; i < ILP; i++) o[i] = x[(i + 1) % ILP];
#pragma unroll
    for (int i = 0; i < ILP; i++) x[i] = OP::f(x[i], o[i] ^ y);
  }
  uint64_t s = 0;
#pragma unroll
  for (int i = 0; i < ILP; i++) s ^= x[i];
  if (s == 0x12345678u) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// whole radix-16 butterflies on registers: the current DIF network against the lazy DIT one
template <int VARIANT, bool INV>
__global__ void __launch_bounds__(256) k_radix16(uint64_t* out, uint64_t seed, uint64_t* check) {
  uint64_t a[16];
#pragma unroll
  for (int i = 0; i < 16; i++) {
    a[i] = seed * (2 * i + 1) + threadIdx.x * 0x9E3779B97F4A7C15ULL + blockIdx.x;
    if (a[i] >= Gl64::P) a[i] -= Gl64::P;
  }
#pragma unroll 1
  for (int it = 0; it < ITERS / 16; it++) {
    if (VARIANT == 0) radix16_dif_cur<INV>(a, seed >> 1);
    else radix16_dit_lazy<INV>(a, seed >> 1);
  }
  if (check && blockIdx.x == 0 && threadIdx.x < 32) {
#pragma unroll
    for (int i = 0; i < 16; i++) check[threadIdx.x * 16 + i] = a[i];
  }
  uint64_t s = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) s ^= a[i];
  if (s == 0x12345678u) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// ------------------------------------------------------------------------------------------- harness
static int g_sms = 148;
static double g_clock_mhz = 0;

template <class K, class... Args>
static double time_kernel(K kern, int blocks_per_sm, Args... args) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  const int grid = g_sms * blocks_per_sm;
  for (int w = 0; w < 2; w++) kern<<<grid, 256>>>(args...);
  CK(cudaDeviceSynchronize());
  float best = 1e30f;
  for (int r = 0; r < 5; r++) {
    CK(cudaEventRecord(e0));
    kern<<<grid, 256>>>(args...);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    if (ms < best) best = ms;
  }
  CK(cudaGetLastError());
  return best * 1e-3;
}

static void report(const char* name, double secs, int blocks_per_sm, double ops_per_thread, int sass_per_op) {
  const double threads = (double)g_sms * blocks_per_sm * 256;
  const double ops = threads * ops_per_thread;
  const double warp_ops = ops / 32.0;
  printf("{\"test\": \"%s\", \"ms\": %.4f, \"blocks_per_sm\": %d, \"thread_ops_per_s\": %.4e, "
         "\"warp_ops_per_clk_per_smsp\": %.4f, \"sass_per_op\": %d, \"warp_inst_per_clk_per_smsp\": %.4f}\n",
         name, secs * 1e3, blocks_per_sm, ops / secs, warp_ops / secs / (g_sms * 4.0) / (g_clock_mhz * 1e6), sass_per_op,
         sass_per_op * warp_ops / secs / (g_sms * 4.0) / (g_clock_mhz * 1e6));
  fflush(stdout);
}

int main(int argc, char** argv) {
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, 0));
  g_sms = prop.multiProcessorCount;
  int khz = 0;
  CK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0));
  g_clock_mhz = khz / 1000.0;
  printf("{\"device\": \"%s\", \"sms\": %d, \"clock_mhz_attr\": %.0f, \"note\": \"per-clk figures assume the attr clock\"}\n",
         prop.name, g_sms, g_clock_mhz);
  uint64_t* out;
  CK(cudaMalloc(&out, (size_t)g_sms * 8 * 256 * 8));
  uint64_t* check;
  CK(cudaMalloc(&check, 2 * 32 * 16 * 8));
  const int B = 8;  // 8 x 256 threads = 64 warps per SM
  const double n32 = (double)ITERS * 4 * ILP;
  report("iadd3", time_kernel(k_raw32<OpIadd3>, B, (uint32_t*)out, 12345u), B, n32, 1);
  report("lop3", time_kernel(k_raw32<OpLop3>, B, (uint32_t*)out, 12345u), B, n32, 1);
  report("shf", time_kernel(k_raw32<OpShf>, B, (uint32_t*)out, 12345u), B, n32, 1);
  report("prmt", time_kernel(k_raw32<OpPrmt>, B, (uint32_t*)out, 12345u), B, n32, 1);
  report("imad", time_kernel(k_raw32<OpImad>, B, (uint32_t*)out, 12345u), B, n32, 1);
  report("imad_hi", time_kernel(k_raw32<OpImadHi>, B, (uint32_t*)out, 12345u), B, n32, 1);
  report("imad_wide64", time_kernel(k_imad_wide, B, out, 12345u), B, n32, 1);
  report("add64_carry", time_kernel(k_add64, B, out, 12345u), B, (double)ITERS * 2 * ILP, 2);
  report("mix_1alu_1fma", time_kernel(k_mix<1, 1>, B, (uint32_t*)out, 12345u), B, (double)ITERS * ILP, 2);
  report("mix_2alu_1fma", time_kernel(k_mix<2, 1>, B, (uint32_t*)out, 12345u), B, (double)ITERS * ILP, 3);
  report("mix_1alu_2fma", time_kernel(k_mix<1, 2>, B, (uint32_t*)out, 12345u), B, (double)ITERS * ILP, 3);
  report("mix_9alu_8fma", time_kernel(k_mix<9, 8>, B, (uint32_t*)out, 12345u), B, (double)ITERS * ILP, 17);
  const double nf = (double)ITERS * ILP;
  for (int bps : {4, 8}) {
    report("gl_add_cur", time_kernel(k_field<FAddCur>, bps, out, 0x9E3779B97F4A7C15ULL), bps, nf, 9);
    report("gl_sub_cur", time_kernel(k_field<FSubCur>, bps, out, 0x9E3779B97F4A7C15ULL), bps, nf, 6);
    report("gl_mul_cur", time_kernel(k_field<FMulCur>, bps, out, 0x9E3779B97F4A7C15ULL), bps, nf, 19);
    report("gl_shift28_cur", time_kernel(k_field<FSh28Cur>, bps, out, 0x9E3779B97F4A7C15ULL), bps, nf, 13);
    report("gl_shift52_cur", time_kernel(k_field<FSh52Cur>, bps, out, 0x9E3779B97F4A7C15ULL), bps, nf, 11);
    report("gl_add_lazy", time_kernel(k_field<FAddLazy>, bps, out, 0x9E3779B97F4A7C15ULL), bps, nf, 4);
    report("gl_sub_lazy", time_kernel(k_field<FSubLazy>, bps, out, 0x9E3779B97F4A7C15ULL), bps, nf, 5);
    report("gl_add_lazy_b", time_kernel(k_field<FAddLazyB>, bps, out, 0x9E3779B97F4A7C15ULL), bps, nf, 4);
    report("gl_sub_lazy_b", time_kernel(k_field<FSubLazyB>, bps, out, 0x9E3779B97F4A7C15ULL), bps, nf, 5);
    report("gl_canon", time_kernel(k_field<FCanon>, bps, out, 0x9E3779B97F4A7C15ULL), bps, nf, 5);
    report("gl_mul_w", time_kernel(k_field<FMulW>, bps, out, 0x9E3779B97F4A7C15ULL), bps, nf, 17);
    report("gl_mul_w2", time_kernel(k_field<FMulW2>, bps, out, 0x9E3779B97F4A7C15ULL), bps, nf, 17);
    report("gl_shift28_lazy", time_kernel(k_field<FSh28L>, bps, out, 0x9E3779B97F4A7C15ULL), bps, nf, 11);
    report("gl_shift52_lazy", time_kernel(k_field<FSh52L>, bps, out, 0x9E3779B97F4A7C15ULL), bps, nf, 9);
  }
  const double nr = (double)(ITERS / 16);
  for (int bps : {2, 4}) {
    report("radix16_dif_cur_fwd", time_kernel(k_radix16<0, false>, bps, out, 0x9E3779B97F4A7C15ULL, check), bps, nr, 0);
    report("radix16_dit_lazy_fwd", time_kernel(k_radix16<1, false>, bps, out, 0x9E3779B97F4A7C15ULL, check + 512), bps, nr, 0);
    report("radix16_dif_cur_inv", time_kernel(k_radix16<0, true>, bps, out, 0x9E3779B97F4A7C15ULL, check), bps, nr, 0);
    report("radix16_dit_lazy_inv", time_kernel(k_radix16<1, true>, bps, out, 0x9E3779B97F4A7C15ULL, check + 512), bps, nr, 0);
  }
  // the two radix-16 networks must agree (canonical outputs) after the same number of rounds
  {
    static uint64_t h[1024];
    k_radix16<0, false><<<1, 256>>>(out, 0x9E3779B97F4A7C15ULL, check);
    k_radix16<1, false><<<1, 256>>>(out, 0x9E3779B97F4A7C15ULL, check + 512);
    CK(cudaMemcpy(h, check, sizeof(h), cudaMemcpyDeviceToHost));
    int bad = 0;
    for (int i = 0; i < 512; i++) {
      uint64_t x = h[i], y = h[512 + i];
      if (y >= Gl64::P) y -= Gl64::P;
      if (x != y) bad++;
    }
    printf("{\"test\": \"radix16_agree_fwd\", \"mismatches\": %d}\n", bad);
    k_radix16<0, true><<<1, 256>>>(out, 0x9E3779B97F4A7C15ULL, check);
    k_radix16<1, true><<<1, 256>>>(out, 0x9E3779B97F4A7C15ULL, check + 512);
    CK(cudaMemcpy(h, check, sizeof(h), cudaMemcpyDeviceToHost));
    bad = 0;
    for (int i = 0; i < 512; i++) {
      uint64_t x = h[i], y = h[512 + i];
      if (y >= Gl64::P) y -= Gl64::P;
      if (x != y) bad++;
    }
    printf("{\"test\": \"radix16_agree_inv\", \"mismatches\": %d}\n", bad);
  }
  return 0;
}
