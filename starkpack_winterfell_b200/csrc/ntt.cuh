// Batched shared-memory NTT pass for sm_100a.
//
// One kernel template covers every NTT-shaped step of the trace commitment:
//   * ColMatrix::interpolate_columns  (prover/src/matrix/col_matrix.rs:196-206 -> fft::interpolate_poly,
//     math/src/fft/mod.rs:274, serial.rs:64-73)
//   * RowMatrix::evaluate_polys_over  (prover/src/matrix/row_matrix.rs:82-133, segments.rs:53-207)
//   * fft::evaluate_poly_with_offset  (math/src/fft/mod.rs:171, serial.rs:26-57)
// The reference runs an in-place radix-2 network with bit-reversed output, then a permutation pass (and for
// the matrix a transpose). Field arithmetic is exact, so any algorithm gives the same words; here a
// length-N transform is split N = K1*K2 into at most two passes over HBM, each pass a batch of length-S
// (S <= 4096) transforms done entirely on chip:
//     - thread-level radix-16/8/4/2 butterflies in registers (for Goldilocks the inner twiddles are powers of
//       two because w64 = 8, so they are shifts, not multiplications),
//     - digit exchange through shared memory between register steps, twiddles w_S^e read from a shared table,
//     - first step loads straight from HBM, last step stores straight to HBM in NATURAL order, so no
//       permutation pass and no transpose exist.
// A tile is S transform points x TW independent problems. LAYOUT 0 ("strided pass"): the TW problems are
// contiguous in memory (coalesced 64-byte pieces), the transform runs along a stride. LAYOUT 1 ("contiguous
// pass"): the transform runs along contiguous memory, the TW problems are strided on input and contiguous
// on output (this is what writes row-major LDE rows, 8 columns = one 64-byte row segment per store).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "field.cuh"
#include "field_lazy.cuh"

namespace wg {

template <class F>
struct PassArgs {
  typedef typename F::T T;
  const T* in;
  T* out;
  // element offsets: blockIdx.y*ystride + blockIdx.z*zstride + wg*wstride + s*sstride   (wg = global w index)
  long long in_ystride, in_zstride, in_wstride, in_sstride;
  long long out_ystride, out_zstride, out_wstride, out_jstride;
  uint32_t nw_in;   // problems with wg >= nw_in read as zero
  uint32_t nw_out;  // problems with wg >= nw_out are not written
  // in-block twiddles, one sub-table per register step u < NSTEPS-1 laid out [j][m_rest]:
  // tws[off_u + j*P_u + m] = w_{P_{u-1}}^(+-m*j)  (lanes walk m, so shared-memory reads are conflict-free)
  const T* tws;
  // optional pre-multiplication x[s] *= PM^(s*pm_s + wg*pm_w)  (two-level table, per-z tables)
  const T* pm_lo;
  const T* pm_hi;
  long long pm_zstride_lo, pm_zstride_hi;
  uint32_t pm_s, pm_w, pm_lb, pm_mask;  // exponent & pm_mask, lo index = e & ((1<<pm_lb)-1)
  // optional post-multiplication X[j] *= TM^(j * (wg*tm_w + y*tm_y))  (two-level table)
  const T* tm_lo;
  const T* tm_hi;
  uint32_t tm_w, tm_y, tm_lb, tm_mask;
  // optional constant post-scale (n^-1 when no tm table carries it)
  T post_const;
  // has_pm: 0 none, 1 two-level table above, 2 direct: x[s] *= pm_hi[z*pm_zstride_hi + s]
  int has_pm, has_tm, has_const;
  // optional per-block geometric pre-multiplication x[s] *= q^s (LAYOUT 1 only) with
  // q = q_z[z*q_zstride] * TM^(y*q_y): the table q^0..q^(S-1) is built in shared memory once per (y, z) and shared by
  // every tile the block walks along x. This is how the LDE folds the coset shift t_z^m2 and the inter-pass twiddle
  // w_N^(j1*m2) into one multiplication (q = t_z * w_N^j1).
  int has_q;
  const T* q_z;
  long long q_zstride;
  uint32_t q_y;
  // a block walks x = blockIdx.x, blockIdx.x + gridDim.x, ... < nx and y likewise < ny (0: one tile per block);
  // the in-block twiddle table is loaded once per block
  uint32_t nx, ny;
  // zfast = nz > 0: the grid is (gx*nz, gy, 1) and z = blockIdx.x % nz varies fastest, so the nz blocks that read
  // the same input tile (the cosets of one polynomial slice) are co-resident and all but one hit in L2
  uint32_t zfast;
  // n_peers > 0: the output is scattered over n_peers buffers by transform index j: element j goes to
  // out_peers[j >> peer_shift] at index j & (2^peer_shift - 1) (everything else as for `out`). This is the multi-GPU
  // LDE: output row k = j*stride + ... belongs to the GPU that hashes rows [g*L/G, (g+1)*L/G), and the pointers are
  // peer-mapped (CUDA IPC) buffers on those GPUs, so the pass stores its rows over NVLink while it computes and no
  // separate all-to-all exists.
  static constexpr int MAX_PEERS = 8;
  T* out_peers[MAX_PEERS];
  uint32_t n_peers, peer_shift;
};

// ---------------------------------------------------------------------------------------------- radix butterflies
constexpr __host__ __device__ int bitrev_c(int x, int bits) {
  int r = 0;
  for (int i = 0; i < bits; i++) r |= ((x >> i) & 1) << (bits - 1 - i);
  return r;
}

// natural-order in, natural-order out length-R transform of a[] held in registers (R = 2^LR, LR <= 4)
template <class F, int LR, bool INV>
__device__ __forceinline__ void radix_ntt(typename F::T (&a)[1 << LR]) {
  typedef typename F::T T;
  constexpr int R = 1 << LR;
#pragma unroll
  for (int half = R / 2; half >= 1; half >>= 1) {
#pragma unroll
    for (int g = 0; g < R; g += 2 * half) {
#pragma unroll
      for (int k = 0; k < half; k++) {
        T x = a[g + k], y = a[g + k + half];
        a[g + k] = F::add(x, y);
        a[g + k + half] = F::template sub_mul_root16<INV>(x, y, k * (8 / half));
      }
    }
  }
  // DIF leaves X[bitrev(pos)] at pos
  T b[R];
#pragma unroll
  for (int i = 0; i < R; i++) b[bitrev_c(i, LR)] = a[i];
#pragma unroll
  for (int i = 0; i < R; i++) a[i] = b[i];
}

// Which register network a field uses. Goldilocks: the decimation-in-time network on lazily reduced values
// (field_lazy.cuh; inputs canonical, outputs any u64 congruent mod p, so a pass canonicalises whatever it stores without a
// multiplication in between). Everything else: the canonical decimation-in-frequency network above.
template <class F, int LR, bool INV>
struct RadixNet {
  static constexpr bool LAZY = false;
  static __device__ __forceinline__ void run(typename F::T (&a)[1 << LR]) { radix_ntt<F, LR, INV>(a); }
  static __device__ __forceinline__ typename F::T canon(typename F::T v) { return v; }
};
#if defined(WG_DIT_LAZY) && defined(__CUDACC__)
template <int LR, bool INV>
struct RadixNet<Gl64, LR, INV> {
  static constexpr bool LAZY = true;
  static __device__ __forceinline__ void run(uint64_t (&a)[1 << LR]) { radix_dit_lazy<LR, INV>(a); }
  static __device__ __forceinline__ uint64_t canon(uint64_t v) { return GlLazy::canon(v); }
};
#endif

// ---------------------------------------------------------------------------------------------- compile-time plan
template <class F>
struct FieldCfg;
template <>
struct FieldCfg<Gl64> {
  static constexpr int MAXLR = 4;  // radix-16 register butterflies
#ifndef WG_GL_LOG_TILE
#define WG_GL_LOG_TILE 13
#endif
#ifndef WG_GL_MINB
#define WG_GL_MINB 2
#endif
  static constexpr int LOG_TILE = WG_GL_LOG_TILE;  // S*TW <= 2^LOG_TILE elements (64 KB at 13)
#ifndef WG_GL_LOG_TILE12
#define WG_GL_LOG_TILE12 14
#endif
  // 2^12-point passes (2^23, 2^24 rows) get 128 KB tiles: 4 problems per tile keep the global pieces at 32 bytes
  static constexpr __host__ __device__ int log_tile(int log_s) { return log_s >= 12 ? WG_GL_LOG_TILE12 : LOG_TILE; }
  static constexpr int MAX_LOG_TW = 3;
  // Shared-memory padding. A 64-bit access is served per half-warp, conflict-free when its 16 lanes fall in 16 different
  // 8-byte bank pairs. The last register step reads element 8q + r of column w in lane (w, q) = (lane & 3, lane >> 2):
  //   LAYOUT 0 (index s*TW + w, TW = 4): 4 pad elements per 32 make the q stride 36 = 4 (mod 16)  -> {0,4,8,12} + w
  //   LAYOUT 1 (index w*SP + s):        1 pad element per 16 and SP = 2 (mod 16) -> {0,8,1,9} + 2w
  // (8 pad elements per 32 / per 128, the earlier choice, left a 2-way conflict on that step: 20 % of all wavefronts.)
#ifndef WG_GL_PAD0
#define WG_GL_PAD0 2
#endif
#ifndef WG_GL_L1FINE
#define WG_GL_L1FINE 1
#endif
  static constexpr int PAD_LOG = WG_GL_PAD0;
  static constexpr bool L1_FINE = WG_GL_L1FINE;
  static constexpr bool SKIP_DEAD = false;
  static constexpr int MIN_BLOCKS = WG_GL_MINB;  // resident 512-thread blocks per SM the kernels are compiled for
};
template <>
struct FieldCfg<F128> {
  static constexpr int MAXLR = 3;
  static constexpr int LOG_TILE = 13;
  static constexpr __host__ __device__ int log_tile(int) { return LOG_TILE; }
  static constexpr int MAX_LOG_TW = 2;
  static constexpr int PAD_LOG = 2;  // 4 elements = 64 B
  static constexpr bool L1_FINE = false;
  static constexpr bool SKIP_DEAD = true;  // 2-column fib traces fill half a tile (BASELINE configs 1 and 4)
  // two resident blocks per SM: the 64-register cap costs ~48 bytes of spills per thread and hides the block barriers of
  // the half-filled tiles of 2-column traces (cfg1 2.85 -> 2.72 ms, profiles/r2_ab_f128_minb2.jsonl); 2^12-point passes
  // still run one block per SM (128 KB tiles)
#ifndef WG_F128_MINB
#define WG_F128_MINB 2
#endif
  static constexpr int MIN_BLOCKS = WG_F128_MINB;
};

template <class F, int LOG_S, int LAYOUT>
struct PassCfg {
  typedef FieldCfg<F> FC;
  static constexpr int MAXLR = FC::MAXLR;
  static constexpr int NSTEPS = (LOG_S + MAXLR - 1) / MAXLR;
  static constexpr int LOG_TW = (FC::log_tile(LOG_S) - LOG_S) < FC::MAX_LOG_TW ? (FC::log_tile(LOG_S) - LOG_S) : FC::MAX_LOG_TW;
  static constexpr int TW = 1 << LOG_TW;
  static constexpr int S = 1 << LOG_S;
  static constexpr __host__ __device__ int lr(int u) { return u < NSTEPS - 1 ? MAXLR : LOG_S - MAXLR * (NSTEPS - 1); }
  // log2 of the sub-problem size AFTER step u
  static constexpr __host__ __device__ int logp(int u) {
    int l = LOG_S;
    for (int v = 0; v <= u; v++) l -= lr(v);
    return l;
  }
  static constexpr int LR_LAST = LOG_S - MAXLR * (NSTEPS - 1);
  // threads: one item (R elements) per thread per step where possible
  static constexpr int ITEMS0 = (S >> (NSTEPS > 1 ? MAXLR : LOG_S)) * TW;
  static constexpr int NT = ITEMS0 >= 512 ? 512 : (ITEMS0 <= 32 ? 32 : ITEMS0);
  // shared-memory index of logical (s, w); padded against bank conflicts of the last register step
  static constexpr int PAD_SHIFT = LAYOUT == 0 ? (LR_LAST + LOG_TW) : 7;
  static constexpr int SP = FC::L1_FINE ? S + (S >> 4) + 2 : S + ((S >> 7) << FC::PAD_LOG) + 1;  // LAYOUT 1 row pitch
  static __host__ __device__ __forceinline__ int phys(int s, int w) {
    if (LAYOUT == 0) {
      int idx = s * TW + w;
      return idx + ((idx >> PAD_SHIFT) << FC::PAD_LOG);
    } else if (FC::L1_FINE) {
      return w * SP + s + (s >> 4);
    } else {
      return w * SP + s + ((s >> 7) << FC::PAD_LOG);
    }
  }
  static constexpr int TILE_ELEMS =
      NSTEPS == 1 ? 0 : (LAYOUT == 0 ? (S * TW + (((S * TW) >> PAD_SHIFT) << FC::PAD_LOG)) : TW * SP);
  // offset of step u's twiddle sub-table; sub-table u has R_u * P_u = 2^(logp(u) + lr(u)) entries
  static constexpr __host__ __device__ int tws_off(int u) {
    int o = 0;
    for (int v = 0; v < u; v++) o += 1 << (logp(v) + lr(v));
    return o;
  }
  static constexpr int TWS_ELEMS = NSTEPS == 1 ? 0 : tws_off(NSTEPS - 1);
  static constexpr size_t SMEM_BYTES = (size_t)(TILE_ELEMS + TWS_ELEMS) * sizeof(typename F::T);
};

template <class F>
__device__ __forceinline__ typename F::T table2(const typename F::T* __restrict__ lo, const typename F::T* __restrict__ hi,
                                                uint32_t lb, uint32_t e) {
  typename F::T a = lo[e & ((1u << lb) - 1u)];
  uint32_t h = e >> lb;
  return h ? F::mul(a, hi[h]) : a;
}

// ---------------------------------------------------------------------------------------------- one register step
template <class F, int LOG_S, int LAYOUT, bool INV, int U>
__device__ __forceinline__ void pass_step(const PassArgs<F>& A, typename F::T* tile, const typename F::T* tws,
                                          const typename F::T* qtab, long long in_base, long long out_base, uint32_t w0,
                                          uint32_t y, uint32_t z, bool tws_sync = false) {
  typedef typename F::T T;
  typedef PassCfg<F, LOG_S, LAYOUT> C;
  constexpr int LR = C::lr(U);
  constexpr int R = 1 << LR;
  constexpr int LOGP = C::logp(U);       // sub-problem size after this step
  constexpr int LOGPP = LOGP + LR;       // before this step
  constexpr int LOGQ = LOG_S - LR;       // items per problem
  constexpr bool FIRST = (U == 0), LAST = (U == C::NSTEPS - 1);
  constexpr bool W_FASTEST = (LAYOUT == 0) || (LAST && !FIRST);
  constexpr int NITEMS = (1 << LOGQ) * C::TW;

  // fewer items than threads (tiny transforms): some threads never enter the loop, so the table barrier goes here
  if (FIRST && !LAST && tws_sync && NITEMS < C::NT) __syncthreads();
  for (int item = threadIdx.x; item < NITEMS; item += C::NT) {
    int q, w;
    if (W_FASTEST) {
      w = item & (C::TW - 1);
      q = item >> C::LOG_TW;
    } else {
      q = item & ((1 << LOGQ) - 1);
      w = item >> LOGQ;
    }
    const int blk = q >> LOGP;
    const int m_rest = q & ((1 << LOGP) - 1);
    const int s0 = (blk << LOGPP) + m_rest;  // element r lives at s0 + (r << LOGP)
    const uint32_t wg = w0 + (uint32_t)w;
    // Problems past nw_in are all-zero inputs (the LDE's padding columns inside a partly filled tile: the 2-column
    // f128 traces fill half a tile): their transform is zero, so they are neither loaded nor transformed nor exchanged;
    // the last step just stores zeros. In every step but the last of a contiguous pass a warp works on one problem,
    // so the skipped work is skipped by whole warps.
    // (Compiled in per field, FieldCfg::SKIP_DEAD: the Goldilocks configurations fill their tiles, and the extra
    // predicate costs the 2^22 x 100 commit 2.8 %.)
    const bool in_range = wg < A.nw_in;
    const bool live = C::FC::SKIP_DEAD ? in_range : true;
    T a[R];
    if (FIRST ? !in_range : (C::FC::SKIP_DEAD && !in_range)) {
#pragma unroll
      for (int r = 0; r < R; r++) a[r] = F::zero();
    } else if (FIRST) {
      {
        const T* src = A.in + in_base + (long long)wg * A.in_wstride;
#pragma unroll
        for (int r = 0; r < R; r++) a[r] = src[(long long)(s0 + (r << LOGP)) * A.in_sstride];
        if (A.has_pm == 1) {
          const T* lo = A.pm_lo + (long long)z * A.pm_zstride_lo;
          const T* hi = A.pm_hi + (long long)z * A.pm_zstride_hi;
#pragma unroll
          for (int r = 0; r < R; r++) {
            uint32_t e = ((uint32_t)(s0 + (r << LOGP)) * A.pm_s + wg * A.pm_w) & A.pm_mask;
            if (e) a[r] = F::mul(a[r], table2<F>(lo, hi, A.pm_lb, e));
          }
        } else if (A.has_pm == 2) {
          const T* hi = A.pm_hi + (long long)z * A.pm_zstride_hi;
          T f[R];
#pragma unroll
          for (int r = 0; r < R; r++) f[r] = hi[s0 + (r << LOGP)];
#pragma unroll
          for (int r = 0; r < R; r++) a[r] = F::mul(a[r], f[r]);
        }
        if (LAYOUT == 1 && A.has_q) {
#pragma unroll
          for (int r = 0; r < R; r++) a[r] = F::mul(a[r], qtab[s0 + (r << LOGP)]);
        }
      }
    } else {
#pragma unroll
      for (int r = 0; r < R; r++) a[r] = tile[C::phys(s0 + (r << LOGP), w)];
    }

    typedef RadixNet<F, LR, INV> Net;
    if (live) Net::run(a);

    if (!LAST) {
      // The block's first step 0 runs while the shared twiddle table is still being filled (its loads overlap the
      // tile loads and the first butterflies); everyone's table stores must have landed before the first read.
      // With NITEMS >= NT every thread runs the same number of item iterations, so the barrier is uniform.
      if (FIRST && tws_sync && NITEMS >= C::NT && item == (int)threadIdx.x) __syncthreads();
      // twiddle w_{2^LOGPP}^(m_rest*j) from this step's [j][m_rest] sub-table; then scatter digit j
      if (live) {
#pragma unroll
        for (int j = 0; j < R; j++) {
          T v = a[j];
          if (j > 0) v = F::mul(v, tws[C::tws_off(U) + (j << LOGP) + m_rest]);
          else if (Net::LAZY) v = Net::canon(v);  // the next step's butterflies want canonical inputs
          tile[C::phys(s0 + (j << LOGP), w)] = v;
        }
      }
    } else {
      // natural output index: digits j_0 (most significant in blk) ... j_{U-1}, then this step's j
      int nat0 = 0, shift_out = 0, rem = blk;
      {
        // blk = sum_v j_v << (bits below v); walk from the least significant digit (step U-1) upwards
        int widths[4] = {0, 0, 0, 0};
#pragma unroll
        for (int v = 0; v < U; v++) widths[v] = C::lr(v);
        int pos[4] = {0, 0, 0, 0};
        int acc = 0;
#pragma unroll
        for (int v = 0; v < U; v++) {
          pos[v] = acc;  // weight of j_v in the natural index
          acc += widths[v];
        }
        shift_out = acc;
#pragma unroll
        for (int v = U - 1; v >= 0; v--) {
          int d = rem & ((1 << widths[v]) - 1);
          rem >>= widths[v];
          nat0 |= d << pos[v];
        }
      }
      if (wg < A.nw_out) {
        const long long dst_off = out_base + (long long)wg * A.out_wstride;
        T* dst = A.out + dst_off;
        uint32_t ebase = 0, estep = 0;
        if (A.has_tm) {
          uint32_t f = wg * A.tm_w + y * A.tm_y;
          ebase = (uint32_t)nat0 * f;
          estep = f << shift_out;
        }
#pragma unroll
        for (int j = 0; j < R; j++) {
          T v = a[j];
          if (!live) {
            // zero stays zero
          } else if (A.has_tm) {
            uint32_t e = (ebase + (uint32_t)j * estep) & A.tm_mask;
            v = F::mul(v, table2<F>(A.tm_lo, A.tm_hi, A.tm_lb, e));
          } else if (A.has_const) {
            v = F::mul(v, A.post_const);
          } else if (Net::LAZY) {
            v = Net::canon(v);  // results leave the chip canonical, as the reference's are
          }
          const uint32_t jj = (uint32_t)(nat0 + (j << shift_out));
          if (A.n_peers) {
            A.out_peers[jj >> A.peer_shift][dst_off + (long long)(jj & ((1u << A.peer_shift) - 1u)) * A.out_jstride] = v;
          } else {
            dst[(long long)jj * A.out_jstride] = v;
          }
        }
      }
    }
  }
}

template <class F, int LOG_S, int LAYOUT, bool INV>
__global__ void __launch_bounds__(PassCfg<F, LOG_S, LAYOUT>::NT, FieldCfg<F>::MIN_BLOCKS) ntt_pass_kernel(const PassArgs<F> A) {
  typedef typename F::T T;
  typedef PassCfg<F, LOG_S, LAYOUT> C;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* tile = reinterpret_cast<T*>(smem_raw);
  T* tws = tile + C::TILE_ELEMS;
  T* qtab = tws + C::TWS_ELEMS;  // S entries, present when has_q
  if (C::NSTEPS > 1) {
    for (int i = threadIdx.x; i < C::TWS_ELEMS; i += C::NT) tws[i] = A.tws[i];
  }
  uint32_t bx = blockIdx.x, gx = gridDim.x, z = blockIdx.z;
  if (A.zfast) {
    z = bx % A.zfast;
    bx /= A.zfast;
    gx /= A.zfast;
  }
  const uint32_t nx = A.nx ? A.nx : gx, ny = A.ny ? A.ny : gridDim.y;
  bool tws_ready = false;  // becomes true at the first barrier after the table load above
  for (uint32_t y = blockIdx.y; y < ny; y += gridDim.y) {
    if (LAYOUT == 1 && A.has_q) {
      // qtab[i] = q^i by doubling: level k appends q^(2^k) * qtab[0 .. 2^k)
      T qp = table2<F>(A.tm_lo, A.tm_hi, A.tm_lb, (y * A.q_y) & A.tm_mask);
      if (A.q_z) qp = F::mul(qp, A.q_z[(long long)z * A.q_zstride]);
      __syncthreads();  // previous y's readers are done
      if (threadIdx.x == 0) qtab[0] = F::one();
      __syncthreads();
#pragma unroll 1
      for (int k = 0; k < LOG_S; k++) {
        for (int i = threadIdx.x; i < (1 << k); i += C::NT) qtab[(1 << k) + i] = F::mul(qtab[i], qp);
        qp = F::mul(qp, qp);
        __syncthreads();
      }
      tws_ready = true;
    }
    const long long in_base = (long long)y * A.in_ystride + (long long)z * A.in_zstride;
    const long long out_base = (long long)y * A.out_ystride + (long long)z * A.out_zstride;
    for (uint32_t x = bx; x < nx; x += gx) {
      const uint32_t w0 = x * C::TW;
      if (w0 >= A.nw_in) {
        // a tile of zero problems (the LDE's padding columns): the transform of zeros is zeros
        for (int i = threadIdx.x; i < C::S * C::TW; i += C::NT) {
          const uint32_t wg = w0 + (uint32_t)(i & (C::TW - 1));
          const uint32_t jj = (uint32_t)(i >> C::LOG_TW);
          if (wg >= A.nw_out) continue;
          if (A.n_peers)
            A.out_peers[jj >> A.peer_shift][out_base + (long long)wg * A.out_wstride +
                                            (long long)(jj & ((1u << A.peer_shift) - 1u)) * A.out_jstride] = F::zero();
          else
            A.out[out_base + (long long)wg * A.out_wstride + (long long)jj * A.out_jstride] = F::zero();
        }
        continue;
      }
      pass_step<F, LOG_S, LAYOUT, INV, 0>(A, tile, tws, qtab, in_base, out_base, w0, y, z, !tws_ready);
      tws_ready = true;
      if (C::NSTEPS > 1) {
        __syncthreads();
        pass_step<F, LOG_S, LAYOUT, INV, (C::NSTEPS > 1 ? 1 : 0)>(A, tile, tws, qtab, in_base, out_base, w0, y, z);
      }
      if (C::NSTEPS > 2) {
        __syncthreads();
        pass_step<F, LOG_S, LAYOUT, INV, (C::NSTEPS > 2 ? 2 : 0)>(A, tile, tws, qtab, in_base, out_base, w0, y, z);
      }
      if (C::NSTEPS > 3) {
        __syncthreads();
        pass_step<F, LOG_S, LAYOUT, INV, (C::NSTEPS > 3 ? 3 : 0)>(A, tile, tws, qtab, in_base, out_base, w0, y, z);
      }
      if (C::NSTEPS > 1) __syncthreads();  // the tile is rewritten by the next walk step
    }
  }
}

}  // namespace wg
