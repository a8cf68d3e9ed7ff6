// C ABI (include/winter_gpu.h) and host-side orchestration of the trace-commitment path:
//   Prover::build_trace_commitment (prover/src/lib.rs:615-670)
//     = ColMatrix::interpolate_columns (col_matrix.rs:196-206)
//     + RowMatrix::evaluate_polys_over::<8> (row_matrix.rs:82-133)
//     + RowMatrix::commit_to_comb_rows (row_matrix.rs:204-238) -> MerkleTree::new (crypto/src/merkle/mod.rs:117-136)
// Host code only builds small twiddle tables, picks pass shapes and launches kernels. There is no CPU
// implementation of the path in this library.
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <map>
#include <mutex>
#include <string>
#include <tuple>
#include <vector>

#include "../../include/winter_gpu.h"
#include "field.cuh"
#include "hash.cuh"
#include "hash_kernels.cuh"
#include "ntt_launch.cuh"
#include "split.cuh"
#include "fri.cuh"

namespace wg {

// ------------------------------------------------------------------------------------------------ errors
static thread_local std::string g_last_error;
static std::atomic<uint64_t> g_launches{0};
static int g_device = -1;
static std::mutex g_mu;

static int fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_last_error = buf;
  return code;
}
#define CUDA_TRY(expr)                                                                           \
  do {                                                                                           \
    cudaError_t e__ = (expr);                                                                    \
    if (e__ != cudaSuccess) return fail(WG_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(e__));  \
  } while (0)
#define WG_TRY(expr)          \
  do {                        \
    int rc__ = (expr);        \
    if (rc__ != WG_OK) return rc__; \
  } while (0)

// memory the library parks between calls (defined with the commit entry points below)
static void cache_release_all();
static void arena_release_if_idle();
// cudaMalloc that gives parked memory back before reporting failure
static cudaError_t dev_malloc(void** p, size_t bytes) {
  cudaError_t e = cudaMalloc(p, bytes);
  if (e == cudaSuccess) return e;
  cudaGetLastError();
  cache_release_all();
  e = cudaMalloc(p, bytes);
  if (e == cudaSuccess) return e;
  cudaGetLastError();
  arena_release_if_idle();
  return cudaMalloc(p, bytes);
}

static inline bool is_pow2(uint64_t n) { return n && !(n & (n - 1)); }
static inline size_t dlen(int hash) { return (size_t)hash_digest_bytes(hash); }  // 24 for Blake3_192, else 32
static inline bool hash_ok(int hash) { return hash == WG_HASH_BLAKE3_256 || hash == WG_HASH_SHA3_256 || hash == WG_HASH_BLAKE3_192; }
static inline int ilog2(uint64_t n) { return 63 - __builtin_clzll(n); }

static int ensure_device() {
  if (g_device < 0) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
      return fail(WG_ERR_CUDA, "no CUDA device (%s); this library has no CPU fallback", cudaGetErrorString(e));
    std::lock_guard<std::mutex> lk(g_mu);
    if (g_device < 0) g_device = 0;
  }
  CUDA_TRY(cudaSetDevice(g_device));
  return WG_OK;
}

// ------------------------------------------------------------------------------------------------ tables
// Small device-resident tables, built on the host with the same field code the kernels use, cached forever.
template <class F>
struct Tables {
  typedef typename F::T T;
  std::mutex mu;
  std::map<std::pair<int, int>, T*> tws;                       // (log_s, inv) -> w_S^(+-e), e < S
  struct Two {
    T* lo = nullptr;
    T* hi = nullptr;
    int lb = 0;
  };
  std::map<std::pair<int, int>, Two> tm;                       // (log_n, inv) -> w_N^(+-e) two-level (inv: times n^-1)
  std::map<std::tuple<int, int, uint64_t, uint64_t>, Two> pm;  // (log_n, log_b, offset) -> per-coset powers
  std::map<std::tuple<int, int, int, uint64_t, uint64_t>, Two> pw;  // (log_root, inv, log_len, scale) -> split.cuh PowTab

  static int upload(const std::vector<T>& h, T** out) {
    T* d = nullptr;
    CUDA_TRY(cudaMalloc(&d, h.size() * sizeof(T)));
    CUDA_TRY(cudaMemcpy(d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice));
    *out = d;
    return WG_OK;
  }
  static T root(int log_n, bool inv) {
    T r = f_root_of_unity<F>(log_n);
    return inv ? f_inv<F>(r) : r;
  }

  // in-block twiddles in the per-step [j][m_rest] layout of PassCfg::tws_off (ntt.cuh)
  int get_tws(int log_s, bool inv, const T** out) {
    std::lock_guard<std::mutex> lk(mu);
    auto key = std::make_pair(log_s, (int)inv);
    auto it = tws.find(key);
    if (it == tws.end()) {
      const int maxlr = FieldCfg<F>::MAXLR;
      const int nsteps = (log_s + maxlr - 1) / maxlr;
      std::vector<T> pw((size_t)1 << log_s);  // w_S^e
      T w = root(log_s, inv), acc = F::one();
      for (auto& x : pw) {
        x = acc;
        acc = F::mul(acc, w);
      }
      std::vector<T> h;
      int logpp = log_s;
      for (int u = 0; u + 1 < nsteps; u++) {
        const int lr = maxlr, logp = logpp - lr;
        for (int j = 0; j < (1 << lr); j++)
          for (int m = 0; m < (1 << logp); m++) h.push_back(pw[((size_t)m * j) << (log_s - logpp)]);
        logpp = logp;
      }
      if (h.empty()) h.push_back(F::one());
      T* d;
      WG_TRY(upload(h, &d));
      it = tws.emplace(key, d).first;
    }
    *out = it->second;
    return WG_OK;
  }

  // lookup only (never builds, never allocates): are the tables of a length-2^log_n inverse two-pass transform resident?
  bool have_tws(int log_s, bool inv) {
    std::lock_guard<std::mutex> lk(mu);
    return tws.find(std::make_pair(log_s, (int)inv)) != tws.end();
  }
  bool have_tm(int log_n, bool inv) {
    std::lock_guard<std::mutex> lk(mu);
    return tm.find(std::make_pair(log_n, (int)inv)) != tm.end();
  }

  // two-level power table of `base`: lo[e] = scale*base^e (e < 2^lb), hi[h] = base^(h << lb) (h < 2^(log_n-lb))
  static void fill_two(T base, T scale, int log_n, int lb, T* lo, T* hi) {
    T acc = scale;
    for (size_t e = 0; e < ((size_t)1 << lb); e++) {
      lo[e] = acc;
      acc = F::mul(acc, base);
    }
    T step = f_pow<F>(base, (uint64_t)1 << lb);
    acc = F::one();
    for (size_t h = 0; h < ((size_t)1 << (log_n - lb)); h++) {
      hi[h] = acc;
      acc = F::mul(acc, step);
    }
  }

  int get_tm(int log_n, bool inv, Two* out) {
    std::lock_guard<std::mutex> lk(mu);
    auto key = std::make_pair(log_n, (int)inv);
    auto it = tm.find(key);
    if (it == tm.end()) {
      Two t;
      t.lb = (log_n + 1) / 2;
      std::vector<T> lo((size_t)1 << t.lb), hi((size_t)1 << (log_n - t.lb));
      T scale = inv ? f_inv<F>(F::from_u64((uint64_t)1 << log_n)) : F::one();
      fill_two(root(log_n, inv), scale, log_n, t.lb, lo.data(), hi.data());
      WG_TRY(upload(lo, &t.lo));
      WG_TRY(upload(hi, &t.hi));
      it = tm.emplace(key, t).first;
    }
    *out = it->second;
    return WG_OK;
  }

  // coset z of the LDE domain is offset * w_{N*B}^z; table z holds powers e < N of that base
  int get_pm(int log_n, int log_b, T offset, Two* out, long long* zstride_lo, long long* zstride_hi) {
    std::lock_guard<std::mutex> lk(mu);
    uint64_t o_lo, o_hi;
    memcpy(&o_lo, &offset, 8);
    o_hi = 0;
    if (sizeof(T) == 16) memcpy(&o_hi, (const char*)&offset + 8, 8);
    auto key = std::make_tuple(log_n, log_b, o_lo, o_hi);
    auto it = pm.find(key);
    const int lb = log_n <= MAX_LOG_S ? log_n : (log_n + 1) / 2;
    const size_t nlo = (size_t)1 << lb, nhi = (size_t)1 << (log_n - lb);
    if (it == pm.end()) {
      // keyed by the caller's offset: a caller that walks through many offsets must not grow device memory for ever
      if (pm.size() >= 32) {
        cudaDeviceSynchronize();  // kernels that still read the old tables
        for (auto& kv : pm) {
          cudaFree(kv.second.lo);
          cudaFree(kv.second.hi);
        }
        pm.clear();
      }
      Two t;
      t.lb = lb;
      const size_t B = (size_t)1 << log_b;
      std::vector<T> lo(nlo * B), hi(nhi * B);
      T g = f_root_of_unity<F>(log_n + log_b);
      T base = offset;
      for (size_t z = 0; z < B; z++) {
        fill_two(base, F::one(), log_n, lb, lo.data() + z * nlo, hi.data() + z * nhi);
        base = F::mul(base, g);
      }
      WG_TRY(upload(lo, &t.lo));
      WG_TRY(upload(hi, &t.hi));
      it = pm.emplace(key, t).first;
    }
    *out = it->second;
    *zstride_lo = (long long)nlo;
    *zstride_hi = (long long)nhi;
    return WG_OK;
  }

  // scale * w^k, k < 2^log_len, w = the root of order 2^log_root (inverted if inv), as a two-level table (split.cuh)
  int get_pow(int log_root, bool inv, int log_len, T scale, Two* out) {
    std::lock_guard<std::mutex> lk(mu);
    uint64_t s_lo, s_hi = 0;
    memcpy(&s_lo, &scale, 8);
    if (sizeof(T) == 16) memcpy(&s_hi, (const char*)&scale + 8, 8);
    auto key = std::make_tuple(log_root, (int)inv, log_len, s_lo, s_hi);
    auto it = pw.find(key);
    if (it == pw.end()) {
      if (pw.size() >= 32) {  // keyed by the caller's offset, like pm
        cudaDeviceSynchronize();
        for (auto& kv : pw) {
          cudaFree(kv.second.lo);
          cudaFree(kv.second.hi);
        }
        pw.clear();
      }
      Two t;
      t.lb = (log_len + 1) / 2;
      std::vector<T> lo((size_t)1 << t.lb), hi((size_t)1 << (log_len - t.lb));
      fill_two(root(log_root, inv), scale, log_len, t.lb, lo.data(), hi.data());
      WG_TRY(upload(lo, &t.lo));
      WG_TRY(upload(hi, &t.hi));
      it = pw.emplace(key, t).first;
    }
    *out = it->second;
    return WG_OK;
  }
};

template <class F>
static Tables<F>& tables() {
  static Tables<F> t;
  return t;
}

template <class F>
static int launch_pass(int layout, bool inv, int log_s, const PassArgs<F>& a, dim3 grid, cudaStream_t stream) {
  cudaError_t e = launch_ntt_pass<F>(layout, inv, log_s, a, grid, stream);
  if (e != cudaSuccess) return fail(WG_ERR_CUDA, "ntt pass launch (layout %d log_s %d): %s", layout, log_s, cudaGetErrorString(e));
  g_launches++;
  return WG_OK;
}
static inline unsigned cdiv(uint64_t a, uint64_t b) { return (unsigned)((a + b - 1) / b); }
// Tuning knobs, read from the environment ONCE (first use), never on the per-call path.
struct Tunables {
  int lde_q, walk_a, walk_b, zfast;
  Tunables() {
    auto env_int = [](const char* name, int dflt) {
      const char* s = getenv(name);
      return s && *s ? atoi(s) : dflt;
    };
    lde_q = env_int("WG_LDE_Q", 2);
    walk_a = env_int("WG_WALK_A", 1);
    walk_b = env_int("WG_WALK_B", 13);
    zfast = env_int("WG_ZFAST", 1);
  }
};
static const Tunables& tunables() {
  static const Tunables t;
  return t;
}

// ------------------------------------------------------------------------------------------------ phase timing
// Optional CUDA-event timing of the phases of a commit, on the launching stream (bench.py's roofline figure).
enum { PH_INTERP_A = 0, PH_INTERP_B, PH_LDE_A, PH_LDE_B, PH_HASH, PH_MERKLE, PH_COUNT };
struct ProfRec {
  int phase;
  cudaEvent_t a, b;
};
static bool g_prof_on = false;
static std::vector<ProfRec> g_prof;
static std::vector<cudaEvent_t> g_prof_pool;
static cudaEvent_t prof_event() {
  cudaEvent_t e;
  if (!g_prof_pool.empty()) {
    e = g_prof_pool.back();
    g_prof_pool.pop_back();
  } else {
    cudaEventCreate(&e);
  }
  return e;
}
struct ProfScope {
  int idx = -1;
  cudaStream_t s;
  cudaEvent_t end = nullptr;
  ProfScope(int phase, cudaStream_t stream) : s(stream) {
    if (!g_prof_on) return;
    std::lock_guard<std::mutex> lk(g_mu);  // g_prof / g_prof_pool are shared with wg_profile_read and other callers
    ProfRec r{phase, prof_event(), prof_event()};
    cudaEventRecord(r.a, s);
    g_prof.push_back(r);
    idx = (int)g_prof.size() - 1;
    end = r.b;
  }
  ~ProfScope() {
    if (idx >= 0) cudaEventRecord(end, s);
  }
};

template <class F>
static PassArgs<F> blank_args() {
  PassArgs<F> a;
  memset(&a, 0, sizeof a);
  return a;
}

// split of a length-2^log_n transform: strided pass of 2^na points, then contiguous pass of 2^nb points
static inline void split_passes(int log_n, int* na, int* nb) {
  if (log_n <= MAX_LOG_S) {
    *na = 0;
    *nb = log_n;
  } else {
    *na = log_n / 2;
    *nb = log_n - *na;
  }
}

template <class F>
static bool elem_is_canonical(const typename F::T& v);
template <>
bool elem_is_canonical<Gl64>(const uint64_t& v) { return v < Gl64::P; }
template <>
bool elem_is_canonical<F128>(const u128& v) { return u128_lt(v, F128::modulus()); }

template <class F>
static int check_domain(uint64_t n, uint64_t blowup) {
  if (!is_pow2(n) || !is_pow2(blowup)) return fail(WG_ERR_NOT_POW2, "size %llu / blowup %llu must be powers of two", (unsigned long long)n, (unsigned long long)blowup);
  if (n < 2) return fail(WG_ERR_ARG, "transform length must be at least 2");
  if (ilog2(n) + ilog2(blowup) > F::kTwoAdicity) return fail(WG_ERR_TOO_LARGE, "domain 2^%d exceeds the field's two-adicity", ilog2(n) + ilog2(blowup));
  return WG_OK;
}

// Transforms longer than 2^g_split_log points take an outer decimation step by R = 2^lr <= 16 around 2^(log_n - lr)-point
// transforms (split.cuh); the inner call splits again when it is still too long. Default: the two-pass plan's limit.
static std::atomic<int> g_split_log{2 * MAX_LOG_S};
// Inner length the split aims at. 2^12-point passes run one block per SM at 1.3x the per-element cost of 2^11-point ones,
// so (a) a transform that must split takes the radix that leaves 2^22-point inner transforms (2^25 -> 8 x 2^22, not
// 2 x 2^24), and (b) the LDE of a 2^24-row trace, whose two passes would both be 2^12-point, PREFERS radix 4 around 2^22
// when its scratch is there (2^24 x 32 columns on one GPU: LDE 92.5 -> 82.4 ms; a 2^23-row trace has one 2^11-point pass
// already and loses with radix 2: 78.8 -> 83.7 ms, so a preferred split needs radix >= 4; profiles/r2_split_1gpu.jsonl).
// The interpolation gains nothing from (b) (12.5 ms either way) and only splits above g_split_log.
static std::atomic<int> g_split_pref{22};
static inline int split_target(int log_n) {
  const int hard = g_split_log.load(), pref = g_split_pref.load();
  const int over = log_n - (pref < hard ? pref : hard);
  return over <= 0 ? 0 : (over < 4 ? over : 4);
}
static inline int split_lr(uint64_t n) { return ilog2(n) > g_split_log.load() ? split_target(ilog2(n)) : 0; }
static inline int split_lr_pref(uint64_t n) {
  const int lr = split_target(ilog2(n));
  return split_lr(n) ? lr : (lr >= 2 ? lr : 0);
}

// ------------------------------------------------------------------------------------------------ interpolate
// c_j = n^-1 sum_k v_k w_n^(-jk), natural order in and out (math/src/fft/serial.rs:64-73), for n_cols
// contiguous columns.
template <class F>
static uint64_t interpolate_scratch_elems(uint32_t n_cols, uint64_t n) {
  if (split_lr(n)) return 2 * (uint64_t)n_cols * n;  // the R*C sub-columns and the inner transform's scratch
  return ilog2(n) <= MAX_LOG_S ? 0 : (uint64_t)n_cols * n;
}

// n <= 2^24 (one or two on-chip passes). d_out == d_in is allowed for two-pass lengths (pass A reads d_in into the scratch).
template <class F>
static int interpolate_columns_base(const typename F::T* d_in, uint32_t n_cols, uint64_t n, typename F::T* d_out,
                                    typename F::T* d_scratch, uint64_t scratch_elems, cudaStream_t stream) {
  typedef typename F::T T;
  WG_TRY(check_domain<F>(n, 1));
  if (ilog2(n) > 2 * MAX_LOG_S) return fail(WG_ERR_TOO_LARGE, "transform length above 2^%d needs the split plan", 2 * MAX_LOG_S);
  if (n_cols == 0) return WG_OK;
  if (n_cols > 65535) return fail(WG_ERR_UNSUPPORTED, "more than 65535 columns per call");
  const int log_n = ilog2(n);
  int na, nb;
  split_passes(log_n, &na, &nb);
  Tables<F>& tb = tables<F>();
  if (na == 0) {
    PassArgs<F> a = blank_args<F>();
    a.in = d_in;
    a.out = d_out;
    a.in_wstride = (long long)n;
    a.in_sstride = 1;
    a.out_wstride = (long long)n;
    a.out_jstride = 1;
    a.nw_in = a.nw_out = n_cols;
    WG_TRY(tb.get_tws(nb, true, &a.tws));
    a.has_const = 1;
    a.post_const = f_inv<F>(F::from_u64(n));
    const int tw = 1 << pass_log_tw<F>(nb);
    ProfScope ps(PH_INTERP_B, stream);
    return launch_pass<F>(1, true, nb, a, dim3(cdiv(n_cols, tw), 1, 1), stream);
  }
  if (scratch_elems < (uint64_t)n_cols * n || !d_scratch) return fail(WG_ERR_NOMEM, "interpolate scratch too small");
  const uint64_t K1 = (uint64_t)1 << na, K2 = (uint64_t)1 << nb;
  typename Tables<F>::Two tm;
  WG_TRY(tb.get_tm(log_n, true, &tm));
  {  // pass A: for each k2, transform over k1 (stride K2); Y[mA*K2 + k2] *= n^-1 w_n^(-mA*k2)
    ProfScope ps(PH_INTERP_A, stream);
    PassArgs<F> a = blank_args<F>();
    a.in = d_in;
    a.out = d_scratch;
    a.in_ystride = a.out_ystride = (long long)n;
    a.in_wstride = a.out_wstride = 1;
    a.in_sstride = a.out_jstride = (long long)K2;
    a.nw_in = a.nw_out = (uint32_t)K2;
    WG_TRY(tb.get_tws(na, true, &a.tws));
    a.has_tm = 1;
    a.tm_lo = tm.lo;
    a.tm_hi = tm.hi;
    a.tm_lb = tm.lb;
    a.tm_mask = (uint32_t)(n - 1);
    a.tm_w = 1;
    const int tw = 1 << pass_log_tw<F>(na);
    WG_TRY(launch_pass<F>(0, true, na, a, dim3(cdiv(K2, tw), n_cols, 1), stream));
  }
  {  // pass B: for each mA, transform over k2 (contiguous); c[mA + K1*mB]
    ProfScope ps(PH_INTERP_B, stream);
    PassArgs<F> a = blank_args<F>();
    a.in = d_scratch;
    a.out = d_out;
    a.in_ystride = a.out_ystride = (long long)n;
    a.in_wstride = (long long)K2;
    a.in_sstride = 1;
    a.out_wstride = 1;
    a.out_jstride = (long long)K1;
    a.nw_in = a.nw_out = (uint32_t)K1;
    WG_TRY(tb.get_tws(nb, true, &a.tws));
    const int tw = 1 << pass_log_tw<F>(nb);
    WG_TRY(launch_pass<F>(1, true, nb, a, dim3(cdiv(K1, tw), n_cols, 1), stream));
  }
  return WG_OK;
}

// ------------------------------------------------------------------------------------------------ LDE
// out[(k)*row_width + c] = P_c(offset * w_{nB}^k), k = j*B + z: coset z (shift offset*w_{nB}^z) evaluated at
// w_n^j. Natural row order, row-major, pad columns zero (row_matrix.rs:82-133; segments.rs:53-207).
template <class F>
static uint64_t lde_scratch_elems(uint32_t n_cols, uint64_t n, uint64_t blowup) {
  // split plan: the de-interleaved sub-polynomials (n), their LDE (n*B) and the inner LDE's scratch (n*B) per column
  if (split_lr_pref(n)) return (uint64_t)n_cols * n * (2 * blowup + 1);
  return ilog2(n) <= MAX_LOG_S ? 0 : (uint64_t)n_cols * n * blowup;
}
// the least the LDE can work with (two-pass plan where it applies): for callers that ran out of memory
template <class F>
static uint64_t lde_scratch_elems_min(uint32_t n_cols, uint64_t n, uint64_t blowup) {
  if (split_lr(n)) return (uint64_t)n_cols * n * (2 * blowup + 1);
  return ilog2(n) <= MAX_LOG_S ? 0 : (uint64_t)n_cols * n * blowup;
}

template <class F>
static int lde_base(const typename F::T* d_polys, uint32_t n_cols, uint64_t n, uint64_t blowup, typename F::T offset,
                    typename F::T* d_lde, uint64_t row_width, typename F::T* d_scratch, uint64_t scratch_elems,
                    cudaStream_t stream, typename F::T* const* d_peers = nullptr, uint32_t n_peers = 0) {
  // d_peers (multi-GPU): row block g of the result, rows [g*L/G, (g+1)*L/G), is written to d_peers[g] (a buffer of
  // L/G x row_width elements, usually peer-mapped memory of GPU g) instead of d_lde; see PassArgs::out_peers.
  typedef typename F::T T;
  WG_TRY(check_domain<F>(n, blowup));
  if (ilog2(n) > 2 * MAX_LOG_S) return fail(WG_ERR_TOO_LARGE, "transform length above 2^%d needs the split plan", 2 * MAX_LOG_S);
  if (n_peers) {
    if (!d_peers || !is_pow2(n_peers) || n_peers > (uint32_t)PassArgs<F>::MAX_PEERS)
      return fail(WG_ERR_ARG, "number of row-block owners must be a power of two <= %d", PassArgs<F>::MAX_PEERS);
    int sa, sb;
    split_passes(ilog2(n), &sa, &sb);
    if (((uint64_t)1 << sb) < n_peers) return fail(WG_ERR_UNSUPPORTED, "trace too short for %u row blocks", n_peers);
  }
  if (F::is_zero(offset)) return fail(WG_ERR_ZERO_OFFSET, "domain offset cannot be zero");
  if (!elem_is_canonical<F>(offset)) return fail(WG_ERR_ARG, "domain offset is not a reduced field element");
  if (row_width < n_cols) return fail(WG_ERR_ARG, "row_width < n_cols");
  if (n_cols == 0) return WG_OK;
  if (blowup > 65535) return fail(WG_ERR_UNSUPPORTED, "blowup too large");
  const int log_n = ilog2(n), log_b = ilog2(blowup);
  int na, nb;
  split_passes(log_n, &na, &nb);
  Tables<F>& tb = tables<F>();
  typename Tables<F>::Two pm;
  long long pzl, pzh;
  WG_TRY(tb.get_pm(log_n, log_b, offset, &pm, &pzl, &pzh));
  const uint32_t B = (uint32_t)blowup;
  if (na == 0) {
    PassArgs<F> a = blank_args<F>();
    a.in = d_polys;
    a.out = d_lde;
    a.in_wstride = (long long)n;
    a.in_sstride = 1;
    a.out_wstride = 1;
    a.out_zstride = (long long)row_width;
    a.out_jstride = (long long)B * row_width;
    a.nw_in = n_cols;
    a.nw_out = (uint32_t)row_width;
    WG_TRY(tb.get_tws(nb, false, &a.tws));
    a.has_pm = 1;
    a.pm_lo = pm.lo;
    a.pm_hi = pm.hi;
    a.pm_zstride_lo = pzl;
    a.pm_zstride_hi = pzh;
    a.pm_s = 1;
    a.pm_w = 0;
    a.pm_lb = pm.lb;
    a.pm_mask = (uint32_t)(n - 1);
    if (n_peers) {
      a.n_peers = n_peers;
      a.peer_shift = (uint32_t)(nb - ilog2(n_peers));
      for (uint32_t g = 0; g < n_peers; g++) a.out_peers[g] = d_peers[g];
    }
    const int tw = 1 << pass_log_tw<F>(nb);
    ProfScope ps(PH_LDE_B, stream);
    return launch_pass<F>(1, false, nb, a, dim3(cdiv(row_width, tw), 1, B), stream);
  }
  const uint64_t N1 = (uint64_t)1 << na, N2 = (uint64_t)1 << nb;
  // columns per chunk limited by the scratch (B*n elements per column)
  uint64_t chunk = scratch_elems / (n * blowup);
  if (!d_scratch || chunk == 0) return fail(WG_ERR_NOMEM, "lde scratch too small");
  if (chunk >= n_cols && row_width <= 65528) {
    chunk = row_width;  // everything (pad columns included) in one sweep
  } else {
    if (chunk > 65528) chunk = 65528;
    chunk = chunk / 8 * 8;
    if (chunk == 0) return fail(WG_ERR_NOMEM, "lde scratch must hold at least 8 columns (8*n*blowup elements)");
  }
  typename Tables<F>::Two tm;
  WG_TRY(tb.get_tm(log_n, false, &tm));
  // With room for a 2^nb-entry table next to the tile (at the kernel's resident-block count), pass B' multiplies by
  // q^m2, q = t_z * w_n^j1, which carries both the low half of the coset shift and the inter-pass twiddle; pass A' is
  // then left with the direct table t_z^(N2*m1). Otherwise pass A' applies shift and twiddle through two-level tables.
  const size_t smem_b = ntt_pass_smem<F>(1, nb) + (sizeof(T) << nb);
  // 0: never, 1: only when the table fits at the compiled residency, 2 (default): whenever it fits. At 2^12-point passes
  // (2^23, 2^24 rows) the table leaves room for one block per SM only, and that is still the faster plan: 2^24 x 32
  // columns 175.5 -> 148.2 ms (strided pass 74.2 -> 51.3 ms: one table product per element instead of four).
  const int q_mode = tunables().lde_q;
  const bool use_q = pm.lb == nb && q_mode && smem_b * (q_mode == 2 ? 1 : FieldCfg<F>::MIN_BLOCKS) <= SMEM_PER_SM;
  const unsigned walk_a = (unsigned)tunables().walk_a, walk_b = (unsigned)tunables().walk_b;
  for (uint64_t c0 = 0; c0 < row_width; c0 += chunk) {
    const uint32_t nc_out = (uint32_t)((row_width - c0) < chunk ? (row_width - c0) : chunk);  // incl. pad columns
    const uint32_t nc_in = c0 >= n_cols ? 0 : (uint32_t)((n_cols - c0) < chunk ? (n_cols - c0) : chunk);
    if (nc_in > 0) {
      // pass A': coefficient m = m1*N2 + m2; scale by t_z^m, transform over m1 (stride N2),
      // Z_z[c][j1*N2 + m2] *= w_n^(m2*j1)
      PassArgs<F> a = blank_args<F>();
      a.in = d_polys + c0 * n;
      a.out = d_scratch;
      a.in_ystride = (long long)n;
      a.in_zstride = 0;
      a.in_wstride = 1;
      a.in_sstride = (long long)N2;
      a.out_ystride = (long long)n;
      a.out_zstride = (long long)nc_in * (long long)n;
      a.out_wstride = 1;
      a.out_jstride = (long long)N2;
      a.nw_in = a.nw_out = (uint32_t)N2;
      WG_TRY(tb.get_tws(na, false, &a.tws));
      a.pm_lo = pm.lo;
      a.pm_hi = pm.hi;
      a.pm_zstride_lo = pzl;
      a.pm_zstride_hi = pzh;
      if (use_q) {
        a.has_pm = 2;
      } else {
        a.has_pm = 1;
        a.pm_s = (uint32_t)N2;
        a.pm_w = 1;
        a.pm_lb = pm.lb;
        a.pm_mask = (uint32_t)(n - 1);
        a.has_tm = 1;
        a.tm_lo = tm.lo;
        a.tm_hi = tm.hi;
        a.tm_lb = tm.lb;
        a.tm_mask = (uint32_t)(n - 1);
        a.tm_w = 1;
      }
      const int tw = 1 << pass_log_tw<F>(na);
      a.ny = nc_in;
      const unsigned gy = cdiv(nc_in, walk_a ? walk_a : 1);
      ProfScope ps(PH_LDE_A, stream);
      if (tunables().zfast) {
        a.zfast = B;
        WG_TRY(launch_pass<F>(0, false, na, a, dim3(cdiv(N2, tw) * B, gy, 1), stream));
      } else {
        WG_TRY(launch_pass<F>(0, false, na, a, dim3(cdiv(N2, tw), gy, B), stream));
      }
    }
    {
      // pass B': for each (z, j1) transform over m2 (contiguous) -> rows (j1 + N1*j2)*B + z, columns c0..
      PassArgs<F> a = blank_args<F>();
      a.in = d_scratch;
      a.out = d_lde + c0;
      a.in_ystride = (long long)N2;
      a.in_zstride = (long long)nc_in * (long long)n;
      a.in_wstride = (long long)n;
      a.in_sstride = 1;
      a.out_ystride = (long long)B * row_width;
      a.out_zstride = (long long)row_width;
      a.out_wstride = 1;
      a.out_jstride = (long long)N1 * B * row_width;
      a.nw_in = nc_in;
      a.nw_out = nc_out;
      WG_TRY(tb.get_tws(nb, false, &a.tws));
      if (use_q && nc_in > 0) {
        a.has_q = 1;
        a.q_z = pm.lo + 1;  // t_z = first power of coset z's table
        a.q_zstride = pzl;
        a.q_y = 1;
        a.tm_lo = tm.lo;
        a.tm_hi = tm.hi;
        a.tm_lb = tm.lb;
        a.tm_mask = (uint32_t)(n - 1);
      }
      if (n_peers) {
        a.n_peers = n_peers;
        a.peer_shift = (uint32_t)(nb - ilog2(n_peers));
        for (uint32_t g = 0; g < n_peers; g++) a.out_peers[g] = d_peers[g] + c0;
      }
      const int tw = 1 << pass_log_tw<F>(nb);
      a.nx = cdiv(nc_out, tw);
      const unsigned gx = cdiv(a.nx, walk_b ? walk_b : 1);
      ProfScope ps(PH_LDE_B, stream);
      WG_TRY(launch_pass<F>(1, false, nb, a, dim3(gx, (unsigned)N1, B), stream));
    }
  }
  return WG_OK;
}

// ------------------------------------------------------------------------------------------------ split plan (split.cuh)
template <class F>
static PowTab<F> powtab_of(const typename Tables<F>::Two& t) {
  PowTab<F> p;
  p.lo = t.lo;
  p.hi = t.hi;
  p.lb = (uint32_t)t.lb;
  return p;
}
#define WG_SPLIT_LAUNCH(KERNEL, LR, GRID, BLOCK, ...)                                   \
  do {                                                                                  \
    switch (LR) {                                                                       \
      case 1: KERNEL<F, 1><<<GRID, BLOCK, 0, stream>>>(__VA_ARGS__); break;             \
      case 2: KERNEL<F, 2><<<GRID, BLOCK, 0, stream>>>(__VA_ARGS__); break;             \
      case 3: KERNEL<F, 3><<<GRID, BLOCK, 0, stream>>>(__VA_ARGS__); break;             \
      default: KERNEL<F, 4><<<GRID, BLOCK, 0, stream>>>(__VA_ARGS__); break;            \
    }                                                                                   \
    CUDA_TRY(cudaGetLastError());                                                       \
    g_launches++;                                                                       \
  } while (0)

template <class F>
static int interpolate_columns_dev(const typename F::T* d_in, uint32_t n_cols, uint64_t n, typename F::T* d_out,
                                   typename F::T* d_scratch, uint64_t scratch_elems, cudaStream_t stream) {
  typedef typename F::T T;
  WG_TRY(check_domain<F>(n, 1));
  const int lr = split_lr(n);
  if (lr == 0) return interpolate_columns_base<F>(d_in, n_cols, n, d_out, d_scratch, scratch_elems, stream);
  if (n_cols == 0) return WG_OK;
  const uint64_t R = (uint64_t)1 << lr, M = n >> lr;
  uint64_t chunk = d_scratch ? scratch_elems / (2 * n) : 0;
  if (chunk == 0) return fail(WG_ERR_NOMEM, "interpolation of 2^%d points needs scratch for 2*n elements per column", ilog2(n));
  if (chunk > 65535 / R) chunk = 65535 / R;
  typename Tables<F>::Two tw;
  WG_TRY(tables<F>().get_pow(ilog2(n), true, ilog2(M), F::one(), &tw));
  const PowTab<F> tab = powtab_of<F>(tw);
  const T r_inv = f_inv<F>(F::from_u64(R));
  for (uint64_t c0 = 0; c0 < n_cols; c0 += chunk) {
    const uint32_t nc = (uint32_t)((n_cols - c0) < chunk ? (n_cols - c0) : chunk);
    T* e = d_scratch;                 // nc*R sub-columns of M values, transformed in place
    T* s = d_scratch + (uint64_t)nc * n;
    const dim3 grid(cdiv(M, 256), nc, 1);
    {
      ProfScope ps(PH_INTERP_A, stream);
      WG_SPLIT_LAUNCH(split_interp_pre_kernel, lr, grid, 256, d_in + c0 * n, e, M, tab, r_inv);
    }
    WG_TRY(interpolate_columns_dev<F>(e, (uint32_t)(nc * R), M, e, s, scratch_elems - (uint64_t)nc * n, stream));
    {
      ProfScope ps(PH_INTERP_B, stream);
      WG_SPLIT_LAUNCH(split_interleave_kernel, lr, grid, 256, (const T*)e, d_out + c0 * n, M);
    }
  }
  return WG_OK;
}

template <class F>
static int lde_dev(const typename F::T* d_polys, uint32_t n_cols, uint64_t n, uint64_t blowup, typename F::T offset,
                   typename F::T* d_lde, uint64_t row_width, typename F::T* d_scratch, uint64_t scratch_elems,
                   cudaStream_t stream, typename F::T* const* d_peers = nullptr, uint32_t n_peers = 0) {
  typedef typename F::T T;
  WG_TRY(check_domain<F>(n, blowup));
  int lr = split_lr(n);
  if (lr == 0) {
    // preferred, not required: only with scratch for at least min(n_cols, 8) columns of the split plan
    const int lp = split_lr_pref(n);
    const uint64_t need = n * (2 * blowup + 1) * (n_cols < 8 ? n_cols : 8);
    if (lp == 0 || n_peers || !d_scratch || scratch_elems < need)
      return lde_base<F>(d_polys, n_cols, n, blowup, offset, d_lde, row_width, d_scratch, scratch_elems, stream, d_peers, n_peers);
    lr = lp;
  }
  if (n_peers) return fail(WG_ERR_UNSUPPORTED, "row-block scatter is not available for transforms above 2^%d points", g_split_log.load());
  if (F::is_zero(offset)) return fail(WG_ERR_ZERO_OFFSET, "domain offset cannot be zero");
  if (!elem_is_canonical<F>(offset)) return fail(WG_ERR_ARG, "domain offset is not a reduced field element");
  if (row_width < n_cols) return fail(WG_ERR_ARG, "row_width < n_cols");
  if (n_cols == 0) return WG_OK;
  const uint64_t R = (uint64_t)1 << lr, M = n >> lr, MB = M * blowup;
  uint64_t chunk = d_scratch ? scratch_elems / (n * (2 * blowup + 1)) : 0;
  if (chunk == 0) return fail(WG_ERR_NOMEM, "LDE of 2^%d coefficients needs scratch for (2*blowup + 1)*n elements per column", ilog2(n));
  if (chunk > 65528 / R) chunk = 65528 / R;
  typename Tables<F>::Two tw;
  WG_TRY(tables<F>().get_pow(ilog2(n) + ilog2(blowup), false, ilog2(MB), offset, &tw));
  const PowTab<F> tab = powtab_of<F>(tw);
  const T offset_r = f_pow<F>(offset, R);
  for (uint64_t c0 = 0; c0 < n_cols; c0 += chunk) {
    const uint32_t nc = (uint32_t)((n_cols - c0) < chunk ? (n_cols - c0) : chunk);
    const bool last = c0 + nc >= n_cols;
    const uint32_t nc_out = last ? (uint32_t)(row_width - c0) : nc;  // the last sweep also zeroes the pad columns
    T* d = d_scratch;                          // nc*R sub-polynomials of M coefficients
    T* e = d + (uint64_t)nc * n;               // their LDE: [MB][nc*R]
    T* s = e + (uint64_t)nc * n * blowup;
    {
      ProfScope ps(PH_LDE_A, stream);
      WG_SPLIT_LAUNCH(split_deinterleave_kernel, lr, dim3(cdiv(M, 256), nc, 1), 256, d_polys + c0 * n, d, M);
    }
    WG_TRY(lde_dev<F>(d, (uint32_t)(nc * R), M, blowup, offset_r, e, (uint64_t)nc * R, s,
                      scratch_elems - (uint64_t)nc * n * (blowup + 1), stream));
    {
      ProfScope ps(PH_LDE_B, stream);
      // block = bx columns x (256 / bx) rows; narrow sweeps (an 8-column chunk of a sharded commit) keep every thread busy
      const unsigned bx = nc_out <= 4 ? 4 : (nc_out <= 8 ? 8 : (nc_out <= 16 ? 16 : 32)), by = 256 / bx;
      WG_SPLIT_LAUNCH(split_lde_post_kernel, lr, dim3(cdiv(MB, by), cdiv(nc_out, bx), 1), dim3(bx, by, 1), (const T*)e,
                      (uint64_t)nc * R, d_lde + c0, row_width, nc, nc_out, MB, tab);
    }
  }
  return WG_OK;
}

// ------------------------------------------------------------------------------------------------ LDE of several traces
// StarkPack commits several NARROW traces together (cfg4: 8 fib traces of 2 columns). Extended one by one, each runs partly
// filled 4-column tiles and its own pair of launches; all columns extended together as one [L][W] matrix fill the tiles, and
// a memory-bound copy then lays every trace out as its own row-major matrix (what RowMatrix / TraceLde hold,
// row_matrix.rs:38-45). Taken when it saves more tile columns than the copy costs (the copy moves 2 x 16..32 bytes per
// element, the LDE costs ~5x that in time) and the scratch holds the combined matrix next to the LDE's own scratch.
static inline bool traces_prefer_combined(uint32_t n_traces, const uint32_t* cols) {
  if (n_traces < 2) return false;
  uint64_t total = 0, tiles = 0;
  for (uint32_t t = 0; t < n_traces; t++) {
    total += cols[t];
    tiles += (cols[t] + 3) / 4;
  }
  const uint64_t saved_cols = 4 * tiles - (total + 3) / 4 * 4;
  return saved_cols * 5 > total;
}
template <class F>
static uint64_t traces_combined_elems(uint32_t n_traces, const uint32_t* cols, uint64_t n, uint64_t blowup) {
  uint64_t total = 0;
  for (uint32_t t = 0; t < n_traces; t++) total += cols[t];
  const uint64_t W = (total + 7) / 8 * 8;
  return n * blowup * W + lde_scratch_elems<F>((uint32_t)total, n, blowup);
}
// scratch for a whole commit of these traces at its fastest plan
template <class F>
static uint64_t commit_scratch_elems(uint32_t n_traces, const uint32_t* cols, uint64_t n, uint64_t blowup) {
  uint64_t total = 0, max_cols = 0;
  for (uint32_t t = 0; t < n_traces; t++) {
    total += cols[t];
    if (cols[t] > max_cols) max_cols = cols[t];
  }
  uint64_t se = interpolate_scratch_elems<F>((uint32_t)(total < 65535 ? total : 65535), n);
  const uint64_t se2 = lde_scratch_elems<F>((uint32_t)max_cols, n, blowup);
  if (se2 > se) se = se2;
  if (traces_prefer_combined(n_traces, cols)) {
    const uint64_t se3 = traces_combined_elems<F>(n_traces, cols, n, blowup);
    if (se3 > se) se = se3;
  }
  return se;
}
// dst[row*rw + c] = c < cols ? src[row*W + c] : 0
template <class T>
__global__ void __launch_bounds__(256) unpack_trace_kernel(const T* __restrict__ src, uint64_t W, T* __restrict__ dst, uint32_t rw,
                                                           uint32_t cols, uint64_t rows) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= rows * rw) return;
  const uint64_t row = i / rw;
  const uint32_t c = (uint32_t)(i - row * rw);
  T v;
  memset(&v, 0, sizeof v);
  if (c < cols) v = src[row * W + c];
  dst[i] = v;
}
// every trace's LDE into ldes[t] ([L][rws[t]], pad columns zero); after(t) runs once trace t's matrix is enqueued
template <class F, class After>
static int lde_traces_dev(const typename F::T* d_polys, uint32_t n_traces, const uint32_t* cols, uint64_t n, uint64_t blowup,
                          typename F::T offset, void* const* ldes, const uint64_t* rws, typename F::T* d_scratch,
                          uint64_t scratch_elems, cudaStream_t stream, After after) {
  typedef typename F::T T;
  const uint64_t L = n * blowup;
  uint64_t total = 0;
  for (uint32_t t = 0; t < n_traces; t++) total += cols[t];
  const uint64_t W = (total + 7) / 8 * 8, comb = L * W;
  const uint64_t lde_min = lde_scratch_elems_min<F>((uint32_t)(total < 8 ? total : 8), n, blowup);
  if (traces_prefer_combined(n_traces, cols) && d_scratch && scratch_elems >= comb + lde_min && total <= 65528) {
    T* cm = d_scratch;
    WG_TRY(lde_dev<F>(d_polys, (uint32_t)total, n, blowup, offset, cm, W, d_scratch + comb, scratch_elems - comb, stream));
    uint64_t c0 = 0;
    for (uint32_t t = 0; t < n_traces; t++) {
      ProfScope ps(PH_LDE_B, stream);
      unpack_trace_kernel<T><<<cdiv(L * rws[t], 256), 256, 0, stream>>>(cm + c0, W, (T*)ldes[t], (uint32_t)rws[t], cols[t], L);
      CUDA_TRY(cudaGetLastError());
      g_launches++;
      c0 += cols[t];
      WG_TRY(after(t));
    }
    return WG_OK;
  }
  uint64_t c0 = 0;
  for (uint32_t t = 0; t < n_traces; t++) {
    WG_TRY(lde_dev<F>(d_polys + c0 * n, cols[t], n, blowup, offset, (T*)ldes[t], rws[t], d_scratch, scratch_elems, stream));
    c0 += cols[t];
    WG_TRY(after(t));
  }
  return WG_OK;
}

// ------------------------------------------------------------------------------------------------ hashing / Merkle
template <class F>
static int hash_rows_dev(int hash, uint32_t n_traces, const void* const* d_ldes, const uint64_t* row_widths,
                         const uint64_t* elems, uint64_t rows, uint8_t* d_leaves, cudaStream_t stream) {
  if (n_traces == 0 || n_traces > HASH_MAX_TRACES) return fail(WG_ERR_UNSUPPORTED, "1..%d aggregated traces supported", HASH_MAX_TRACES);
  if (!hash_ok(hash)) return fail(WG_ERR_ARG, "unknown hash %d", hash);
  if (rows == 0) return WG_OK;
  RowHashArgs a;
  memset(&a, 0, sizeof a);
  a.n_traces = n_traces;
  a.rows = rows;
  a.leaves = d_leaves;
  a.one = 1;
  uint64_t off = 0;
  for (uint32_t t = 0; t < n_traces; t++) {
    // elems == 0 is legal: hash_elements(&[]) is the digest of the empty input (crypto/src/hash/blake/mod.rs:46-59)
    if (elems[t] > row_widths[t]) return fail(WG_ERR_ARG, "elems_per_row must not exceed row_width");
    a.lde[t] = d_ldes[t];
    a.row_width[t] = (uint32_t)row_widths[t];
    a.unit_off[t] = (uint32_t)off;
    off += elems[t] * (F::kElemBytes / 8);
  }
  if (off * 8 > (1u << 22)) return fail(WG_ERR_UNSUPPORTED, "rows longer than 4 MiB");
  a.unit_off[n_traces] = (uint32_t)off;
  const unsigned blocks = cdiv(rows, 32 * HASH_WARPS);
  ProfScope ps(PH_HASH, stream);
  const bool multi = off * 8 > 1024, one = n_traces == 1;
#define WG_HASH_LAUNCH(H, M, O) hash_rows_kernel<F, H, M, O><<<blocks, HASH_WARPS * 32, 0, stream>>>(a)
#define WG_HASH_PICK(H)                                  \
  do {                                                   \
    if (multi && one) WG_HASH_LAUNCH(H, true, true);     \
    else if (multi) WG_HASH_LAUNCH(H, true, false);      \
    else if (one) WG_HASH_LAUNCH(H, false, true);        \
    else WG_HASH_LAUNCH(H, false, false);                \
  } while (0)
  if (hash == WG_HASH_BLAKE3_256) {
    WG_HASH_PICK(HASH_BLAKE3_256);
  } else if (hash == WG_HASH_BLAKE3_192) {
    WG_HASH_PICK(HASH_BLAKE3_192);
  } else if (one) {
    WG_HASH_LAUNCH(HASH_SHA3_256, false, true);
  } else {
    WG_HASH_LAUNCH(HASH_SHA3_256, false, false);
  }
#undef WG_HASH_PICK
#undef WG_HASH_LAUNCH
  CUDA_TRY(cudaGetLastError());
  g_launches++;
  return WG_OK;
}

template <int HASH>
static int merkle_dev_t(const uint8_t* d_leaves, uint64_t n_leaves, uint8_t* d_nodes, cudaStream_t stream) {
  constexpr int U = DigestCfg<HASH>::U;  // 8-byte units per digest
  uint2* nodes = reinterpret_cast<uint2*>(d_nodes);
  uint64_t n = n_leaves / 2;
  ProfScope ps(PH_MERKLE, stream);
  merkle_level_kernel<HASH><<<cdiv(n, 256), 256, 0, stream>>>(reinterpret_cast<const uint2*>(d_leaves), nodes + U * n, n, 1u);
  CUDA_TRY(cudaGetLastError());
  g_launches++;
  while (n > 1024) {
    merkle_level_kernel<HASH><<<cdiv(n / 2, 256), 256, 0, stream>>>(nodes + U * n, nodes + U * (n / 2), n / 2, 1u);
    CUDA_TRY(cudaGetLastError());
    g_launches++;
    n /= 2;
  }
  merkle_top_kernel<HASH><<<1, 512, 0, stream>>>(nodes, (uint32_t)n, 1u);
  CUDA_TRY(cudaGetLastError());
  g_launches++;
  return WG_OK;
}

static int merkle_dev(int hash, const uint8_t* d_leaves, uint64_t n_leaves, uint8_t* d_nodes, cudaStream_t stream) {
  if (n_leaves < 2) return fail(WG_ERR_TOO_FEW, "a Merkle tree needs at least 2 leaves, got %llu", (unsigned long long)n_leaves);
  if (!is_pow2(n_leaves)) return fail(WG_ERR_NOT_POW2, "number of leaves %llu is not a power of two", (unsigned long long)n_leaves);
  if (hash == WG_HASH_BLAKE3_256) return merkle_dev_t<HASH_BLAKE3_256>(d_leaves, n_leaves, d_nodes, stream);
  if (hash == WG_HASH_SHA3_256) return merkle_dev_t<HASH_SHA3_256>(d_leaves, n_leaves, d_nodes, stream);
  if (hash == WG_HASH_BLAKE3_192) return merkle_dev_t<HASH_BLAKE3_192>(d_leaves, n_leaves, d_nodes, stream);
  return fail(WG_ERR_ARG, "unknown hash %d", hash);
}

// ------------------------------------------------------------------------------------------------ twiddles
// get_twiddles / get_inv_twiddles (math/src/fft/mod.rs:466-482,505-522):
// out[bitrev_{n/2}(i)] = internal(w_n^(+-i)), i < n/2; w_n^-i is read as w_n^(n-i) from the forward table.
template <class F>
__global__ void twiddles_kernel(typename F::T* out, uint32_t n, int log_half, int inverse, const typename F::T* lo,
                                const typename F::T* hi, uint32_t lb) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n / 2) return;
  const uint32_t e = inverse ? ((n - i) & (n - 1)) : i;
  typename F::T v = table2<F>(lo, hi, lb, e);
  const uint32_t j = log_half == 0 ? 0 : (__brev(i) >> (32 - log_half));
  out[j] = v;  // tables already hold the internal representation
}

template <class F>
static int get_twiddles_dev(uint64_t n, bool inverse, typename F::T* d_out, cudaStream_t stream) {
  WG_TRY(check_domain<F>(n, 1));
  const int log_n = ilog2(n);
  typename Tables<F>::Two tm;
  WG_TRY(tables<F>().get_tm(log_n, false, &tm));
  twiddles_kernel<F><<<cdiv(n / 2, 256), 256, 0, stream>>>(d_out, (uint32_t)n, log_n - 1, inverse ? 1 : 0, tm.lo, tm.hi, tm.lb);
  CUDA_TRY(cudaGetLastError());
  g_launches++;
  return WG_OK;
}

struct DevBufFwd {
  void* p = nullptr;
  ~DevBufFwd() {
    if (p) cudaFree(p);
  }
  int alloc(size_t bytes) {
    cudaError_t e = dev_malloc(&p, bytes ? bytes : 16);
    if (e != cudaSuccess) {
      p = nullptr;
      return fail(WG_ERR_NOMEM, "cudaMalloc(%zu): %s", bytes, cudaGetErrorString(e));
    }
    return WG_OK;
  }
};

// ------------------------------------------------------------------------------------------------ device arithmetic probe
// Element-wise field ops on the device (tests/test_gpu_parity.py checks them against Python big ints):
// op 0 add, 1 sub, 2 mul, 16+i mul_root16<false>(a, i), 32+i mul_root16<true>(a, i), 48 as_int (f64 only),
// 64+i / 80+i sub_mul_root16<false / true>(a, b, i)
template <class F>
__global__ void field_op_kernel(int op, const typename F::T* a, const typename F::T* b, typename F::T* out, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  typename F::T x = a[i], y = b[i], r = x;
  if (op == 0) r = F::add(x, y);
  else if (op == 1) r = F::sub(x, y);
  else if (op == 2) r = F::mul(x, y);
  else if (op >= 16 && op < 32) {
#pragma unroll
    for (int k = 0; k < 16; k++)
      if (op - 16 == k) r = F::template mul_root16<false>(x, k);
  } else if (op >= 32 && op < 48) {
#pragma unroll
    for (int k = 0; k < 16; k++)
      if (op - 32 == k) r = F::template mul_root16<true>(x, k);
  } else if (op >= 64 && op < 80) {  // (a - b) * w16^(op-64), the butterfly form the NTT kernels use
#pragma unroll
    for (int k = 0; k < 16; k++)
      if (op - 64 == k) r = F::template sub_mul_root16<false>(x, y, k);
  } else if (op >= 80 && op < 96) {
#pragma unroll
    for (int k = 0; k < 16; k++)
      if (op - 80 == k) r = F::template sub_mul_root16<true>(x, y, k);
  } else if (op == 48) {
    if (F::kFieldId == 0) {
      uint64_t w;
      memcpy(&w, &x, 8);
      w = Gl64::as_int(w);
      memcpy(&r, &w, 8);
    }
  } else if (op >= 96 && op < 104) {
    // lazily reduced Goldilocks ops (field_lazy.cuh): first operand ANY u64, second <= p; results compared mod p by the test
    if (F::kFieldId == 0) {
      uint64_t xa, yb, w = 0;
      memcpy(&xa, &x, 8);
      memcpy(&yb, &y, 8);
      if (op == 96) w = GlLazy::add(xa, yb);
      else if (op == 97) w = GlLazy::add_b(xa, yb);
      else if (op == 98) w = GlLazy::sub(xa, yb);
      else if (op == 99) w = GlLazy::canon(xa);
      else if (op == 100) w = Gl64::mul(xa, yb);            // lazy first operand through the canonical product
      else if (op == 101) w = Gl64::mont_shift<28>(xa);
      else if (op == 102) w = Gl64::mont_shift<52>(xa);
      else w = GlLazy::mul_w2(xa, yb);
      memcpy(&r, &w, 8);
    }
  } else if (op >= 112 && op < 116) {
    // 16-point register networks on (x, y, x+1, y+1, ...): 112/113 canonical DIF fwd/inv, 114/115 lazy DIT fwd/inv; output
    // element (i & 15), canonicalised
    if (F::kFieldId == 0) {
      uint64_t xa, yb, v[16];
      memcpy(&xa, &x, 8);
      memcpy(&yb, &y, 8);
#pragma unroll
      for (int k = 0; k < 16; k++) {
        const uint64_t t = ((k & 1) ? yb : xa) + (uint64_t)(k >> 1) * 0x9E3779B97F4A7C15ULL;
        v[k] = t >= Gl64::P ? t - Gl64::P : t;
      }
      if (op == 112) radix_ntt<Gl64, 4, false>(v);
      else if (op == 113) radix_ntt<Gl64, 4, true>(v);
      else if (op == 114) radix_dit_lazy<4, false>(v);
      else radix_dit_lazy<4, true>(v);
      uint64_t w = 0;
#pragma unroll
      for (int k = 0; k < 16; k++)
        if ((int)(i & 15) == k) w = GlLazy::canon(v[k]);
      memcpy(&r, &w, 8);
    }
  }
  out[i] = r;
}
template <class F>
static int h_field_op(int op, const void* a, const void* b, void* out, uint32_t n) {
  typedef typename F::T T;
  DevBufFwd da, db, dout;
  WG_TRY(da.alloc(n * sizeof(T)));
  WG_TRY(db.alloc(n * sizeof(T)));
  WG_TRY(dout.alloc(n * sizeof(T)));
  CUDA_TRY(cudaMemcpy(da.p, a, n * sizeof(T), cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(db.p, b, n * sizeof(T), cudaMemcpyHostToDevice));
  field_op_kernel<F><<<cdiv(n, 128), 128>>>(op, (const T*)da.p, (const T*)db.p, (T*)dout.p, n);
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaMemcpy(out, dout.p, n * sizeof(T), cudaMemcpyDeviceToHost));
  return WG_OK;
}

// ------------------------------------------------------------------------------------------------ whole commit
template <class F>
static typename F::T generator() {
  return F::from_u64(F::GENERATOR_CANON);
}
template <class F>
static typename F::T offset_from_internal(const void* p) {
  typename F::T v;  // kernels compute on the reference's internal representation directly
  memcpy(&v, p, sizeof v);
  return v;
}

template <class F>
static int commit_trace_dev(int hash, uint64_t n, uint64_t blowup, uint32_t n_traces, const uint32_t* cols_per_trace,
                            const typename F::T* d_cols, typename F::T* d_polys, void* const* d_ldes, uint8_t* d_leaves,
                            uint8_t* d_nodes, typename F::T* d_scratch, uint64_t scratch_elems, cudaStream_t stream) {
  typedef typename F::T T;
  WG_TRY(check_domain<F>(n, blowup));
  if (n_traces == 0 || n_traces > HASH_MAX_TRACES) return fail(WG_ERR_UNSUPPORTED, "1..%d aggregated traces supported", HASH_MAX_TRACES);
  uint64_t total_cols = 0;
  for (uint32_t t = 0; t < n_traces; t++) total_cols += cols_per_trace[t];
  // interpolate_columns for every trace at once (columns are independent; prover/src/lib.rs:631-640)
  for (uint64_t c0 = 0; c0 < total_cols; c0 += 65535) {
    uint32_t nc = (uint32_t)((total_cols - c0) < 65535 ? (total_cols - c0) : 65535);
    WG_TRY(interpolate_columns_dev<F>(d_cols + c0 * n, nc, n, d_polys + c0 * n, d_scratch, scratch_elems, stream));
  }
  // evaluate_polys_over::<8> per trace (prover/src/lib.rs:641-650), offset = GENERATOR (air/src/options.rs:199-201)
  // (several narrow traces are extended together when the scratch allows: lde_traces_dev)
  const void* ldes[HASH_MAX_TRACES];
  uint64_t rws[HASH_MAX_TRACES], els[HASH_MAX_TRACES];
  for (uint32_t t = 0; t < n_traces; t++) {
    ldes[t] = d_ldes[t];
    rws[t] = (uint64_t)(cols_per_trace[t] + 7) / 8 * 8;
    els[t] = cols_per_trace[t];
  }
  WG_TRY(lde_traces_dev<F>(d_polys, n_traces, cols_per_trace, n, blowup, generator<F>(), d_ldes, rws, d_scratch, scratch_elems, stream,
                           [](uint32_t) { return WG_OK; }));
  // commit_to_comb_rows (row_matrix.rs:204-238)
  WG_TRY(hash_rows_dev<F>(hash, n_traces, ldes, rws, els, n * blowup, d_leaves, stream));
  return merkle_dev(hash, d_leaves, n * blowup, d_nodes, stream);
}

// drains the legacy stream on every exit path of a host-pointer entry point (async copies against caller buffers)
struct LegacyStreamDrainFwd {
  ~LegacyStreamDrainFwd() { cudaStreamSynchronize(0); }
};

// RAII device buffer for the host-pointer entry points
struct DevBuf {
  void* p = nullptr;
  ~DevBuf() {
    if (p) cudaFree(p);
  }
  int alloc(size_t bytes) {
    if (bytes == 0) bytes = 16;
    cudaError_t e = dev_malloc(&p, bytes);
    if (e != cudaSuccess) {
      p = nullptr;
      return fail(WG_ERR_NOMEM, "cudaMalloc(%zu): %s", bytes, cudaGetErrorString(e));
    }
    return WG_OK;
  }
  template <class T>
  T* as() { return reinterpret_cast<T*>(p); }
};

template <class F>
static int h_get_twiddles(uint64_t n, int inverse, void* out) {
  typedef typename F::T T;
  WG_TRY(check_domain<F>(n, 1));
  DevBuf d;
  WG_TRY(d.alloc(n / 2 * sizeof(T)));
  WG_TRY(get_twiddles_dev<F>(n, inverse != 0, d.as<T>(), 0));
  CUDA_TRY(cudaMemcpy(out, d.p, n / 2 * sizeof(T), cudaMemcpyDeviceToHost));
  return WG_OK;
}

// Extension-field element layouts: a QuadExtension / CubeExtension element is `deg` consecutive base
// elements (math/src/field/extensions/quadratic.rs:28, cubic.rs:28) and every fft:: function multiplies it by base-field
// twiddles only (fft_inputs.rs:101-144 `mul_base`), i.e. acts per coordinate. A column of E elements [n][deg] is therefore
// split into deg base columns on the device, transformed by the base-field kernels, and merged back.
template <class T>
__global__ void __launch_bounds__(256) deinterleave_kernel(const T* __restrict__ in, T* __restrict__ out, uint64_t n, uint32_t deg) {
  const uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;  // out [deg][n]
  if (idx >= n * deg) return;
  const uint64_t d = idx / n, i = idx - d * n;
  out[idx] = in[i * deg + d];
}
template <class T>
__global__ void __launch_bounds__(256) interleave_kernel(const T* __restrict__ in, T* __restrict__ out, uint64_t n, uint32_t deg) {
  const uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;  // out [n][deg]
  if (idx >= n * deg) return;
  const uint64_t i = idx / deg, d = idx - i * deg;
  out[idx] = in[d * n + i];
}
static inline int check_ext_degree(uint32_t deg) {
  if (deg < 1 || deg > 3) return fail(WG_ERR_UNSUPPORTED, "extension degree %u is not supported (1, 2, 3)", deg);
  return WG_OK;
}

// Uploads n_cols host columns of n extension elements each and leaves them on the device as n_cols*deg base columns
// [c*deg + d][n] in `dst` (deg = 1: a plain copy per column). `stg` is scratch of the same size.
template <class T>
static int upload_columns(const void* const* cols, uint32_t n_cols, uint64_t n, uint32_t deg, T* dst, T* stg, cudaStream_t s) {
  const size_t colb = n * deg * sizeof(T);
  char* to = reinterpret_cast<char*>(deg == 1 ? dst : stg);
  for (uint32_t c = 0; c < n_cols; c++) CUDA_TRY(cudaMemcpyAsync(to + c * colb, cols[c], colb, cudaMemcpyHostToDevice, s));
  if (deg > 1) {
    for (uint32_t c = 0; c < n_cols; c++) {
      deinterleave_kernel<T><<<cdiv(n * deg, 256), 256, 0, s>>>(stg + (size_t)c * n * deg, dst + (size_t)c * n * deg, n, deg);
      CUDA_TRY(cudaGetLastError());
      g_launches++;
    }
  }
  return WG_OK;
}
template <class T>
static int download_columns(void* const* outs, uint32_t n_cols, uint64_t n, uint32_t deg, const T* src, T* stg, cudaStream_t s) {
  const size_t colb = n * deg * sizeof(T);
  if (deg > 1) {
    for (uint32_t c = 0; c < n_cols; c++) {
      interleave_kernel<T><<<cdiv(n * deg, 256), 256, 0, s>>>(src + (size_t)c * n * deg, stg + (size_t)c * n * deg, n, deg);
      CUDA_TRY(cudaGetLastError());
      g_launches++;
    }
  }
  const char* from = reinterpret_cast<const char*>(deg == 1 ? src : stg);
  for (uint32_t c = 0; c < n_cols; c++) CUDA_TRY(cudaMemcpyAsync(outs[c], from + c * colb, colb, cudaMemcpyDeviceToHost, s));
  return WG_OK;
}

template <class F>
static int h_interpolate_columns(const void* const* cols, uint32_t n_cols, uint64_t n, void* const* outs, uint32_t deg = 1) {
  typedef typename F::T T;
  WG_TRY(check_domain<F>(n, 1));
  WG_TRY(check_ext_degree(deg));
  DevBuf in, out, scr;
  const uint32_t nb = n_cols * deg;  // base columns
  const size_t colb = n * sizeof(T);
  WG_TRY(in.alloc(colb * nb));
  WG_TRY(out.alloc(colb * nb));
  const uint64_t se = interpolate_scratch_elems<F>(nb, n);
  WG_TRY(scr.alloc(se * sizeof(T)));
  WG_TRY(upload_columns<T>(cols, n_cols, n, deg, in.as<T>(), out.as<T>(), 0));
  for (uint32_t c0 = 0; c0 < nb; c0 += 65535) {
    uint32_t nc = (nb - c0) < 65535 ? (nb - c0) : 65535;
    WG_TRY(interpolate_columns_dev<F>(in.as<T>() + (size_t)c0 * n, nc, n, out.as<T>() + (size_t)c0 * n, scr.as<T>(), se, 0));
  }
  WG_TRY(download_columns<T>(outs, n_cols, n, deg, out.as<T>(), in.as<T>(), 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return WG_OK;
}

// x[i] *= base^i (two-level table of `base`), the coefficient rescale of fft::interpolate_poly_with_offset
template <class F>
__global__ void scale_powers_kernel(typename F::T* x, uint64_t n, const typename F::T* lo, const typename F::T* hi, uint32_t lb) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) x[i] = F::mul(x[i], table2<F>(lo, hi, lb, (uint32_t)i));
}

// fft::interpolate_poly_with_offset (math/src/fft/mod.rs:362-400; serial.rs:78-93): inverse transform, then
// coefficient i times offset^-i
template <class F>
static int h_interpolate_poly_with_offset(void* evals, uint64_t n, const void* offset, uint32_t deg = 1) {
  typedef typename F::T T;
  WG_TRY(check_domain<F>(n, 1));
  WG_TRY(check_ext_degree(deg));
  const T off = offset_from_internal<F>(offset);
  if (F::is_zero(off)) return fail(WG_ERR_ZERO_OFFSET, "domain offset cannot be zero");
  DevBuf in, out, scr;
  WG_TRY(in.alloc(n * deg * sizeof(T)));
  WG_TRY(out.alloc(n * deg * sizeof(T)));
  const uint64_t se = interpolate_scratch_elems<F>(deg, n);
  WG_TRY(scr.alloc(se * sizeof(T)));
  const void* cols[1] = {evals};
  void* outs[1] = {evals};
  WG_TRY(upload_columns<T>(cols, 1, n, deg, in.as<T>(), out.as<T>(), 0));
  WG_TRY(interpolate_columns_dev<F>(in.as<T>(), deg, n, out.as<T>(), scr.as<T>(), se, 0));
  typename Tables<F>::Two pm;
  long long zl, zh;
  WG_TRY(tables<F>().get_pm(ilog2(n), 0, f_inv<F>(off), &pm, &zl, &zh));
  for (uint32_t d = 0; d < deg; d++) {
    scale_powers_kernel<F><<<cdiv(n, 256), 256, 0, 0>>>(out.as<T>() + (size_t)d * n, n, pm.lo, pm.hi, (uint32_t)pm.lb);
    CUDA_TRY(cudaGetLastError());
    g_launches++;
  }
  WG_TRY(download_columns<T>(outs, 1, n, deg, out.as<T>(), in.as<T>(), 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return WG_OK;
}

template <class F>
static int h_evaluate_polys_over(const void* const* polys, uint32_t n_cols, uint64_t n, uint64_t blowup, const void* offset,
                                 void* out, uint64_t row_width, uint32_t deg = 1) {
  typedef typename F::T T;
  WG_TRY(check_domain<F>(n, blowup));
  WG_TRY(check_ext_degree(deg));
  DevBuf in, stg, lde, scr;
  const uint32_t nb = n_cols * deg;
  const size_t colb = n * sizeof(T);
  WG_TRY(in.alloc(colb * nb));
  if (deg > 1) WG_TRY(stg.alloc(colb * nb));
  WG_TRY(lde.alloc(n * blowup * row_width * sizeof(T)));
  const uint64_t se = lde_scratch_elems<F>(nb, n, blowup);
  WG_TRY(scr.alloc(se * sizeof(T)));
  WG_TRY(upload_columns<T>(polys, n_cols, n, deg, in.as<T>(), stg.as<T>(), 0));
  WG_TRY(lde_dev<F>(in.as<T>(), nb, n, blowup, offset_from_internal<F>(offset), lde.as<T>(), row_width, scr.as<T>(), se, 0));
  CUDA_TRY(cudaMemcpyAsync(out, lde.p, n * blowup * row_width * sizeof(T), cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return WG_OK;
}

template <class F>
static int h_hash_rows(int hash, const void* rows_host, uint64_t rows, uint64_t row_width, uint64_t elems, uint8_t* leaves) {
  typedef typename F::T T;
  DevBuf m, l;
  WG_TRY(m.alloc(rows * row_width * sizeof(T)));
  WG_TRY(l.alloc(rows * dlen(hash)));
  CUDA_TRY(cudaMemcpyAsync(m.p, rows_host, rows * row_width * sizeof(T), cudaMemcpyHostToDevice, 0));
  const void* ptrs[1] = {m.p};
  WG_TRY(hash_rows_dev<F>(hash, 1, ptrs, &row_width, &elems, rows, l.as<uint8_t>(), 0));
  CUDA_TRY(cudaMemcpyAsync(leaves, l.p, rows * dlen(hash), cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return WG_OK;
}

// ColMatrix::evaluate_columns_at (col_matrix.rs:210-216): one block per column; thread t evaluates its contiguous slice of
// coefficients by Horner, then the 256 partial values are combined as a polynomial in x^m (m = slice length).
template <class F>
__global__ void __launch_bounds__(256) eval_at_kernel(const typename F::T* polys, uint64_t n, typename F::T x, typename F::T* out) {
  typedef typename F::T T;
  __shared__ T part[256];
  const T* p = polys + (size_t)blockIdx.x * n;
  const uint64_t m = (n + 255) / 256;
  const uint64_t lo = (uint64_t)threadIdx.x * m, hi = lo + m < n ? lo + m : n;
  T acc = F::zero();
  for (uint64_t i = hi; i > lo; i--) acc = F::add(F::mul(acc, x), p[i - 1]);
  part[threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    T xm = F::one(), b = x;  // x^m
    for (uint64_t e = m; e; e >>= 1) {
      if (e & 1) xm = F::mul(xm, b);
      b = F::mul(b, b);
    }
    T r = F::zero();
    for (int t = 255; t >= 0; t--) r = F::add(F::mul(r, xm), part[t]);
    out[blockIdx.x] = r;
  }
}
template <class F>
static int h_evaluate_columns_at(const void* const* polys, uint32_t n_cols, uint64_t n, const void* x, void* out) {
  typedef typename F::T T;
  if (n == 0 || n_cols == 0) return WG_OK;
  const T xv = offset_from_internal<F>(x);
  if (!elem_is_canonical<F>(xv)) return fail(WG_ERR_ARG, "evaluation point is not a reduced field element");
  DevBuf in, res;
  LegacyStreamDrainFwd drain;
  WG_TRY(in.alloc((size_t)n_cols * n * sizeof(T)));
  WG_TRY(res.alloc((size_t)n_cols * sizeof(T)));
  for (uint32_t c = 0; c < n_cols; c++) CUDA_TRY(cudaMemcpyAsync(in.as<T>() + (size_t)c * n, polys[c], n * sizeof(T), cudaMemcpyHostToDevice, 0));
  eval_at_kernel<F><<<n_cols, 256, 0, 0>>>(in.as<T>(), n, xv, res.as<T>());
  CUDA_TRY(cudaGetLastError());
  g_launches++;
  CUDA_TRY(cudaMemcpyAsync(out, res.p, (size_t)n_cols * sizeof(T), cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return WG_OK;
}

template <class F>
static int h_hash_comb_rows(int hash, uint32_t n_traces, const void* const* mats, const uint64_t* rws, const uint64_t* els, uint64_t rows,
                            uint8_t* leaves) {
  typedef typename F::T T;
  if (n_traces == 0 || n_traces > HASH_MAX_TRACES) return fail(WG_ERR_UNSUPPORTED, "1..%d aggregated traces supported", HASH_MAX_TRACES);
  std::vector<DevBuf> d(n_traces);
  DevBuf l;
  std::vector<const void*> ptrs(n_traces);
  LegacyStreamDrainFwd drain;
  WG_TRY(l.alloc(rows * dlen(hash)));
  for (uint32_t t = 0; t < n_traces; t++) {
    if (!mats[t]) return fail(WG_ERR_ARG, "null matrix %u", t);
    WG_TRY(d[t].alloc(rows * rws[t] * sizeof(T)));
    CUDA_TRY(cudaMemcpyAsync(d[t].p, mats[t], rows * rws[t] * sizeof(T), cudaMemcpyHostToDevice, 0));
    ptrs[t] = d[t].p;
  }
  WG_TRY(hash_rows_dev<F>(hash, n_traces, ptrs.data(), rws, els, rows, l.as<uint8_t>(), 0));
  CUDA_TRY(cudaMemcpyAsync(leaves, l.p, rows * dlen(hash), cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return WG_OK;
}

// Device arena + two streams kept across host-pointer commits: a 2^22 x 100 commit needs 64 GB of device buffers
// and cudaMalloc/cudaFree of that per call costs tens of milliseconds. Calls are serialised by g_commit_mu.
static std::mutex g_commit_mu;
static char* g_arena = nullptr;
static size_t g_arena_bytes = 0;
constexpr int H2D_CHUNKS = 8;
static cudaStream_t g_s_compute = nullptr, g_s_copy = nullptr, g_s_h2d = nullptr;
static cudaEvent_t g_ev[3] = {nullptr, nullptr, nullptr};
static cudaEvent_t g_ev_h2d[H2D_CHUNKS];

static void arena_release_if_idle() {
  if (!g_commit_mu.try_lock()) return;  // a commit is using it
  if (g_arena) cudaFree(g_arena);
  g_arena = nullptr;
  g_arena_bytes = 0;
  g_commit_mu.unlock();
}
static int arena_reserve(size_t bytes) {
  if (!g_s_compute) {
    CUDA_TRY(cudaStreamCreateWithFlags(&g_s_compute, cudaStreamNonBlocking));
    CUDA_TRY(cudaStreamCreateWithFlags(&g_s_copy, cudaStreamNonBlocking));
    CUDA_TRY(cudaStreamCreateWithFlags(&g_s_h2d, cudaStreamNonBlocking));
    for (auto& e : g_ev) CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    for (auto& e : g_ev_h2d) CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  }
  if (bytes <= g_arena_bytes) return WG_OK;
  if (g_arena) cudaFree(g_arena);
  g_arena = nullptr;
  g_arena_bytes = 0;
  cudaError_t e = cudaMalloc(&g_arena, bytes);
  if (e != cudaSuccess) {  // give back parked commitment buffers and retry once
    cudaGetLastError();
    cache_release_all();
    e = cudaMalloc(&g_arena, bytes);
  }
  if (e != cudaSuccess) {
    g_arena = nullptr;
    return fail(WG_ERR_NOMEM, "cudaMalloc(%zu): %s", bytes, cudaGetErrorString(e));
  }
  g_arena_bytes = bytes;
  return WG_OK;
}

// Error paths of the pipelined host-pointer entry points return early (CUDA_TRY / WG_TRY) while async copies against the
// caller's host buffers and kernels against the shared arena may still be in flight on the three streams; this guard
// drains them before g_commit_mu is released, so a failed call can neither scribble on memory the caller frees next
// nor race the next commit's use of the arena.
struct StreamDrain {
  cudaStream_t s[3];
  ~StreamDrain() {
    for (cudaStream_t x : s)
      if (x) cudaStreamSynchronize(x);
  }
};
struct LegacyStreamDrain {
  ~LegacyStreamDrain() { cudaStreamSynchronize(0); }
};

// Prover::build_trace_commitment through host buffers. Copies ride a second stream so that the D2H of the
// polynomials overlaps the LDE kernels, the D2H of each trace's LDE overlaps the next trace's LDE / the row hashing,
// and only the tree comes back after the last kernel.
template <class F>
static int h_commit_trace(int hash, uint64_t n, uint64_t blowup, uint32_t n_traces, const uint32_t* cols_per_trace,
                          const void* const* column_ptrs, void* const* polys_out, void* const* lde_out, uint8_t* leaves_out,
                          uint8_t* nodes_out, uint8_t* root_out) {
  typedef typename F::T T;
  WG_TRY(check_domain<F>(n, blowup));
  if (n_traces == 0 || n_traces > HASH_MAX_TRACES) return fail(WG_ERR_UNSUPPORTED, "1..%d aggregated traces supported", HASH_MAX_TRACES);
  uint64_t total_cols = 0, max_cols = 0;
  for (uint32_t t = 0; t < n_traces; t++) {
    total_cols += cols_per_trace[t];
    if (cols_per_trace[t] > max_cols) max_cols = cols_per_trace[t];
    if (cols_per_trace[t] == 0) return fail(WG_ERR_ARG, "trace %u has no columns", t);
  }
  const uint64_t L = n * blowup;
  const size_t colb = n * sizeof(T);
  uint64_t se = interpolate_scratch_elems<F>((uint32_t)(total_cols < 65535 ? total_cols : 65535), n);
  const uint64_t se2 = lde_scratch_elems<F>((uint32_t)max_cols, n, blowup);
  if (se2 > se) se = se2;
  auto al = [](size_t b) { return (b + 255) / 256 * 256; };
  size_t fixed = 2 * al(colb * total_cols) + 2 * al(L * dlen(hash));
  for (uint32_t t = 0; t < n_traces; t++) fixed += al(L * ((uint64_t)(cols_per_trace[t] + 7) / 8 * 8) * sizeof(T));
  std::lock_guard<std::mutex> lk(g_commit_mu);
  {
    // several narrow traces: room for the combined LDE matrix when the device has it (lde_traces_dev)
    const uint64_t se_fast = commit_scratch_elems<F>(n_traces, cols_per_trace, n, blowup);
    size_t free_b = 0, total_b = 0;
    if (se_fast > se && cudaMemGetInfo(&free_b, &total_b) == cudaSuccess && fixed + al(se_fast * sizeof(T) + 16) <= free_b + g_arena_bytes)
      se = se_fast;
  }
  const size_t need = fixed + al(se * sizeof(T) + 16);
  WG_TRY(arena_reserve(need));
  char* cur = g_arena;
  auto take = [&](size_t b) {
    char* r = cur;
    cur += al(b);
    return r;
  };
  T* cols = (T*)take(colb * total_cols);
  T* polys = (T*)take(colb * total_cols);
  uint8_t* leaves = (uint8_t*)take(L * dlen(hash));
  uint8_t* nodes = (uint8_t*)take(L * dlen(hash));
  T* scr = (T*)take(se * sizeof(T) + 16);
  std::vector<T*> ldes(n_traces);
  std::vector<uint64_t> rws(n_traces), els(n_traces);
  for (uint32_t t = 0; t < n_traces; t++) {
    rws[t] = (uint64_t)(cols_per_trace[t] + 7) / 8 * 8;
    els[t] = cols_per_trace[t];
    ldes[t] = (T*)take(L * rws[t] * sizeof(T));
  }
  cudaStream_t sc = g_s_compute, sx = g_s_copy, sh = g_s_h2d;
  StreamDrain drain{{sc, sx, sh}};
  // upload in up to H2D_CHUNKS column groups; group i is interpolated while group i+1 is still on the wire
  const uint64_t per = (total_cols + H2D_CHUNKS - 1) / H2D_CHUNKS;
  for (uint64_t c0 = 0, k = 0; c0 < total_cols; c0 += per, k++) {
    const uint32_t nc = (uint32_t)((total_cols - c0) < per ? (total_cols - c0) : per);
    for (uint64_t c = c0; c < c0 + nc; c++)
      CUDA_TRY(cudaMemcpyAsync((char*)cols + c * colb, column_ptrs[c], colb, cudaMemcpyHostToDevice, sh));
    CUDA_TRY(cudaEventRecord(g_ev_h2d[k], sh));
    CUDA_TRY(cudaStreamWaitEvent(sc, g_ev_h2d[k], 0));
    WG_TRY(interpolate_columns_dev<F>(cols + c0 * n, nc, n, polys + c0 * n, scr, se, sc));
  }
  if (polys_out) {
    CUDA_TRY(cudaEventRecord(g_ev[0], sc));
    CUDA_TRY(cudaStreamWaitEvent(sx, g_ev[0], 0));
    for (uint64_t c = 0; c < total_cols; c++)
      CUDA_TRY(cudaMemcpyAsync(polys_out[c], (char*)polys + c * colb, colb, cudaMemcpyDeviceToHost, sx));
  }
  std::vector<const void*> lde_c(ldes.begin(), ldes.end());
  std::vector<void*> lde_v(ldes.begin(), ldes.end());
  WG_TRY(lde_traces_dev<F>(polys, n_traces, cols_per_trace, n, blowup, generator<F>(), lde_v.data(), rws.data(), scr, se, sc,
                           [&](uint32_t t) -> int {
                             if (lde_out && lde_out[t]) {  // this trace's D2H rides the copy stream under the next LDE / the hashing
                               CUDA_TRY(cudaEventRecord(g_ev[1], sc));
                               CUDA_TRY(cudaStreamWaitEvent(sx, g_ev[1], 0));
                               CUDA_TRY(cudaMemcpyAsync(lde_out[t], ldes[t], L * rws[t] * sizeof(T), cudaMemcpyDeviceToHost, sx));
                             }
                             return WG_OK;
                           }));
  WG_TRY(hash_rows_dev<F>(hash, n_traces, lde_c.data(), rws.data(), els.data(), L, leaves, sc));
  WG_TRY(merkle_dev(hash, leaves, L, nodes, sc));
  CUDA_TRY(cudaEventRecord(g_ev[2], sc));
  CUDA_TRY(cudaStreamWaitEvent(sx, g_ev[2], 0));
  if (root_out) CUDA_TRY(cudaMemcpyAsync(root_out, nodes + dlen(hash), dlen(hash), cudaMemcpyDeviceToHost, sx));
  if (leaves_out) CUDA_TRY(cudaMemcpyAsync(leaves_out, leaves, L * dlen(hash), cudaMemcpyDeviceToHost, sx));
  if (nodes_out) CUDA_TRY(cudaMemcpyAsync(nodes_out, nodes, L * dlen(hash), cudaMemcpyDeviceToHost, sx));
  CUDA_TRY(cudaStreamSynchronize(sx));
  CUDA_TRY(cudaStreamSynchronize(sc));
  CUDA_TRY(cudaStreamSynchronize(sh));
  return WG_OK;
}

// Prover::build_trace_commitment::<E> for an extension field E (auxiliary trace segments, prover/src/lib.rs:333-334):
// the same path on n_cols*deg base columns (ColMatrix::get_base_element, col_matrix.rs:113-119; RowMatrix keeps base
// elements, row_matrix.rs:33-41). Not pipelined like h_commit_trace: auxiliary segments are a few columns.
template <class F>
static int h_commit_trace_ext(int hash, uint32_t deg, uint64_t n, uint64_t blowup, uint32_t n_traces, const uint32_t* cols_per_trace,
                              const void* const* column_ptrs, void* const* polys_out, void* const* lde_out, uint8_t* leaves_out,
                              uint8_t* nodes_out, uint8_t* root_out) {
  typedef typename F::T T;
  WG_TRY(check_domain<F>(n, blowup));
  WG_TRY(check_ext_degree(deg));
  if (n_traces == 0 || n_traces > HASH_MAX_TRACES) return fail(WG_ERR_UNSUPPORTED, "1..%d aggregated traces supported", HASH_MAX_TRACES);
  uint64_t total = 0, max_cols = 0;
  for (uint32_t t = 0; t < n_traces; t++) {
    if (cols_per_trace[t] == 0) return fail(WG_ERR_ARG, "trace %u has no columns", t);
    total += cols_per_trace[t];
    if (cols_per_trace[t] > max_cols) max_cols = cols_per_trace[t];
  }
  if (total * deg > 65535) return fail(WG_ERR_UNSUPPORTED, "too many columns");
  const uint32_t nb = (uint32_t)(total * deg);
  const uint64_t L = n * blowup;
  DevBuf cols, polys, scr, leaves, nodes;
  WG_TRY(cols.alloc((size_t)nb * n * sizeof(T)));
  WG_TRY(polys.alloc((size_t)nb * n * sizeof(T)));
  uint64_t se = interpolate_scratch_elems<F>(nb, n);
  const uint64_t se2 = lde_scratch_elems<F>((uint32_t)(max_cols * deg), n, blowup);
  if (se2 > se) se = se2;
  WG_TRY(scr.alloc(se * sizeof(T)));
  WG_TRY(leaves.alloc(L * dlen(hash)));
  WG_TRY(nodes.alloc(L * dlen(hash)));
  std::vector<DevBuf> ldes(n_traces);
  std::vector<uint64_t> rws(n_traces), els(n_traces);
  std::vector<const void*> lde_c(n_traces);
  for (uint32_t t = 0; t < n_traces; t++) {
    els[t] = (uint64_t)cols_per_trace[t] * deg;
    rws[t] = (els[t] + 7) / 8 * 8;
    WG_TRY(ldes[t].alloc(L * rws[t] * sizeof(T)));
    lde_c[t] = ldes[t].p;
  }
  LegacyStreamDrain drain;
  WG_TRY(upload_columns<T>(column_ptrs, (uint32_t)total, n, deg, cols.as<T>(), polys.as<T>(), 0));
  WG_TRY(interpolate_columns_dev<F>(cols.as<T>(), nb, n, polys.as<T>(), scr.as<T>(), se, 0));
  uint64_t c0 = 0;
  for (uint32_t t = 0; t < n_traces; t++) {
    WG_TRY(lde_dev<F>(polys.as<T>() + c0 * n, (uint32_t)els[t], n, blowup, generator<F>(), ldes[t].as<T>(), rws[t], scr.as<T>(), se, 0));
    c0 += els[t];
    if (lde_out && lde_out[t]) CUDA_TRY(cudaMemcpyAsync(lde_out[t], ldes[t].p, L * rws[t] * sizeof(T), cudaMemcpyDeviceToHost, 0));
  }
  WG_TRY(hash_rows_dev<F>(hash, n_traces, lde_c.data(), rws.data(), els.data(), L, leaves.as<uint8_t>(), 0));
  WG_TRY(merkle_dev(hash, leaves.as<uint8_t>(), L, nodes.as<uint8_t>(), 0));
  if (polys_out) WG_TRY(download_columns<T>(polys_out, (uint32_t)total, n, deg, polys.as<T>(), cols.as<T>(), 0));
  if (root_out) CUDA_TRY(cudaMemcpyAsync(root_out, nodes.as<uint8_t>() + dlen(hash), dlen(hash), cudaMemcpyDeviceToHost, 0));
  if (leaves_out) CUDA_TRY(cudaMemcpyAsync(leaves_out, leaves.p, L * dlen(hash), cudaMemcpyDeviceToHost, 0));
  if (nodes_out) CUDA_TRY(cudaMemcpyAsync(nodes_out, nodes.p, L * dlen(hash), cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return WG_OK;
}

// ------------------------------------------------------------------------------------------------ resident commitments
// A commitment whose LDE matrices, leaves and tree stay in HBM: the prover's later steps only need the root, the rows at
// the ~30-100 query positions and their authentication paths (TraceCommitment::query, prover/src/trace/commitment.rs:
// 87-190; MerkleTree::prove_batch, crypto/src/merkle/mod.rs:222-284), so the 33 GB D2H of the host-pointer commit is
// replaced by a few kilobytes. wg_commitment_download() still returns everything for a CPU prover that wants it.
// Freed commitment buffers are parked here and handed to the next commitment of the same shape: cudaMalloc / cudaFree
// of a 28 GB matrix cost tens of milliseconds each. wg_shutdown() (or a failed allocation) empties the cache.
static std::mutex g_cache_mu;
static std::vector<std::pair<void*, size_t>> g_block_cache;
static void cache_release_all() {
  std::lock_guard<std::mutex> lk(g_cache_mu);
  for (auto& b : g_block_cache) cudaFree(b.first);
  g_block_cache.clear();
}
static int cached_malloc(void** p, size_t bytes) {
  {
    std::lock_guard<std::mutex> lk(g_cache_mu);
    for (size_t i = 0; i < g_block_cache.size(); i++)
      if (g_block_cache[i].second == bytes) {
        *p = g_block_cache[i].first;
        g_block_cache.erase(g_block_cache.begin() + i);
        return WG_OK;
      }
  }
  cudaError_t e = dev_malloc(p, bytes);
  if (e != cudaSuccess) {
    *p = nullptr;
    return fail(WG_ERR_NOMEM, "cudaMalloc(%zu): %s", bytes, cudaGetErrorString(e));
  }
  return WG_OK;
}
static void cached_free(void* p, size_t bytes) {
  if (!p) return;
  std::lock_guard<std::mutex> lk(g_cache_mu);
  if (g_block_cache.size() >= 40) {
    cudaFree(p);
    return;
  }
  g_block_cache.emplace_back(p, bytes);
}

struct Commitment {
  int field = 0, hash = 0;
  uint64_t n = 0, blowup = 0, L = 0;
  std::vector<uint32_t> cols;
  std::vector<uint64_t> rws;
  std::vector<void*> ldes;
  uint8_t* leaves = nullptr;
  uint8_t* nodes = nullptr;
  size_t lde_bytes(size_t t) const { return L * rws[t] * (field == WG_FIELD_F64 ? 8 : 16); }
  ~Commitment() {
    for (size_t t = 0; t < ldes.size(); t++) cached_free(ldes[t], lde_bytes(t));
    cached_free(leaves, L * dlen(hash));
    cached_free(nodes, L * dlen(hash));
  }
};

template <class F>
static int commitment_build(Commitment* cm, const void* const* column_ptrs, bool inputs_are_polys, void* const* polys_out,
                            uint8_t* root_out) {
  typedef typename F::T T;
  const uint64_t n = cm->n, blowup = cm->blowup, L = cm->L;
  const uint32_t n_traces = (uint32_t)cm->cols.size();
  WG_TRY(check_domain<F>(n, blowup));
  if (n_traces == 0 || n_traces > HASH_MAX_TRACES) return fail(WG_ERR_UNSUPPORTED, "1..%d aggregated traces supported", HASH_MAX_TRACES);
  if (!hash_ok(cm->hash)) return fail(WG_ERR_ARG, "unknown hash %d", cm->hash);
  uint64_t total_cols = 0, max_cols = 0;
  for (uint32_t c : cm->cols) {
    if (c == 0) return fail(WG_ERR_ARG, "a trace has no columns");
    total_cols += c;
    if (c > max_cols) max_cols = c;
  }
  const size_t colb = n * sizeof(T);
  // persistent buffers
  auto dmalloc = [&](void** p, size_t bytes) -> int { return cached_malloc(p, bytes); };
  cm->ldes.assign(n_traces, nullptr);
  cm->rws.resize(n_traces);
  for (uint32_t t = 0; t < n_traces; t++) {
    cm->rws[t] = (uint64_t)(cm->cols[t] + 7) / 8 * 8;
    WG_TRY(dmalloc(&cm->ldes[t], L * cm->rws[t] * sizeof(T)));
  }
  WG_TRY(dmalloc((void**)&cm->leaves, L * dlen(cm->hash)));
  WG_TRY(dmalloc((void**)&cm->nodes, L * dlen(cm->hash)));
  // temporaries from the shared arena
  uint64_t se = inputs_are_polys ? 0 : interpolate_scratch_elems<F>((uint32_t)(total_cols < 65535 ? total_cols : 65535), n);
  // LDE scratch: whole matrix when it fits next to everything else, else 8-column sweeps
  size_t free_b = 0, total_b = 0;
  CUDA_TRY(cudaMemGetInfo(&free_b, &total_b));
  auto al = [](size_t b) { return (b + 255) / 256 * 256; };
  std::lock_guard<std::mutex> lk(g_commit_mu);
  uint64_t lde_cols = max_cols;
  const size_t fixed = (inputs_are_polys ? 1 : 2) * al(colb * total_cols) + 4096;
  while (lde_cols > 8 && fixed + lde_scratch_elems<F>((uint32_t)lde_cols, n, blowup) * sizeof(T) > free_b + g_arena_bytes) lde_cols = (lde_cols / 2 + 7) / 8 * 8;
  const uint64_t se2 = lde_scratch_elems<F>((uint32_t)lde_cols, n, blowup);
  if (se2 > se) se = se2;
  {
    // several narrow traces: room for the combined LDE matrix when the device has it (lde_traces_dev)
    const uint64_t se_fast = commit_scratch_elems<F>(n_traces, cm->cols.data(), n, blowup);
    if (se_fast > se && fixed + al(se_fast * sizeof(T) + 16) <= free_b + g_arena_bytes) se = se_fast;
  }
  WG_TRY(arena_reserve(fixed + al(se * sizeof(T) + 16)));
  char* cur = g_arena;
  auto take = [&](size_t b) {
    char* r = cur;
    cur += al(b);
    return r;
  };
  T* polys = (T*)take(colb * total_cols);
  T* cols = inputs_are_polys ? nullptr : (T*)take(colb * total_cols);
  T* scr = (T*)take(se * sizeof(T) + 16);
  cudaStream_t sc = g_s_compute, sx = g_s_copy, sh = g_s_h2d;
  StreamDrain drain{{sc, sx, sh}};
  const uint64_t per = (total_cols + H2D_CHUNKS - 1) / H2D_CHUNKS;
  for (uint64_t c0 = 0, k = 0; c0 < total_cols; c0 += per, k++) {
    const uint32_t nc = (uint32_t)((total_cols - c0) < per ? (total_cols - c0) : per);
    T* dst = inputs_are_polys ? polys : cols;
    for (uint64_t c = c0; c < c0 + nc; c++)
      CUDA_TRY(cudaMemcpyAsync((char*)dst + c * colb, column_ptrs[c], colb, cudaMemcpyHostToDevice, sh));
    CUDA_TRY(cudaEventRecord(g_ev_h2d[k], sh));
    CUDA_TRY(cudaStreamWaitEvent(sc, g_ev_h2d[k], 0));
    if (!inputs_are_polys) WG_TRY(interpolate_columns_dev<F>(cols + c0 * n, nc, n, polys + c0 * n, scr, se, sc));
  }
  if (polys_out && !inputs_are_polys) {
    CUDA_TRY(cudaEventRecord(g_ev[0], sc));
    CUDA_TRY(cudaStreamWaitEvent(sx, g_ev[0], 0));
    for (uint64_t c = 0; c < total_cols; c++)
      CUDA_TRY(cudaMemcpyAsync(polys_out[c], (char*)polys + c * colb, colb, cudaMemcpyDeviceToHost, sx));
  }
  std::vector<const void*> lde_c(cm->ldes.begin(), cm->ldes.end());
  std::vector<uint64_t> els(cm->cols.begin(), cm->cols.end());
  WG_TRY(lde_traces_dev<F>(polys, n_traces, cm->cols.data(), n, blowup, generator<F>(), cm->ldes.data(), cm->rws.data(), scr, se, sc,
                           [](uint32_t) { return WG_OK; }));
  WG_TRY(hash_rows_dev<F>(cm->hash, n_traces, lde_c.data(), cm->rws.data(), els.data(), L, cm->leaves, sc));
  WG_TRY(merkle_dev(cm->hash, cm->leaves, L, cm->nodes, sc));
  if (root_out) CUDA_TRY(cudaMemcpyAsync(root_out, cm->nodes + dlen(cm->hash), dlen(cm->hash), cudaMemcpyDeviceToHost, sc));
  CUDA_TRY(cudaStreamSynchronize(sc));
  CUDA_TRY(cudaStreamSynchronize(sx));
  CUDA_TRY(cudaStreamSynchronize(sh));
  return WG_OK;
}

static int commitment_create(int field, int hash, uint64_t n, uint64_t blowup, uint32_t n_traces, const uint32_t* cols_per_trace,
                             const void* const* column_ptrs, bool inputs_are_polys, void* const* polys_out, uint8_t* root_out,
                             void** handle) {
  if (!cols_per_trace || !column_ptrs || !handle) return fail(WG_ERR_ARG, "null pointer");
  if (field != WG_FIELD_F64 && field != WG_FIELD_F128) return fail(WG_ERR_ARG, "unknown field id %d", field);
  WG_TRY(ensure_device());
  Commitment* cm = new Commitment;
  cm->field = field;
  cm->hash = hash;
  cm->n = n;
  cm->blowup = blowup;
  cm->L = n * blowup;
  cm->cols.assign(cols_per_trace, cols_per_trace + n_traces);
  int rc = field == WG_FIELD_F64 ? commitment_build<Gl64>(cm, column_ptrs, inputs_are_polys, polys_out, root_out)
                                 : commitment_build<F128>(cm, column_ptrs, inputs_are_polys, polys_out, root_out);
  if (rc != WG_OK) {
    delete cm;
    *handle = nullptr;
    return rc;
  }
  *handle = cm;
  return WG_OK;
}

// out[p*elems + e] = lde[pos[p]*rw + e] in 8-byte words
__global__ void gather_rows_kernel(const uint64_t* lde, uint64_t rw_words, uint32_t elem_words, const uint64_t* pos, uint32_t n_pos,
                                   uint64_t* out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (uint64_t)n_pos * elem_words) return;
  const uint32_t p = (uint32_t)(i / elem_words), e = (uint32_t)(i % elem_words);
  out[i] = lde[pos[p] * rw_words + e];
}
// digests of `units` 8-byte words each (4 for the 32-byte hashes, 3 for Blake3_192)
__global__ void gather_digests_kernel(const uint2* src, const uint64_t* idx, uint32_t n, uint32_t units, uint2* out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= units * n) return;
  out[i] = src[units * idx[i / units] + (i % units)];
}

// ------------------------------------------------------------------------------------------------ FRI layers
// FriProver::build_layer (fri/src/prover/mod.rs:185-211): transpose_slice -> hash_values -> MerkleTree::new, then
// folding::apply_drp (fri/src/folding/mod.rs:115-146). Base field only.
static inline bool fri_folding_ok(uint32_t ff) { return ff == 2 || ff == 4 || ff == 8 || ff == 16; }

template <class F>
static int fri_commit_layer_dev(int hash, const typename F::T* d_evals, uint64_t n, uint32_t ff, typename F::T* d_transposed,
                                uint8_t* d_leaves, uint8_t* d_nodes, cudaStream_t stream) {
  typedef typename F::T T;
  if (!fri_folding_ok(ff)) return fail(WG_ERR_UNSUPPORTED, "folding factor %u is not supported", ff);  // prover/mod.rs:186
  if (n == 0 || n % ff) return fail(WG_ERR_ARG, "source length must be divisible by %u, but was %llu", ff, (unsigned long long)n);
  const uint64_t rows = n / ff;
  if (rows < 2) return fail(WG_ERR_TOO_FEW, "a Merkle tree needs at least 2 leaves, got %llu", (unsigned long long)rows);
  if (!is_pow2(rows)) return fail(WG_ERR_NOT_POW2, "number of leaves %llu is not a power of two", (unsigned long long)rows);
  {
    ProfScope ps(PH_HASH, stream);
    fri_transpose_kernel<T><<<cdiv(n, 256), 256, 0, stream>>>(d_evals, d_transposed, rows, (uint32_t)ilog2(ff));
    CUDA_TRY(cudaGetLastError());
    g_launches++;
  }
  const void* ptrs[1] = {d_transposed};
  const uint64_t rw = ff;
  WG_TRY(hash_rows_dev<F>(hash, 1, ptrs, &rw, &rw, rows, d_leaves, stream));
  return merkle_dev(hash, d_leaves, rows, d_nodes, stream);
}

template <class F>
static int fri_apply_drp_dev(const typename F::T* d_transposed, uint64_t rows, uint32_t ff, typename F::T offset,
                             typename F::T alpha, typename F::T* d_out, cudaStream_t stream) {
  typedef typename F::T T;
  if (!fri_folding_ok(ff)) return fail(WG_ERR_UNSUPPORTED, "folding factor %u is not supported", ff);
  if (rows == 0) return WG_OK;
  if (!is_pow2(rows)) return fail(WG_ERR_NOT_POW2, "number of rows %llu is not a power of two", (unsigned long long)rows);
  const int log_n = ilog2(rows * ff);
  if (log_n > F::kTwoAdicity || log_n > FRI_MAX_LOG_DOMAIN)
    return fail(WG_ERR_TOO_LARGE, "multiplicative subgroup of size 2^%d does not exist in the specified base field", log_n);
  if (F::is_zero(offset)) return fail(WG_ERR_ZERO_OFFSET, "domain offset cannot be zero");
  if (!elem_is_canonical<F>(alpha) || !elem_is_canonical<F>(offset)) return fail(WG_ERR_ARG, "alpha / offset not a reduced field element");
  FriFoldArgs<F> a;
  memset(&a, 0, sizeof a);
  a.rows_in = d_transposed;
  a.out = d_out;
  a.rows = rows;
  a.log_rows = (uint32_t)ilog2(rows);
  a.alpha_sinv0 = F::mul(alpha, f_inv<F>(offset));
  a.n_inv = f_inv<F>(F::from_u64(ff));
  T gi = f_inv<F>(f_root_of_unity<F>(log_n));
  for (uint32_t b = 0; b < a.log_rows; b++) {
    a.ginv_pow[b] = gi;
    gi = F::mul(gi, gi);
  }
  ProfScope ps(PH_LDE_B, stream);
  const unsigned blocks = cdiv(rows, 128);
  switch (ff) {
    case 2: fri_fold_kernel<F, 1><<<blocks, 128, 0, stream>>>(a); break;
    case 4: fri_fold_kernel<F, 2><<<blocks, 128, 0, stream>>>(a); break;
    case 8: fri_fold_kernel<F, 3><<<blocks, 128, 0, stream>>>(a); break;
    default: fri_fold_kernel<F, 4><<<blocks, 128, 0, stream>>>(a); break;
  }
  CUDA_TRY(cudaGetLastError());
  g_launches++;
  return WG_OK;
}

template <class F>
static int h_fri_commit_layer(int hash, const void* evals, uint64_t n, uint32_t ff, void* transposed_out, uint8_t* leaves_out,
                              uint8_t* nodes_out) {
  typedef typename F::T T;
  if (!fri_folding_ok(ff)) return fail(WG_ERR_UNSUPPORTED, "folding factor %u is not supported", ff);
  if (n == 0 || n % ff) return fail(WG_ERR_ARG, "source length must be divisible by %u, but was %llu", ff, (unsigned long long)n);
  const uint64_t rows = n / ff;
  DevBuf e, t, l, nd;
  WG_TRY(e.alloc(n * sizeof(T)));
  WG_TRY(t.alloc(n * sizeof(T)));
  WG_TRY(l.alloc(rows * dlen(hash)));
  WG_TRY(nd.alloc(rows * dlen(hash)));
  CUDA_TRY(cudaMemcpyAsync(e.p, evals, n * sizeof(T), cudaMemcpyHostToDevice, 0));
  WG_TRY(fri_commit_layer_dev<F>(hash, e.as<T>(), n, ff, t.as<T>(), l.as<uint8_t>(), nd.as<uint8_t>(), 0));
  if (transposed_out) CUDA_TRY(cudaMemcpyAsync(transposed_out, t.p, n * sizeof(T), cudaMemcpyDeviceToHost, 0));
  if (leaves_out) CUDA_TRY(cudaMemcpyAsync(leaves_out, l.p, rows * dlen(hash), cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaMemcpyAsync(nodes_out, nd.p, rows * dlen(hash), cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return WG_OK;
}

template <class F>
static int h_fri_apply_drp(const void* transposed, uint64_t rows, uint32_t ff, const void* offset, const void* alpha, void* out) {
  typedef typename F::T T;
  if (!fri_folding_ok(ff)) return fail(WG_ERR_UNSUPPORTED, "folding factor %u is not supported", ff);
  DevBuf t, o;
  WG_TRY(t.alloc(rows * ff * sizeof(T)));
  WG_TRY(o.alloc(rows * sizeof(T)));
  CUDA_TRY(cudaMemcpyAsync(t.p, transposed, rows * ff * sizeof(T), cudaMemcpyHostToDevice, 0));
  WG_TRY(fri_apply_drp_dev<F>(t.as<T>(), rows, ff, offset_from_internal<F>(offset), offset_from_internal<F>(alpha), o.as<T>(), 0));
  CUDA_TRY(cudaMemcpyAsync(out, o.p, rows * sizeof(T), cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return WG_OK;
}

// ------------------------------------------------------------------------------------------------ no-allocation variant
// Scratch of wg_interpolate_columns_noalloc_dev: a file-scope device array instead of a caller / cudaMalloc buffer. Every
// GPU that loads the library holds these 16 MiB until exit. In HOST code the name below is only a stand-in for the symbol,
// not a device address: the address is fetched once (per process = per device, the library is bound to one device) with
// cudaGetSymbolAddress while that device is current, and only that pointer is handed to kernels.
constexpr size_t NOALLOC_SCRATCH_BYTES = (size_t)WG_NOALLOC_SCRATCH_BYTES;
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];
static void* g_noalloc_ptr = nullptr;
static int g_noalloc_dev = -1;
static std::mutex g_noalloc_mu;

static int noalloc_scratch(void** out) {
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  int dev = -1;
  CUDA_TRY(cudaGetDevice(&dev));
  if (!g_noalloc_ptr || g_noalloc_dev != dev) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p) return fail(WG_ERR_CUDA, "cudaGetSymbolAddress(g_noalloc_scratch): %s", cudaGetErrorString(e));
    g_noalloc_ptr = p;
    g_noalloc_dev = dev;
  }
  *out = g_noalloc_ptr;
  return WG_OK;
}

template <class F>
static int interpolate_columns_noalloc(const typename F::T* d_in, uint32_t n_cols, uint64_t n, typename F::T* d_out, cudaStream_t stream) {
  typedef typename F::T T;
  WG_TRY(check_domain<F>(n, 1));
  if (n_cols == 0) return WG_OK;
  const uint64_t max_elems = NOALLOC_SCRATCH_BYTES / sizeof(T);
  if ((uint64_t)n_cols * n > max_elems)
    return fail(WG_ERR_TOO_LARGE, "no-allocation interpolation holds at most %llu elements (%zu bytes of static scratch); asked for %u x %llu",
                (unsigned long long)max_elems, NOALLOC_SCRATCH_BYTES, n_cols, (unsigned long long)n);
  // the twiddle tables are cached device allocations built on first use of a length by the ordinary entry points; this
  // entry point only looks them up
  const int log_n = ilog2(n);
  int na, nb;
  split_passes(log_n, &na, &nb);
  Tables<F>& tb = tables<F>();
  const bool ok = tb.have_tws(nb, true) && (na == 0 || (tb.have_tws(na, true) && tb.have_tm(log_n, true)));
  if (!ok)
    return fail(WG_ERR_ARG, "twiddle tables for length 2^%d are not resident: run wg_interpolate_columns_dev once with this length "
                "outside the no-allocation region", log_n);
  void* scr = nullptr;
  WG_TRY(noalloc_scratch(&scr));
  // pass A writes every one of the n_cols*n scratch elements before pass B reads them: nothing has to start at zero
  // (the base plan: the split plan of longer transforms builds a table on first use)
  return interpolate_columns_base<F>(d_in, n_cols, n, d_out, (T*)scr, max_elems, stream);
}

}  // namespace wg

#include "dist.cuh"

// ================================================================================================ extern "C"
using namespace wg;

#define WG_FIELD_DISPATCH(field, CALL64, CALL128)                 \
  do {                                                            \
    if ((field) == WG_FIELD_F64) return CALL64;                   \
    if ((field) == WG_FIELD_F128) return CALL128;                 \
    return fail(WG_ERR_ARG, "unknown field id %d", (int)(field)); \
  } while (0)

extern "C" {

int wg_init(int device) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0)
    return fail(WG_ERR_CUDA, "no CUDA device (%s); this library has no CPU fallback", cudaGetErrorString(e));
  if (device < 0 || device >= n) return fail(WG_ERR_ARG, "device %d out of range (have %d)", device, n);
  {
    std::lock_guard<std::mutex> lk(g_mu);
    if (g_device >= 0 && g_device != device)
      return fail(WG_ERR_ARG, "library already bound to device %d (one process per GPU)", g_device);
    g_device = device;
  }
  CUDA_TRY(cudaSetDevice(device));
  return WG_OK;
}

// releases the transfer arena and parked commitment buffers; twiddle tables live for the process lifetime
int wg_shutdown(void) {
  cache_release_all();
  std::lock_guard<std::mutex> lk(g_commit_mu);
  if (g_arena) cudaFree(g_arena);
  g_arena = nullptr;
  g_arena_bytes = 0;
  return WG_OK;
}

const char* wg_last_error(void) { return g_last_error.c_str(); }
uint32_t wg_digest_bytes(int hash) { return hash_ok(hash) ? (uint32_t)dlen(hash) : 0u; }
uint64_t wg_launch_count(void) { return g_launches.load(); }

int wg_set_split_log(int log_len) {
  const int prev = g_split_log.load();
  if (log_len == 0) return prev;
  if (log_len < MAX_LOG_S + 1 || log_len > 2 * MAX_LOG_S)
    return fail(WG_ERR_ARG, "split threshold 2^%d outside [2^%d, 2^%d]", log_len, MAX_LOG_S + 1, 2 * MAX_LOG_S);
  g_split_log.store(log_len);
  return prev;
}

int wg_set_split_preferred_log(int log_len) {
  const int prev = g_split_pref.load();
  if (log_len == 0) return prev;
  if (log_len < MAX_LOG_S + 1 || log_len > 2 * MAX_LOG_S)
    return fail(WG_ERR_ARG, "preferred split threshold 2^%d outside [2^%d, 2^%d]", log_len, MAX_LOG_S + 1, 2 * MAX_LOG_S);
  g_split_pref.store(log_len);
  return prev;
}

int wg_debug_field_op(int field, int op, const void* a, const void* b, void* out, uint32_t n) {
  if (!a || !b || !out) return fail(WG_ERR_ARG, "null pointer");
  WG_TRY(ensure_device());
  WG_FIELD_DISPATCH(field, h_field_op<Gl64>(op, a, b, out, n), h_field_op<F128>(op, a, b, out, n));
}

int wg_profile_enable(int on) {
  std::lock_guard<std::mutex> lk(g_mu);
  g_prof_on = on != 0;
  return WG_OK;
}
int wg_profile_read(float* ms_out, uint32_t* count_out) {
  std::lock_guard<std::mutex> lk(g_mu);
  for (int i = 0; i < PH_COUNT; i++) {
    ms_out[i] = 0.f;
    count_out[i] = 0;
  }
  for (auto& r : g_prof) {
    CUDA_TRY(cudaEventSynchronize(r.b));
    float ms = 0.f;
    CUDA_TRY(cudaEventElapsedTime(&ms, r.a, r.b));
    ms_out[r.phase] += ms;
    count_out[r.phase]++;
    g_prof_pool.push_back(r.a);
    g_prof_pool.push_back(r.b);
  }
  g_prof.clear();
  return WG_OK;
}

int wg_get_twiddles(int field, uint64_t n, int inverse, void* out) {
  if (!out) return fail(WG_ERR_ARG, "null output");
  WG_TRY(ensure_device());
  WG_FIELD_DISPATCH(field, h_get_twiddles<Gl64>(n, inverse, out), h_get_twiddles<F128>(n, inverse, out));
}

int wg_interpolate_poly(int field, void* evals, uint64_t n, const void* inv_twiddles) {
  (void)inv_twiddles;  // the kernels keep their own on-chip twiddle tables
  if (!evals) return fail(WG_ERR_ARG, "null input");
  WG_TRY(ensure_device());
  const void* in[1] = {evals};
  void* out[1] = {evals};
  WG_FIELD_DISPATCH(field, h_interpolate_columns<Gl64>(in, 1, n, out), h_interpolate_columns<F128>(in, 1, n, out));
}

int wg_interpolate_poly_with_offset(int field, void* evals, uint64_t n, const void* inv_twiddles, const void* offset) {
  (void)inv_twiddles;
  if (!evals || !offset) return fail(WG_ERR_ARG, "null pointer");
  WG_TRY(ensure_device());
  WG_FIELD_DISPATCH(field, h_interpolate_poly_with_offset<Gl64>(evals, n, offset), h_interpolate_poly_with_offset<F128>(evals, n, offset));
}

int wg_evaluate_poly(int field, void* p, uint64_t n, const void* twiddles) {
  (void)twiddles;
  if (!p) return fail(WG_ERR_ARG, "null input");
  WG_TRY(ensure_device());
  const void* in[1] = {p};
  if (field == WG_FIELD_F64) {
    const Gl64::T one = Gl64::one();
    return h_evaluate_polys_over<Gl64>(in, 1, n, 1, &one, p, 1);
  } else if (field == WG_FIELD_F128) {
    const F128::T one = F128::one();
    return h_evaluate_polys_over<F128>(in, 1, n, 1, &one, p, 1);
  }
  return fail(WG_ERR_ARG, "unknown field %d", field);
}

int wg_evaluate_poly_with_offset(int field, const void* p, uint64_t n, const void* twiddles, const void* offset,
                                 uint64_t blowup, void* out) {
  (void)twiddles;
  if (!p || !offset || !out) return fail(WG_ERR_ARG, "null pointer");
  WG_TRY(ensure_device());
  const void* in[1] = {p};
  WG_FIELD_DISPATCH(field, h_evaluate_polys_over<Gl64>(in, 1, n, blowup, offset, out, 1),
                    h_evaluate_polys_over<F128>(in, 1, n, blowup, offset, out, 1));
}

int wg_interpolate_columns(int field, const void* const* columns, uint32_t n_cols, uint64_t n, void* const* polys_out) {
  if (!columns || !polys_out) return fail(WG_ERR_ARG, "null pointer");
  WG_TRY(ensure_device());
  WG_FIELD_DISPATCH(field, h_interpolate_columns<Gl64>(columns, n_cols, n, polys_out),
                    h_interpolate_columns<F128>(columns, n_cols, n, polys_out));
}

int wg_evaluate_polys_over(int field, const void* const* polys, uint32_t n_cols, uint64_t n, uint64_t blowup,
                           const void* offset, void* out) {
  if (!polys || !offset || !out) return fail(WG_ERR_ARG, "null pointer");
  WG_TRY(ensure_device());
  const uint64_t rw = (uint64_t)(n_cols + 7) / 8 * 8;
  WG_FIELD_DISPATCH(field, h_evaluate_polys_over<Gl64>(polys, n_cols, n, blowup, offset, out, rw),
                    h_evaluate_polys_over<F128>(polys, n_cols, n, blowup, offset, out, rw));
}

int wg_hash_rows(int field, int hash, const void* row_major, uint64_t rows, uint64_t row_width, uint64_t elems_per_row,
                 uint8_t* leaves_out) {
  if (!row_major || !leaves_out) return fail(WG_ERR_ARG, "null pointer");
  WG_TRY(ensure_device());
  WG_FIELD_DISPATCH(field, h_hash_rows<Gl64>(hash, row_major, rows, row_width, elems_per_row, leaves_out),
                    h_hash_rows<F128>(hash, row_major, rows, row_width, elems_per_row, leaves_out));
}

int wg_evaluate_columns_at(int field, const void* const* polys, uint32_t n_cols, uint64_t n, const void* x, void* out) {
  if (!polys || !x || !out) return fail(WG_ERR_ARG, "null pointer");
  WG_TRY(ensure_device());
  WG_FIELD_DISPATCH(field, h_evaluate_columns_at<Gl64>(polys, n_cols, n, x, out), h_evaluate_columns_at<F128>(polys, n_cols, n, x, out));
}

int wg_hash_comb_rows(int field, int hash, uint32_t n_traces, const void* const* matrices, const uint64_t* row_widths,
                      const uint64_t* elems_per_row, uint64_t rows, uint8_t* leaves_out) {
  if (!matrices || !row_widths || !elems_per_row || !leaves_out) return fail(WG_ERR_ARG, "null pointer");
  WG_TRY(ensure_device());
  WG_FIELD_DISPATCH(field, h_hash_comb_rows<Gl64>(hash, n_traces, matrices, row_widths, elems_per_row, rows, leaves_out),
                    h_hash_comb_rows<F128>(hash, n_traces, matrices, row_widths, elems_per_row, rows, leaves_out));
}

int wg_merkle_new(int hash, const uint8_t* leaves, uint64_t n_leaves, uint8_t* nodes_out) {
  if (!leaves || !nodes_out) return fail(WG_ERR_ARG, "null pointer");
  if (n_leaves < 2) return fail(WG_ERR_TOO_FEW, "a Merkle tree needs at least 2 leaves, got %llu", (unsigned long long)n_leaves);
  if (!is_pow2(n_leaves)) return fail(WG_ERR_NOT_POW2, "number of leaves %llu is not a power of two", (unsigned long long)n_leaves);
  WG_TRY(ensure_device());
  DevBuf l, nd;
  WG_TRY(l.alloc(n_leaves * dlen(hash)));
  WG_TRY(nd.alloc(n_leaves * dlen(hash)));
  CUDA_TRY(cudaMemcpyAsync(l.p, leaves, n_leaves * dlen(hash), cudaMemcpyHostToDevice, 0));
  WG_TRY(merkle_dev(hash, l.as<uint8_t>(), n_leaves, nd.as<uint8_t>(), 0));
  CUDA_TRY(cudaMemcpyAsync(nodes_out, nd.p, n_leaves * dlen(hash), cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return WG_OK;
}

int wg_commit_trace(int field, int hash, uint64_t n_rows, uint64_t blowup, uint32_t n_traces,
                    const uint32_t* cols_per_trace, const void* const* column_ptrs, void* const* polys_out,
                    void* const* lde_out, uint8_t* leaves_out, uint8_t* nodes_out, uint8_t* root_out) {
  if (!cols_per_trace || !column_ptrs) return fail(WG_ERR_ARG, "null pointer");
  WG_TRY(ensure_device());
  WG_FIELD_DISPATCH(field,
                    h_commit_trace<Gl64>(hash, n_rows, blowup, n_traces, cols_per_trace, column_ptrs, polys_out, lde_out,
                                         leaves_out, nodes_out, root_out),
                    h_commit_trace<F128>(hash, n_rows, blowup, n_traces, cols_per_trace, column_ptrs, polys_out, lde_out,
                                         leaves_out, nodes_out, root_out));
}

int wg_commit_polys(int field, int hash, uint64_t n_coeffs, uint64_t blowup, uint32_t n_cols, const void* const* poly_ptrs,
                    void* lde_out, uint8_t* leaves_out, uint8_t* nodes_out, uint8_t* root_out) {
  void* h = nullptr;
  WG_TRY(commitment_create(field, hash, n_coeffs, blowup, 1, &n_cols, poly_ptrs, true, nullptr, root_out, &h));
  void* ldes[1] = {lde_out};
  int rc = wg_commitment_download(h, ldes, leaves_out, nodes_out);
  wg_commitment_free(h);
  return rc;
}

// ---- resident commitments
int wg_commitment_from_trace(int field, int hash, uint64_t n_rows, uint64_t blowup, uint32_t n_traces,
                             const uint32_t* cols_per_trace, const void* const* column_ptrs, void* const* polys_out,
                             uint8_t* root_out, void** handle) {
  return commitment_create(field, hash, n_rows, blowup, n_traces, cols_per_trace, column_ptrs, false, polys_out, root_out, handle);
}

int wg_commitment_from_polys(int field, int hash, uint64_t n_coeffs, uint64_t blowup, uint32_t n_cols, const void* const* poly_ptrs,
                             uint8_t* root_out, void** handle) {
  return commitment_create(field, hash, n_coeffs, blowup, 1, &n_cols, poly_ptrs, true, nullptr, root_out, handle);
}

int wg_commitment_rows(const void* handle, uint32_t trace, uint32_t n_pos, const uint64_t* positions, void* rows_out) {
  const Commitment* cm = (const Commitment*)handle;
  if (!cm || !positions || !rows_out) return fail(WG_ERR_ARG, "null pointer");
  if (trace >= cm->cols.size()) return fail(WG_ERR_ARG, "trace index %u out of range", trace);
  if (n_pos == 0) return WG_OK;
  for (uint32_t i = 0; i < n_pos; i++)
    if (positions[i] >= cm->L) return fail(WG_ERR_ARG, "position %llu out of range (%llu rows)", (unsigned long long)positions[i], (unsigned long long)cm->L);
  WG_TRY(ensure_device());
  const uint32_t w = cm->field == WG_FIELD_F64 ? 1 : 2;
  const uint32_t ew = cm->cols[trace] * w;
  DevBuf dpos, dout;
  WG_TRY(dpos.alloc(n_pos * 8));
  WG_TRY(dout.alloc((size_t)n_pos * ew * 8));
  CUDA_TRY(cudaMemcpyAsync(dpos.p, positions, n_pos * 8, cudaMemcpyHostToDevice, 0));
  gather_rows_kernel<<<cdiv((uint64_t)n_pos * ew, 256), 256, 0, 0>>>((const uint64_t*)cm->ldes[trace], cm->rws[trace] * w, ew,
                                                                   dpos.as<uint64_t>(), n_pos, dout.as<uint64_t>());
  CUDA_TRY(cudaGetLastError());
  g_launches++;
  CUDA_TRY(cudaMemcpyAsync(rows_out, dout.p, (size_t)n_pos * ew * 8, cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return WG_OK;
}

int wg_commitment_digests(const void* handle, int which, uint32_t n, const uint64_t* indexes, uint8_t* out) {
  const Commitment* cm = (const Commitment*)handle;
  if (!cm || !indexes || !out) return fail(WG_ERR_ARG, "null pointer");
  if (which != 0 && which != 1) return fail(WG_ERR_ARG, "which must be 0 (leaves) or 1 (nodes)");
  if (n == 0) return WG_OK;
  for (uint32_t i = 0; i < n; i++)
    if (indexes[i] >= cm->L) return fail(WG_ERR_ARG, "index %llu out of range (%llu digests)", (unsigned long long)indexes[i], (unsigned long long)cm->L);
  WG_TRY(ensure_device());
  DevBuf didx, dout;
  WG_TRY(didx.alloc(n * 8));
  WG_TRY(dout.alloc((size_t)n * dlen(cm->hash)));
  CUDA_TRY(cudaMemcpyAsync(didx.p, indexes, n * 8, cudaMemcpyHostToDevice, 0));
  const uint32_t units = (uint32_t)dlen(cm->hash) / 8;
  gather_digests_kernel<<<cdiv(units * (uint64_t)n, 256), 256, 0, 0>>>((const uint2*)(which ? cm->nodes : cm->leaves), didx.as<uint64_t>(), n,
                                                                      units, dout.as<uint2>());
  CUDA_TRY(cudaGetLastError());
  g_launches++;
  CUDA_TRY(cudaMemcpyAsync(out, dout.p, (size_t)n * dlen(cm->hash), cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return WG_OK;
}

int wg_commitment_info(const void* handle, uint64_t* lde_rows, uint32_t* n_traces, uint32_t* cols_per_trace, uint32_t* row_widths) {
  const Commitment* cm = (const Commitment*)handle;
  if (!cm) return fail(WG_ERR_ARG, "null handle");
  if (lde_rows) *lde_rows = cm->L;
  if (n_traces) *n_traces = (uint32_t)cm->cols.size();
  for (size_t t = 0; t < cm->cols.size(); t++) {
    if (cols_per_trace) cols_per_trace[t] = cm->cols[t];
    if (row_widths) row_widths[t] = (uint32_t)cm->rws[t];
  }
  return WG_OK;
}

int wg_commitment_download(const void* handle, void* const* lde_out, uint8_t* leaves_out, uint8_t* nodes_out) {
  const Commitment* cm = (const Commitment*)handle;
  if (!cm) return fail(WG_ERR_ARG, "null handle");
  WG_TRY(ensure_device());
  const size_t es = cm->field == WG_FIELD_F64 ? 8 : 16;
  for (size_t t = 0; lde_out && t < cm->cols.size(); t++)
    if (lde_out[t]) CUDA_TRY(cudaMemcpyAsync(lde_out[t], cm->ldes[t], cm->L * cm->rws[t] * es, cudaMemcpyDeviceToHost, 0));
  if (leaves_out) CUDA_TRY(cudaMemcpyAsync(leaves_out, cm->leaves, cm->L * dlen(cm->hash), cudaMemcpyDeviceToHost, 0));
  if (nodes_out) CUDA_TRY(cudaMemcpyAsync(nodes_out, cm->nodes, cm->L * dlen(cm->hash), cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return WG_OK;
}

int wg_commitment_free(void* handle) {
  delete (Commitment*)handle;
  return WG_OK;
}

// ---- device-resident
uint64_t wg_interpolate_scratch_elems(int field, uint32_t n_cols, uint64_t n) {
  if (!is_pow2(n)) return 0;
  return field == WG_FIELD_F64 ? interpolate_scratch_elems<Gl64>(n_cols, n) : interpolate_scratch_elems<F128>(n_cols, n);
}
uint64_t wg_lde_scratch_elems(int field, uint32_t n_cols, uint64_t n, uint64_t blowup) {
  if (!is_pow2(n)) return 0;
  return field == WG_FIELD_F64 ? lde_scratch_elems<Gl64>(n_cols, n, blowup) : lde_scratch_elems<F128>(n_cols, n, blowup);
}
uint64_t wg_commit_scratch_elems(int field, uint64_t n_rows, uint64_t blowup, uint32_t n_traces, const uint32_t* cols_per_trace) {
  if (!cols_per_trace || n_traces == 0 || !is_pow2(n_rows) || !is_pow2(blowup)) return 0;
  return field == WG_FIELD_F64 ? commit_scratch_elems<Gl64>(n_traces, cols_per_trace, n_rows, blowup)
                               : commit_scratch_elems<F128>(n_traces, cols_per_trace, n_rows, blowup);
}

int wg_interpolate_columns_dev(int field, const void* d_cols, uint32_t n_cols, uint64_t n, void* d_polys, void* d_scratch,
                               uint64_t scratch_elems, void* stream) {
  WG_TRY(ensure_device());
  cudaStream_t s = (cudaStream_t)stream;
  WG_FIELD_DISPATCH(field,
                    interpolate_columns_dev<Gl64>((const uint64_t*)d_cols, n_cols, n, (uint64_t*)d_polys, (uint64_t*)d_scratch, scratch_elems, s),
                    interpolate_columns_dev<F128>((const u128*)d_cols, n_cols, n, (u128*)d_polys, (u128*)d_scratch, scratch_elems, s));
}

int wg_interpolate_columns_noalloc_dev(int field, const void* d_cols, uint32_t n_cols, uint64_t n, void* d_polys, void* stream) {
  if (!d_cols || !d_polys) return fail(WG_ERR_ARG, "null pointer");
  WG_TRY(ensure_device());
  cudaStream_t s = (cudaStream_t)stream;
  WG_FIELD_DISPATCH(field, interpolate_columns_noalloc<Gl64>((const uint64_t*)d_cols, n_cols, n, (uint64_t*)d_polys, s),
                    interpolate_columns_noalloc<F128>((const u128*)d_cols, n_cols, n, (u128*)d_polys, s));
}

int wg_lde_dev(int field, const void* d_polys, uint32_t n_cols, uint64_t n, uint64_t blowup, const void* offset_host,
               void* d_lde, uint64_t row_width, void* d_scratch, uint64_t scratch_elems, void* stream) {
  if (!offset_host) return fail(WG_ERR_ARG, "null offset");
  WG_TRY(ensure_device());
  cudaStream_t s = (cudaStream_t)stream;
  WG_FIELD_DISPATCH(field,
                    lde_dev<Gl64>((const uint64_t*)d_polys, n_cols, n, blowup, offset_from_internal<Gl64>(offset_host), (uint64_t*)d_lde, row_width, (uint64_t*)d_scratch, scratch_elems, s),
                    lde_dev<F128>((const u128*)d_polys, n_cols, n, blowup, offset_from_internal<F128>(offset_host), (u128*)d_lde, row_width, (u128*)d_scratch, scratch_elems, s));
}

int wg_hash_rows_dev(int field, int hash, uint32_t n_traces, const void* const* d_ldes, const uint64_t* row_widths,
                     const uint64_t* elems_per_row, uint64_t rows, void* d_leaves, void* stream) {
  WG_TRY(ensure_device());
  cudaStream_t s = (cudaStream_t)stream;
  WG_FIELD_DISPATCH(field, hash_rows_dev<Gl64>(hash, n_traces, d_ldes, row_widths, elems_per_row, rows, (uint8_t*)d_leaves, s),
                    hash_rows_dev<F128>(hash, n_traces, d_ldes, row_widths, elems_per_row, rows, (uint8_t*)d_leaves, s));
}

int wg_merkle_dev(int hash, const void* d_leaves, uint64_t n_leaves, void* d_nodes, void* stream) {
  WG_TRY(ensure_device());
  return merkle_dev(hash, (const uint8_t*)d_leaves, n_leaves, (uint8_t*)d_nodes, (cudaStream_t)stream);
}

int wg_commit_trace_dev(int field, int hash, uint64_t n_rows, uint64_t blowup, uint32_t n_traces,
                        const uint32_t* cols_per_trace, const void* d_cols, void* d_polys, void* const* d_ldes,
                        void* d_leaves, void* d_nodes, void* d_scratch, uint64_t scratch_elems, void* stream) {
  if (!cols_per_trace || !d_cols || !d_polys || !d_ldes || !d_leaves || !d_nodes) return fail(WG_ERR_ARG, "null pointer");
  WG_TRY(ensure_device());
  cudaStream_t s = (cudaStream_t)stream;
  WG_FIELD_DISPATCH(field,
                    commit_trace_dev<Gl64>(hash, n_rows, blowup, n_traces, cols_per_trace, (const uint64_t*)d_cols, (uint64_t*)d_polys, d_ldes, (uint8_t*)d_leaves, (uint8_t*)d_nodes, (uint64_t*)d_scratch, scratch_elems, s),
                    commit_trace_dev<F128>(hash, n_rows, blowup, n_traces, cols_per_trace, (const u128*)d_cols, (u128*)d_polys, d_ldes, (uint8_t*)d_leaves, (uint8_t*)d_nodes, (u128*)d_scratch, scratch_elems, s));
}

// ---- FRI layers
int wg_fri_commit_layer(int field, int hash, const void* evaluations, uint64_t n, uint32_t folding_factor, void* transposed_out,
                        uint8_t* leaves_out, uint8_t* nodes_out) {
  if (!evaluations || !nodes_out) return fail(WG_ERR_ARG, "null pointer");
  WG_TRY(ensure_device());
  WG_FIELD_DISPATCH(field, h_fri_commit_layer<Gl64>(hash, evaluations, n, folding_factor, transposed_out, leaves_out, nodes_out),
                    h_fri_commit_layer<F128>(hash, evaluations, n, folding_factor, transposed_out, leaves_out, nodes_out));
}

int wg_fri_apply_drp(int field, const void* transposed, uint64_t rows, uint32_t folding_factor, const void* domain_offset,
                     const void* alpha, void* out) {
  if (!transposed || !domain_offset || !alpha || !out) return fail(WG_ERR_ARG, "null pointer");
  WG_TRY(ensure_device());
  WG_FIELD_DISPATCH(field, h_fri_apply_drp<Gl64>(transposed, rows, folding_factor, domain_offset, alpha, out),
                    h_fri_apply_drp<F128>(transposed, rows, folding_factor, domain_offset, alpha, out));
}

int wg_fri_commit_layer_dev(int field, int hash, const void* d_evaluations, uint64_t n, uint32_t folding_factor,
                            void* d_transposed, void* d_leaves, void* d_nodes, void* stream) {
  if (!d_evaluations || !d_transposed || !d_leaves || !d_nodes) return fail(WG_ERR_ARG, "null pointer");
  WG_TRY(ensure_device());
  cudaStream_t s = (cudaStream_t)stream;
  WG_FIELD_DISPATCH(field,
                    fri_commit_layer_dev<Gl64>(hash, (const uint64_t*)d_evaluations, n, folding_factor, (uint64_t*)d_transposed, (uint8_t*)d_leaves, (uint8_t*)d_nodes, s),
                    fri_commit_layer_dev<F128>(hash, (const u128*)d_evaluations, n, folding_factor, (u128*)d_transposed, (uint8_t*)d_leaves, (uint8_t*)d_nodes, s));
}

int wg_fri_apply_drp_dev(int field, const void* d_transposed, uint64_t rows, uint32_t folding_factor, const void* domain_offset_host,
                         const void* alpha_host, void* d_out, void* stream) {
  if (!d_transposed || !domain_offset_host || !alpha_host || !d_out) return fail(WG_ERR_ARG, "null pointer");
  WG_TRY(ensure_device());
  cudaStream_t s = (cudaStream_t)stream;
  WG_FIELD_DISPATCH(field,
                    fri_apply_drp_dev<Gl64>((const uint64_t*)d_transposed, rows, folding_factor, offset_from_internal<Gl64>(domain_offset_host), offset_from_internal<Gl64>(alpha_host), (uint64_t*)d_out, s),
                    fri_apply_drp_dev<F128>((const u128*)d_transposed, rows, folding_factor, offset_from_internal<F128>(domain_offset_host), offset_from_internal<F128>(alpha_host), (u128*)d_out, s));
}

// ---- extension-field element layouts
int wg_interpolate_poly_ext(int field, uint32_t ext_degree, void* evals, uint64_t n, const void* inv_twiddles) {
  (void)inv_twiddles;
  if (!evals) return fail(WG_ERR_ARG, "null input");
  WG_TRY(ensure_device());
  const void* in[1] = {evals};
  void* out[1] = {evals};
  WG_FIELD_DISPATCH(field, h_interpolate_columns<Gl64>(in, 1, n, out, ext_degree), h_interpolate_columns<F128>(in, 1, n, out, ext_degree));
}

int wg_interpolate_poly_with_offset_ext(int field, uint32_t ext_degree, void* evals, uint64_t n, const void* inv_twiddles,
                                        const void* offset) {
  (void)inv_twiddles;
  if (!evals || !offset) return fail(WG_ERR_ARG, "null pointer");
  WG_TRY(ensure_device());
  WG_FIELD_DISPATCH(field, h_interpolate_poly_with_offset<Gl64>(evals, n, offset, ext_degree),
                    h_interpolate_poly_with_offset<F128>(evals, n, offset, ext_degree));
}

int wg_evaluate_poly_ext(int field, uint32_t ext_degree, void* p, uint64_t n, const void* twiddles) {
  (void)twiddles;
  if (!p) return fail(WG_ERR_ARG, "null input");
  WG_TRY(ensure_device());
  const void* in[1] = {p};
  if (field == WG_FIELD_F64) {
    const Gl64::T one = Gl64::one();
    return h_evaluate_polys_over<Gl64>(in, 1, n, 1, &one, p, ext_degree, ext_degree);
  } else if (field == WG_FIELD_F128) {
    const F128::T one = F128::one();
    return h_evaluate_polys_over<F128>(in, 1, n, 1, &one, p, ext_degree, ext_degree);
  }
  return fail(WG_ERR_ARG, "unknown field %d", field);
}

int wg_evaluate_poly_with_offset_ext(int field, uint32_t ext_degree, const void* p, uint64_t n, const void* twiddles,
                                     const void* offset, uint64_t blowup, void* out) {
  (void)twiddles;
  if (!p || !offset || !out) return fail(WG_ERR_ARG, "null pointer");
  WG_TRY(ensure_device());
  const void* in[1] = {p};
  WG_FIELD_DISPATCH(field, h_evaluate_polys_over<Gl64>(in, 1, n, blowup, offset, out, ext_degree, ext_degree),
                    h_evaluate_polys_over<F128>(in, 1, n, blowup, offset, out, ext_degree, ext_degree));
}

int wg_interpolate_columns_ext(int field, uint32_t ext_degree, const void* const* columns, uint32_t n_cols, uint64_t n,
                               void* const* polys_out) {
  if (!columns || !polys_out) return fail(WG_ERR_ARG, "null pointer");
  WG_TRY(ensure_device());
  WG_FIELD_DISPATCH(field, h_interpolate_columns<Gl64>(columns, n_cols, n, polys_out, ext_degree),
                    h_interpolate_columns<F128>(columns, n_cols, n, polys_out, ext_degree));
}

int wg_evaluate_polys_over_ext(int field, uint32_t ext_degree, const void* const* polys, uint32_t n_cols, uint64_t n,
                               uint64_t blowup, const void* offset, void* out) {
  if (!polys || !offset || !out) return fail(WG_ERR_ARG, "null pointer");
  WG_TRY(ensure_device());
  WG_TRY(check_ext_degree(ext_degree));
  const uint64_t rw = ((uint64_t)n_cols * ext_degree + 7) / 8 * 8;
  WG_FIELD_DISPATCH(field, h_evaluate_polys_over<Gl64>(polys, n_cols, n, blowup, offset, out, rw, ext_degree),
                    h_evaluate_polys_over<F128>(polys, n_cols, n, blowup, offset, out, rw, ext_degree));
}

int wg_commit_trace_ext(int field, int hash, uint32_t ext_degree, uint64_t n_rows, uint64_t blowup, uint32_t n_traces,
                        const uint32_t* cols_per_trace, const void* const* column_ptrs, void* const* polys_out,
                        void* const* lde_out, uint8_t* leaves_out, uint8_t* nodes_out, uint8_t* root_out) {
  if (!cols_per_trace || !column_ptrs) return fail(WG_ERR_ARG, "null pointer");
  WG_TRY(ensure_device());
  WG_FIELD_DISPATCH(field,
                    h_commit_trace_ext<Gl64>(hash, ext_degree, n_rows, blowup, n_traces, cols_per_trace, column_ptrs, polys_out, lde_out, leaves_out, nodes_out, root_out),
                    h_commit_trace_ext<F128>(hash, ext_degree, n_rows, blowup, n_traces, cols_per_trace, column_ptrs, polys_out, lde_out, leaves_out, nodes_out, root_out));
}

// ---- multi-GPU: LDE that stores its rows straight into the row-block owners' buffers over NVLink
int wg_peer_alloc(uint64_t bytes, void** d_ptr, uint8_t* handle_out) {
  if (!d_ptr || !handle_out) return fail(WG_ERR_ARG, "null pointer");
  WG_TRY(ensure_device());
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, bytes ? bytes : 256);  // a whole allocation of its own: IPC handles name allocations
  if (e != cudaSuccess) return fail(WG_ERR_NOMEM, "cudaMalloc(%llu): %s", (unsigned long long)bytes, cudaGetErrorString(e));
  cudaIpcMemHandle_t h;
  e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    return fail(WG_ERR_CUDA, "cudaIpcGetMemHandle: %s", cudaGetErrorString(e));
  }
  static_assert(sizeof(h) == WG_PEER_HANDLE_BYTES, "IPC handle size");
  memcpy(handle_out, &h, sizeof h);
  *d_ptr = p;
  return WG_OK;
}

int wg_peer_open(const uint8_t* handle, void** d_ptr) {
  if (!handle || !d_ptr) return fail(WG_ERR_ARG, "null pointer");
  WG_TRY(ensure_device());
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof h);
  CUDA_TRY(cudaIpcOpenMemHandle(d_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return WG_OK;
}

int wg_peer_close(void* d_ptr) {
  if (!d_ptr) return WG_OK;
  WG_TRY(ensure_device());
  CUDA_TRY(cudaIpcCloseMemHandle(d_ptr));
  return WG_OK;
}

int wg_peer_free(void* d_ptr) {
  if (!d_ptr) return WG_OK;
  WG_TRY(ensure_device());
  CUDA_TRY(cudaFree(d_ptr));
  return WG_OK;
}

int wg_lde_scatter_dev(int field, const void* d_polys, uint32_t n_cols, uint64_t n, uint64_t blowup, const void* offset_host,
                       uint32_t n_owners, void* const* d_row_blocks, uint64_t row_width, void* d_scratch, uint64_t scratch_elems,
                       void* stream) {
  if (!offset_host || !d_row_blocks || !d_polys) return fail(WG_ERR_ARG, "null pointer");
  if (n_owners == 0) return fail(WG_ERR_ARG, "no row-block owners");
  WG_TRY(ensure_device());
  cudaStream_t s = (cudaStream_t)stream;
  WG_FIELD_DISPATCH(field,
                    lde_dev<Gl64>((const uint64_t*)d_polys, n_cols, n, blowup, offset_from_internal<Gl64>(offset_host), nullptr, row_width, (uint64_t*)d_scratch, scratch_elems, s, (uint64_t* const*)d_row_blocks, n_owners),
                    lde_dev<F128>((const u128*)d_polys, n_cols, n, blowup, offset_from_internal<F128>(offset_host), nullptr, row_width, (u128*)d_scratch, scratch_elems, s, (u128* const*)d_row_blocks, n_owners));
}

// ---- sharded commit (dist.cuh)
int wg_dist_unique_id(uint8_t* id_out) {
  if (!id_out) return fail(WG_ERR_ARG, "null pointer");
  WG_TRY(nccl_load());
  static_assert(sizeof(ncclUniqueId) == WG_DIST_ID_BYTES, "ncclUniqueId size");
  ncclUniqueId id;
  NCCL_TRY(g_nccl.GetUniqueId(&id));
  memcpy(id_out, &id, sizeof id);
  return WG_OK;
}

int wg_dist_init(int rank, int world, const uint8_t* id) {
  if (world < 1 || rank < 0 || rank >= world) return fail(WG_ERR_ARG, "rank %d of %d", rank, world);
  if (!is_pow2((uint64_t)world)) return fail(WG_ERR_ARG, "number of ranks %d must be a power of two", world);
  WG_TRY(ensure_device());
  std::lock_guard<std::mutex> lk(g_dist_mu);
  if (g_dist.on) return fail(WG_ERR_ARG, "wg_dist_init called twice (rank %d of %d is active)", g_dist.rank, g_dist.world);
  WG_TRY(dist_streams());
  if (world > 1) {
    if (!id) return fail(WG_ERR_ARG, "null NCCL id");
    WG_TRY(nccl_load());
    ncclUniqueId nid;
    memcpy(&nid, id, sizeof nid);
    NCCL_TRY(g_nccl.CommInitRank(&g_dist.comm, world, nid, rank));
  }
  g_dist.rank = rank;
  g_dist.world = world;
  g_dist.on = true;
  return WG_OK;
}

int wg_dist_shutdown(void) {
  std::lock_guard<std::mutex> lk(g_dist_mu);
  if (!g_dist.on) return WG_OK;
  cudaDeviceSynchronize();
  if (g_dist.world > 1 && g_dist.comm) {
    // collective: every rank stops using the peers' buffers before anyone unmaps or frees
    uint8_t* tok = g_dist.d_small;
    NCCL_TRY(g_nccl.AllGather(tok + 64 * g_dist.world, tok, 4, ncclUint8, g_dist.comm, g_dist.s_comm));
    CUDA_TRY(cudaStreamSynchronize(g_dist.s_comm));
    for (auto& pb : g_dist.peer_bufs)
      for (int g = 0; g < (int)pb.peers.size(); g++)
        if (g != g_dist.rank && pb.peers[g]) {
          cudaIpcCloseMemHandle(pb.peers[g]);
          pb.peers[g] = nullptr;
        }
    NCCL_TRY(g_nccl.AllGather(tok + 64 * g_dist.world, tok, 4, ncclUint8, g_dist.comm, g_dist.s_comm));
    CUDA_TRY(cudaStreamSynchronize(g_dist.s_comm));
    for (auto& pb : g_dist.peer_bufs) cudaFree(pb.ptr);
    g_dist.peer_bufs.clear();
    g_nccl.CommDestroy(g_dist.comm);
    g_dist.comm = nullptr;
  }
  g_dist.on = false;
  g_dist.rank = 0;
  g_dist.world = 1;
  return WG_OK;
}

int wg_dist_info(int* rank, int* world) {
  if (rank) *rank = g_dist.rank;
  if (world) *world = g_dist.world;
  return WG_OK;
}

int wg_shard_columns(uint32_t total_cols, uint32_t world, uint32_t rank, uint32_t* start, uint32_t* count) {
  if (!start || !count || world == 0 || rank >= world) return fail(WG_ERR_ARG, "bad shard query");
  shard_columns(total_cols, world, rank, start, count);
  return WG_OK;
}
int wg_shard_chunks(uint32_t cols, uint32_t k, uint32_t idx, uint32_t* start, uint32_t* count) {
  if (!start || !count || k == 0 || idx >= k) return fail(WG_ERR_ARG, "bad chunk query");
  shard_chunk(cols, k, idx, start, count);
  return WG_OK;
}
int wg_shard_plan(uint32_t total_cols, uint32_t world, int exchange, uint32_t chunks, int* exchange_out, uint32_t* chunks_out) {
  if (!exchange_out || !chunks_out || world == 0) return fail(WG_ERR_ARG, "bad plan query");
  if (exchange < WG_EXCHANGE_AUTO || exchange > WG_EXCHANGE_P2P_COPY) return fail(WG_ERR_ARG, "unknown exchange form %d", exchange);
  if (exchange == WG_EXCHANGE_FUSED_P2P && world > (uint32_t)PassArgs<Gl64>::MAX_PEERS)
    return fail(WG_ERR_UNSUPPORTED, "fused exchange supports at most %d ranks", PassArgs<Gl64>::MAX_PEERS);
  shard_plan(total_cols, world, exchange, chunks, exchange_out, chunks_out);
  return WG_OK;
}
int wg_shard_node_owner(uint64_t n_leaves, uint32_t world, uint64_t index, int* rank_out, uint64_t* local_index_out) {
  if (!rank_out || !local_index_out || world == 0 || !is_pow2(world) || !is_pow2(n_leaves) || index >= n_leaves || n_leaves < 2 * (uint64_t)world)
    return fail(WG_ERR_ARG, "bad node query");
  shard_node_owner(n_leaves, world, index, rank_out, local_index_out);
  return WG_OK;
}

int wg_sharded_commit_dev(int field, int hash, uint64_t n_rows, uint64_t blowup, uint32_t total_cols, const void* d_local_cols,
                          void* d_local_polys, int exchange, uint32_t chunks, uint8_t* root_out, void** handle, void* stream) {
  int ex;
  uint32_t k;
  WG_TRY(wg_shard_plan(total_cols, (uint32_t)g_dist.world, exchange, chunks, &ex, &k));
  return sharded_commit_create(field, hash, n_rows, blowup, total_cols, d_local_cols, d_local_polys, exchange, chunks, root_out, handle,
                               (cudaStream_t)stream);
}

int wg_sharded_commit(int field, int hash, uint64_t n_rows, uint64_t blowup, uint32_t total_cols, const void* const* local_column_ptrs,
                      void* const* local_polys_out, int exchange, uint32_t chunks, uint8_t* root_out, void** handle) {
  if (!local_column_ptrs || !handle) return fail(WG_ERR_ARG, "null pointer");
  *handle = nullptr;
  WG_TRY(ensure_device());
  int ex;
  uint32_t k;
  WG_TRY(wg_shard_plan(total_cols, (uint32_t)g_dist.world, exchange, chunks, &ex, &k));
  WG_TRY(shard_check_shape(field, hash, n_rows, blowup, total_cols, (uint32_t)g_dist.world));
  uint32_t s, c;
  shard_columns(total_cols, (uint32_t)g_dist.world, (uint32_t)g_dist.rank, &s, &c);
  const size_t colb = n_rows * (field == WG_FIELD_F64 ? 8 : 16);
  void *d_cols = nullptr, *d_polys = nullptr;
  WG_TRY(cached_malloc(&d_cols, colb * c));
  int rc = cached_malloc(&d_polys, colb * c);
  if (rc == WG_OK) {
    LegacyStreamDrain drain;
    for (uint32_t i = 0; i < c && rc == WG_OK; i++)
      if (cudaMemcpyAsync((char*)d_cols + i * colb, local_column_ptrs[i], colb, cudaMemcpyHostToDevice, 0) != cudaSuccess)
        rc = fail(WG_ERR_CUDA, "H2D of column %u", i);
    if (rc == WG_OK) rc = sharded_commit_create(field, hash, n_rows, blowup, total_cols, d_cols, d_polys, exchange, chunks, root_out, handle, 0);
    for (uint32_t i = 0; local_polys_out && i < c && rc == WG_OK; i++)
      if (cudaMemcpyAsync(local_polys_out[i], (char*)d_polys + i * colb, colb, cudaMemcpyDeviceToHost, 0) != cudaSuccess)
        rc = fail(WG_ERR_CUDA, "D2H of polynomial %u", i);
  }
  cached_free(d_cols, colb * c);
  cached_free(d_polys, colb * c);
  return rc;
}

int wg_sharded_info(const void* handle, uint64_t* lde_rows, uint64_t* row_start, uint64_t* row_count, uint32_t* n_blocks,
                    uint32_t* block_cols, uint32_t* block_row_widths, int* exchange, uint32_t* chunks) {
  const ShardedCommit* h = (const ShardedCommit*)handle;
  if (!h) return fail(WG_ERR_ARG, "null handle");
  if (lde_rows) *lde_rows = h->L;
  if (row_start) *row_start = h->rows * h->rank;
  if (row_count) *row_count = h->rows;
  if (n_blocks) *n_blocks = (uint32_t)h->blocks.size();
  for (size_t i = 0; i < h->blocks.size(); i++) {
    if (block_cols) block_cols[i] = h->blocks[i].cols;
    if (block_row_widths) block_row_widths[i] = (uint32_t)h->blocks[i].rw;
  }
  if (exchange) *exchange = h->exchange;
  if (chunks) *chunks = h->K;
  return WG_OK;
}

int wg_sharded_phase_ms(const void* handle, float* ms_out, uint64_t* bytes_sent) {
  const ShardedCommit* h = (const ShardedCommit*)handle;
  if (!h) return fail(WG_ERR_ARG, "null handle");
  for (int i = 0; ms_out && i < 6; i++) ms_out[i] = h->phase_ms[i];
  if (bytes_sent) *bytes_sent = h->bytes_sent;
  return WG_OK;
}

int wg_sharded_rows(const void* handle, uint32_t n_pos, const uint64_t* positions, void* rows_out) {
  const ShardedCommit* h = (const ShardedCommit*)handle;
  if (!h || !positions || !rows_out) return fail(WG_ERR_ARG, "null pointer");
  if (n_pos == 0) return WG_OK;
  const uint64_t r0 = h->rows * h->rank;
  for (uint32_t i = 0; i < n_pos; i++)
    if (positions[i] < r0 || positions[i] >= r0 + h->rows)
      return fail(WG_ERR_ARG, "position %llu is not in this rank's rows [%llu, %llu)", (unsigned long long)positions[i], (unsigned long long)r0,
                  (unsigned long long)(r0 + h->rows));
  WG_TRY(ensure_device());
  const uint32_t w = h->field == WG_FIELD_F64 ? 1 : 2;
  std::vector<const uint64_t*> ptrs;
  std::vector<uint32_t> rww, uoff;
  uint32_t off = 0;
  for (const ShardBlock& b : h->blocks) {
    if (!b.cols) continue;
    ptrs.push_back((const uint64_t*)h->recv + b.off * w);
    rww.push_back((uint32_t)b.rw * w);
    uoff.push_back(off);
    off += b.cols * w;
  }
  uoff.push_back(off);
  const uint32_t nb = (uint32_t)ptrs.size();
  DevBuf dp, dr, du, dpos, dout;
  WG_TRY(dp.alloc(nb * 8));
  WG_TRY(dr.alloc(nb * 4));
  WG_TRY(du.alloc((nb + 1) * 4));
  WG_TRY(dpos.alloc(n_pos * 8));
  WG_TRY(dout.alloc((size_t)n_pos * off * 8));
  LegacyStreamDrain drain;
  CUDA_TRY(cudaMemcpyAsync(dp.p, ptrs.data(), nb * 8, cudaMemcpyHostToDevice, 0));
  CUDA_TRY(cudaMemcpyAsync(dr.p, rww.data(), nb * 4, cudaMemcpyHostToDevice, 0));
  CUDA_TRY(cudaMemcpyAsync(du.p, uoff.data(), (nb + 1) * 4, cudaMemcpyHostToDevice, 0));
  CUDA_TRY(cudaMemcpyAsync(dpos.p, positions, n_pos * 8, cudaMemcpyHostToDevice, 0));
  shard_gather_rows_kernel<<<cdiv((uint64_t)n_pos * off, 256), 256, 0, 0>>>(dp.as<const uint64_t*>(), dr.as<uint32_t>(), du.as<uint32_t>(), nb, off,
                                                                          dpos.as<uint64_t>(), n_pos, r0, dout.as<uint64_t>());
  CUDA_TRY(cudaGetLastError());
  g_launches++;
  CUDA_TRY(cudaMemcpyAsync(rows_out, dout.p, (size_t)n_pos * off * 8, cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return WG_OK;
}

int wg_sharded_digests(const void* handle, int which, uint32_t n, const uint64_t* indexes, uint8_t* out) {
  const ShardedCommit* h = (const ShardedCommit*)handle;
  if (!h || !indexes || !out) return fail(WG_ERR_ARG, "null pointer");
  if (which < 0 || which > 2) return fail(WG_ERR_ARG, "which must be 0 (leaves), 1 (subtree nodes) or 2 (top nodes)");
  if (n == 0) return WG_OK;
  const uint64_t lim = which == 2 ? (h->world < 2 ? 2 : h->world) : h->rows;
  for (uint32_t i = 0; i < n; i++)
    if (indexes[i] >= lim) return fail(WG_ERR_ARG, "index %llu out of range (%llu digests)", (unsigned long long)indexes[i], (unsigned long long)lim);
  if (which == 2) {
    for (uint32_t i = 0; i < n; i++) memcpy(out + (size_t)i * 32, h->top_host.data() + indexes[i] * 32, 32);
    return WG_OK;
  }
  WG_TRY(ensure_device());
  DevBuf didx, dout;
  WG_TRY(didx.alloc(n * 8));
  WG_TRY(dout.alloc((size_t)n * 32));
  LegacyStreamDrain drain;
  CUDA_TRY(cudaMemcpyAsync(didx.p, indexes, n * 8, cudaMemcpyHostToDevice, 0));
  gather_digests_kernel<<<cdiv(4 * (uint64_t)n, 256), 256, 0, 0>>>((const uint2*)(which ? h->nodes : h->leaves), didx.as<uint64_t>(), n, 4u, dout.as<uint2>());
  CUDA_TRY(cudaGetLastError());
  g_launches++;
  CUDA_TRY(cudaMemcpyAsync(out, dout.p, (size_t)n * 32, cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return WG_OK;
}

int wg_sharded_gather(const void* handle, uint8_t* leaves_out, uint8_t* nodes_out) {
  const ShardedCommit* h = (const ShardedCommit*)handle;
  if (!h) return fail(WG_ERR_ARG, "null handle");
  WG_TRY(ensure_device());
  const uint32_t G = h->world;
  const size_t blk = (size_t)h->rows * 32;
  if (G == 1) {
    if (leaves_out) CUDA_TRY(cudaMemcpy(leaves_out, h->leaves, blk, cudaMemcpyDeviceToHost));
    if (nodes_out) CUDA_TRY(cudaMemcpy(nodes_out, h->nodes, blk, cudaMemcpyDeviceToHost));
    return WG_OK;
  }
  std::lock_guard<std::mutex> lk(g_dist_mu);
  void* all = nullptr;
  WG_TRY(cached_malloc(&all, blk * G));
  int rc = WG_OK;
  cudaStream_t sx = g_dist.s_comm;
  auto gather = [&](const uint8_t* src, uint8_t* host_dst) -> int {
    NCCL_TRY(g_nccl.AllGather(src, all, blk, ncclUint8, g_dist.comm, sx));
    CUDA_TRY(cudaMemcpyAsync(host_dst, all, blk * G, cudaMemcpyDeviceToHost, sx));
    CUDA_TRY(cudaStreamSynchronize(sx));
    return WG_OK;
  };
  if (leaves_out) rc = gather(h->leaves, leaves_out);
  if (rc == WG_OK && nodes_out) {
    std::vector<uint8_t> locals(blk * G);
    rc = gather(h->nodes, locals.data());
    if (rc == WG_OK) stitch_nodes_host(locals.data(), h->top_host.data(), G, h->rows, nodes_out);
  }
  cudaStreamSynchronize(sx);
  cached_free(all, blk * G);
  return rc;
}

int wg_sharded_download(const void* handle, void* const* blocks_out, uint8_t* leaves_out, uint8_t* nodes_out, uint8_t* top_out) {
  const ShardedCommit* h = (const ShardedCommit*)handle;
  if (!h) return fail(WG_ERR_ARG, "null handle");
  WG_TRY(ensure_device());
  LegacyStreamDrain drain;
  const size_t es = h->esize();
  for (size_t i = 0; blocks_out && i < h->blocks.size(); i++)
    if (blocks_out[i] && h->blocks[i].cols)
      CUDA_TRY(cudaMemcpyAsync(blocks_out[i], (const char*)h->recv + h->blocks[i].off * es, (size_t)h->rows * h->blocks[i].rw * es,
                               cudaMemcpyDeviceToHost, 0));
  if (leaves_out) CUDA_TRY(cudaMemcpyAsync(leaves_out, h->leaves, h->rows * 32, cudaMemcpyDeviceToHost, 0));
  if (nodes_out) CUDA_TRY(cudaMemcpyAsync(nodes_out, h->nodes, h->rows * 32, cudaMemcpyDeviceToHost, 0));
  if (top_out) memcpy(top_out, h->top_host.data(), (size_t)(h->world < 2 ? 2 : h->world) * 32);
  CUDA_TRY(cudaStreamSynchronize(0));
  return WG_OK;
}

int wg_sharded_free(void* handle) {
  std::lock_guard<std::mutex> lk(g_dist_mu);
  delete (ShardedCommit*)handle;
  return WG_OK;
}

int wg_sharded_commit_emulated(int field, int hash, uint64_t n_rows, uint64_t blowup, uint32_t total_cols, uint32_t world, const void* d_cols,
                               int exchange, uint32_t chunks, uint8_t* root_out, uint8_t* leaves_out, uint8_t* nodes_out) {
  if (!d_cols) return fail(WG_ERR_ARG, "null pointer");
  WG_TRY(ensure_device());
  if (world == 0) return fail(WG_ERR_ARG, "world must be positive");
  int ex;
  uint32_t k;
  WG_TRY(wg_shard_plan(total_cols, world, exchange, chunks, &ex, &k));
  WG_TRY(shard_check_shape(field, hash, n_rows, blowup, total_cols, world));
  std::lock_guard<std::mutex> lk(g_dist_mu);
  WG_FIELD_DISPATCH(field,
                    sharded_emulate<Gl64>(hash, n_rows, blowup, total_cols, world, (const uint64_t*)d_cols, exchange, chunks, root_out, leaves_out, nodes_out),
                    sharded_emulate<F128>(hash, n_rows, blowup, total_cols, world, (const u128*)d_cols, exchange, chunks, root_out, leaves_out, nodes_out));
}

}  // extern "C"
