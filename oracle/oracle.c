/* ORACLE — TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement (plain C + OpenMP) of the reference's trace-commitment hot path. It exists so that
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg can CHECK and
 * TIME-BESIDE the CUDA path. Nothing under starkpack_winterfell_b200/ may import, link or call it.
 *
 * Reference = /root/reference (NethermindEth/starkpack-winterfell, pure Rust). It cannot be built on
 * this image (no cargo/rustc, crates not vendored), so there is no oracle/_ref; this restatement is
 * the oracle, kind "port". Parity pinning: see tests/test_oracle_*.py — checked against Python
 * big-int O(n^2) polynomial evaluation (the reference's own test oracle, math/src/fft/tests.rs:19-60,
 * prover/src/matrix/tests.rs:14-40), the reference's fixed fixtures (fib trace, LEAVES4/LEAVES8,
 * root-of-unity constants) and independent hash implementations (Python blake3 / hashlib).
 *
 * Field ids: 0 = f64 (Goldilocks, Montgomery words), 1 = f128 (canonical u128 LE).
 * Hash ids : 0 = Blake3_256, 1 = Sha3_256.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "oracle_field.h"
#include "oracle_hash.h"

#define FE uint64_t
#define FADD f64_add
#define FSUB f64_sub
#define FMUL f64_mul
#define FEXP f64_exp
#define NAME(x) f64_##x
#include "oracle_ntt_impl.h"
#undef FEXP
#undef FE
#undef FADD
#undef FSUB
#undef FMUL
#undef NAME

#define FE u128
#define FADD f128_add
#define FSUB f128_sub
#define FMUL f128_mul
#define FEXP f128_exp
#define NAME(x) f128_##x
#include "oracle_ntt_impl.h"
#undef FEXP
#undef FE
#undef FADD
#undef FSUB
#undef FMUL
#undef NAME

enum { OR_OK = 0, OR_ERR_NOT_POW2 = 1, OR_ERR_TOO_LARGE = 2, OR_ERR_ZERO_OFFSET = 3, OR_ERR_TOO_FEW = 4, OR_ERR_ARG = 5 };

static inline int is_pow2(uint64_t n) { return n && !(n & (n - 1)); }
static inline unsigned ilog2(uint64_t n) { return 63u - (unsigned)__builtin_clzll(n); }
static inline size_t esize(int field) { return field == 0 ? 8 : 16; }

/* math/src/field/traits.rs:254-263 */
static uint64_t f64_root_of_unity(unsigned n) { return f64_exp(f64_new(F64_ROOT_CANON), 1ULL << (F64_TWO_ADICITY - n)); }
static u128 f128_root_of_unity(unsigned n) { return f128_exp(f128_root(), (u128)1 << (F128_TWO_ADICITY - n)); }

int or_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
/* torchrun exports OMP_NUM_THREADS=1; the timed CPU baseline sets its thread count explicitly */
void or_set_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

/* ---- scalar helpers exported for the field tests */
uint64_t or_f64_new(uint64_t v) { return f64_new(v); }
uint64_t or_f64_as_int(uint64_t v) { return f64_as_int(v); }
uint64_t or_f64_add(uint64_t a, uint64_t b) { return f64_add(a, b); }
uint64_t or_f64_sub(uint64_t a, uint64_t b) { return f64_sub(a, b); }
uint64_t or_f64_mul(uint64_t a, uint64_t b) { return f64_mul(a, b); }
uint64_t or_f64_inv(uint64_t a) { return f64_inv(a); }
uint64_t or_f64_exp(uint64_t a, uint64_t e) { return f64_exp(a, e); }
void or_f128_op(int op, const void* a, const void* b, void* out) {
  u128 x, y, r;
  memcpy(&x, a, 16);
  memcpy(&y, b, 16);
  switch (op) {
    case 0: r = f128_add(x, y); break;
    case 1: r = f128_sub(x, y); break;
    case 2: r = f128_mul(x, y); break;
    case 3: r = f128_inv(x); break;
    default: r = f128_exp(x, y); break;
  }
  memcpy(out, &r, 16);
}
void or_root_of_unity(int field, unsigned log_n, void* out) {
  if (field == 0) {
    uint64_t r = f64_root_of_unity(log_n);
    memcpy(out, &r, 8);
  } else {
    u128 r = f128_root_of_unity(log_n);
    memcpy(out, &r, 16);
  }
}
void or_hash_bytes(int hash, const uint8_t* in, uint64_t len, uint8_t* out) { oracle_hash(hash, in, len, out); }

/* ---- math/src/fft/mod.rs:466-482 (get_twiddles), :505-522 (get_inv_twiddles) */
int or_get_twiddles(int field, uint64_t n, int inverse, void* out) {
  if (!is_pow2(n)) return OR_ERR_NOT_POW2;
  if (ilog2(n) > (field == 0 ? F64_TWO_ADICITY : F128_TWO_ADICITY)) return OR_ERR_TOO_LARGE;
  if (n < 2) return OR_OK;
  if (field == 0) {
    uint64_t root = f64_root_of_unity(ilog2(n));
    if (inverse) root = f64_exp(root, n - 1);
    f64_power_series(root, n / 2, f64_new(1), (uint64_t*)out);
    f64_permute((uint64_t*)out, n / 2, 1);
  } else {
    u128 root = f128_root_of_unity(ilog2(n));
    if (inverse) root = f128_exp(root, n - 1);
    f128_power_series(root, n / 2, 1, (u128*)out);
    f128_permute((u128*)out, n / 2, 1);
  }
  return OR_OK;
}

/* ---- math/src/fft/serial.rs:64-73: fft_in_place(inv_tw); shift_by(1/n); permute */
int or_interpolate_poly(int field, void* evals, uint64_t n, const void* inv_tw) {
  if (!is_pow2(n)) return OR_ERR_NOT_POW2;
  if (ilog2(n) > (field == 0 ? F64_TWO_ADICITY : F128_TWO_ADICITY)) return OR_ERR_TOO_LARGE;
  if (field == 0) {
    uint64_t* v = (uint64_t*)evals;
    uint64_t inv_len = f64_inv(f64_new(n));
    f64_fft_in_place(v, n, 1, (const uint64_t*)inv_tw);
    for (uint64_t i = 0; i < n; i++) v[i] = f64_mul(v[i], inv_len);
    f64_permute(v, n, 1);
  } else {
    u128* v = (u128*)evals;
    u128 inv_len = f128_inv((u128)n);
    f128_fft_in_place(v, n, 1, (const u128*)inv_tw);
    for (uint64_t i = 0; i < n; i++) v[i] = f128_mul(v[i], inv_len);
    f128_permute(v, n, 1);
  }
  return OR_OK;
}

/* ---- math/src/fft/serial.rs:26-57 */
int or_evaluate_poly_with_offset(int field, const void* p, uint64_t n, const void* tw, const void* offset,
                                 uint64_t blowup, void* out) {
  if (!is_pow2(n) || !is_pow2(blowup)) return OR_ERR_NOT_POW2;
  uint64_t dom = n * blowup;
  if (ilog2(dom) > (field == 0 ? F64_TWO_ADICITY : F128_TWO_ADICITY)) return OR_ERR_TOO_LARGE;
  if (field == 0) {
    uint64_t off;
    memcpy(&off, offset, 8);
    if (off == 0) return OR_ERR_ZERO_OFFSET;
    uint64_t g = f64_root_of_unity(ilog2(dom));
    uint64_t* r = (uint64_t*)out;
    const uint64_t* c = (const uint64_t*)p;
#pragma omp parallel for schedule(dynamic)
    for (uint64_t i = 0; i < blowup; i++) {
      uint64_t idx = f64_permute_index(blowup, i);
      uint64_t o = f64_mul(f64_exp(g, idx), off), factor = f64_new(1);
      uint64_t* chunk = r + i * n;
      for (uint64_t k = 0; k < n; k++) {
        chunk[k] = f64_mul(c[k], factor);
        factor = f64_mul(factor, o);
      }
      f64_fft_in_place(chunk, n, 1, (const uint64_t*)tw);
    }
    f64_permute(r, dom, 1);
  } else {
    u128 off;
    memcpy(&off, offset, 16);
    if (off == 0) return OR_ERR_ZERO_OFFSET;
    u128 g = f128_root_of_unity(ilog2(dom));
    u128* r = (u128*)out;
    const u128* c = (const u128*)p;
#pragma omp parallel for schedule(dynamic)
    for (uint64_t i = 0; i < blowup; i++) {
      uint64_t idx = f128_permute_index(blowup, i);
      u128 o = f128_mul(f128_exp(g, idx), off), factor = 1;
      u128* chunk = r + i * n;
      for (uint64_t k = 0; k < n; k++) {
        chunk[k] = f128_mul(c[k], factor);
        factor = f128_mul(factor, o);
      }
      f128_fft_in_place(chunk, n, 1, (const u128*)tw);
    }
    f128_permute(r, dom, 1);
  }
  return OR_OK;
}

/* ---- prover/src/matrix/col_matrix.rs:196-206: per column clone + interpolate_poly (rayon over columns) */
int or_interpolate_columns(int field, const void* const* cols, uint32_t ncols, uint64_t n, void* const* polys_out) {
  if (!is_pow2(n)) return OR_ERR_NOT_POW2;
  void* itw = malloc((n / 2 + 1) * esize(field));
  int rc = or_get_twiddles(field, n, 1, itw);
  if (rc) {
    free(itw);
    return rc;
  }
#pragma omp parallel for schedule(dynamic)
  for (uint32_t c = 0; c < ncols; c++) {
    if (polys_out[c] != cols[c]) memcpy(polys_out[c], cols[c], n * esize(field));
    or_interpolate_poly(field, polys_out[c], n, itw);
  }
  free(itw);
  return OR_OK;
}

/* ---- prover/src/matrix/row_matrix.rs:248-287 get_evaluation_offsets (f64 shown; f128 twin below) */
static void f64_eval_offsets(uint64_t n, uint64_t blowup, uint64_t domain_offset, uint64_t* offsets) {
  uint64_t g = f64_root_of_unity(ilog2(n * blowup));
#pragma omp parallel for schedule(dynamic)
  for (uint64_t i = 0; i < blowup; i++) {
    uint64_t idx = f64_permute_index(blowup, i);
    uint64_t o = f64_mul(f64_exp(g, idx), domain_offset), factor = f64_new(1);
    for (uint64_t k = 0; k < n; k++) {
      offsets[i * n + k] = factor;
      factor = f64_mul(factor, o);
    }
  }
}
static void f128_eval_offsets(uint64_t n, uint64_t blowup, u128 domain_offset, u128* offsets) {
  u128 g = f128_root_of_unity(ilog2(n * blowup));
#pragma omp parallel for schedule(dynamic)
  for (uint64_t i = 0; i < blowup; i++) {
    uint64_t idx = f128_permute_index(blowup, i);
    u128 o = f128_mul(f128_exp(g, idx), domain_offset), factor = 1;
    for (uint64_t k = 0; k < n; k++) {
      offsets[i * n + k] = factor;
      factor = f128_mul(factor, o);
    }
  }
}

/* ---- RowMatrix::evaluate_polys_over::<8> (prover/src/matrix/row_matrix.rs:82-133):
 * per 8-column Segment (segments.rs:53-155): for each coset chunk copy_polys (coeff * offsets,
 * segments.rs:175-207) + fft_in_place over [B;8] rows; permute the whole segment; then transpose the
 * segments into one row-major matrix (row_matrix.rs:314-359). out: (n*blowup) x row_width elements,
 * row_width = 8*ceil(ncols/8), pad columns zero. twiddles = get_twiddles(n). */
#define SEGW 8
int or_evaluate_polys_over(int field, const void* const* polys, uint32_t ncols, uint64_t n, uint64_t blowup,
                           const void* domain_offset, const void* twiddles, void* out) {
  if (!is_pow2(n) || !is_pow2(blowup)) return OR_ERR_NOT_POW2;
  uint64_t L = n * blowup;
  if (ilog2(L) > (field == 0 ? F64_TWO_ADICITY : F128_TWO_ADICITY)) return OR_ERR_TOO_LARGE;
  uint32_t nseg = (ncols + SEGW - 1) / SEGW;
  uint64_t row_width = (uint64_t)nseg * SEGW;
  size_t es = esize(field);
  void* offsets = malloc(L * es);
  void* seg = malloc(L * SEGW * es);
  if (!offsets || !seg) return OR_ERR_ARG;
  if (field == 0) {
    uint64_t off;
    memcpy(&off, domain_offset, 8);
    if (off == 0) return OR_ERR_ZERO_OFFSET;
    f64_eval_offsets(n, blowup, off, (uint64_t*)offsets);
  } else {
    u128 off;
    memcpy(&off, domain_offset, 16);
    if (off == 0) return OR_ERR_ZERO_OFFSET;
    f128_eval_offsets(n, blowup, off, (u128*)offsets);
  }
  /* Segments one after the other (row_matrix.rs:304-306 is a serial `map`); inside Segment::new_with_buffer the B coset
   * chunks run in parallel, each copy_polys + concurrent::split_radix_fft whose inner / outer FFTs are parallel again
   * (segments.rs:125-139, 232-272; nested rayon -> nested OpenMP tasks), then the concurrent permute (:276-298), and
   * finally the batched transpose of row_matrix.rs:314-359 (here per segment: same bytes moved). */
  int nthreads = or_threads();
  size_t batches = 1;
  while ((int)batches < nthreads) batches *= 2;
  batches *= 2; /* rayon::current_num_threads().next_power_of_two() * 2 */
#pragma omp parallel
#pragma omp single
  for (uint32_t s = 0; s < nseg; s++) {
    uint32_t c0 = s * SEGW, np = ncols - c0 < SEGW ? ncols - c0 : SEGW;
    if (np < SEGW) memset(seg, 0, L * SEGW * es);
#pragma omp taskloop default(shared) grainsize(1)
    for (uint64_t i = 0; i < blowup; i++) {
      if (field == 0) {
        uint64_t* d = (uint64_t*)seg + i * n * SEGW;
        const uint64_t* o = (const uint64_t*)offsets + i * n;
        for (uint64_t r = 0; r < n; r++)
          for (uint32_t c = 0; c < np; c++) d[r * SEGW + c] = f64_mul(((const uint64_t*)polys[c0 + c])[r], o[r]);
        f64_split_radix_fft(d, n, SEGW, (const uint64_t*)twiddles);
      } else {
        u128* d = (u128*)seg + i * n * SEGW;
        const u128* o = (const u128*)offsets + i * n;
        for (uint64_t r = 0; r < n; r++)
          for (uint32_t c = 0; c < np; c++) d[r * SEGW + c] = f128_mul(((const u128*)polys[c0 + c])[r], o[r]);
        f128_split_radix_fft(d, n, SEGW, (const u128*)twiddles);
      }
    }
    if (field == 0) f64_permute_par((uint64_t*)seg, L, SEGW, batches);
    else f128_permute_par((u128*)seg, L, SEGW, batches);
    /* transpose step: segment s occupies columns [8s, 8s+8) of every row */
#pragma omp taskloop default(shared) grainsize(4096)
    for (uint64_t r = 0; r < L; r++)
      memcpy((uint8_t*)out + (r * row_width + c0) * es, (uint8_t*)seg + r * SEGW * es, SEGW * es);
  }
  free(seg);
  free(offsets);
  return OR_OK;
}

/* ---- hash_elements byte layout (crypto/src/hash/blake/mod.rs:46-59, sha/mod.rs:43-56):
 * f128 (IS_CANONICAL): raw LE memory; f64: each element as_int().to_le_bytes() (f64/mod.rs:605-610). */
static size_t row_bytes(int field, const void* row, uint64_t elems, uint8_t* dst) {
  if (field == 0) {
    const uint64_t* r = (const uint64_t*)row;
    for (uint64_t i = 0; i < elems; i++) {
      uint64_t v = f64_as_int(r[i]);
      memcpy(dst + 8 * i, &v, 8);
    }
    return 8 * elems;
  }
  memcpy(dst, row, 16 * elems);
  return 16 * elems;
}

/* ---- RowMatrix::commit_to_comb_rows leaves (row_matrix.rs:204-238); n_traces = 1 gives commit_to_rows
 * (row_matrix.rs:183-203). ldes[t] is row-major rows x row_widths[t]; only elems[t] are hashed. */
int or_hash_comb_rows(int field, int hash, uint32_t n_traces, const void* const* ldes, const uint64_t* row_widths,
                      const uint64_t* elems, uint64_t rows, uint8_t* leaves_out) {
  size_t es = esize(field), total = 0, dl = oracle_digest_len(hash);
  for (uint32_t t = 0; t < n_traces; t++) total += elems[t] * es;
#pragma omp parallel
  {
    uint8_t* buf = (uint8_t*)malloc(total + 16);
#pragma omp for schedule(static)
    for (uint64_t r = 0; r < rows; r++) {
      size_t pos = 0;
      for (uint32_t t = 0; t < n_traces; t++)
        pos += row_bytes(field, (const uint8_t*)ldes[t] + r * row_widths[t] * es, elems[t], buf + pos);
      oracle_hash(hash, buf, pos, leaves_out + dl * r);
    }
    free(buf);
  }
  return OR_OK;
}
int or_hash_rows(int field, int hash, const void* row_major, uint64_t rows, uint64_t row_width, uint64_t elems_per_row,
                 uint8_t* leaves_out) {
  const void* l[1] = {row_major};
  return or_hash_comb_rows(field, hash, 1, l, &row_width, &elems_per_row, rows, leaves_out);
}

/* ---- crypto/src/merkle/mod.rs:117-136 (new: validation), :350-374 (build_merkle_nodes) */
int or_merkle_nodes(int hash, const uint8_t* leaves, uint64_t n_leaves, uint8_t* nodes) {
  if (n_leaves < 2) return OR_ERR_TOO_FEW;
  if (!is_pow2(n_leaves)) return OR_ERR_NOT_POW2;
  uint64_t n = n_leaves / 2;
  size_t dl = oracle_digest_len(hash); /* merge = hash of the 2*dl bytes of the two children (blake/mod.rs:31-33,85-88) */
  memset(nodes, 0, dl);
#pragma omp parallel for schedule(static)
  for (uint64_t i = 0; i < n; i++) oracle_hash(hash, leaves + 2 * dl * i, 2 * dl, nodes + dl * (n + i));
  for (uint64_t lvl = n / 2; lvl >= 1; lvl /= 2) {
#pragma omp parallel for schedule(static) if (lvl >= 1024)
    for (uint64_t i = lvl; i < 2 * lvl; i++) oracle_hash(hash, nodes + 2 * dl * i, 2 * dl, nodes + dl * i);
  }
  return OR_OK;
}

/* ---- Prover::build_trace_commitment (prover/src/lib.rs:615-670): interpolate_columns per trace,
 * evaluate_polys_over::<8> per trace (StarkDomain: twiddles = get_twiddles(n), offset = GENERATOR,
 * prover/src/domain.rs:38-52, air/src/options.rs:199-201), commit_to_comb_rows.
 * Any of polys_out / lde_out / leaves_out / nodes_out may be NULL. */
int or_commit_trace(int field, int hash, uint64_t n, uint64_t blowup, uint32_t n_traces, const uint32_t* cols_per_trace,
                    const void* const* column_ptrs, void* const* polys_out, void* const* lde_out, uint8_t* leaves_out,
                    uint8_t* nodes_out, uint8_t root_out[32]) {
  if (!is_pow2(n) || !is_pow2(blowup)) return OR_ERR_NOT_POW2;
  size_t es = esize(field);
  uint64_t L = n * blowup;
  uint32_t total_cols = 0;
  for (uint32_t t = 0; t < n_traces; t++) total_cols += cols_per_trace[t];
  void** polys = (void**)malloc(sizeof(void*) * total_cols);
  for (uint32_t c = 0; c < total_cols; c++) polys[c] = polys_out ? polys_out[c] : malloc(n * es);
  int rc = or_interpolate_columns(field, column_ptrs, total_cols, n, polys);
  void* tw = malloc((n / 2 + 1) * es);
  if (!rc) rc = or_get_twiddles(field, n, 0, tw);
  void** ldes = (void**)calloc(n_traces, sizeof(void*));
  uint64_t* rw = (uint64_t*)malloc(8 * n_traces);
  uint64_t* el = (uint64_t*)malloc(8 * n_traces);
  uint64_t gen64 = f64_new(F64_GENERATOR_CANON);
  u128 gen128 = 3;
  uint32_t c0 = 0;
  for (uint32_t t = 0; t < n_traces && !rc; t++) {
    el[t] = cols_per_trace[t];
    rw[t] = (uint64_t)((cols_per_trace[t] + 7) / 8) * 8;
    ldes[t] = lde_out ? lde_out[t] : malloc(L * rw[t] * es);
    rc = or_evaluate_polys_over(field, (const void* const*)(polys + c0), cols_per_trace[t], n, blowup,
                                field == 0 ? (const void*)&gen64 : (const void*)&gen128, tw, ldes[t]);
    c0 += cols_per_trace[t];
  }
  uint8_t* leaves = leaves_out ? leaves_out : (uint8_t*)malloc(32 * L); /* 32 >= any digest length */
  uint8_t* nodes = nodes_out ? nodes_out : (uint8_t*)malloc(32 * L);
  if (!rc) rc = or_hash_comb_rows(field, hash, n_traces, (const void* const*)ldes, rw, el, L, leaves);
  if (!rc) rc = or_merkle_nodes(hash, leaves, L, nodes);
  if (!rc) memcpy(root_out, nodes + oracle_digest_len(hash), oracle_digest_len(hash));
  if (!leaves_out) free(leaves);
  if (!nodes_out) free(nodes);
  for (uint32_t t = 0; t < n_traces; t++)
    if (!lde_out && ldes[t]) free(ldes[t]);
  if (!polys_out)
    for (uint32_t c = 0; c < total_cols; c++) free(polys[c]);
  free(polys);
  free(ldes);
  free(rw);
  free(el);
  free(tw);
  return rc;
}

/* ================================================================================================ FRI layers
 * (SURVEY 8(f) rank 3). Base field (E = B). */

/* ---- FriProver::build_layer, first half (fri/src/prover/mod.rs:185-199):
 *   transpose_slice::<E, N> (utils/core/src/lib.rs:206-225): result[i][j] = source[i + j * row_count]
 *   hash_values::<H, E, N> (fri/src/utils.rs): leaf_i = H::hash_elements(&values[i])
 *   MerkleTree::new(leaves) */
int or_fri_commit_layer(int field, int hash, const void* evals, uint64_t n, uint32_t ff, void* transposed_out,
                        uint8_t* leaves_out, uint8_t* nodes_out) {
  if (ff != 2 && ff != 4 && ff != 8 && ff != 16) return OR_ERR_ARG;
  if (n == 0 || n % ff) return OR_ERR_ARG;
  uint64_t rows = n / ff;
  size_t es = esize(field);
  uint8_t* tr = (uint8_t*)transposed_out;
  uint8_t* tmp = NULL;
  if (!tr) tr = tmp = (uint8_t*)malloc(n * es);
  const uint8_t* src = (const uint8_t*)evals;
  for (uint64_t i = 0; i < rows; i++)
    for (uint32_t j = 0; j < ff; j++) memcpy(tr + (i * ff + j) * es, src + (i + (uint64_t)j * rows) * es, es);
  uint8_t* leaves = leaves_out ? leaves_out : (uint8_t*)malloc(rows * 32);
  uint64_t rw = ff;
  int rc = or_hash_rows(field, hash, tr, rows, rw, rw, leaves);
  if (rc == OR_OK) rc = or_merkle_nodes(hash, leaves, rows, nodes_out);
  if (!leaves_out) free(leaves);
  free(tmp);
  return rc;
}

/* ---- folding::apply_drp (fri/src/folding/mod.rs:115-146) with get_inv_offsets (:186-193):
 *   inv_offsets = get_power_series_with_offset(g^-1, domain_offset^-1, rows), g = root of unity of order rows * N
 *   per row: poly = row; serial_fft(poly, get_inv_twiddles(N)) (= fft_in_place + permute, math/src/fft/mod.rs:416-440);
 *            coeff_k *= (1/N) * inv_offset_i^k; result_i = polynom::eval(poly, alpha) (Horner from the top coefficient) */
int or_fri_apply_drp(int field, const void* transposed, uint64_t rows, uint32_t ff, const void* domain_offset,
                     const void* alpha, void* out) {
  if (ff != 2 && ff != 4 && ff != 8 && ff != 16) return OR_ERR_ARG;
  if (rows == 0) return OR_OK;
  if (!is_pow2(rows)) return OR_ERR_NOT_POW2;
  unsigned log_n = ilog2(rows * ff);
  if (log_n > (unsigned)(field == 0 ? F64_TWO_ADICITY : F128_TWO_ADICITY)) return OR_ERR_TOO_LARGE;
  if (field == 0) {
    uint64_t inv_tw[8];
    or_get_twiddles(0, ff, 1, inv_tw);
    uint64_t g_inv = f64_inv(f64_root_of_unity(log_n));
    uint64_t off_inv = f64_inv(*(const uint64_t*)domain_offset), a = *(const uint64_t*)alpha;
    uint64_t len_offset = f64_inv(f64_new(ff));
    const uint64_t* v = (const uint64_t*)transposed;
    uint64_t* res = (uint64_t*)out;
    uint64_t s = off_inv; /* inv_offsets[i] */
    for (uint64_t i = 0; i < rows; i++) {
      uint64_t poly[16];
      memcpy(poly, v + i * ff, ff * 8);
      f64_fft_in_place(poly, ff, 1, inv_tw);
      f64_permute(poly, ff, 1);
      uint64_t offset = len_offset;
      for (uint32_t k = 0; k < ff; k++) {
        poly[k] = f64_mul(poly[k], offset);
        offset = f64_mul(offset, s);
      }
      uint64_t acc = 0;
      for (int k = (int)ff - 1; k >= 0; k--) acc = f64_add(f64_mul(acc, a), poly[k]);
      res[i] = acc;
      s = f64_mul(s, g_inv);
    }
  } else {
    u128 inv_tw[8];
    or_get_twiddles(1, ff, 1, inv_tw);
    u128 g_inv = f128_inv(f128_root_of_unity(log_n));
    u128 off, a;
    memcpy(&off, domain_offset, 16);
    memcpy(&a, alpha, 16);
    u128 off_inv = f128_inv(off);
    u128 len_offset = f128_inv((u128)ff);
    const u128* v = (const u128*)transposed;
    u128* res = (u128*)out;
    u128 s = off_inv;
    for (uint64_t i = 0; i < rows; i++) {
      u128 poly[16];
      memcpy(poly, v + i * ff, ff * 16);
      f128_fft_in_place(poly, ff, 1, inv_tw);
      f128_permute(poly, ff, 1);
      u128 offset = len_offset;
      for (uint32_t k = 0; k < ff; k++) {
        poly[k] = f128_mul(poly[k], offset);
        offset = f128_mul(offset, s);
      }
      u128 acc = 0;
      for (int k = (int)ff - 1; k >= 0; k--) acc = f128_add(f128_mul(acc, a), poly[k]);
      res[i] = acc;
      s = f128_mul(s, g_inv);
    }
  }
  return OR_OK;
}


/* ================================================================================================ extension fields
 * (SURVEY 8(f) rank 4). QuadExtension<B> / CubeExtension<B> elements are 2 / 3 consecutive base elements
 * (math/src/field/extensions/quadratic.rs:28, cubic.rs:28) and the FFT functions multiply them by BASE-field twiddles and
 * offsets only (`FftInputs for [E]`, math/src/fft/fft_inputs.rs:101-144: `mul_base`), which acts on every coordinate
 * separately. The generic NTT above already works on rows of w elements, so E = degree-w extension is the w-wide case. */
int or_interpolate_poly_ext(int field, uint32_t deg, void* evals, uint64_t n, const void* inv_tw, const void* offset) {
  if (!is_pow2(n)) return OR_ERR_NOT_POW2;
  if (ilog2(n) > (field == 0 ? F64_TWO_ADICITY : F128_TWO_ADICITY)) return OR_ERR_TOO_LARGE;
  if (field == 0) {
    uint64_t off_inv = f64_new(1);
    if (offset) {
      if (*(const uint64_t*)offset == 0) return OR_ERR_ZERO_OFFSET;
      off_inv = f64_inv(*(const uint64_t*)offset);
    }
    f64_interpolate_w((uint64_t*)evals, n, deg, (const uint64_t*)inv_tw, f64_inv(f64_new(n)), off_inv);
  } else {
    u128 off_inv = 1;
    if (offset) {
      u128 o;
      memcpy(&o, offset, 16);
      if (o == 0) return OR_ERR_ZERO_OFFSET;
      off_inv = f128_inv(o);
    }
    f128_interpolate_w((u128*)evals, n, deg, (const u128*)inv_tw, f128_inv((u128)n), off_inv);
  }
  return OR_OK;
}

int or_evaluate_poly_with_offset_ext(int field, uint32_t deg, const void* p, uint64_t n, const void* tw, const void* offset,
                                     uint64_t blowup, void* out) {
  if (!is_pow2(n) || !is_pow2(blowup)) return OR_ERR_NOT_POW2;
  uint64_t dom = n * blowup;
  if (ilog2(dom) > (field == 0 ? F64_TWO_ADICITY : F128_TWO_ADICITY)) return OR_ERR_TOO_LARGE;
  if (field == 0) {
    uint64_t off = *(const uint64_t*)offset;
    if (off == 0) return OR_ERR_ZERO_OFFSET;
    f64_evaluate_offset_w((const uint64_t*)p, n, deg, (const uint64_t*)tw, f64_root_of_unity(ilog2(dom)), off, f64_new(1), blowup,
                          (uint64_t*)out);
  } else {
    u128 off;
    memcpy(&off, offset, 16);
    if (off == 0) return OR_ERR_ZERO_OFFSET;
    f128_evaluate_offset_w((const u128*)p, n, deg, (const u128*)tw, f128_root_of_unity(ilog2(dom)), off, 1, blowup, (u128*)out);
  }
  return OR_OK;
}
