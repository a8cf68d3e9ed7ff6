/* ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.c). Included once per field with
 *   FE    element type            FADD/FSUB/FMUL  field ops on the INTERNAL representation
 *   NAME(x) symbol prefixing
 * "values" is a slice of rows, each row = w consecutive field elements (w = 1: a plain [E];
 * w = 8: the reference's [[B; 8]] segment rows, math/src/fft/fft_inputs.rs:151-207).
 */

/* math/src/fft/mod.rs:592-600 */
static inline size_t NAME(permute_index)(size_t size, size_t index) {
  unsigned bits = (unsigned)__builtin_ctzll((unsigned long long)size);
  if (bits == 0) return 0;
  size_t r = 0;
  for (unsigned i = 0; i < bits; i++) r |= ((index >> i) & 1) << (bits - 1 - i);
  return r;
}

/* math/src/fft/fft_inputs.rs:56-64 (permute): swap i <-> bitrev(i) when j > i */
static void NAME(permute)(FE* v, size_t n, size_t w) {
  for (size_t i = 0; i < n; i++) {
    size_t j = NAME(permute_index)(n, i);
    if (j > i)
      for (size_t t = 0; t < w; t++) {
        FE tmp = v[i * w + t];
        v[i * w + t] = v[j * w + t];
        v[j * w + t] = tmp;
      }
  }
}

/* math/src/fft/fft_inputs.rs:215-256: in-place recursive radix-2, natural in -> bit-reversed out,
 * twiddles bit-reversed. n = number of rows. MAX_LOOP = 256 (fft_inputs.rs:10). */
static void NAME(fft_rec)(FE* values, size_t n, size_t w, const FE* tw, size_t count, size_t stride, size_t offset) {
  size_t size = n / stride;
  if (size > 2) {
    if (stride == count && count < 256) {
      NAME(fft_rec)(values, n, w, tw, 2 * count, 2 * stride, offset);
    } else {
      NAME(fft_rec)(values, n, w, tw, count, 2 * stride, offset);
      NAME(fft_rec)(values, n, w, tw, count, 2 * stride, offset + stride);
    }
  }
  /* butterflies without twiddle (fft_inputs.rs:241-243, :106-112) */
  for (size_t o = offset; o < offset + count; o++) {
    FE* a = values + o * w;
    FE* b = values + (o + stride) * w;
    for (size_t t = 0; t < w; t++) {
      FE x = a[t], y = b[t];
      a[t] = FADD(x, y);
      b[t] = FSUB(x, y);
    }
  }
  /* butterflies with twiddles (fft_inputs.rs:246-255, :115-122) */
  size_t last = offset + size * stride;
  size_t i = 0;
  for (size_t o = offset; o < last; o += 2 * stride, i++) {
    if (i == 0) continue;
    FE twd = tw[i];
    for (size_t j = o; j < o + count; j++) {
      FE* a = values + j * w;
      FE* b = values + (j + stride) * w;
      for (size_t t = 0; t < w; t++) {
        FE x = a[t], y = FMUL(b[t], twd);
        a[t] = FADD(x, y);
        b[t] = FSUB(x, y);
      }
    }
  }
}
static inline void NAME(fft_in_place)(FE* values, size_t n, size_t w, const FE* tw) {
  if (n >= 2) NAME(fft_rec)(values, n, w, tw, 1, 1, 0);
}

/* ---- the reference's `concurrent` transform of a segment (prover/src/matrix/segments.rs:232-345; the [E] twin is
 * math/src/fft/concurrent.rs:121-210): sqrt(n) decomposition, natural in -> bit-reversed out, i.e. the same output as
 * fft_in_place. transpose (serial, as in the reference: "TODO: investigate concurrent implementation"), inner FFTs over
 * the rows in parallel (rayon par_chunks_mut -> OpenMP taskloop; must run inside a parallel region), transpose, twiddle
 * scaling + outer FFTs over the rows in parallel. */
static void NAME(transpose_square_stretch)(FE* d, size_t size, size_t stretch, size_t w) {
  /* segments.rs:308-345: blocks of `stretch` consecutive rows-of-w are the matrix entries */
  size_t bw = stretch * w;
  for (size_t row = 0; row < size; row++)
    for (size_t col = row + 1; col < size; col++) {
      FE* a = d + (row * size + col) * bw;
      FE* b = d + (col * size + row) * bw;
      for (size_t t = 0; t < bw; t++) {
        FE tmp = a[t];
        a[t] = b[t];
        b[t] = tmp;
      }
    }
}
static void NAME(split_radix_fft)(FE* data, size_t n, size_t w, const FE* tw) {
  unsigned log_n = (unsigned)__builtin_ctzll((unsigned long long)n);
  if (n < 1024) { /* MIN_CONCURRENT_SIZE (math/src/fft/mod.rs:34): the serial transform */
    NAME(fft_in_place)(data, n, w, tw);
    return;
  }
  FE g = tw[n / 4]; /* twiddles[twiddles.len() / 2] = w_n (segments.rs:236) */
  size_t inner_len = (size_t)1 << (log_n / 2), outer_len = n / inner_len, stretch = outer_len / inner_len;
  NAME(transpose_square_stretch)(data, inner_len, stretch, w);
#pragma omp taskloop default(shared) grainsize(4)
  for (size_t i = 0; i < inner_len; i++) NAME(fft_rec)(data + i * outer_len * w, outer_len, w, tw, stretch, stretch, 0);
  NAME(transpose_square_stretch)(data, inner_len, stretch, w);
#pragma omp taskloop default(shared) grainsize(4)
  for (size_t i = 0; i < inner_len; i++) {
    FE* row = data + i * outer_len * w;
    if (i > 0) {
      FE inner_tw = FEXP(g, NAME(permute_index)(inner_len, i)), outer_tw = inner_tw;
      for (size_t e = 1; e < outer_len; e++) {
        for (size_t t = 0; t < w; t++) row[e * w + t] = FMUL(row[e * w + t], outer_tw);
        outer_tw = FMUL(outer_tw, inner_tw);
      }
    }
    NAME(fft_in_place)(row, outer_len, w, tw);
  }
}
/* concurrent permute (segments.rs:276-298): batches of the index range, swap when j > i. Inside a parallel region. */
static void NAME(permute_par)(FE* v, size_t n, size_t w, size_t batches) {
  if (batches < 1 || n < 1024) batches = 1;
  size_t per = n / batches;
#pragma omp taskloop default(shared) grainsize(1)
  for (size_t b = 0; b < batches; b++) {
    size_t end = b + 1 == batches ? n : (b + 1) * per;
    for (size_t i = b * per; i < end; i++) {
      size_t j = NAME(permute_index)(n, i);
      if (j > i)
        for (size_t t = 0; t < w; t++) {
          FE tmp = v[i * w + t];
          v[i * w + t] = v[j * w + t];
          v[j * w + t] = tmp;
        }
    }
  }
}

/* math/src/utils/mod.rs:36-46: [1, b, b^2, ...] */
static void NAME(power_series)(FE b, size_t n, FE one, FE* out) {
  FE acc = one;
  for (size_t i = 0; i < n; i++) {
    out[i] = acc;
    acc = FMUL(acc, b);
  }
}

/* math/src/fft/serial.rs:64-73 (interpolate_poly) and :78-93 (interpolate_poly_with_offset) on rows of w coordinates:
 * fft_in_place(inv_twiddles); coefficient k *= inv_len * offset_inv^k (offset_inv = 1 for the plain variant); permute.
 * The reference scales before permuting (shift_by / the offset loop run on the bit-reversed vector in the plain
 * variant and after permute in the offset variant); products commute, so one order is restated. */
static void NAME(interpolate_w)(FE* v, size_t n, size_t w, const FE* inv_tw, FE inv_len, FE off_inv) {
  NAME(fft_in_place)(v, n, w, inv_tw);
  NAME(permute)(v, n, w);
  FE f = inv_len;
  for (size_t k = 0; k < n; k++) {
    for (size_t t = 0; t < w; t++) v[k * w + t] = FMUL(v[k * w + t], f);
    f = FMUL(f, off_inv);
  }
}

/* math/src/fft/serial.rs:26-57 (evaluate_poly_with_offset) on rows of w coordinates: per coset i (bit-reversed order)
 * scale coefficient k by (g^rev(i) * offset)^k, transform, then one global permute */
static void NAME(evaluate_offset_w)(const FE* p, size_t n, size_t w, const FE* tw, FE g, FE off, FE one, size_t blowup, FE* out) {
  for (size_t i = 0; i < blowup; i++) {
    size_t idx = NAME(permute_index)(blowup, i);
    FE o = FMUL(FEXP(g, idx), off), factor = one;
    FE* chunk = out + i * n * w;
    for (size_t k = 0; k < n; k++) {
      for (size_t t = 0; t < w; t++) chunk[k * w + t] = FMUL(p[k * w + t], factor);
      factor = FMUL(factor, o);
    }
    NAME(fft_in_place)(chunk, n, w, tw);
  }
  NAME(permute)(out, n * blowup, w);
}
