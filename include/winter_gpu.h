/* winter_gpu.h — C ABI of the B200 trace-commitment library (libwinter_gpu.so).
 *
 * Drop-in boundary for ONE path of NethermindEth/starkpack-winterfell (reference = pure Rust, no FFI of
 * its own): interpolate trace columns -> coset LDE -> hash LDE rows -> Merkle tree -> root. Each entry point
 * replaces the body of the reference function cited beside it; a Rust `-sys` crate binds these symbols
 * (see INTEGRATION.md). Plain pointers and sizes only; no torch / CUDA types in signatures (streams are
 * passed as void*).
 *
 * Element encodings (exactly the reference's in-memory representation):
 *   field 0 = f64 Goldilocks, one uint64 per element holding the MONTGOMERY word x*2^64 mod p
 *             (math/src/field/f64/mod.rs:48-53); field 1 = f128, 16 bytes little-endian canonical
 *             (math/src/field/f128/mod.rs:46-51).
 *   hash  0 = Blake3_256 (crypto/src/hash/blake/mod.rs:20-60), 1 = Sha3_256 (crypto/src/hash/sha/mod.rs:17-57),
 *         2 = Blake3_192 (blake/mod.rs:68-114). Digests are 32 bytes (24 for Blake3_192).
 * All functions return WG_OK (0) or a negative error code and are thread-safe. Precondition failures map to
 * the reference's assert!/MerkleTreeError conditions. There is NO CPU fallback: without a CUDA device every
 * compute entry point returns WG_ERR_CUDA.
 */
#ifndef WINTER_GPU_H
#define WINTER_GPU_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum {
  WG_OK = 0,
  WG_ERR_NOT_POW2 = -1,     /* fft/mod.rs:90-105,279-295 "must be a power of 2"; MerkleTreeError::NumberOfLeavesNotPowerOfTwo */
  WG_ERR_TOO_LARGE = -2,    /* log2(domain) > TWO_ADICITY (fft/mod.rs:291-295) (domain = n * blowup) */
  WG_ERR_ZERO_OFFSET = -3,  /* fft/mod.rs:197-200 "domain offset cannot be zero" */
  WG_ERR_TOO_FEW = -4,      /* MerkleTreeError::TooFewLeaves (crypto/src/merkle/mod.rs:118-120) */
  WG_ERR_ARG = -5,          /* null pointer / bad enum / twiddle length mismatch (fft/mod.rs:283-289) */
  WG_ERR_CUDA = -6,         /* CUDA runtime error or no device; wg_last_error() has the text */
  WG_ERR_NOMEM = -7,        /* device allocation failed / scratch too small */
  WG_ERR_UNSUPPORTED = -8
};

enum { WG_FIELD_F64 = 0, WG_FIELD_F128 = 1 };
/* WG_HASH_BLAKE3_192 (crypto/src/hash/blake/mod.rs:68-114): BLAKE3 truncated to 24 bytes; its digests, leaves and nodes are
 * 24 bytes each wherever this header says "32 bytes" / "x 32". Not supported by the sharded (multi-GPU) entry points. */
enum { WG_HASH_BLAKE3_256 = 0, WG_HASH_SHA3_256 = 1, WG_HASH_BLAKE3_192 = 2 };
/* digest length in bytes of a hash id (0 for an unknown id) */
uint32_t wg_digest_bytes(int hash);

/* Select the CUDA device for this process (one process per GPU). Idempotent. */
int wg_init(int device);
int wg_shutdown(void);
/* Text of the last error on the calling thread. */
const char* wg_last_error(void);
/* Number of kernels this library has launched since wg_init (bench.py's gpu_launches). */
uint64_t wg_launch_count(void);

/* Phase timing (CUDA events on the launching stream) of the device-resident entry points. Phases:
 * 0 interpolate strided pass, 1 interpolate contiguous pass, 2 LDE strided pass, 3 LDE contiguous pass,
 * 4 row hashing, 5 Merkle. wg_profile_read waits for the recorded work, writes summed milliseconds and the
 * number of timed launches-groups per phase into 6-entry arrays, and clears the records. */
int wg_profile_enable(int on);
int wg_profile_read(float* ms_out, uint32_t* count_out);

/* Transforms longer than 2^log_len points are computed with one more decimation step (radix 2..16) around 2^log_len-point
 * two-pass transforms (csrc/split.cuh); the default, 24, is the two-pass plan's own limit, so this is how lengths up to
 * the field's two-adicity are reached (fft/mod.rs:291-295). 13 <= log_len <= 24 moves the threshold (a 2^24-row trace
 * with log_len 22 runs 2^11-point passes instead of 2^12-point ones); 0 only reads it. Returns the previous value or a
 * negative error code. Process-wide; not meant to be changed while a call is in flight. */
int wg_set_split_log(int log_len);
/* Inner transform length the split plan aims at (default 2^22: both passes 2^11-point). A transform that must split takes
 * the radix (<= 16) that leaves inner transforms of this length. In addition the LDE (not the interpolation) of a trace
 * whose preferred radix is >= 4 -- with the defaults exactly the 2^24-row traces, both of whose passes would be 2^12-point
 * -- PREFERS the split plan when the scratch it is given holds (2*blowup + 1)*n elements for at least min(n_cols, 8)
 * columns, and uses the two-pass plan otherwise (2^24 x 32 columns: LDE 92.5 -> 82.4 ms). wg_lde_scratch_elems returns the
 * preferred size. 24 switches the preference off; 0 only reads. Returns the previous value or a negative error code. */
int wg_set_split_preferred_log(int log_len);

/* Test probe: out[i] = op(a[i], b[i]) computed by the device field code (op 0 add, 1 sub, 2 mul in the internal
 * representation, 16+i / 32+i multiply by the order-16 root to the +i / -i, 48 f64 as_int). Host buffers. */
int wg_debug_field_op(int field, int op, const void* a, const void* b, void* out, uint32_t n);

/* ---- host-buffer entry points (H2D + kernels + D2H inside) ------------------------------------------- */

/* fft::get_twiddles / get_inv_twiddles (math/src/fft/mod.rs:466-482,505-522): n/2 elements, bit-reversed,
 * internal representation. */
int wg_get_twiddles(int field, uint64_t n, int inverse, void* out);

/* fft::interpolate_poly (math/src/fft/mod.rs:274-305): in place, natural order in and out.
 * inv_twiddles may be NULL; if given only its length contract (n/2) is assumed. */
int wg_interpolate_poly(int field, void* evals, uint64_t n, const void* inv_twiddles);

/* fft::interpolate_poly_with_offset (math/src/fft/mod.rs:362-400): in place; coefficient i of the inverse transform is
 * multiplied by offset^-i, so the result interpolates evaluations taken over the coset offset*<w_n>. */
int wg_interpolate_poly_with_offset(int field, void* evals, uint64_t n, const void* inv_twiddles, const void* offset);

/* fft::evaluate_poly (math/src/fft/mod.rs:85-116) and fft::serial_fft (:416-440; fft_in_place + permute gives the same
 * natural-order result): in place, p[k] <- P(w_n^k). */
int wg_evaluate_poly(int field, void* p, uint64_t n, const void* twiddles);

/* fft::evaluate_poly_with_offset (math/src/fft/mod.rs:171-220): out has n*blowup elements,
 * out[k] = P(offset * w_{n*blowup}^k). offset points at one element in internal representation. */
int wg_evaluate_poly_with_offset(int field, const void* p, uint64_t n, const void* twiddles, const void* offset,
                                 uint64_t blowup, void* out);

/* ColMatrix::interpolate_columns (prover/src/matrix/col_matrix.rs:196-206): columns are separate allocations. */
int wg_interpolate_columns(int field, const void* const* columns, uint32_t n_cols, uint64_t n, void* const* polys_out);

/* RowMatrix::evaluate_polys_over::<8> (prover/src/matrix/row_matrix.rs:82-133): out is row-major
 * (n*blowup) x row_width, row_width = 8*ceil(n_cols/8), padding columns zero. offset = domain offset. */
int wg_evaluate_polys_over(int field, const void* const* polys, uint32_t n_cols, uint64_t n, uint64_t blowup,
                           const void* offset, void* out);

/* RowMatrix::commit_to_rows leaves (row_matrix.rs:183-203; ElementHasher::hash_elements): one digest per row
 * over the first elems_per_row elements of each row. */
int wg_hash_rows(int field, int hash, const void* row_major, uint64_t rows, uint64_t row_width, uint64_t elems_per_row,
                 uint8_t* leaves_out);

/* RowMatrix::commit_to_comb_rows leaves (prover/src/matrix/row_matrix.rs:204-238) over n_traces ALREADY-BUILT row-major LDE
 * matrices in host memory: leaf_k = hash_elements(row_k(t0)[..e0] || row_k(t1)[..e1] || ...), without the per-row Vec the
 * reference concatenates (:217-231). matrices[t] is rows x row_widths[t]; only elems_per_row[t] elements of each row are
 * hashed. Feed the result to wg_merkle_new for the tree. */
int wg_hash_comb_rows(int field, int hash, uint32_t n_traces, const void* const* matrices, const uint64_t* row_widths,
                      const uint64_t* elems_per_row, uint64_t rows, uint8_t* leaves_out);

/* MerkleTree::new -> build_merkle_nodes (crypto/src/merkle/mod.rs:117-136,350-374): nodes_out has n_leaves
 * digests in heap layout (nodes[0] = 0, root = nodes[1]); adopt with MerkleTree::from_raw_parts (:149-161). */
int wg_merkle_new(int hash, const uint8_t* leaves, uint64_t n_leaves, uint8_t* nodes_out);

/* Prover::build_trace_commitment (prover/src/lib.rs:615-670): T traces of n_rows rows, trace t has
 * cols_per_trace[t] columns; column_ptrs lists all columns trace by trace. Domain offset = GENERATOR
 * (air/src/options.rs:199-201). Outputs (each may be NULL): polys_out (same shape as column_ptrs),
 * lde_out[t] row-major (n_rows*blowup) x 8*ceil(cols_t/8), leaves_out / nodes_out n_rows*blowup digests,
 * root_out 32 bytes. */
int wg_commit_trace(int field, int hash, uint64_t n_rows, uint64_t blowup, uint32_t n_traces,
                    const uint32_t* cols_per_trace, const void* const* column_ptrs, void* const* polys_out,
                    void* const* lde_out, uint8_t* leaves_out, uint8_t* nodes_out, uint8_t* root_out);

/* ColMatrix::evaluate_columns_at (prover/src/matrix/col_matrix.rs:210-216 -> polynom::eval, Horner) for a base-field point:
 * out[c] = sum_i polys[c][i] * x^i. This is what TracePolyTable::evaluate_at / get_ood_frame call
 * (prover/src/trace/poly_table.rs:66-80) when E = B. x points at one element in internal representation. */
int wg_evaluate_columns_at(int field, const void* const* polys, uint32_t n_cols, uint64_t n, const void* x, void* out);

/* Prover::build_constraint_commitment (prover/src/lib.rs:680-715) = RowMatrix::evaluate_polys_over::<8>(composition_poly
 * columns, domain) + commit_to_rows: n_cols coefficient columns of n_coeffs elements each (an extension-field composition
 * polynomial is passed as its base-field columns, col_matrix.rs:113-119). Outputs as for wg_commit_trace, each may be NULL. */
int wg_commit_polys(int field, int hash, uint64_t n_coeffs, uint64_t blowup, uint32_t n_cols, const void* const* poly_ptrs,
                    void* lde_out, uint8_t* leaves_out, uint8_t* nodes_out, uint8_t* root_out);

/* ---- resident commitments: the LDE matrices, leaves and tree stay in HBM --------------------------------------------
 * What the prover needs after committing is the root, then the rows at the query positions with their authentication
 * paths (TraceCommitment::query -> build_segment_queries -> MerkleTree::prove_batch, prover/src/trace/commitment.rs:
 * 87-190, crypto/src/merkle/mod.rs:222-284; same for ConstraintCommitment::query, prover/src/constraints/commitment.rs).
 * A handle owns the device buffers until wg_commitment_free. */
int wg_commitment_from_trace(int field, int hash, uint64_t n_rows, uint64_t blowup, uint32_t n_traces,
                             const uint32_t* cols_per_trace, const void* const* column_ptrs, void* const* polys_out,
                             uint8_t* root_out, void** handle);
int wg_commitment_from_polys(int field, int hash, uint64_t n_coeffs, uint64_t blowup, uint32_t n_cols,
                             const void* const* poly_ptrs, uint8_t* root_out, void** handle);
/* shape: LDE rows, number of traces, and per trace the real columns and the stored row width (arrays of n_traces) */
int wg_commitment_info(const void* handle, uint64_t* lde_rows, uint32_t* n_traces, uint32_t* cols_per_trace,
                       uint32_t* row_widths);
/* RowMatrix::row(pos) for each position (row_matrix.rs:160-165): rows_out is n_pos x cols_per_trace[trace] elements */
int wg_commitment_rows(const void* handle, uint32_t trace, uint32_t n_pos, const uint64_t* positions, void* rows_out);
/* digests by index: which = 0 -> tree.leaves[i], 1 -> tree.nodes[i] (heap layout); out is n x 32 bytes. These are the
 * reads MerkleTree::prove / prove_batch perform (crypto/src/merkle/mod.rs:194-284). */
int wg_commitment_digests(const void* handle, int which, uint32_t n, const uint64_t* indexes, uint8_t* out);
/* full copy to the host (lde_out[t] may be NULL per trace; leaves_out / nodes_out may be NULL) */
int wg_commitment_download(const void* handle, void* const* lde_out, uint8_t* leaves_out, uint8_t* nodes_out);
int wg_commitment_free(void* handle);

/* ---- FRI layer commitments (base field, E = B) --------------------------------------------------
 * wg_fri_commit_layer replaces the first half of FriProver::build_layer (fri/src/prover/mod.rs:185-199):
 *   transposed = transpose_slice::<E, N>(evaluations)   (utils/core/src/lib.rs:206-225: row i = evals[i + j*rows], j < N)
 *   leaves     = hash_values::<H, E, N>(&transposed)    (fri/src/utils.rs: hash_elements of each row)
 *   tree       = MerkleTree::new(leaves)                (crypto/src/merkle/mod.rs:117-136)
 * n = evaluations.len(), folding_factor N in {2,4,8,16} (prover/mod.rs:178-184, anything else: WG_ERR_UNSUPPORTED).
 * transposed_out (n elements, the FriLayer's stored evaluations) and leaves_out (n/N x 32) may be NULL;
 * nodes_out is n/N x 32 bytes in heap layout, root = nodes_out[32..64).
 * wg_fri_apply_drp replaces folding::apply_drp (fri/src/folding/mod.rs:115-146): out[i] = the degree-respecting
 * projection of row i at alpha over the coset with offset domain_offset (elements in the internal representation). */
int wg_fri_commit_layer(int field, int hash, const void* evaluations, uint64_t n, uint32_t folding_factor, void* transposed_out,
                        uint8_t* leaves_out, uint8_t* nodes_out);
int wg_fri_apply_drp(int field, const void* transposed, uint64_t rows, uint32_t folding_factor, const void* domain_offset,
                     const void* alpha, void* out);

/* ---- extension-field element layouts ------------------------------------------------------------
 * The fft:: functions are generic over E: FieldElement<BaseField = B> (math/src/fft/mod.rs:85,171,274,362) with BASE-field
 * twiddles and offsets; callers with E != B are DEEP composition (prover/src/composer/mod.rs:199), the constraint evaluation
 * table (prover/src/constraints/evaluation_table.rs:180-181,216), FRI remainders and the auxiliary trace segments
 * (prover/src/lib.rs:333-334). An E element is ext_degree (1, 2 or 3) consecutive base elements in memory
 * (math/src/field/extensions/quadratic.rs:28, cubic.rs:28); lengths n count E elements. Same conventions and error
 * codes as the base-field functions above. In wg_evaluate_polys_over_ext / wg_commit_trace_ext a column holds n E
 * elements, n_cols / cols_per_trace count E columns, and the row-major LDE has n_cols*ext_degree base elements per row
 * (row width 8*ceil(n_cols*ext_degree/8)), exactly RowMatrix<E> (prover/src/matrix/row_matrix.rs:33-41). */
int wg_interpolate_poly_ext(int field, uint32_t ext_degree, void* evals, uint64_t n, const void* inv_twiddles);
int wg_interpolate_poly_with_offset_ext(int field, uint32_t ext_degree, void* evals, uint64_t n, const void* inv_twiddles,
                                        const void* offset);
int wg_evaluate_poly_ext(int field, uint32_t ext_degree, void* p, uint64_t n, const void* twiddles);
int wg_evaluate_poly_with_offset_ext(int field, uint32_t ext_degree, const void* p, uint64_t n, const void* twiddles,
                                     const void* offset, uint64_t blowup, void* out);
int wg_interpolate_columns_ext(int field, uint32_t ext_degree, const void* const* columns, uint32_t n_cols, uint64_t n,
                               void* const* polys_out);
int wg_evaluate_polys_over_ext(int field, uint32_t ext_degree, const void* const* polys, uint32_t n_cols, uint64_t n,
                               uint64_t blowup, const void* offset, void* out);
int wg_commit_trace_ext(int field, int hash, uint32_t ext_degree, uint64_t n_rows, uint64_t blowup, uint32_t n_traces,
                        const uint32_t* cols_per_trace, const void* const* column_ptrs, void* const* polys_out,
                        void* const* lde_out, uint8_t* leaves_out, uint8_t* nodes_out, uint8_t* root_out);

/* ---- device-resident entry points (no host transfers; what bench.py's `value` times) -------------------- */
/* All pointers are device pointers on the wg_init device; `stream` is a cudaStream_t (NULL = default stream).
 * Column matrices are contiguous: element (c, r) at [c*n + r]. */

/* scratch elements (not bytes) needed by wg_interpolate_columns_dev / wg_lde_dev for these shapes */
uint64_t wg_interpolate_scratch_elems(int field, uint32_t n_cols, uint64_t n);
uint64_t wg_lde_scratch_elems(int field, uint32_t n_cols, uint64_t n, uint64_t blowup);

int wg_interpolate_columns_dev(int field, const void* d_cols, uint32_t n_cols, uint64_t n, void* d_polys, void* d_scratch,
                               uint64_t scratch_elems, void* stream);
/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it). Same result as
 * wg_interpolate_columns_dev, but the two-pass NTT's intermediate lives in a file-scope static __device__ array of
 * WG_NOALLOC_SCRATCH_BYTES (16 MiB, held by every GPU that loads the library until exit) instead of a caller buffer, and
 * the call performs NO device or host allocation: n_cols * n elements must fit in that array (2^21 f64 / 2^20 f128
 * elements, e.g. 8 columns x 2^18 rows), larger requests return WG_ERR_TOO_LARGE; the cached twiddle tables of length n
 * must already be resident (any earlier wg_interpolate_columns_dev of that length builds them), else WG_ERR_ARG. Every
 * call shares the one array, which keeps the previous call's contents (each call overwrites all it reads, nothing needs
 * clearing): calls that overlap on the GPU (different streams / threads) race. */
#define WG_NOALLOC_SCRATCH_BYTES (16u << 20)
int wg_interpolate_columns_noalloc_dev(int field, const void* d_cols, uint32_t n_cols, uint64_t n, void* d_polys, void* stream);

/* d_lde row-major (n*blowup) x row_width; columns n_cols..row_width-1 are written as zero. A smaller scratch
 * than wg_lde_scratch_elems makes the call run in column chunks (minimum: 8 columns' worth). */
int wg_lde_dev(int field, const void* d_polys, uint32_t n_cols, uint64_t n, uint64_t blowup, const void* offset_host,
               void* d_lde, uint64_t row_width, void* d_scratch, uint64_t scratch_elems, void* stream);
/* leaves over the concatenation of n_traces row-major matrices (commit_to_comb_rows, row_matrix.rs:204-238) */
int wg_hash_rows_dev(int field, int hash, uint32_t n_traces, const void* const* d_ldes, const uint64_t* row_widths,
                     const uint64_t* elems_per_row, uint64_t rows, void* d_leaves, void* stream);
int wg_merkle_dev(int hash, const void* d_leaves, uint64_t n_leaves, void* d_nodes, void* stream);
/* whole commit on device: d_cols holds all columns of all traces contiguously ([sum cols][n_rows]);
 * d_polys same shape; d_ldes[t] per trace ([n_rows*blowup][8*ceil(cols_t/8)]); scratch at least the max of the two queries
 * above. With wg_commit_scratch_elems elements of scratch several NARROW traces (StarkPack: 8 traces of 2 columns) are
 * extended together as one matrix with full column tiles and then laid out per trace (cfg4 on one GPU: 19.0 -> 15.9 ms);
 * same outputs either way. */
uint64_t wg_commit_scratch_elems(int field, uint64_t n_rows, uint64_t blowup, uint32_t n_traces, const uint32_t* cols_per_trace);
int wg_commit_trace_dev(int field, int hash, uint64_t n_rows, uint64_t blowup, uint32_t n_traces,
                        const uint32_t* cols_per_trace, const void* d_cols, void* d_polys, void* const* d_ldes,
                        void* d_leaves, void* d_nodes, void* d_scratch, uint64_t scratch_elems, void* stream);

/* FRI layer on device buffers: d_evaluations n elements -> d_transposed n elements, d_leaves / d_nodes (n/N) x 32 bytes;
 * fold: d_transposed rows x N -> d_out rows elements (the next layer's evaluations). offset and alpha are host elements. */
int wg_fri_commit_layer_dev(int field, int hash, const void* d_evaluations, uint64_t n, uint32_t folding_factor,
                            void* d_transposed, void* d_leaves, void* d_nodes, void* stream);
int wg_fri_apply_drp_dev(int field, const void* d_transposed, uint64_t rows, uint32_t folding_factor,
                         const void* domain_offset_host, const void* alpha_host, void* d_out, void* stream);

/* ---- multi-GPU LDE with the row-block exchange fused into the kernel ---------------------------------------
 * One process per GPU. Leaves need whole rows, so the column-sharded LDE has to become row blocks: GPU g hashes rows
 * [g*L/G, (g+1)*L/G) of every shard. Instead of LDE -> all-to-all, wg_lde_scatter_dev's last pass stores each row
 * directly into its owner's buffer through peer-mapped pointers (NVLink P2P stores), overlapping the transfer with the
 * butterflies tile by tile: d_row_blocks[g] is the (L/G) x row_width matrix that receives THIS rank's columns on GPU g
 * (g = own rank: ordinary local memory). n_owners is a power of two <= 8. The caller orders the hash after the scatter
 * with any cross-GPU barrier on the stream (starkpack_winterfell_b200/parallel.py uses a 4-byte NCCL all-reduce).
 * wg_peer_alloc / wg_peer_open export and map such buffers between processes (CUDA IPC; the 64-byte handle travels over
 * any channel, e.g. torch.distributed.all_gather). */
#define WG_PEER_HANDLE_BYTES 64
int wg_peer_alloc(uint64_t bytes, void** d_ptr, uint8_t* handle_out /* WG_PEER_HANDLE_BYTES */);
int wg_peer_open(const uint8_t* handle, void** d_ptr);
int wg_peer_close(void* d_ptr);
int wg_peer_free(void* d_ptr);
int wg_lde_scatter_dev(int field, const void* d_polys, uint32_t n_cols, uint64_t n, uint64_t blowup, const void* offset_host,
                       uint32_t n_owners, void* const* d_row_blocks, uint64_t row_width, void* d_scratch,
                       uint64_t scratch_elems, void* stream);

/* ---- one tree over a column-sharded trace on G GPUs, behind the ABI (SURVEY 8(e), 8(b) last row) ------------------------
 * The reference's entry is one in-process call, Prover::build_trace_commitment (prover/src/lib.rs:615-626), and a leaf
 * needs the whole row of every column of every aggregated trace (RowMatrix::commit_to_comb_rows,
 * prover/src/matrix/row_matrix.rs:204-238). Here the GLOBAL column list (trace 0 column 0 ... trace T-1 last column) is cut
 * into G contiguous blocks (wg_shard_columns); one process per GPU calls wg_sharded_commit* with ITS block. Inside the
 * library: interpolate + LDE of the local columns, the one exchange that turns column shards into row blocks (rank g
 * receives rows [g*L/G, (g+1)*L/G) of every shard), row hashing over all columns, the local subtree, an all-gather of the
 * G subtree roots and the top log2(G) levels. Every rank returns the same root.
 *
 * Communicator: rank 0 calls wg_dist_unique_id and ships the WG_DIST_ID_BYTES to the other ranks over any channel; every
 * rank then calls wg_dist_init (collective; ncclCommInitRank on the wg_init device). NCCL is loaded at run time
 * (dlopen "libnccl.so.2"); without it wg_dist_* return WG_ERR_UNSUPPORTED and the single-GPU entry points are unaffected.
 *
 * Exchange forms (identical results):
 *   WG_EXCHANGE_ALLTOALL   LDE, then one grouped ncclSend/ncclRecv all-to-all;
 *   WG_EXCHANGE_PIPELINED  the shard is LDE'd in `chunks` column chunks (multiples of 8 columns) and chunk k's
 *                          all-to-all runs on its own stream under chunk k+1's LDE;
 *   WG_EXCHANGE_FUSED_P2P  the LDE's last pass stores each row straight into its owner's receive buffer through
 *                          peer-mapped pointers (CUDA IPC, NVLink P2P stores): no send buffer, no all-to-all;
 *   WG_EXCHANGE_P2P_COPY   chunked like PIPELINED, but every rank PUSHES its slices into the owners' receive buffers
 *                          (peer-mapped with CUDA IPC) with cudaMemcpyAsync: copy engines over NVLink, no SM used by the
 *                          exchange, one 4-byte NCCL all-gather as the barrier behind the copies (single node);
 *   WG_EXCHANGE_AUTO       fused up to 8 columns per rank (G <= 8), P2P_COPY in 8-column chunks otherwise (at most 4 chunks; 8 from
 *                          2^26 LDE rows, which wg_shard_plan cannot know: wg_sharded_info reports the plan a commit used). */
#define WG_DIST_ID_BYTES 128
enum { WG_EXCHANGE_AUTO = 0, WG_EXCHANGE_ALLTOALL = 1, WG_EXCHANGE_PIPELINED = 2, WG_EXCHANGE_FUSED_P2P = 3, WG_EXCHANGE_P2P_COPY = 4 };
#define WG_SHARD_MAX_BLOCKS 64 /* received column blocks (ranks x chunks) the leaf kernel concatenates */

int wg_dist_unique_id(uint8_t* id_out /* WG_DIST_ID_BYTES */);
int wg_dist_init(int rank, int world, const uint8_t* id);
int wg_dist_shutdown(void);
int wg_dist_info(int* rank, int* world); /* world 1 / rank 0 before wg_dist_init */

/* Host-only partition logic (no GPU needed; tests/test_cabi_cpu.py and the gloo tests call these):
 * wg_shard_columns: block of rank `rank` in a list of total_cols columns (first total%world ranks get one more);
 * wg_shard_chunks : chunk `idx` of `k` of a block of `cols` columns (multiples of 8 except the last non-empty one);
 * wg_shard_plan   : resolves WG_EXCHANGE_AUTO / clamps `chunks` (ranks x chunks <= WG_SHARD_MAX_BLOCKS);
 * wg_shard_node_owner: global heap index i of an L-leaf tree over `world` subtrees -> rank -1 (top levels, index into the
 *                   G-entry top array) or the owning rank and its local heap index. */
int wg_shard_columns(uint32_t total_cols, uint32_t world, uint32_t rank, uint32_t* start, uint32_t* count);
int wg_shard_chunks(uint32_t cols, uint32_t k, uint32_t idx, uint32_t* start, uint32_t* count);
int wg_shard_plan(uint32_t total_cols, uint32_t world, int exchange, uint32_t chunks, int* exchange_out, uint32_t* chunks_out);
int wg_shard_node_owner(uint64_t n_leaves, uint32_t world, uint64_t index, int* rank_out, uint64_t* local_index_out);

/* Collective. d_local_cols: this rank's columns, contiguous [count][n_rows] on the device; d_local_polys (same shape) may
 * be NULL (the handle then owns the coefficients). `chunks` = 0 lets the library choose. root_out: 32 host bytes.
 * The handle keeps this rank's row block (all columns of rows [rank*L/G, (rank+1)*L/G)), leaves and subtree in HBM. */
int wg_sharded_commit_dev(int field, int hash, uint64_t n_rows, uint64_t blowup, uint32_t total_cols, const void* d_local_cols,
                          void* d_local_polys, int exchange, uint32_t chunks, uint8_t* root_out, void** handle, void* stream);
/* The same through host buffers: local_column_ptrs / local_polys_out list this rank's columns (separate allocations,
 * as ColMatrix holds them); polys_out may be NULL. */
int wg_sharded_commit(int field, int hash, uint64_t n_rows, uint64_t blowup, uint32_t total_cols,
                      const void* const* local_column_ptrs, void* const* local_polys_out, int exchange, uint32_t chunks,
                      uint8_t* root_out, void** handle);
/* shape of the handle: global LDE rows, first row / number of rows held here, received column blocks (global column
 * order) with their real columns and stored row widths (arrays of WG_SHARD_MAX_BLOCKS; an exchanged block is stored without
 * pad columns: width = columns), exchange form and chunks used */
int wg_sharded_info(const void* handle, uint64_t* lde_rows, uint64_t* row_start, uint64_t* row_count, uint32_t* n_blocks,
                    uint32_t* block_cols, uint32_t* block_row_widths, int* exchange, uint32_t* chunks);
/* milliseconds on the commit's stream (CUDA events): 0 interpolate, 1 LDE (exchange of earlier chunks runs underneath),
 * 2 exchange left exposed after the last LDE kernel, 3 row hashing, 4 subtree, 5 root all-gather + top levels;
 * bytes_sent = LDE bytes this rank shipped to OTHER ranks */
int wg_sharded_phase_ms(const void* handle, float* ms_out /* 6 */, uint64_t* bytes_sent);
/* rows at GLOBAL LDE positions inside [row_start, row_start + row_count): n_pos x total_cols elements (pad columns
 * dropped: RowMatrix::row of every shard, concatenated in global column order) */
int wg_sharded_rows(const void* handle, uint32_t n_pos, const uint64_t* positions, void* rows_out);
/* digests by index: which = 0 local leaves, 1 local subtree nodes (heap layout over row_count leaves), 2 top nodes (heap
 * layout over the G subtree roots; entry 1 = root) */
int wg_sharded_digests(const void* handle, int which, uint32_t n, const uint64_t* indexes, uint8_t* out);
/* Collective: the reference's full arrays on every rank (host): leaves_out L x 32 (tree.leaves) and nodes_out L x 32 in heap
 * layout (crypto/src/merkle/mod.rs:350-374), ready for MerkleTree::from_raw_parts. Either may be NULL. */
int wg_sharded_gather(const void* handle, uint8_t* leaves_out, uint8_t* nodes_out);
/* This rank's part to the host: blocks_out[i] (may be NULL per block) receives received column block i as a row_count x
 * block_row_widths[i] row-major matrix (the rows of RowMatrix shard i this rank owns); leaves_out / nodes_out row_count x 32
 * (local subtree, heap layout); top_out world x 32. Any may be NULL. */
int wg_sharded_download(const void* handle, void* const* blocks_out, uint8_t* leaves_out, uint8_t* nodes_out, uint8_t* top_out);
int wg_sharded_free(void* handle);
/* Test / single-GPU boxes: the SAME per-rank code for all `world` ranks on one device, device copies standing in for the
 * exchange (NCCL cannot put two ranks on one GPU). d_cols holds the whole column list [total_cols][n_rows]. Outputs may be
 * NULL: root_out 32 bytes, leaves_out / nodes_out L x 32 host bytes (stitched reference layout). */
int wg_sharded_commit_emulated(int field, int hash, uint64_t n_rows, uint64_t blowup, uint32_t total_cols, uint32_t world,
                               const void* d_cols, int exchange, uint32_t chunks, uint8_t* root_out, uint8_t* leaves_out,
                               uint8_t* nodes_out);

#ifdef __cplusplus
}
#endif
#endif
