// One Merkle tree over a column-sharded trace on G GPUs, behind the C ABI. Included by api.cu.
//
// The reference has no distributed code: Prover::build_trace_commitment (prover/src/lib.rs:615-670) is one in-process
// call, and RowMatrix::commit_to_comb_rows (prover/src/matrix/row_matrix.rs:204-238) needs whole rows. Interpolation and
// LDE are independent per column, so every rank runs them on ITS block of the global column list, the column shards
// become row blocks in ONE exchange, each rank hashes its rows and builds its subtree, and the G subtree roots are
// all-gathered for the top log2(G) levels (redundantly on every rank, G - 1 compressions).
//
// NCCL is bound at run time (dlopen) so the library has no link-time dependency on it; types come from <nccl.h>.
#pragma once
#include <dlfcn.h>
#include <nccl.h>

namespace wg {

// ------------------------------------------------------------------------------------------------ partition (host only)
static inline void shard_columns(uint32_t total, uint32_t world, uint32_t rank, uint32_t* start, uint32_t* count) {
  const uint32_t base = total / world, extra = total % world;
  *count = base + (rank < extra ? 1 : 0);
  *start = rank * base + (rank < extra ? rank : extra);
}
// chunk idx of k: sizes are multiples of 8 columns except the last non-empty one (an LDE matrix is padded to a multiple
// of 8 columns and the pad columns are not hashed, so only the LAST chunk of a block may carry padding)
static inline void shard_chunk(uint32_t cols, uint32_t k, uint32_t idx, uint32_t* start, uint32_t* count) {
  const uint32_t per = ((cols + k - 1) / k + 7) / 8 * 8;
  const uint64_t s = (uint64_t)per * idx;
  *start = s < cols ? (uint32_t)s : cols;
  *count = s < cols ? (cols - (uint32_t)s < per ? cols - (uint32_t)s : per) : 0;
}
static inline void shard_plan(uint32_t total_cols, uint32_t world, int exchange, uint32_t chunks, int* ex_out, uint32_t* k_out,
                              uint64_t lde_rows = 0) {
  const uint32_t per_rank = (total_cols + world - 1) / world;
  int ex = exchange;
  uint32_t k = chunks;
  if (world <= 1) {
    ex = WG_EXCHANGE_ALLTOALL;
    k = 1;
  } else if (ex == WG_EXCHANGE_AUTO) {
    // narrow shards (a StarkPack trace per GPU) finish their last LDE pass in a few milliseconds and the fused stores
    // save the whole all-to-all; wider shards would be NVLink-bound at 32-byte stores and pipeline NCCL instead
    // (measured, profiles/r2_forms_*gpu.jsonl: 25 f64 columns per rank on 4 GPUs: fused 43.3 ms, pipelined x4 29.4 ms;
    //  2 - 8 f128 columns per rank, cfg4: fused 3.9 / 7.1 ms against 5.8 / 8.3 ms)
    //  13 f64 columns per rank on 8 GPUs: NCCL send/recv moves ~350 GB/s per GPU and its kernels share the SMs with the LDE,
    //  copy-engine pushes into the peers' mapped buffers leave the SMs alone: WG_EXCHANGE_P2P_COPY is the wide-shard default)
    if (per_rank <= 8 && world <= (uint32_t)PassArgs<Gl64>::MAX_PEERS) ex = WG_EXCHANGE_FUSED_P2P;
    else ex = WG_EXCHANGE_P2P_COPY;
  }
  if (ex == WG_EXCHANGE_FUSED_P2P || ex == WG_EXCHANGE_ALLTOALL) k = 1;
  if (ex == WG_EXCHANGE_PIPELINED || ex == WG_EXCHANGE_P2P_COPY) {
    // 8-column chunks; the last chunk's exchange is the exposed one, every chunk costs the LDE a little. At most 4 chunks
    // (headline on 2 GPUs, 50 columns per rank: x4 44.1 ms, x7 45.3 ms), 8 for traces of 2^26 LDE rows and more, where a
    // 16-column chunk is 13 GB on the wire (cfg5 on 4 GPUs: 20 ms exposed with x4)
    if (k == 0) {
      const uint32_t cap = lde_rows >= ((uint64_t)1 << 26) ? 8 : 4;
      k = (per_rank + 7) / 8 > cap ? cap : (per_rank + 7) / 8;
    }
    const uint32_t max_k = (per_rank + 7) / 8;  // at least 8 columns per chunk
    if (k > max_k) k = max_k;
    while (k > 1 && world * k > WG_SHARD_MAX_BLOCKS) k--;
    if (k <= 1) {
      k = 1;
      if (ex == WG_EXCHANGE_PIPELINED) ex = WG_EXCHANGE_ALLTOALL;
    }
  }
  *ex_out = ex;
  *k_out = k ? k : 1;
}
// stored row width of an exchanged column block: exactly its columns. The single-GPU LDE pads rows to a multiple of 8
// columns because the reference's RowMatrix does (row_matrix.rs:110-133); pad columns are never hashed, and here they
// would only be bytes on NVLink (13 columns per rank at 8 GPUs: 16 stored = +23 %).
static inline uint64_t shard_rw(uint32_t cols) { return cols; }
// heap index of the global tree (n_leaves leaves, `world` subtrees) -> owner. Levels above the subtree roots (index <
// world) live in the top array; node i at level m = 2^floor(log2 i) >= world belongs to rank (i - m) / (m / world) with
// local heap index m/world + (i - m) % (m / world).
static inline void shard_node_owner(uint64_t n_leaves, uint32_t world, uint64_t i, int* rank, uint64_t* local) {
  (void)n_leaves;
  if (i < world) {
    *rank = -1;
    *local = i;
    return;
  }
  const uint64_t m = (uint64_t)1 << ilog2(i);
  const uint64_t per = m / world;
  *rank = (int)((i - m) / per);
  *local = per + (i - m) % per;
}

// ------------------------------------------------------------------------------------------------ NCCL binding
struct NcclApi {
  void* lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
static NcclApi g_nccl;
static std::mutex g_dist_mu;

static int nccl_load() {
  if (g_nccl.lib) return WG_OK;
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!h) return fail(WG_ERR_UNSUPPORTED, "NCCL not found (dlopen libnccl.so.2: %s)", dlerror());
#define WG_NCCL_SYM(name)                                                                       \
  *(void**)(&g_nccl.name) = dlsym(h, "nccl" #name);                                             \
  if (!g_nccl.name) return fail(WG_ERR_UNSUPPORTED, "NCCL symbol nccl" #name " missing")
  WG_NCCL_SYM(GetUniqueId);
  WG_NCCL_SYM(CommInitRank);
  WG_NCCL_SYM(CommDestroy);
  WG_NCCL_SYM(GroupStart);
  WG_NCCL_SYM(GroupEnd);
  WG_NCCL_SYM(Send);
  WG_NCCL_SYM(Recv);
  WG_NCCL_SYM(AllGather);
  WG_NCCL_SYM(GetErrorString);
#undef WG_NCCL_SYM
  g_nccl.lib = h;
  return WG_OK;
}
#define NCCL_TRY(expr)                                                                                        \
  do {                                                                                                        \
    ncclResult_t r__ = (expr);                                                                                \
    if (r__ != ncclSuccess) return fail(WG_ERR_CUDA, "%s: NCCL %s", #expr, g_nccl.GetErrorString(r__));        \
  } while (0)

// ------------------------------------------------------------------------------------------------ process-wide state
struct PeerBuf {  // a receive buffer other ranks can store into (fused exchange); mappings are kept across commits
  void* ptr = nullptr;
  size_t bytes = 0;
  bool busy = false;
  std::vector<void*> peers;  // peers[g]: this buffer's counterpart on rank g, mapped here (own rank: ptr)
};
struct DistState {
  bool on = false;
  int rank = 0, world = 1;
  ncclComm_t comm = nullptr;
  cudaStream_t s_comm = nullptr;
  static constexpr int N_COPY = 4;           // peer copies of one chunk run on several streams = several copy engines
  cudaStream_t s_copy[N_COPY] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t ev_copy[N_COPY] = {nullptr, nullptr, nullptr, nullptr};
  std::vector<PeerBuf> peer_bufs;
  uint8_t* d_small = nullptr;  // 64-byte slots: IPC handles / barrier tokens
};
static DistState g_dist;

static int dist_streams() {
  if (!g_dist.s_comm) CUDA_TRY(cudaStreamCreateWithFlags(&g_dist.s_comm, cudaStreamNonBlocking));
  for (int i = 0; i < DistState::N_COPY; i++) {
    if (!g_dist.s_copy[i]) CUDA_TRY(cudaStreamCreateWithFlags(&g_dist.s_copy[i], cudaStreamNonBlocking));
    if (!g_dist.ev_copy[i]) CUDA_TRY(cudaEventCreateWithFlags(&g_dist.ev_copy[i], cudaEventDisableTiming));
  }
  if (!g_dist.d_small) CUDA_TRY(cudaMalloc(&g_dist.d_small, 64 * (PassArgs<Gl64>::MAX_PEERS + 64 + 1)));
  return WG_OK;
}

// ------------------------------------------------------------------------------------------------ the handle
struct ShardBlock {
  uint32_t cols = 0;    // real columns
  uint64_t rw = 0;      // stored row width (multiple of 8)
  size_t off = 0;       // element offset of the rows x rw matrix inside recv
};
struct ShardedCommit {
  int field = 0, hash = 0, exchange = 0;
  uint32_t world = 1, rank = 0, K = 1, total_cols = 0;
  uint64_t n = 0, blowup = 0, L = 0, rows = 0;
  std::vector<ShardBlock> blocks;  // global column order: rank-major, then chunk
  void* recv = nullptr;
  size_t recv_bytes = 0;
  int peer_slot = -1;  // recv is g_dist.peer_bufs[peer_slot] (fused exchange)
  void* polys = nullptr;
  size_t polys_bytes = 0;
  uint8_t* leaves = nullptr;
  uint8_t* nodes = nullptr;
  uint8_t* top = nullptr;  // world x 32, heap layout over the subtree roots (device)
  std::vector<uint8_t> top_host;
  cudaEvent_t ev[8] = {};
  float phase_ms[6] = {0, 0, 0, 0, 0, 0};
  bool timed = false;
  uint64_t bytes_sent = 0;
  size_t esize() const { return field == WG_FIELD_F64 ? 8 : 16; }
  ~ShardedCommit() {
    if (peer_slot >= 0) g_dist.peer_bufs[peer_slot].busy = false;
    else cached_free(recv, recv_bytes);
    cached_free(polys, polys_bytes);
    cached_free(leaves, rows * 32);
    cached_free(nodes, rows * 32);
    cached_free(top, (size_t)(world < 2 ? 2 : world) * 32);
    for (auto& e : ev)
      if (e) cudaEventDestroy(e);
  }
};

// Receive-buffer layout shared by every exchange form: block (g, k) — chunk k of rank g's columns — is a rows x rw
// matrix; blocks follow each other in global column order.
static void shard_layout(ShardedCommit* h) {
  h->blocks.clear();
  size_t off = 0;
  for (uint32_t g = 0; g < h->world; g++) {
    uint32_t s, c;
    shard_columns(h->total_cols, h->world, g, &s, &c);
    for (uint32_t k = 0; k < h->K; k++) {
      uint32_t cs, cc;
      shard_chunk(c, h->K, k, &cs, &cc);
      ShardBlock b;
      b.cols = cc;
      b.rw = shard_rw(cc);
      b.off = off;
      off += (size_t)h->rows * b.rw;
      h->blocks.push_back(b);
    }
  }
  h->recv_bytes = off * h->esize();
  if (h->recv_bytes == 0) h->recv_bytes = 16;
}

// fused exchange: a receive buffer of this size that every rank has mapped (collective on first use of a size)
static int peer_buf_acquire(size_t bytes, int* slot_out) {
  for (size_t i = 0; i < g_dist.peer_bufs.size(); i++)
    if (!g_dist.peer_bufs[i].busy && g_dist.peer_bufs[i].bytes == bytes) {
      g_dist.peer_bufs[i].busy = true;
      *slot_out = (int)i;
      return WG_OK;
    }
  const int G = g_dist.world, r = g_dist.rank;
  PeerBuf pb;
  pb.bytes = bytes;
  cudaError_t e = dev_malloc(&pb.ptr, bytes);
  if (e != cudaSuccess) return fail(WG_ERR_NOMEM, "cudaMalloc(%zu): %s", bytes, cudaGetErrorString(e));
  cudaIpcMemHandle_t mine;
  CUDA_TRY(cudaIpcGetMemHandle(&mine, pb.ptr));
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  uint8_t* slots = g_dist.d_small;  // [G] gathered + [1] mine
  CUDA_TRY(cudaMemcpyAsync(slots + 64 * G, &mine, 64, cudaMemcpyHostToDevice, g_dist.s_comm));
  NCCL_TRY(g_nccl.AllGather(slots + 64 * G, slots, 64, ncclUint8, g_dist.comm, g_dist.s_comm));
  std::vector<cudaIpcMemHandle_t> all(G);
  CUDA_TRY(cudaMemcpyAsync(all.data(), slots, 64 * (size_t)G, cudaMemcpyDeviceToHost, g_dist.s_comm));
  CUDA_TRY(cudaStreamSynchronize(g_dist.s_comm));
  pb.peers.assign(G, nullptr);
  for (int g = 0; g < G; g++) {
    if (g == r) pb.peers[g] = pb.ptr;
    else CUDA_TRY(cudaIpcOpenMemHandle(&pb.peers[g], all[g], cudaIpcMemLazyEnablePeerAccess));
  }
  pb.busy = true;
  g_dist.peer_bufs.push_back(pb);
  *slot_out = (int)g_dist.peer_bufs.size() - 1;
  return WG_OK;
}
// scratch + send buffers of one rank's LDE, sized for the chunk plan
struct ShardScratch {
  void* scr = nullptr;
  size_t scr_bytes = 0;
  uint64_t scr_elems = 0;
  void* send[2] = {nullptr, nullptr};
  size_t send_bytes = 0;
  ~ShardScratch() {
    cached_free(scr, scr_bytes);
    cached_free(send[0], send_bytes);
    cached_free(send[1], send_bytes);
  }
};

template <class F>
static int shard_scratch_alloc(const ShardedCommit* h, uint32_t my_cols, bool need_send, ShardScratch* s) {
  typedef typename F::T T;
  uint32_t c0, maxc = 0;
  for (uint32_t k = 0; k < h->K; k++) {
    uint32_t cc;
    shard_chunk(my_cols, h->K, k, &c0, &cc);
    if (cc > maxc) maxc = cc;
  }
  // the send buffers first, then the transform scratch: at its preferred size (split plan for 2^23 / 2^24-row traces, see
  // api.cu g_split_pref) when that still fits, else the two-pass plan's
  if (need_send) {
    s->send_bytes = (size_t)h->L * shard_rw(maxc) * sizeof(T);
    if (s->send_bytes == 0) s->send_bytes = 16;
    WG_TRY(cached_malloc(&s->send[0], s->send_bytes));
    if (h->K > 1) WG_TRY(cached_malloc(&s->send[1], s->send_bytes));
  }
  const uint64_t se_i = interpolate_scratch_elems<F>(my_cols, h->n);
  const bool scatter = h->exchange == WG_EXCHANGE_FUSED_P2P && h->world > 1;  // the scatter form keeps the two-pass plan
  const uint64_t se_min = lde_scratch_elems_min<F>(maxc, h->n, h->blowup);
  const uint64_t se_pref = scatter ? se_min : lde_scratch_elems<F>(maxc, h->n, h->blowup);
  for (int attempt = 0; attempt < 2; attempt++) {
    const uint64_t se_l = attempt == 0 ? se_pref : se_min;
    s->scr_elems = se_i > se_l ? se_i : se_l;
    s->scr_bytes = s->scr_elems * sizeof(T) + 16;
    if (cached_malloc(&s->scr, s->scr_bytes) == WG_OK) return WG_OK;
    s->scr = nullptr;
    cudaGetLastError();  // the failed allocation must not surface at the next launch check
    if (se_pref == se_min) break;
  }
  return fail(WG_ERR_NOMEM, "no memory for %zu bytes of transform scratch", s->scr_bytes);
  return WG_OK;
}

// leaves over this rank's row block (all received column blocks in global column order) and the local subtree
template <class F>
static int shard_hash_and_tree(ShardedCommit* h, cudaStream_t sc) {
  typedef typename F::T T;
  std::vector<const void*> ptrs;
  std::vector<uint64_t> rws, els;
  for (const ShardBlock& b : h->blocks) {
    if (b.cols == 0) continue;
    ptrs.push_back((const T*)h->recv + b.off);
    rws.push_back(b.rw);
    els.push_back(b.cols);
  }
  WG_TRY(hash_rows_dev<F>(h->hash, (uint32_t)ptrs.size(), ptrs.data(), rws.data(), els.data(), h->rows, h->leaves, sc));
  if (h->timed) CUDA_TRY(cudaEventRecord(h->ev[4], sc));
  WG_TRY(merkle_dev(h->hash, h->leaves, h->rows, h->nodes, sc));
  if (h->timed) CUDA_TRY(cudaEventRecord(h->ev[5], sc));
  return WG_OK;
}

static int shard_alloc_outputs(ShardedCommit* h, bool own_polys, uint32_t my_cols) {
  WG_TRY(cached_malloc((void**)&h->leaves, h->rows * 32));
  WG_TRY(cached_malloc((void**)&h->nodes, h->rows * 32));
  WG_TRY(cached_malloc((void**)&h->top, (size_t)(h->world < 2 ? 2 : h->world) * 32));
  if (own_polys) {
    h->polys_bytes = (size_t)my_cols * h->n * h->esize();
    if (h->polys_bytes == 0) h->polys_bytes = 16;
    WG_TRY(cached_malloc(&h->polys, h->polys_bytes));
  }
  return WG_OK;
}

static int shard_check_shape(int field, int hash, uint64_t n, uint64_t blowup, uint32_t total_cols, uint32_t world) {
  if (field != WG_FIELD_F64 && field != WG_FIELD_F128) return fail(WG_ERR_ARG, "unknown field id %d", field);
  if (hash == WG_HASH_BLAKE3_192) return fail(WG_ERR_UNSUPPORTED, "Blake3_192 is not supported by the sharded entry points");
  if (hash != WG_HASH_BLAKE3_256 && hash != WG_HASH_SHA3_256) return fail(WG_ERR_ARG, "unknown hash %d", hash);
  if (!is_pow2(world)) return fail(WG_ERR_ARG, "number of ranks %u must be a power of two", world);
  if (total_cols < world) return fail(WG_ERR_ARG, "more ranks (%u) than columns (%u)", world, total_cols);
  if (!is_pow2(n) || !is_pow2(blowup)) return fail(WG_ERR_NOT_POW2, "rows %llu / blowup %llu must be powers of two", (unsigned long long)n, (unsigned long long)blowup);
  if (n * blowup / world < 2) return fail(WG_ERR_TOO_FEW, "each rank needs at least 2 leaves");
  return WG_OK;
}

// ------------------------------------------------------------------------------------------------ the per-rank commit
template <class F>
static int sharded_commit_run(ShardedCommit* h, const typename F::T* d_cols, typename F::T* d_polys, uint8_t* root_out,
                              cudaStream_t sc) {
  typedef typename F::T T;
  const uint32_t G = h->world, r = h->rank;
  uint32_t my_start, my_cols;
  shard_columns(h->total_cols, G, r, &my_start, &my_cols);
  WG_TRY(check_domain<F>(h->n, h->blowup));
  shard_layout(h);
  WG_TRY(shard_alloc_outputs(h, d_polys == nullptr, my_cols));
  T* polys = d_polys ? d_polys : (T*)h->polys;
  const bool fused = h->exchange == WG_EXCHANGE_FUSED_P2P && G > 1;
  const bool p2pcopy = h->exchange == WG_EXCHANGE_P2P_COPY && G > 1;
  if (fused || p2pcopy) {
    WG_TRY(peer_buf_acquire(h->recv_bytes, &h->peer_slot));
    h->recv = g_dist.peer_bufs[h->peer_slot].ptr;
  } else {
    WG_TRY(cached_malloc(&h->recv, h->recv_bytes));
  }
  ShardScratch ss;
  WG_TRY(shard_scratch_alloc<F>(h, my_cols, !fused && G > 1, &ss));
  for (auto& e : h->ev) CUDA_TRY(cudaEventCreate(&e));
  h->timed = true;
  cudaStream_t sx = g_dist.s_comm;
  const T offset = generator<F>();

  CUDA_TRY(cudaEventRecord(h->ev[0], sc));
  WG_TRY(interpolate_columns_dev<F>(d_cols, my_cols, h->n, polys, (T*)ss.scr, ss.scr_elems, sc));
  CUDA_TRY(cudaEventRecord(h->ev[1], sc));

  if (G == 1) {
    // one rank: the LDE writes the row block directly
    WG_TRY(lde_dev<F>(polys, my_cols, h->n, h->blowup, offset, (T*)h->recv, h->blocks[0].rw, (T*)ss.scr, ss.scr_elems, sc));
    CUDA_TRY(cudaEventRecord(h->ev[2], sc));
    CUDA_TRY(cudaEventRecord(h->ev[3], sc));
  } else if (fused) {
    // owner o's slot for MY columns is block (r, 0) of ITS receive buffer
    T* peers[PassArgs<F>::MAX_PEERS];
    const ShardBlock& mine = h->blocks[(size_t)r * h->K];
    for (uint32_t o = 0; o < G; o++) peers[o] = (T*)g_dist.peer_bufs[h->peer_slot].peers[o] + mine.off;
    WG_TRY(lde_dev<F>(polys, my_cols, h->n, h->blowup, offset, nullptr, mine.rw, (T*)ss.scr, ss.scr_elems, sc, peers, G));
    CUDA_TRY(cudaEventRecord(h->ev[2], sc));
    // every rank's rows have landed once every rank's LDE kernel has finished: a stream-ordered barrier
    uint8_t* tok = g_dist.d_small + 64 * (PassArgs<Gl64>::MAX_PEERS + 1);
    NCCL_TRY(g_nccl.AllGather(tok + 64 * G, tok, 4, ncclUint8, g_dist.comm, sc));
    CUDA_TRY(cudaEventRecord(h->ev[3], sc));
    h->bytes_sent = (uint64_t)(h->L - h->rows) * mine.rw * sizeof(T);
  } else {
    // LDE of chunk k into a send buffer, its all-to-all on the exchange stream while chunk k+1 is computed
    cudaEvent_t ev_lde, ev_x[2] = {h->ev[6], h->ev[7]};
    CUDA_TRY(cudaEventCreateWithFlags(&ev_lde, cudaEventDisableTiming));
    int rc = WG_OK;
    for (uint32_t k = 0; k < h->K && rc == WG_OK; k++) {
      uint32_t cs, cc;
      shard_chunk(my_cols, h->K, k, &cs, &cc);
      const uint64_t rw_k = shard_rw(cc);
      T* send = (T*)ss.send[k & 1];
      if (k >= 2) {  // the exchange that last read this send buffer must be over
        if (cudaStreamWaitEvent(sc, ev_x[k & 1], 0) != cudaSuccess) rc = fail(WG_ERR_CUDA, "cudaStreamWaitEvent");
      }
      if (rc == WG_OK && cc) rc = lde_dev<F>(polys + (size_t)cs * h->n, cc, h->n, h->blowup, offset, send, rw_k, (T*)ss.scr, ss.scr_elems, sc);
      if (rc != WG_OK) break;
      cudaEventRecord(ev_lde, sc);
      cudaStreamWaitEvent(sx, ev_lde, 0);
      if (p2pcopy) {
        // push my slice of every owner's row block into ITS receive buffer (peer-mapped, CUDA IPC) with the copy engines:
        // each slice is contiguous on both sides; owners are visited round-robin from my right-hand neighbour
        const size_t bytes = (size_t)h->rows * rw_k * sizeof(T);
        const ShardBlock& mine_k = h->blocks[(size_t)r * h->K + k];
        // the G copies of a chunk go out on N_COPY streams (one copy engine each) and join on the exchange stream
        for (int c = 0; c < DistState::N_COPY; c++) cudaStreamWaitEvent(g_dist.s_copy[c], ev_lde, 0);
        for (uint32_t i = 0; i < G && cc && rc == WG_OK; i++) {
          const uint32_t o = (r + 1 + i) % G;
          T* dst = (T*)g_dist.peer_bufs[h->peer_slot].peers[o] + mine_k.off;
          if (cudaMemcpyAsync(dst, send + (size_t)o * h->rows * rw_k, bytes, cudaMemcpyDeviceToDevice,
                              g_dist.s_copy[i % DistState::N_COPY]) != cudaSuccess)
            rc = fail(WG_ERR_CUDA, "peer copy of chunk %u to rank %u: %s", k, o, cudaGetErrorString(cudaGetLastError()));
          if (o != r) h->bytes_sent += bytes;
        }
        for (int c = 0; c < DistState::N_COPY; c++) {
          cudaEventRecord(g_dist.ev_copy[c], g_dist.s_copy[c]);
          cudaStreamWaitEvent(sx, g_dist.ev_copy[c], 0);
        }
        cudaEventRecord(ev_x[k & 1], sx);
        continue;
      }
      ncclResult_t nr = g_nccl.GroupStart();
      for (uint32_t g = 0; g < G && nr == ncclSuccess; g++) {
        const ShardBlock& from_g = h->blocks[(size_t)g * h->K + k];
        const size_t send_bytes = (size_t)h->rows * rw_k * sizeof(T);
        if (cc) nr = g_nccl.Send(send + (size_t)g * h->rows * rw_k, send_bytes, ncclUint8, (int)g, g_dist.comm, sx);
        if (nr == ncclSuccess && from_g.cols)
          nr = g_nccl.Recv((T*)h->recv + from_g.off, (size_t)h->rows * from_g.rw * sizeof(T), ncclUint8, (int)g, g_dist.comm, sx);
        if (g != r) h->bytes_sent += cc ? send_bytes : 0;
      }
      ncclResult_t ne = g_nccl.GroupEnd();
      if (nr == ncclSuccess) nr = ne;
      if (nr != ncclSuccess) rc = fail(WG_ERR_CUDA, "NCCL all-to-all of chunk %u: %s", k, g_nccl.GetErrorString(nr));
      cudaEventRecord(ev_x[k & 1], sx);
    }
    CUDA_TRY(cudaEventRecord(h->ev[2], sc));
    if (rc == WG_OK && p2pcopy) {
      // every rank's pushes have landed once every rank's copies are complete: a stream-ordered barrier behind them
      uint8_t* tok = g_dist.d_small + 64 * (PassArgs<Gl64>::MAX_PEERS + 1);
      ncclResult_t nr = g_nccl.AllGather(tok + 64 * G, tok, 4, ncclUint8, g_dist.comm, sx);
      if (nr != ncclSuccess) rc = fail(WG_ERR_CUDA, "NCCL barrier after the peer copies: %s", g_nccl.GetErrorString(nr));
      cudaEventRecord(ev_x[(h->K - 1) & 1], sx);
    }
    if (rc == WG_OK) {
      cudaStreamWaitEvent(sc, ev_x[0], 0);
      if (h->K > 1) cudaStreamWaitEvent(sc, ev_x[1], 0);
    }
    CUDA_TRY(cudaEventRecord(h->ev[3], sc));
    cudaEventDestroy(ev_lde);
    if (rc != WG_OK) {
      cudaStreamSynchronize(sx);
      cudaStreamSynchronize(sc);
      return rc;
    }
  }
  WG_TRY(shard_hash_and_tree<F>(h, sc));
  // top log2(G) levels over the G subtree roots
  h->top_host.assign((size_t)(G < 2 ? 2 : G) * 32, 0);
  if (G > 1) {
    uint8_t* roots = g_dist.d_small + 64 * (PassArgs<Gl64>::MAX_PEERS + 1);  // G x 32 <= 64 slots
    if ((size_t)G * 32 > 64 * 64) return fail(WG_ERR_UNSUPPORTED, "more than 128 ranks");
    NCCL_TRY(g_nccl.AllGather(h->nodes + 32, roots, 32, ncclUint8, g_dist.comm, sc));
    WG_TRY(merkle_dev(h->hash, roots, G, h->top, sc));
    CUDA_TRY(cudaMemcpyAsync(h->top_host.data(), h->top, (size_t)G * 32, cudaMemcpyDeviceToHost, sc));
  } else {
    CUDA_TRY(cudaMemcpyAsync(h->top_host.data(), h->nodes, 64, cudaMemcpyDeviceToHost, sc));
    CUDA_TRY(cudaMemcpyAsync(h->top, h->nodes, 64, cudaMemcpyDeviceToDevice, sc));
  }
  CUDA_TRY(cudaEventRecord(h->ev[6], sc));
  CUDA_TRY(cudaStreamSynchronize(sc));
  CUDA_TRY(cudaStreamSynchronize(sx));
  if (root_out) memcpy(root_out, h->top_host.data() + 32, 32);
  const int pairs[6][2] = {{0, 1}, {1, 2}, {2, 3}, {3, 4}, {4, 5}, {5, 6}};
  for (int i = 0; i < 6; i++) CUDA_TRY(cudaEventElapsedTime(&h->phase_ms[i], h->ev[pairs[i][0]], h->ev[pairs[i][1]]));
  return WG_OK;
}

static int sharded_commit_create(int field, int hash, uint64_t n, uint64_t blowup, uint32_t total_cols, const void* d_cols,
                                 void* d_polys, int exchange, uint32_t chunks, uint8_t* root_out, void** handle, cudaStream_t sc) {
  if (!handle || !d_cols) return fail(WG_ERR_ARG, "null pointer");
  *handle = nullptr;
  WG_TRY(ensure_device());
  const uint32_t G = (uint32_t)g_dist.world;
  WG_TRY(shard_check_shape(field, hash, n, blowup, total_cols, G));
  std::lock_guard<std::mutex> lk(g_dist_mu);
  WG_TRY(dist_streams());
  ShardedCommit* h = new ShardedCommit;
  h->field = field;
  h->hash = hash;
  h->world = G;
  h->rank = (uint32_t)g_dist.rank;
  h->n = n;
  h->blowup = blowup;
  h->L = n * blowup;
  h->rows = h->L / G;
  h->total_cols = total_cols;
  int ex;
  uint32_t k;
  shard_plan(total_cols, G, exchange, chunks, &ex, &k, h->L);
  h->exchange = ex;
  h->K = k;
  int rc = field == WG_FIELD_F64 ? sharded_commit_run<Gl64>(h, (const uint64_t*)d_cols, (uint64_t*)d_polys, root_out, sc)
                                 : sharded_commit_run<F128>(h, (const u128*)d_cols, (u128*)d_polys, root_out, sc);
  if (rc != WG_OK) {
    cudaStreamSynchronize(sc);
    if (g_dist.s_comm) cudaStreamSynchronize(g_dist.s_comm);
    delete h;
    return rc;
  }
  *handle = h;
  return WG_OK;
}

// ------------------------------------------------------------------------------------------------ stitching
// local subtree heaps [g][rows x 32] + top heap [G x 32] -> the reference's heap layout over L leaves
// (crypto/src/merkle/mod.rs:350-374): level m >= G of the global tree is the concatenation of every rank's level m/G.
static void stitch_nodes_host(const uint8_t* locals, const uint8_t* top, uint32_t G, uint64_t rows, uint8_t* out) {
  const uint64_t L = rows * G;
  memset(out, 0, 32);
  if (G > 1) memcpy(out + 32, top + 32, (size_t)(G - 1) * 32);
  for (uint64_t m = G; m < L; m *= 2) {
    const uint64_t per = m / G;
    for (uint32_t g = 0; g < G; g++)
      memcpy(out + (m + g * per) * 32, locals + ((size_t)g * rows + per) * 32, (size_t)per * 32);
  }
}

// ------------------------------------------------------------------------------------------------ one-device emulation
// Runs the per-rank code of all G ranks on the current device; the exchange is a device copy of the slice each
// receiver would get (or, fused, the peers' buffers are simply the other ranks' buffers on this device).
template <class F>
static int sharded_emulate(int hash, uint64_t n, uint64_t blowup, uint32_t total_cols, uint32_t G, const typename F::T* d_cols,
                           int exchange, uint32_t chunks, uint8_t* root_out, uint8_t* leaves_out, uint8_t* nodes_out) {
  typedef typename F::T T;
  WG_TRY(check_domain<F>(n, blowup));
  cudaStream_t sc = 0;
  std::vector<ShardedCommit*> hs(G, nullptr);
  struct Cleanup {
    std::vector<ShardedCommit*>& v;
    ~Cleanup() {
      cudaDeviceSynchronize();
      for (auto* p : v) delete p;
    }
  } cleanup{hs};
  int ex;
  uint32_t K;
  shard_plan(total_cols, G, exchange, chunks, &ex, &K);
  for (uint32_t r = 0; r < G; r++) {
    ShardedCommit* h = hs[r] = new ShardedCommit;
    h->field = F::kFieldId;
    h->hash = hash;
    h->world = G;
    h->rank = r;
    h->n = n;
    h->blowup = blowup;
    h->L = n * blowup;
    h->rows = h->L / G;
    h->total_cols = total_cols;
    h->exchange = ex;
    h->K = K;
    shard_layout(h);
    uint32_t s, c;
    shard_columns(total_cols, G, r, &s, &c);
    WG_TRY(shard_alloc_outputs(h, true, c));
    WG_TRY(cached_malloc(&h->recv, h->recv_bytes));
  }
  const T offset = generator<F>();
  for (uint32_t r = 0; r < G; r++) {  // "rank r" produces its shard and delivers it
    ShardedCommit* h = hs[r];
    uint32_t s, c;
    shard_columns(total_cols, G, r, &s, &c);
    const bool fused = ex == WG_EXCHANGE_FUSED_P2P && G > 1;
    ShardScratch ss;
    WG_TRY(shard_scratch_alloc<F>(h, c, !fused, &ss));
    T* polys = (T*)h->polys;
    WG_TRY(interpolate_columns_dev<F>(d_cols + (size_t)s * n, c, n, polys, (T*)ss.scr, ss.scr_elems, sc));
    if (fused) {
      T* peers[PassArgs<F>::MAX_PEERS];
      const ShardBlock& mine = h->blocks[(size_t)r * K];
      for (uint32_t o = 0; o < G; o++) peers[o] = (T*)hs[o]->recv + mine.off;
      WG_TRY(lde_dev<F>(polys, c, n, blowup, offset, nullptr, mine.rw, (T*)ss.scr, ss.scr_elems, sc, peers, G));
    } else {
      for (uint32_t k = 0; k < K; k++) {
        uint32_t cs, cc;
        shard_chunk(c, K, k, &cs, &cc);
        if (!cc) continue;
        const uint64_t rw_k = shard_rw(cc);
        T* send = (T*)ss.send[k & 1];
        WG_TRY(lde_dev<F>(polys + (size_t)cs * n, cc, n, blowup, offset, send, rw_k, (T*)ss.scr, ss.scr_elems, sc));
        for (uint32_t o = 0; o < G; o++) {  // what ncclSend(o) / ncclRecv(r) on rank o would move
          const ShardBlock& dst = hs[o]->blocks[(size_t)r * K + k];
          CUDA_TRY(cudaMemcpyAsync((T*)hs[o]->recv + dst.off, send + (size_t)o * h->rows * rw_k, (size_t)h->rows * rw_k * sizeof(T),
                                   cudaMemcpyDeviceToDevice, sc));
        }
      }
    }
    CUDA_TRY(cudaStreamSynchronize(sc));
  }
  const uint64_t rows = hs[0]->rows;
  std::vector<uint8_t> roots((size_t)(G < 2 ? 2 : G) * 32), top((size_t)(G < 2 ? 2 : G) * 32);
  for (uint32_t r = 0; r < G; r++) {
    WG_TRY(shard_hash_and_tree<F>(hs[r], sc));
    CUDA_TRY(cudaMemcpyAsync(roots.data() + (size_t)r * 32, hs[r]->nodes + 32, 32, cudaMemcpyDeviceToHost, sc));
  }
  CUDA_TRY(cudaStreamSynchronize(sc));
  if (G > 1) {
    CUDA_TRY(cudaMemcpyAsync(hs[0]->top, roots.data(), (size_t)G * 32, cudaMemcpyHostToDevice, sc));
    uint8_t* d_top2 = nullptr;
    WG_TRY(cached_malloc((void**)&d_top2, (size_t)G * 32));
    int rc = merkle_dev(hash, hs[0]->top, G, d_top2, sc);
    if (rc == WG_OK && cudaMemcpyAsync(top.data(), d_top2, (size_t)G * 32, cudaMemcpyDeviceToHost, sc) != cudaSuccess) rc = fail(WG_ERR_CUDA, "copy of the top levels");
    cudaStreamSynchronize(sc);
    cached_free(d_top2, (size_t)G * 32);
    WG_TRY(rc);
  } else {
    memcpy(top.data() + 32, roots.data(), 32);
  }
  if (root_out) memcpy(root_out, top.data() + 32, 32);
  if (leaves_out)
    for (uint32_t r = 0; r < G; r++) CUDA_TRY(cudaMemcpy(leaves_out + (size_t)r * rows * 32, hs[r]->leaves, rows * 32, cudaMemcpyDeviceToHost));
  if (nodes_out) {
    std::vector<uint8_t> locals((size_t)G * rows * 32);
    for (uint32_t r = 0; r < G; r++) CUDA_TRY(cudaMemcpy(locals.data() + (size_t)r * rows * 32, hs[r]->nodes, rows * 32, cudaMemcpyDeviceToHost));
    stitch_nodes_host(locals.data(), top.data(), G, rows, nodes_out);
  }
  return WG_OK;
}

// rows_out[p][c] over all real columns, for positions inside this rank's row range
__global__ void shard_gather_rows_kernel(const uint64_t* const* blocks, const uint32_t* rw_words, const uint32_t* unit_off,
                                         uint32_t n_blocks, uint32_t units_total, const uint64_t* pos, uint32_t n_pos, uint64_t row_start,
                                         uint64_t* out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (uint64_t)n_pos * units_total) return;
  const uint32_t p = (uint32_t)(i / units_total), u = (uint32_t)(i % units_total);
  uint32_t b = 0;
  while (b + 1 < n_blocks && u >= unit_off[b + 1]) b++;
  out[i] = blocks[b][(pos[p] - row_start) * rw_words[b] + (u - unit_off[b])];
}

}  // namespace wg
