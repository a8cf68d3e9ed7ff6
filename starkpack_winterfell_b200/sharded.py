"""One Merkle tree over a column-sharded trace on G GPUs — ctypes front end of the C ABI's wg_dist_* / wg_sharded_*
(include/winter_gpu.h). Everything on the data path — interpolation, LDE, the column-shard -> row-block
exchange (NCCL all-to-all, pipelined, or fused P2P stores), row hashing, the subtree, the root all-gather and the top
levels — runs inside libwinter_gpu.so; this module only carries pointers. One process per GPU.

Reference behaviour matched: Prover::build_trace_commitment (prover/src/lib.rs:615-670) with
RowMatrix::commit_to_comb_rows (prover/src/matrix/row_matrix.rs:204-238) over the GLOBAL column list; the stitched
leaves / nodes are the arrays MerkleTree::from_raw_parts (crypto/src/merkle/mod.rs:149-161) adopts."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import (ELEM_BYTES, EXCHANGE_AUTO, EXCHANGE_NAMES, SHARD_MAX_BLOCKS, check, lib)

PHASE_NAMES = ["interpolate", "lde", "exchange_exposed", "hash", "merkle", "top"]


# ---- host-only partition logic (no GPU): thin wrappers over the library's own functions ---------------------------------
def shard_columns(total_cols, world, rank):
    s, c = C.c_uint32(), C.c_uint32()
    check(lib().wg_shard_columns(total_cols, world, rank, C.byref(s), C.byref(c)))
    return s.value, c.value


def shard_chunks(cols, k):
    out = []
    for i in range(k):
        s, c = C.c_uint32(), C.c_uint32()
        check(lib().wg_shard_chunks(cols, k, i, C.byref(s), C.byref(c)))
        out.append((s.value, c.value))
    return out


def shard_plan(total_cols, world, exchange=EXCHANGE_AUTO, chunks=0):
    ex, k = C.c_int(), C.c_uint32()
    check(lib().wg_shard_plan(total_cols, world, exchange, chunks, C.byref(ex), C.byref(k)))
    return ex.value, k.value


def node_owner(n_leaves, world, index):
    """global heap index -> (rank or -1 for the top levels, local heap index)"""
    r, li = C.c_int(), C.c_uint64()
    check(lib().wg_shard_node_owner(n_leaves, world, index, C.byref(r), C.byref(li)))
    return r.value, li.value


# ---- communicator ----------------------------------------------------------------------------------------------------------
def dist_init(rank, world, device=0, broadcast=None):
    """wg_init + wg_dist_init. `broadcast(bytes_or_None) -> bytes` ships rank 0's NCCL id to every rank; by default
    torch.distributed (any backend) is used, which must then be initialised."""
    _lib.init(device)
    ident = None
    if world > 1:
        buf = (C.c_uint8 * _lib.DIST_ID_BYTES)()
        if rank == 0:
            check(lib().wg_dist_unique_id(buf))
        if broadcast is None:
            import torch.distributed as dist

            box = [bytes(buf) if rank == 0 else None]
            dist.broadcast_object_list(box, src=0)
            ident = box[0]
        else:
            ident = broadcast(bytes(buf) if rank == 0 else None)
        ident = (C.c_uint8 * _lib.DIST_ID_BYTES).from_buffer_copy(ident)
    check(lib().wg_dist_init(rank, world, ident))


def dist_shutdown():
    check(lib().wg_dist_shutdown())


def dist_info():
    r, w = C.c_int(), C.c_int()
    check(lib().wg_dist_info(C.byref(r), C.byref(w)))
    return r.value, w.value


class ShardedCommitment:
    """Handle of one sharded commit: this rank's row block, leaves and subtree stay in HBM."""

    def __init__(self, handle, field, root):
        self._h, self.field, self._root = handle, field, root
        L, r0, rc, nb, ex, k = C.c_uint64(), C.c_uint64(), C.c_uint64(), C.c_uint32(), C.c_int(), C.c_uint32()
        bc = (C.c_uint32 * SHARD_MAX_BLOCKS)()
        bw = (C.c_uint32 * SHARD_MAX_BLOCKS)()
        check(lib().wg_sharded_info(self._h, C.byref(L), C.byref(r0), C.byref(rc), C.byref(nb), bc, bw, C.byref(ex), C.byref(k)))
        self.lde_rows, self.row_start, self.row_count = L.value, r0.value, rc.value
        self.block_cols = [bc[i] for i in range(nb.value)]
        self.block_row_widths = [bw[i] for i in range(nb.value)]
        self.exchange, self.chunks = EXCHANGE_NAMES[ex.value], k.value
        self.total_cols = sum(self.block_cols)

    def root(self):
        return self._root

    def phases_ms(self):
        ms = (C.c_float * 6)()
        sent = C.c_uint64()
        check(lib().wg_sharded_phase_ms(self._h, ms, C.byref(sent)))
        return {PHASE_NAMES[i]: float(ms[i]) for i in range(6)}, sent.value

    def rows(self, positions):
        """RowMatrix::row of every column shard, concatenated in global column order, for GLOBAL positions held here."""
        pos = np.ascontiguousarray(positions, dtype=np.uint64)
        w = ELEM_BYTES[self.field] // 8
        out = np.empty((len(pos), self.total_cols * w), dtype=np.uint64)
        check(lib().wg_sharded_rows(self._h, len(pos), C.c_void_p(pos.ctypes.data), C.c_void_p(out.ctypes.data)))
        return out

    def digests(self, which, indexes):
        """which: 0 local leaves, 1 local subtree nodes (heap layout), 2 top nodes (heap layout over the subtree roots)"""
        idx = np.ascontiguousarray(indexes, dtype=np.uint64)
        out = np.empty((len(idx), 32), dtype=np.uint8)
        check(lib().wg_sharded_digests(self._h, which, len(idx), C.c_void_p(idx.ctypes.data), C.c_void_p(out.ctypes.data)))
        return out

    def gather(self, leaves=True, nodes=True):
        """Collective: the reference's full leaves / heap-layout nodes arrays on every rank."""
        lv = np.empty((self.lde_rows, 32), dtype=np.uint8) if leaves else None
        nd = np.empty((self.lde_rows, 32), dtype=np.uint8) if nodes else None
        check(lib().wg_sharded_gather(self._h, C.c_void_p(lv.ctypes.data) if leaves else None, C.c_void_p(nd.ctypes.data) if nodes else None))
        return lv, nd

    def download(self, blocks=True):
        """This rank's row blocks ([row_count, rw_i(, 2)] per received column block), local leaves and subtree nodes."""
        w = ELEM_BYTES[self.field] // 8
        mats = [np.empty((self.row_count, rw * w), dtype=np.uint64) if (blocks and c) else None for c, rw in zip(self.block_cols, self.block_row_widths)]
        ptrs = (C.c_void_p * len(mats))(*[m.ctypes.data if m is not None else None for m in mats])
        lv = np.empty((self.row_count, 32), dtype=np.uint8)
        nd = np.empty((self.row_count, 32), dtype=np.uint8)
        check(lib().wg_sharded_download(self._h, ptrs, C.c_void_p(lv.ctypes.data), C.c_void_p(nd.ctypes.data), None))
        return mats, lv, nd

    def close(self):
        if self._h:
            check(lib().wg_sharded_free(self._h))
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def commit_dev(field, hash_id, n_rows, blowup, total_cols, local_cols, local_polys=None, exchange=EXCHANGE_AUTO, chunks=0, stream=None):
    """local_cols / local_polys: torch int64 CUDA tensors [my_cols, n_rows(, 2)] (this rank's block of the column list)."""
    import torch

    s = torch.cuda.current_stream(local_cols.device) if stream is None else stream
    root = (C.c_uint8 * 32)()
    h = C.c_void_p()
    check(lib().wg_sharded_commit_dev(field, hash_id, n_rows, blowup, total_cols, C.c_void_p(local_cols.data_ptr()),
                                      C.c_void_p(local_polys.data_ptr()) if local_polys is not None else None, exchange, chunks, root,
                                      C.byref(h), C.c_void_p(s.cuda_stream)))
    return ShardedCommitment(h, field, bytes(root))


def commit_host(field, hash_id, n_rows, blowup, total_cols, local_columns, polys_out=None, exchange=EXCHANGE_AUTO, chunks=0):
    """local_columns: list of numpy column arrays (separate allocations, as ColMatrix holds them); polys_out likewise."""
    n = len(local_columns)
    colp = (C.c_void_p * max(n, 1))(*[c.ctypes.data for c in local_columns])
    polp = (C.c_void_p * max(n, 1))(*[c.ctypes.data for c in polys_out]) if polys_out is not None else None
    root = (C.c_uint8 * 32)()
    h = C.c_void_p()
    check(lib().wg_sharded_commit(field, hash_id, n_rows, blowup, total_cols, colp, polp, exchange, chunks, root, C.byref(h)))
    return ShardedCommitment(h, field, bytes(root))


def commit_emulated(field, hash_id, n_rows, blowup, total_cols, world, cols, exchange=EXCHANGE_AUTO, chunks=0, want_leaves=True, want_nodes=True):
    """All `world` ranks' code on ONE device (tests / 1-GPU boxes). cols: torch int64 CUDA tensor [total_cols, n_rows(, 2)].
    -> (root, leaves [L, 32] or None, nodes [L, 32] or None) in the reference's layout."""
    L = n_rows * blowup
    root = (C.c_uint8 * 32)()
    lv = np.empty((L, 32), dtype=np.uint8) if want_leaves else None
    nd = np.empty((L, 32), dtype=np.uint8) if want_nodes else None
    check(lib().wg_sharded_commit_emulated(field, hash_id, n_rows, blowup, total_cols, world, C.c_void_p(cols.data_ptr()), exchange, chunks,
                                           root, C.c_void_p(lv.ctypes.data) if want_leaves else None,
                                           C.c_void_p(nd.ctypes.data) if want_nodes else None))
    return bytes(root), lv, nd
