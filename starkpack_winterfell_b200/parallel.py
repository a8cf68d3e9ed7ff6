"""One Merkle tree over an LDE that is sharded by columns across G GPUs (BASELINE configs[3], [4]).

The reference has no distributed code; its `commit_to_comb_rows` (prover/src/matrix/row_matrix.rs:204-238) needs whole
rows (all columns of all aggregated traces) to form a leaf. Interpolation and LDE are independent per column, so:

  1. every rank interpolates + LDEs ITS columns (contiguous block of the global column list: trace 0 col 0, ...,
     trace T-1 col last) into a local row-major matrix [L, rw_g];
  2. ONE all-to-all (torch.distributed, NCCL over NVLink) turns column shards into row blocks: rank g receives rows
     [g*L/G, (g+1)*L/G) of every shard — a contiguous slice on the sender;
  3. rank g hashes its rows (the leaf kernel concatenates the G received matrices, skipping pad columns, so the bytes
     hashed are exactly the reference's) and builds the subtree over its L/G leaves;
  4. the G subtree roots are all-gathered and the top log2(G) levels are computed (redundantly on every rank; the
     result on rank 0 is the commitment). `gather_nodes()` stitches the reference's heap-layout node array.

`DistributedCommitter(..., fused=True)` replaces steps 1b + 2 by ONE kernel: the LDE's last pass stores every row straight
into its owner's receive buffer through peer-mapped pointers (CUDA IPC, NVLink P2P stores; wg_lde_scatter_dev), so the
transfer overlaps the butterflies tile by tile, no send buffer exists and NCCL only carries a 4-byte barrier and the
32-byte roots. The receive buffer has the all-to-all's layout, so steps 3-4 are unchanged and the results are identical.

One process per GPU. All compute goes through a backend with four methods; the product backend is the C ABI
(`DeviceBackend`); tests/ substitutes a CPU checker backend to exercise the partition / exchange / stitching logic under
gloo without a GPU.
"""
import ctypes as C

import torch
import torch.distributed as dist

from . import _lib
from ._lib import ELEM_BYTES, check, lib


def split_columns(total_cols, world):
    """contiguous, as even as possible: first (total % world) ranks get one more. -> list of (start, count)."""
    base, extra = divmod(total_cols, world)
    out, s = [], 0
    for g in range(world):
        c = base + (1 if g < extra else 0)
        out.append((s, c))
        s += c
    return out


def split_chunks(cols, k):
    """cut `cols` columns into k chunks whose sizes are multiples of 8 except possibly the last non-empty one (the LDE
    pads every matrix to a multiple of 8 columns and pad columns are not hashed). -> list of k (start, count)."""
    per = ((cols + k - 1) // k + 7) // 8 * 8
    out, s = [], 0
    for _ in range(k):
        c = min(per, cols - s)
        out.append((s, max(c, 0)))
        s += max(c, 0)
    return out


def stitch_nodes(local_nodes, top_nodes, world):
    """local_nodes[g]: [L/G, 32] heap layout of rank g's subtree; top_nodes: [G, 32] heap layout of the tree over the G
    subtree roots. Returns the global [L, 32] heap-layout array of crypto/src/merkle/mod.rs:350-374."""
    lg = local_nodes[0].shape[0]
    L = lg * world
    out = torch.zeros((L, 32), dtype=torch.uint8)
    out[1:world] = top_nodes[1:world].cpu() if world > 1 else out[1:world]
    m = world
    while m < L:
        per = m // world  # nodes of this level owned by each rank: local level `per`
        for g in range(world):
            out[m + g * per: m + (g + 1) * per] = local_nodes[g][per: 2 * per].cpu()
        m *= 2
    return out


class DeviceBackend:
    """Compute through libwinter_gpu.so on the current CUDA device."""

    def __init__(self, field, hash_id, device):
        self.field, self.hash_id, self.device = field, hash_id, torch.device(device)
        _lib.init(self.device.index or 0)
        self.w = ELEM_BYTES[field] // 8
        from .matrix import GENERATOR_INTERNAL

        self._offset = GENERATOR_INTERNAL[field]

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def empty(self, shape, dtype=torch.int64):
        return torch.empty(shape, dtype=dtype, device=self.device)

    def interpolate(self, cols, n):
        c = cols.shape[0]
        polys = torch.empty_like(cols)
        se = int(lib().wg_interpolate_scratch_elems(self.field, c, n))
        scratch = self.empty((max(se, 1), self.w))
        check(lib().wg_interpolate_columns_dev(self.field, C.c_void_p(cols.data_ptr()), c, n, C.c_void_p(polys.data_ptr()),
                                               C.c_void_p(scratch.data_ptr()), se, self._stream()))
        return polys

    def lde(self, polys, n, blowup, out=None, scratch=None):
        """polys [c, n, w] -> row-major LDE [n*blowup, 8*ceil(c/8), w] (into `out` when given)"""
        c = polys.shape[0]
        rw = (c + 7) // 8 * 8
        lde = self.empty((n * blowup, rw, self.w)) if out is None else out
        se = int(lib().wg_lde_scratch_elems(self.field, c, n, blowup))
        if scratch is None:
            scratch = self.empty((max(se, 1), self.w))
        check(lib().wg_lde_dev(self.field, C.c_void_p(polys.data_ptr()), c, n, blowup, C.c_void_p(self._offset.ctypes.data),
                               C.c_void_p(lde.data_ptr()), rw, C.c_void_p(scratch.data_ptr()), se, self._stream()))
        return lde

    def interpolate_lde(self, cols, n, blowup):
        """cols: [c, n, w] int64 -> (polys [c, n, w], lde [n*blowup, rw, w])"""
        c = cols.shape[0]
        rw = (c + 7) // 8 * 8
        polys = torch.empty_like(cols)
        lde = self.empty((n * blowup, rw, self.w))
        se = max(lib().wg_interpolate_scratch_elems(self.field, c, n), lib().wg_lde_scratch_elems(self.field, c, n, blowup))
        scratch = self.empty((max(int(se), 1), self.w))
        check(lib().wg_interpolate_columns_dev(self.field, C.c_void_p(cols.data_ptr()), c, n, C.c_void_p(polys.data_ptr()),
                                               C.c_void_p(scratch.data_ptr()), int(se), self._stream()))
        check(lib().wg_lde_dev(self.field, C.c_void_p(polys.data_ptr()), c, n, blowup, C.c_void_p(self._offset.ctypes.data),
                               C.c_void_p(lde.data_ptr()), rw, C.c_void_p(scratch.data_ptr()), int(se), self._stream()))
        return polys, lde

    # ---- fused LDE + exchange ------------------------------------------------------------------------------------
    def peer_alloc(self, nbytes):
        """-> (device pointer, 64-byte IPC handle) of a fresh allocation that other ranks can map."""
        ptr = C.c_void_p()
        handle = (C.c_uint8 * 64)()
        check(lib().wg_peer_alloc(int(nbytes), C.byref(ptr), handle))
        return ptr.value, bytes(handle)

    def peer_open(self, handle):
        ptr = C.c_void_p()
        check(lib().wg_peer_open((C.c_uint8 * 64).from_buffer_copy(handle), C.byref(ptr)))
        return ptr.value

    def peer_close(self, ptr):
        check(lib().wg_peer_close(C.c_void_p(ptr)))

    def peer_free(self, ptr):
        check(lib().wg_peer_free(C.c_void_p(ptr)))

    def wrap(self, ptr, shape, dtype=torch.int64):
        """torch view of raw device memory (no copy, no ownership)."""
        n = 1
        for d in shape:
            n *= d

        class _Raw:
            __cuda_array_interface__ = {"shape": (n,), "typestr": "<i8" if dtype == torch.int64 else "|u1", "data": (ptr, False), "version": 2}

        return torch.as_tensor(_Raw(), device=self.device).view(shape)

    def interpolate_lde_scatter(self, cols, n, blowup, row_block_ptrs):
        """cols [c, n, w] -> polys; LDE rows go to row_block_ptrs[g] (owner g's (L/G) x rw matrix for this rank)."""
        c = cols.shape[0]
        rw = (c + 7) // 8 * 8
        polys = torch.empty_like(cols)
        se = max(lib().wg_interpolate_scratch_elems(self.field, c, n), lib().wg_lde_scratch_elems(self.field, c, n, blowup))
        scratch = self.empty((max(int(se), 1), self.w))
        check(lib().wg_interpolate_columns_dev(self.field, C.c_void_p(cols.data_ptr()), c, n, C.c_void_p(polys.data_ptr()),
                                               C.c_void_p(scratch.data_ptr()), int(se), self._stream()))
        ptrs = (C.c_void_p * len(row_block_ptrs))(*row_block_ptrs)
        check(lib().wg_lde_scatter_dev(self.field, C.c_void_p(polys.data_ptr()), c, n, blowup, C.c_void_p(self._offset.ctypes.data),
                                       len(row_block_ptrs), ptrs, rw, C.c_void_p(scratch.data_ptr()), int(se), self._stream()))
        return polys

    def hash_rows(self, mats, elems, rows):
        """mats: list of [rows, rw_i, w] int64 row blocks; elems[i] real columns of block i -> leaves [rows, 32]"""
        k = len(mats)
        ptrs = (C.c_void_p * k)(*[m.data_ptr() for m in mats])
        rws = (C.c_uint64 * k)(*[m.shape[1] for m in mats])
        els = (C.c_uint64 * k)(*elems)
        leaves = self.empty((rows, 32), torch.uint8)
        check(lib().wg_hash_rows_dev(self.field, self.hash_id, k, ptrs, rws, els, rows, C.c_void_p(leaves.data_ptr()), self._stream()))
        return leaves

    def merkle(self, leaves):
        nodes = self.empty((leaves.shape[0], 32), torch.uint8)
        check(lib().wg_merkle_dev(self.hash_id, C.c_void_p(leaves.data_ptr()), leaves.shape[0], C.c_void_p(nodes.data_ptr()), self._stream()))
        return nodes


class DistributedCommitter:
    def __init__(self, backend, n_rows, blowup, total_cols, group=None, fused=False, pipeline=0):
        """fused: the LDE stores rows into the owners' buffers itself (P2P). pipeline = K > 1: the shard is LDE'd in K
        column chunks and chunk k's all-to-all (NCCL, its own stream) runs under chunk k+1's LDE."""
        self.b, self.n, self.blowup, self.total_cols, self.group = backend, n_rows, blowup, total_cols, group
        self.fused = fused
        self.pipeline = pipeline
        self._recv_ptr = None
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        assert self.world & (self.world - 1) == 0, "number of ranks must be a power of two"
        # fail before any work: the leaf kernel concatenates at most SHARD_MAX_BLOCKS matrices, the fused LDE has 8 peer slots
        assert self.world * max(pipeline, 1) <= _lib.SHARD_MAX_BLOCKS, "ranks x pipeline chunks exceeds the leaf kernel's matrix limit"
        assert not fused or self.world <= 8, "the fused exchange supports at most 8 ranks"
        self.L = n_rows * blowup
        assert self.L % self.world == 0 and self.L // self.world >= 2
        self.shards = split_columns(total_cols, self.world)
        assert all(c > 0 for _, c in self.shards), "more ranks than columns"
        self.col_start, self.col_count = self.shards[self.rank]

    @classmethod
    def auto(cls, backend, n_rows, blowup, total_cols, group=None):
        """Picks the form of the exchange from the shard shape, following the 2-GPU measurements in DESIGN.md: narrow
        shards (a StarkPack trace per GPU: <= 16 columns per rank) finish their LDE pass in a few milliseconds, where
        the fused P2P stores save the whole all-to-all; wide shards (>= 64 columns) would make that pass NVLink-bound
        at 32-byte stores, so they pipeline NCCL all-to-alls of column chunks under the LDE instead. A backend without
        peer memory (the CPU checker of the tests) gets the plain all-to-all."""
        world = dist.get_world_size(group)
        per_rank = (total_cols + world - 1) // world
        can_fuse = hasattr(backend, "peer_alloc") and world <= 8
        if world > 1 and per_rank <= 16 and can_fuse:
            return cls(backend, n_rows, blowup, total_cols, group, fused=True)
        if world > 1 and per_rank >= 64 and hasattr(backend, "lde"):
            k = 4
            while world * k > _lib.SHARD_MAX_BLOCKS and k > 1:  # the leaf kernel's matrix limit
                k //= 2
            return cls(backend, n_rows, blowup, total_cols, group, pipeline=k if k > 1 else 0)
        return cls(backend, n_rows, blowup, total_cols, group)

    def commit(self, local_cols):
        """local_cols: [col_count, n, w] — this rank's block of the global column list. Returns the 32-byte root
        (identical on every rank); keeps polys / lde shard / leaves / local nodes / top nodes as attributes."""
        G, L = self.world, self.L
        rows = L // G
        assert local_cols.shape[0] == self.col_count
        rws = [(c + 7) // 8 * 8 for _, c in self.shards]
        if self.fused and G > 1:
            w = local_cols.shape[2]
            recv = self._fused_setup(rows, rws, w)
            # owner g's slot for MY columns: after the matrices of ranks 0..rank-1
            off = sum(rows * r * w for r in rws[: self.rank]) * 8
            self.polys = self.b.interpolate_lde_scatter(local_cols, self.n, self.blowup, [p + off for p in self._peer_ptrs])
            self.lde = None
            # every rank's rows have landed once every rank's LDE kernel has finished: a stream-ordered barrier
            dist.all_reduce(self._token, group=self.group)
        elif self.pipeline > 1 and G > 1:
            return self._commit_pipelined(local_cols)
        else:
            self.polys, self.lde = self.b.interpolate_lde(local_cols, self.n, self.blowup)
            w = self.lde.shape[2]
            # --- the one exchange: column shards -> row blocks
            send = self.lde.reshape(-1)
            recv = self.b.empty((sum(rows * r * w for r in rws),))
            in_splits = [rows * rws[self.rank] * w] * G
            out_splits = [rows * r * w for r in rws]
            if G > 1:
                dist.all_to_all_single(recv, send, out_splits, in_splits, group=self.group)
            else:
                recv.copy_(send)
        mats, off = [], 0
        for r in rws:
            mats.append(recv[off: off + rows * r * w].view(rows, r, w))
            off += rows * r * w
        self.row_block = mats
        # --- leaves of my rows over ALL columns, my subtree
        self.leaves = self.b.hash_rows(mats, [c for _, c in self.shards], rows)
        self.local_nodes = self.b.merkle(self.leaves)
        # --- top log2(G) levels over the G subtree roots
        my_root = self.local_nodes[1].contiguous()
        if G > 1:
            roots = [torch.empty_like(my_root) for _ in range(G)]
            dist.all_gather(roots, my_root, group=self.group)
            self.top_nodes = self.b.merkle(torch.stack(roots).contiguous())
            return bytes(self.top_nodes[1].cpu().numpy())
        self.top_nodes = self.local_nodes[:1]
        return bytes(my_root.cpu().numpy())

    def _finish(self, mats, elems, rows):
        G = self.world
        self.row_block = mats
        self.leaves = self.b.hash_rows(mats, elems, rows)
        self.local_nodes = self.b.merkle(self.leaves)
        my_root = self.local_nodes[1].contiguous()
        roots = [torch.empty_like(my_root) for _ in range(G)]
        dist.all_gather(roots, my_root, group=self.group)
        self.top_nodes = self.b.merkle(torch.stack(roots).contiguous())
        return bytes(self.top_nodes[1].cpu().numpy())

    def _commit_pipelined(self, local_cols):
        """Column chunks: LDE of chunk k+1 overlaps the all-to-all of chunk k. Every rank cuts its shard into the same
        number K of chunks (multiples of 8 columns except the last), so the K collectives pair up; the leaf kernel reads
        the G*K received matrices in global column order (rank-major, then chunk)."""
        G, L, K = self.world, self.L, self.pipeline
        rows = L // G
        w = local_cols.shape[2]
        chunks = [split_chunks(c, K) for _, c in self.shards]  # per rank: list of (start, count), count may be 0
        self.polys = self.b.interpolate(local_cols, self.n)
        self.lde = None
        on_gpu = local_cols.is_cuda
        if on_gpu:
            comm = torch.cuda.Stream(device=local_cols.device)
            cur = torch.cuda.current_stream(local_cols.device)
        recvs, keep = {}, []
        for k in range(K):
            if all(chunks[g][k][1] == 0 for g in range(G)):
                continue
            s0, cnt = chunks[self.rank][k]
            rws_k = [(chunks[g][k][1] + 7) // 8 * 8 for g in range(G)]
            send = self.b.lde(self.polys[s0: s0 + cnt], self.n, self.blowup).reshape(-1) if cnt else self.b.empty((0,))
            recv = self.b.empty((sum(rows * r * w for r in rws_k),))
            splits = ([rows * r * w for r in rws_k], [rows * rws_k[self.rank] * w] * G)
            if on_gpu:
                ev = torch.cuda.Event()
                ev.record(cur)
                with torch.cuda.stream(comm):
                    comm.wait_event(ev)
                    dist.all_to_all_single(recv, send, splits[0], splits[1], group=self.group)
                    send.record_stream(comm)
                    recv.record_stream(comm)
            else:
                dist.all_to_all_single(recv, send, splits[0], splits[1], group=self.group)
            keep.append(send)
            recvs[k] = (recv, rws_k)
        if on_gpu:
            cur.wait_stream(comm)
        mats, elems = [], []
        for g in range(G):
            for k in range(K):
                if chunks[g][k][1] == 0:
                    continue
                recv, rws_k = recvs[k]
                off = sum(rows * r * w for r in rws_k[:g])
                mats.append(recv[off: off + rows * rws_k[g] * w].view(rows, rws_k[g], w))
                elems.append(chunks[g][k][1])
        return self._finish(mats, elems, rows)

    def _fused_setup(self, rows, rws, w):
        """receive buffer [sum_r rows*rws[r]*w] int64 in its own allocation, mapped into every rank (done once)."""
        if self._recv_ptr is None:
            G = self.world
            self._recv_elems = sum(rows * r * w for r in rws)
            self._recv_ptr, handle = self.b.peer_alloc(self._recv_elems * 8)
            mine = torch.tensor(list(handle), dtype=torch.uint8, device=self.b.device)
            allh = [torch.empty_like(mine) for _ in range(G)]
            dist.all_gather(allh, mine, group=self.group)
            self._peer_ptrs = []
            for g in range(G):
                self._peer_ptrs.append(self._recv_ptr if g == self.rank else self.b.peer_open(bytes(allh[g].cpu().numpy().tobytes())))
            self._token = torch.zeros(1, dtype=torch.int32, device=self.b.device)
            self._recv = self.b.wrap(self._recv_ptr, (self._recv_elems,))
        return self._recv

    def close(self):
        """unmap the peers' buffers and free this rank's (all ranks first stop using them: collective)."""
        if self._recv_ptr is not None:
            torch.cuda.synchronize(self.b.device)
            dist.barrier(group=self.group)
            for g, p in enumerate(self._peer_ptrs):
                if g != self.rank:
                    self.b.peer_close(p)
            dist.barrier(group=self.group)
            self.b.peer_free(self._recv_ptr)
            self._recv_ptr = None

    def gather_nodes(self):
        """Global heap-layout node array on every rank (verification / handing the tree to a CPU prover)."""
        G = self.world
        if G == 1:
            return self.local_nodes.cpu()
        parts = [torch.empty_like(self.local_nodes) for _ in range(G)]
        dist.all_gather(parts, self.local_nodes.contiguous(), group=self.group)
        return stitch_nodes(parts, self.top_nodes, G)

    def gather_leaves(self):
        G = self.world
        if G == 1:
            return self.leaves.cpu()
        parts = [torch.empty_like(self.leaves) for _ in range(G)]
        dist.all_gather(parts, self.leaves.contiguous(), group=self.group)
        return torch.cat([p.cpu() for p in parts])
