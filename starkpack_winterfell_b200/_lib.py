"""ctypes binding of libwinter_gpu.so (C ABI: include/winter_gpu.h). The library is the product; there is no
Python or CPU implementation behind these calls — if the shared object or a CUDA device is missing, every
compute call raises."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("WG_LIB_PATH") or os.path.join(_HERE, "libwinter_gpu.so")  # override: kernel-variant experiments

F64, F128 = 0, 1
BLAKE3_256, SHA3_256, BLAKE3_192 = 0, 1, 2
DIGEST_BYTES = {BLAKE3_256: 32, SHA3_256: 32, BLAKE3_192: 24}
ELEM_BYTES = {F64: 8, F128: 16}

WG_OK = 0
ERR_NOT_POW2, ERR_TOO_LARGE, ERR_ZERO_OFFSET, ERR_TOO_FEW, ERR_ARG, ERR_CUDA, ERR_NOMEM, ERR_UNSUPPORTED = range(-1, -9, -1)

# every symbol include/winter_gpu.h declares (tests check the .so exports all of them)
SYMBOLS = [
    "wg_init", "wg_shutdown", "wg_last_error", "wg_digest_bytes", "wg_launch_count", "wg_profile_enable", "wg_profile_read", "wg_debug_field_op", "wg_set_split_log", "wg_set_split_preferred_log", "wg_get_twiddles", "wg_interpolate_poly",
    "wg_interpolate_poly_with_offset", "wg_evaluate_poly", "wg_evaluate_poly_with_offset", "wg_interpolate_columns", "wg_evaluate_polys_over", "wg_hash_rows", "wg_hash_comb_rows", "wg_evaluate_columns_at", "wg_merkle_new",
    "wg_commit_trace", "wg_interpolate_scratch_elems", "wg_lde_scratch_elems", "wg_commit_scratch_elems", "wg_interpolate_columns_dev",
    "wg_interpolate_columns_noalloc_dev", "wg_lde_dev", "wg_hash_rows_dev", "wg_merkle_dev", "wg_commit_trace_dev",
    "wg_commit_polys", "wg_commitment_from_trace", "wg_commitment_from_polys", "wg_commitment_info", "wg_commitment_rows",
    "wg_commitment_digests", "wg_commitment_download", "wg_commitment_free",
    "wg_fri_commit_layer", "wg_fri_apply_drp", "wg_fri_commit_layer_dev", "wg_fri_apply_drp_dev",
    "wg_interpolate_poly_ext", "wg_interpolate_poly_with_offset_ext", "wg_evaluate_poly_ext", "wg_evaluate_poly_with_offset_ext",
    "wg_interpolate_columns_ext", "wg_evaluate_polys_over_ext", "wg_commit_trace_ext",
    "wg_peer_alloc", "wg_peer_open", "wg_peer_close", "wg_peer_free", "wg_lde_scatter_dev",
    "wg_dist_unique_id", "wg_dist_init", "wg_dist_shutdown", "wg_dist_info", "wg_shard_columns", "wg_shard_chunks", "wg_shard_plan",
    "wg_shard_node_owner", "wg_sharded_commit_dev", "wg_sharded_commit", "wg_sharded_info", "wg_sharded_phase_ms", "wg_sharded_rows",
    "wg_sharded_digests", "wg_sharded_gather", "wg_sharded_download", "wg_sharded_free", "wg_sharded_commit_emulated",
]


class WinterGpuError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"winter_gpu error {code}: {msg}")
        self.code = code


_lib = None


def build():
    """Compile the CUDA library in-tree (nvcc, sm_100a)."""
    import subprocess

    subprocess.check_call(["make", "-s", "-j8", "-C", os.path.join(_HERE, "csrc")])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise WinterGpuError(ERR_CUDA, f"{LIB_PATH} not built (run __graft_entry__.build()); there is no fallback path")
        L = C.CDLL(LIB_PATH)
        u64, u32, vp, i32 = C.c_uint64, C.c_uint32, C.c_void_p, C.c_int
        L.wg_last_error.restype = C.c_char_p
        L.wg_launch_count.restype = u64
        L.wg_init.argtypes = [i32]
        L.wg_digest_bytes.argtypes = [i32]
        L.wg_digest_bytes.restype = u32
        L.wg_get_twiddles.argtypes = [i32, u64, i32, vp]
        L.wg_debug_field_op.argtypes = [i32, i32, vp, vp, vp, u32]
        L.wg_set_split_log.argtypes = [i32]
        L.wg_set_split_preferred_log.argtypes = [i32]
        L.wg_interpolate_poly.argtypes = [i32, vp, u64, vp]
        L.wg_evaluate_poly_with_offset.argtypes = [i32, vp, u64, vp, vp, u64, vp]
        L.wg_interpolate_poly_with_offset.argtypes = [i32, vp, u64, vp, vp]
        L.wg_evaluate_poly.argtypes = [i32, vp, u64, vp]
        L.wg_interpolate_columns.argtypes = [i32, vp, u32, u64, vp]
        L.wg_evaluate_polys_over.argtypes = [i32, vp, u32, u64, u64, vp, vp]
        L.wg_hash_rows.argtypes = [i32, i32, vp, u64, u64, u64, vp]
        L.wg_hash_comb_rows.argtypes = [i32, i32, u32, vp, vp, vp, u64, vp]
        L.wg_evaluate_columns_at.argtypes = [i32, vp, u32, u64, vp, vp]
        L.wg_merkle_new.argtypes = [i32, vp, u64, vp]
        L.wg_commit_trace.argtypes = [i32, i32, u64, u64, u32, vp, vp, vp, vp, vp, vp, vp]
        L.wg_interpolate_scratch_elems.restype = u64
        L.wg_interpolate_scratch_elems.argtypes = [i32, u32, u64]
        L.wg_lde_scratch_elems.restype = u64
        L.wg_lde_scratch_elems.argtypes = [i32, u32, u64, u64]
        L.wg_commit_scratch_elems.restype = u64
        L.wg_commit_scratch_elems.argtypes = [i32, u64, u64, u32, vp]
        L.wg_interpolate_columns_dev.argtypes = [i32, vp, u32, u64, vp, vp, u64, vp]
        L.wg_interpolate_columns_noalloc_dev.argtypes = [i32, vp, u32, u64, vp, vp]
        L.wg_lde_dev.argtypes = [i32, vp, u32, u64, u64, vp, vp, u64, vp, u64, vp]
        L.wg_hash_rows_dev.argtypes = [i32, i32, u32, vp, vp, vp, u64, vp, vp]
        L.wg_merkle_dev.argtypes = [i32, vp, u64, vp, vp]
        L.wg_commit_trace_dev.argtypes = [i32, i32, u64, u64, u32, vp, vp, vp, vp, vp, vp, vp, u64, vp]
        L.wg_commit_polys.argtypes = [i32, i32, u64, u64, u32, vp, vp, vp, vp, vp]
        L.wg_commitment_from_trace.argtypes = [i32, i32, u64, u64, u32, vp, vp, vp, vp, vp]
        L.wg_commitment_from_polys.argtypes = [i32, i32, u64, u64, u32, vp, vp, vp]
        L.wg_commitment_info.argtypes = [vp, vp, vp, vp, vp]
        L.wg_commitment_rows.argtypes = [vp, u32, u32, vp, vp]
        L.wg_commitment_digests.argtypes = [vp, i32, u32, vp, vp]
        L.wg_commitment_download.argtypes = [vp, vp, vp, vp]
        L.wg_commitment_free.argtypes = [vp]
        L.wg_fri_commit_layer.argtypes = [i32, i32, vp, u64, u32, vp, vp, vp]
        L.wg_fri_apply_drp.argtypes = [i32, vp, u64, u32, vp, vp, vp]
        L.wg_fri_commit_layer_dev.argtypes = [i32, i32, vp, u64, u32, vp, vp, vp, vp]
        L.wg_fri_apply_drp_dev.argtypes = [i32, vp, u64, u32, vp, vp, vp, vp]
        L.wg_interpolate_poly_ext.argtypes = [i32, u32, vp, u64, vp]
        L.wg_interpolate_poly_with_offset_ext.argtypes = [i32, u32, vp, u64, vp, vp]
        L.wg_evaluate_poly_ext.argtypes = [i32, u32, vp, u64, vp]
        L.wg_evaluate_poly_with_offset_ext.argtypes = [i32, u32, vp, u64, vp, vp, u64, vp]
        L.wg_interpolate_columns_ext.argtypes = [i32, u32, vp, u32, u64, vp]
        L.wg_evaluate_polys_over_ext.argtypes = [i32, u32, vp, u32, u64, u64, vp, vp]
        L.wg_commit_trace_ext.argtypes = [i32, i32, u32, u64, u64, u32, vp, vp, vp, vp, vp, vp, vp]
        L.wg_peer_alloc.argtypes = [u64, vp, vp]
        L.wg_peer_open.argtypes = [vp, vp]
        L.wg_peer_close.argtypes = [vp]
        L.wg_peer_free.argtypes = [vp]
        L.wg_lde_scatter_dev.argtypes = [i32, vp, u32, u64, u64, vp, u32, vp, u64, vp, u64, vp]
        L.wg_dist_unique_id.argtypes = [vp]
        L.wg_dist_init.argtypes = [i32, i32, vp]
        L.wg_dist_info.argtypes = [vp, vp]
        L.wg_shard_columns.argtypes = [u32, u32, u32, vp, vp]
        L.wg_shard_chunks.argtypes = [u32, u32, u32, vp, vp]
        L.wg_shard_plan.argtypes = [u32, u32, i32, u32, vp, vp]
        L.wg_shard_node_owner.argtypes = [u64, u32, u64, vp, vp]
        L.wg_sharded_commit_dev.argtypes = [i32, i32, u64, u64, u32, vp, vp, i32, u32, vp, vp, vp]
        L.wg_sharded_commit.argtypes = [i32, i32, u64, u64, u32, vp, vp, i32, u32, vp, vp]
        L.wg_sharded_info.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, vp]
        L.wg_sharded_phase_ms.argtypes = [vp, vp, vp]
        L.wg_sharded_rows.argtypes = [vp, u32, vp, vp]
        L.wg_sharded_digests.argtypes = [vp, i32, u32, vp, vp]
        L.wg_sharded_gather.argtypes = [vp, vp, vp]
        L.wg_sharded_download.argtypes = [vp, vp, vp, vp, vp]
        L.wg_sharded_free.argtypes = [vp]
        L.wg_sharded_commit_emulated.argtypes = [i32, i32, u64, u64, u32, u32, vp, i32, u32, vp, vp, vp]
        _lib = L
    return _lib


def check(rc):
    if rc != WG_OK:
        raise WinterGpuError(rc, lib().wg_last_error().decode())


def init(device=0):
    check(lib().wg_init(device))


EXCHANGE_AUTO, EXCHANGE_ALLTOALL, EXCHANGE_PIPELINED, EXCHANGE_FUSED_P2P, EXCHANGE_P2P_COPY = 0, 1, 2, 3, 4
EXCHANGE_NAMES = {0: "auto", 1: "alltoall", 2: "pipelined", 3: "fused_p2p", 4: "p2p_copy"}
DIST_ID_BYTES = 128
SHARD_MAX_BLOCKS = 64

PHASES = ["interp_strided", "interp_contig", "lde_strided", "lde_contig", "hash_rows", "merkle"]


def profile_enable(on=True):
    check(lib().wg_profile_enable(1 if on else 0))


def profile_read():
    """-> {phase: (ms_total, n_records)} since the last read (synchronises)."""
    ms = (C.c_float * 6)()
    cnt = (C.c_uint32 * 6)()
    check(lib().wg_profile_read(ms, cnt))
    return {PHASES[i]: (float(ms[i]), int(cnt[i])) for i in range(6)}


def launch_count():
    return int(lib().wg_launch_count())
