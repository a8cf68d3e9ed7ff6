"""CPU-side checks of the boundary: the library loads, exports every symbol include/winter_gpu.h declares,
and refuses to compute without a CUDA device (no fallback)."""
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def wg():
    import starkpack_winterfell_b200 as w

    if not os.path.exists(w._lib.LIB_PATH):
        w._lib.build()
    return w


def test_header_symbols_exported(wg):
    hdr = open(os.path.join(ROOT, "include", "winter_gpu.h")).read()
    declared = set(re.findall(r"\b(wg_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(wg._lib.SYMBOLS)
    L = wg._lib.lib()
    for s in declared:
        assert hasattr(L, s), s


def test_no_cpu_fallback(wg):
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(wg.WinterGpuError) as e:
        wg.fft.get_twiddles(wg.F64, 8)
    assert e.value.code == wg._lib.ERR_CUDA
    with pytest.raises(wg.WinterGpuError):
        wg.merkle.MerkleTree.new(wg.BLAKE3_256, np.zeros((4, 32), dtype=np.uint8))


def test_product_does_not_touch_oracle():
    pkg = os.path.join(ROOT, "starkpack_winterfell_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")) or f == "Makefile":
                src = open(os.path.join(dp, f)).read()
                assert "import oracle" not in src and "from oracle" not in src and "liboracle" not in src, f


def test_merkle_argument_errors(wg):
    from starkpack_winterfell_b200.merkle import MerkleTree, MerkleTreeError

    with pytest.raises(MerkleTreeError):  # crypto/src/merkle/mod.rs:118-120
        MerkleTree.new(wg.BLAKE3_256, np.zeros((1, 32), dtype=np.uint8))
    with pytest.raises(MerkleTreeError):  # :121-123
        MerkleTree.new(wg.BLAKE3_256, np.zeros((6, 32), dtype=np.uint8))


def test_shard_partition_helpers_host_only():
    """wg_shard_* are host-only (no GPU): the column partition, the chunking and the node-ownership map the sharded
    commit uses must agree with the Python mirror the gloo tests exercise (starkpack_winterfell_b200/parallel.py)."""
    from starkpack_winterfell_b200 import _lib, sharded
    from starkpack_winterfell_b200.parallel import split_chunks, split_columns

    for total, world in [(100, 1), (100, 2), (100, 4), (100, 8), (256, 8), (16, 8), (21, 4), (9, 8)]:
        assert [sharded.shard_columns(total, world, r) for r in range(world)] == split_columns(total, world)
    for cols, k in [(50, 2), (13, 2), (128, 4), (25, 4), (8, 1), (7, 3), (100, 8)]:
        got = sharded.shard_chunks(cols, k)
        assert got == split_chunks(cols, k)
        assert sum(c for _, c in got) == cols and all(c % 8 == 0 for _, c in got[:-1] if c and got[-1][1])
    # auto plan: fused for narrow shards on <= 8 ranks, pipelined otherwise, never more than 64 received blocks
    assert sharded.shard_plan(16, 8) == (_lib.EXCHANGE_FUSED_P2P, 1)
    assert sharded.shard_plan(100, 2) == (_lib.EXCHANGE_P2P_COPY, 4)
    assert sharded.shard_plan(256, 4) == (_lib.EXCHANGE_P2P_COPY, 4)  # 8 inside a commit of >= 2^26 LDE rows
    assert sharded.shard_plan(100, 8) == (_lib.EXCHANGE_P2P_COPY, 2)
    assert sharded.shard_plan(256, 8) == (_lib.EXCHANGE_P2P_COPY, 4)
    assert sharded.shard_plan(100, 8, _lib.EXCHANGE_PIPELINED) == (_lib.EXCHANGE_PIPELINED, 2)
    assert sharded.shard_plan(9, 1, _lib.EXCHANGE_P2P_COPY) == (_lib.EXCHANGE_ALLTOALL, 1)
    assert sharded.shard_plan(100, 1) == (_lib.EXCHANGE_ALLTOALL, 1)
    ex, k = sharded.shard_plan(4096, 32, _lib.EXCHANGE_PIPELINED, 8)
    assert 32 * k <= _lib.SHARD_MAX_BLOCKS
    # node ownership reproduces the stitching of the heap layout (crypto/src/merkle/mod.rs:350-374)
    L, G = 64, 4
    rows = L // G
    for i in range(1, L):
        r, li = sharded.node_owner(L, G, i)
        if i < G:
            assert (r, li) == (-1, i)
        else:
            m = 1 << (i.bit_length() - 1)
            per = m // G
            assert r == (i - m) // per and li == per + (i - m) % per and 1 <= li < rows


def test_scratch_queries_and_split_thresholds_host_only():
    """The scratch-size queries and the split-plan thresholds are host-only: the sizes a caller allocates from must match
    the plan the library will run (two-pass up to 2^24; split plan preferred for the LDE of 2^24-row traces and required
    above; several narrow traces extended together)."""
    import ctypes as C

    from starkpack_winterfell_b200 import _lib

    L = _lib.lib()
    F64, F128 = 0, 1
    assert L.wg_set_split_log(0) == 24 and L.wg_set_split_preferred_log(0) == 22
    assert L.wg_set_split_log(12) == _lib.ERR_ARG and L.wg_set_split_preferred_log(25) == _lib.ERR_ARG
    n22, n23, n24, n25 = 1 << 22, 1 << 23, 1 << 24, 1 << 25
    # one on-chip pass needs no scratch; two passes one intermediate of the output's size
    assert L.wg_interpolate_scratch_elems(F64, 7, 4096) == 0 and L.wg_lde_scratch_elems(F64, 7, 4096, 8) == 0
    assert L.wg_interpolate_scratch_elems(F64, 100, n22) == 100 * n22
    assert L.wg_lde_scratch_elems(F64, 100, n22, 8) == 100 * n22 * 8
    # 2^23 rows keep the two-pass plan (radix 2 loses), 2^24 rows prefer radix 4 around 2^22: (2B + 1) n per column
    assert L.wg_lde_scratch_elems(F64, 32, n23, 8) == 32 * n23 * 8
    assert L.wg_lde_scratch_elems(F64, 32, n24, 8) == 32 * n24 * 17
    assert L.wg_interpolate_scratch_elems(F64, 32, n24) == 32 * n24  # the interpolation only splits above 2^24
    assert L.wg_interpolate_scratch_elems(F64, 2, n25) == 2 * 2 * n25 and L.wg_lde_scratch_elems(F64, 2, n25, 2) == 2 * n25 * 5
    assert L.wg_set_split_preferred_log(24) == 22
    try:
        assert L.wg_lde_scratch_elems(F64, 32, n24, 8) == 32 * n24 * 8  # preference off
    finally:
        assert L.wg_set_split_preferred_log(22) == 24
    # whole commits: 8 traces of 2 columns are extended together (combined [L][16] matrix + its LDE scratch), one wide
    # trace is not
    n20 = 1 << 20
    cpt = (C.c_uint32 * 8)(*([2] * 8))
    assert L.wg_commit_scratch_elems(F128, n20, 8, 8, cpt) == n20 * 8 * 16 + 16 * n20 * 8
    one = (C.c_uint32 * 1)(100)
    assert L.wg_commit_scratch_elems(F64, n22, 8, 1, one) == 100 * n22 * 8
    assert L.wg_commit_scratch_elems(F64, n22, 8, 0, one) == 0 and L.wg_commit_scratch_elems(F64, n22 + 1, 8, 1, one) == 0
