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
