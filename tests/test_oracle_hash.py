"""Pins oracle.c BLAKE3 / SHA3-256 to the independent Python `blake3` module, hashlib and spec KATs."""
import hashlib
import random

import blake3
import pytest


def test_empty_input_kats(oracle):
    assert oracle.hash_bytes(0, b"").hex() == "af1349b9f5f9a1a6a0404dea36dcc9499bcb25c9adc112b7cc9a93cae41f3262"
    assert oracle.hash_bytes(1, b"").hex() == "a7ffc6f8bf1ed76651c14756a061d662f580ff4de43b49fa82d80a4b80f8434a"
    # BLAKE3 spec test-vector input pattern (i % 251), lengths around block/chunk boundaries
    for n in (1, 63, 64, 65, 1023, 1024, 1025, 2048, 2049, 3072, 3073, 4096, 4097, 5120, 8192):
        data = bytes(i % 251 for i in range(n))
        assert oracle.hash_bytes(0, data) == blake3.blake3(data).digest(), n


@pytest.mark.parametrize("hash_id", [0, 1, 2])
def test_all_lengths(oracle, hash_id):
    """every length 0..300, random lengths up to 9,300 bytes (nine BLAKE3 chunks: every tree shape up to 8 + 1 chunks, so
    the 5,600-byte rows of tests/test_gpu_parity.py::test_hash_rows_block_and_chunk_shapes are inside the pinned range),
    and the block / chunk / rate boundaries. hash 2 = Blake3_192 = the first 24 bytes (blake/mod.rs:71-83)."""
    rnd = random.Random(7)
    lens = list(range(0, 300)) + [rnd.randrange(300, 9300) for _ in range(250)]
    lens += [800, 832, 2048, 135, 136, 137, 272, 4096, 4097, 5119, 5120, 5121, 5600, 6144, 7168, 8191, 8192, 8193, 9216]
    for n in lens:
        data = rnd.randbytes(n)
        ref = hashlib.sha3_256(data).digest() if hash_id == 1 else blake3.blake3(data).digest()[: 24 if hash_id == 2 else 32]
        assert oracle.hash_bytes(hash_id, data) == ref, (hash_id, n)
