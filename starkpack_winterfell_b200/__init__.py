"""starkpack_winterfell_b200 — B200-native trace commitment (interpolate -> coset LDE -> row hash -> Merkle)
behind the reference's own surface. The product is csrc/ (CUDA kernels + C ABI, include/winter_gpu.h); these
modules are the thin host-side mirror of the reference interface used by the tests and the bench."""
from . import _lib, fft, fri, matrix, merkle, prover  # noqa: F401
from ._lib import BLAKE3_256, F64, F128, SHA3_256, WinterGpuError  # noqa: F401
