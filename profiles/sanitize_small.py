"""Small end-to-end run for compute-sanitizer (memcheck / racecheck): every kernel family at tiny sizes, checked against
the oracle. usage: compute-sanitizer --tool racecheck python profiles/sanitize_small.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import oracle
import starkpack_winterfell_b200 as w
from tests import fixtures as fx

oracle.build()
w._lib.init(0)
cases = [(w.F64, w.BLAKE3_256, [10], 1 << 13, 8), (w.F64, w.SHA3_256, [3, 2], 64, 4), (w.F128, w.BLAKE3_256, [3], 1 << 13, 2),
         (w.F128, w.SHA3_256, [2, 2], 256, 8), (w.F64, w.BLAKE3_256, [140], 1 << 13, 2)]
for field, h, shapes, n, B in cases:
    traces = [fx.synth_f64(3 + i, c, n) if field == w.F64 else fx.synth_f128(3 + i, c, n) for i, c in enumerate(shapes)]
    want = oracle.commit_trace(field, h, traces, B, want=("nodes",))
    dom = w.matrix.StarkDomain(field, n, B)
    cms = [w.matrix.ColMatrix(field, [np.ascontiguousarray(c) for c in t]) for t in traces]
    ldes, tree, polys = w.prover.build_trace_commitment(h, cms, dom)
    assert tree.root() == bytes(want["root"]), (field, h, shapes, n)
    rc = w.prover.ResidentCommitment.from_traces(h, cms, dom)
    q = rc.query([1, 5, n * B - 1])
    assert rc.root() == tree.root() and q.trace_proof == tree.prove_batch([1, 5, n * B - 1])
    rc.close()
print("SANITIZE_OK")
