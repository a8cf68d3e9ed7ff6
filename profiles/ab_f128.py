"""A/B timing of library variants on the f128 configs (cfg1: 2^20 x 2 Blake3, cfg3: 2^20 x 4 Sha3, cfg4: 8 traces of 2^20 x 2,
one GPU, device-resident, CUDA events, best of 10). One process per variant (WG_LIB_PATH).
usage: python profiles/ab_f128.py name [name ...]   (names of starkpack_winterfell_b200/variants/libwinter_gpu_<name>.so)"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def child():
    sys.path.insert(0, ROOT)
    import numpy as np
    import torch

    import starkpack_winterfell_b200 as w
    from starkpack_winterfell_b200 import _lib
    from starkpack_winterfell_b200.device import TraceCommitter
    from tests import fixtures as fx

    n = 1 << 20
    out = {"lib": os.path.basename(_lib.LIB_PATH)}
    for name, hash_id, cpt, seed in (("cfg1", w.BLAKE3_256, [2], 1), ("cfg3", w.SHA3_256, [4], 3), ("cfg4", w.BLAKE3_256, [2] * 8, 4)):
        tc = TraceCommitter(w.F128, hash_id, n, 8, cpt, device="cuda:0")
        a = np.ascontiguousarray(fx.synth_f128(seed, sum(cpt), n)).view(np.int64)
        tc.cols.copy_(torch.from_numpy(a))
        for _ in range(3):
            tc.commit()
        torch.cuda.synchronize()
        ts = []
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            tc.commit()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        out[name + "_ms"] = round(min(ts), 3)
        out[name + "_root"] = tc.root().hex()[:16]
        del tc
        torch.cuda.empty_cache()
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "--child":
        child()
    else:
        for name in sys.argv[1:]:
            lib = os.path.join(ROOT, "starkpack_winterfell_b200", "variants", f"libwinter_gpu_{name}.so")
            env = dict(os.environ) if name == "main" else dict(os.environ, WG_LIB_PATH=lib)  # "main": the product library
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--child"], env=env, capture_output=True, text=True)
            line = [x for x in r.stdout.splitlines() if x.startswith("{")]
            print(line[-1] if line else json.dumps({"lib": name, "error": (r.stderr or r.stdout)[-400:]}), flush=True)
