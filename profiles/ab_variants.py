"""A/B timing of kernel variants of the product library on the headline shape (2^22 x 100 f64, blowup 8, Blake3_256).
Each variant is a separately compiled libwinter_gpu_<name>.so (profiles/build_variant.sh) loaded through WG_LIB_PATH in
its own process; prints one JSON line per variant with the CUDA-event time per commit, the per-phase split and the root
(which must not change).   usage: python profiles/ab_variants.py [--log-rows 22] [--cols 100] name [name ...]"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def child(log_rows, cols):
    sys.path.insert(0, ROOT)
    import torch

    import bench
    import starkpack_winterfell_b200 as w
    from starkpack_winterfell_b200 import _lib
    from starkpack_winterfell_b200.device import TraceCommitter

    n = 1 << log_rows
    tc = TraceCommitter(w.F64, w.BLAKE3_256, n, 8, [cols], device="cuda:0")
    tc.cols.copy_(bench.synth_trace_device(torch, cols, n, bench.SEED, tc.device))
    for _ in range(3):
        tc.commit()
    torch.cuda.synchronize()
    _lib.profile_enable(True)
    _lib.profile_read()
    K = 5
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        tc.commit()
    e1.record()
    torch.cuda.synchronize()
    ph = {k: round(v[0] / K, 3) for k, v in _lib.profile_read().items()}
    print(json.dumps({"lib": os.path.basename(_lib.LIB_PATH), "rows": n, "cols": cols, "ms": round(e0.elapsed_time(e1) / K, 3), "phases_ms": ph,
                      "root": tc.root().hex()}), flush=True)


def main():
    args = sys.argv[1:]
    if args and args[0] == "--child":
        return child(int(args[1]), int(args[2]))
    log_rows, cols = 22, 100
    while args and args[0].startswith("--"):
        if args[0] == "--log-rows":
            log_rows = int(args[1])
        elif args[0] == "--cols":
            cols = int(args[1])
        args = args[2:]
    for name in args:
        lib = os.path.join(ROOT, "starkpack_winterfell_b200", "variants", f"libwinter_gpu_{name}.so")
        env = dict(os.environ) if name == "main" else dict(os.environ, WG_LIB_PATH=lib)  # "main": the product library
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--child", str(log_rows), str(cols)], env=env, capture_output=True, text=True)
        line = [x for x in r.stdout.splitlines() if x.startswith("{")]
        print(line[-1] if line else json.dumps({"lib": name, "error": (r.stderr or r.stdout)[-400:]}), flush=True)


if __name__ == "__main__":
    main()
