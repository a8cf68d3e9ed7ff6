"""GPU side of the multi-GPU commit: the C-ABI device backend of parallel.py, checked against the oracle.
On a 1-GPU box the G ranks are emulated one after the other on the same device (same calls, same buffers, the
all-to-all replaced by slicing); with >= 2 GPUs the real NCCL path is launched through torchrun."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from tests import fixtures as fx

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _to_dev(a, field, dev):
    a = np.ascontiguousarray(a).view(np.int64)
    return torch.from_numpy(a[..., None] if field == 0 else a).to(dev)


@pytest.mark.parametrize("G,field,hash_id,n,blowup,total_cols", [
    (2, 0, 0, 8192, 8, 21), (4, 0, 0, 1024, 8, 100), (8, 1, 0, 4096, 8, 16), (2, 0, 1, 256, 4, 17), (8, 0, 0, 2048, 8, 256)])
def test_emulated_ranks_on_one_device(oracle, G, field, hash_id, n, blowup, total_cols):
    from starkpack_winterfell_b200.parallel import DeviceBackend, split_columns, stitch_nodes

    dev = torch.device("cuda:0")
    be = DeviceBackend(field, hash_id, dev)
    full = fx.synth_f64(5, total_cols, n) if field == 0 else fx.synth_f128(5, total_cols, n)
    shards = split_columns(total_cols, G)
    L, rows = n * blowup, n * blowup // G
    ldes = [be.interpolate_lde(_to_dev(full[s: s + c], field, dev), n, blowup)[1] for s, c in shards]
    local_nodes, leaves = [], []
    for g in range(G):
        mats = [l[g * rows: (g + 1) * rows].contiguous() for l in ldes]  # what the all-to-all delivers to rank g
        lv = be.hash_rows(mats, [c for _, c in shards], rows)
        leaves.append(lv.cpu())
        local_nodes.append(be.merkle(lv))
    top = be.merkle(torch.stack([ln[1] for ln in local_nodes]).contiguous())
    nodes = stitch_nodes(local_nodes, top, G).numpy()
    want = oracle.commit_trace(field, hash_id, [full], blowup, want=("leaves", "nodes"))
    assert np.array_equal(torch.cat(leaves).numpy(), want["leaves"])
    assert np.array_equal(nodes, want["nodes"])
    assert bytes(top[1].cpu().numpy()) == bytes(want["root"])


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_nccl_distributed_commit():
    n_gpus = 1 << (torch.cuda.device_count().bit_length() - 1)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n_gpus}",
                        "--master-addr", "127.0.0.1", "--master-port", "29577", os.path.join(ROOT, "tests", "run_distributed_gpu.py")],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "DISTRIBUTED_OK" in r.stdout
