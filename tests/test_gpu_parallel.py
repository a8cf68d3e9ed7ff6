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


@pytest.mark.parametrize("G,field,n,blowup,cols", [(2, 0, 8192, 8, 13), (4, 0, 1 << 14, 4, 8), (8, 1, 1 << 13, 8, 3), (2, 0, 512, 8, 5),
                                                   (4, 1, 64, 2, 2), (8, 0, 1 << 16, 2, 4)])
def test_lde_scatter_into_row_blocks(G, field, n, blowup, cols):
    """wg_lde_scatter_dev (the LDE whose last pass stores each row into its owner's buffer) with the G owner buffers on
    this device: block g must equal rows [g*L/G, (g+1)*L/G) of the plain LDE. Covers single-pass (n <= 4096) and two-pass
    transforms; over NVLink the same kernel receives peer-mapped pointers (tests/run_distributed_gpu.py)."""
    import ctypes as C

    from starkpack_winterfell_b200._lib import check, lib
    from starkpack_winterfell_b200.parallel import DeviceBackend

    dev = torch.device("cuda:0")
    be = DeviceBackend(field, 0, dev)
    full = fx.synth_f64(77, cols, n) if field == 0 else fx.synth_f128(77, cols, n)
    d = _to_dev(full, field, dev)
    polys, lde = be.interpolate_lde(d, n, blowup)
    rows = n * blowup // G
    rw = lde.shape[1]
    blocks = [torch.full((rows, rw, be.w), -1, dtype=torch.int64, device=dev) for _ in range(G)]
    polys2 = be.interpolate_lde_scatter(d, n, blowup, [b.data_ptr() for b in blocks])
    torch.cuda.synchronize()
    assert torch.equal(polys, polys2)
    for g in range(G):
        assert torch.equal(blocks[g], lde[g * rows: (g + 1) * rows]), g
    # argument checks: owners must be a power of two <= 8 and no more than the last pass has points
    ptrs = (C.c_void_p * 3)(*[b.data_ptr() for b in blocks[:1] * 3])
    rc = lib().wg_lde_scatter_dev(field, C.c_void_p(polys.data_ptr()), cols, n, blowup, C.c_void_p(be._offset.ctypes.data), 3, ptrs, rw,
                                  None, 0, None)
    assert rc != 0


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_nccl_distributed_commit():
    n_gpus = 1 << (torch.cuda.device_count().bit_length() - 1)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n_gpus}",
                        "--master-addr", "127.0.0.1", "--master-port", "29577", os.path.join(ROOT, "tests", "run_distributed_gpu.py")],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "DISTRIBUTED_OK" in r.stdout
