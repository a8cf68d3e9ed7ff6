"""torchrun entry: NCCL run of parallel.DistributedCommitter on G GPUs, checked on rank 0 against the oracle."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import fixtures as fx  # noqa: E402


def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from starkpack_winterfell_b200.parallel import DeviceBackend, DistributedCommitter

    for field, hash_id, n, blowup, total_cols in [(0, 0, 8192, 8, 100), (1, 0, 4096, 8, 16), (0, 1, 1024, 8, 21)]:
        if total_cols < dist.get_world_size():
            continue
        full = fx.synth_f64(9, total_cols, n) if field == 0 else fx.synth_f128(9, total_cols, n)
        be = DeviceBackend(field, hash_id, dev)
        # fused: the LDE stores rows into the owners' buffers over NVLink (CUDA IPC peer pointers), no all-to-all
        for fused, pipeline in ((False, 0), (True, 0), (False, 2), (False, 4)):
            run_one(be, field, hash_id, n, blowup, total_cols, full, dev, fused, pipeline)
    dist.barrier()
    if dist.get_rank() == 0:
        print("DISTRIBUTED_OK")
    dist.destroy_process_group()


def run_one(be, field, hash_id, n, blowup, total_cols, full, dev, fused, pipeline):
    from starkpack_winterfell_b200.parallel import DistributedCommitter

    if True:
        dc = DistributedCommitter(be, n, blowup, total_cols, fused=fused, pipeline=pipeline)
        a = np.ascontiguousarray(full[dc.col_start: dc.col_start + dc.col_count]).view(np.int64)
        cols = torch.from_numpy(a[..., None] if field == 0 else a).to(dev)
        root = dc.commit(cols)
        nodes = dc.gather_nodes().numpy()
        if dist.get_rank() == 0:
            import oracle

            want = oracle.commit_trace(field, hash_id, [full], blowup, want=("nodes",))
            assert np.array_equal(nodes, want["nodes"]), "stitched nodes differ from the oracle"
            assert root == bytes(want["root"])
        if fused or pipeline:
            root2 = dc.commit(cols)  # buffers reused: a second commit must give the same tree
            assert root2 == root
        dc.close()


if __name__ == "__main__":
    main()
