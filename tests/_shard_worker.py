"""Worker process for tests/test_sharding_gloo.py (one rank of a gloo group on CPU)."""
import json
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))


def main():
    rank, world, port, N, K, T, wtype = (int(a) for a in sys.argv[1:8])
    import torch
    import torch.distributed as dist
    import helpers as H
    from llamafile_b200.shard import gather_rows, shard_rows
    from oracle import binding as B
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    o = B.oracle()
    W = H.make_weights(wtype, N, K, seed=3, realistic_q4k=False)
    X = H.make_acts(T, K, seed=4)
    lo, hi = shard_rows(N, rank, world)
    rb = B.row_bytes(wtype, K)
    local = o.mul_mat(W[lo * rb:hi * rb], wtype, X) if hi > lo else np.zeros((T, 0), np.float32)
    full = gather_rows(torch.from_numpy(np.ascontiguousarray(local)), N, rank, world).numpy()
    ref = o.mul_mat(W, wtype, X)
    print(json.dumps({"rank": rank, "equal": bool(np.array_equal(full, ref)), "span": [lo, hi]}), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
