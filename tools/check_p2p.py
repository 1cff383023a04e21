#!/usr/bin/env python
"""Multi-GPU parity check of the fused all-gather decode path (run under torchrun, one rank per GPU):
every rank must end up with the full, unsharded result, equal to the oracle's, for Q4_0 and Q4_K, T = 1 and 4,
including an uneven split (vocab 128256 over 128-row tiles), and for prefill-size T (tcgen05 GEMM epilogue) in all four types.
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/check_p2p.py
"""
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))


def main():
    import torch
    import torch.distributed as dist
    import helpers as H
    import llamafile_b200 as tb
    from llamafile_b200.peer import PeerGroup, shard_rows
    from oracle import binding as B
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    assert tb.link()
    o = B.oracle()
    ok_all = True
    # (weight type, K, matrix row counts, token counts); T > 8 goes through the tcgen05 GEMM whose epilogue does the peer stores
    # 61440 rows / rank at K = 256 take the RESIDENT decode kernel (3 CTAs per SM walking row pairs; its last round is part empty)
    cases = ((H.Q4_0, 256, [61440 * world + 256], (1, 2)), (H.Q4_0, 1024, [1000, 384, 130], (1, 4)), (H.Q4_K, 1024, [520, 256], (1, 4)), (H.Q4_0, 512, [128256 // 8], (1, 4)),
             (H.Q4_0, 1024, [1000, 384], (40, 300)), (H.Q8_0, 512, [700], (129,)), (H.Q4_K, 1024, [520], (64,)), (H.F16, 512, [300], (33,)))
    for wt, K, shapes, toks in cases:
        for T in toks:
            X = H.make_acts(T, K, seed=T + K)
            Ws = [H.make_weights(wt, n, K, seed=n, realistic_q4k=False) for n in shapes]
            rb = B.row_bytes(wt, K)
            spans = [shard_rows(n, rank, world) for n in shapes]
            ws = [tb.Weights(W[lo * rb:hi * rb], wt, hi - lo, K) for W, (lo, hi) in zip(Ws, spans) if hi > lo]
            live = [i for i, (lo, hi) in enumerate(spans) if hi > lo]
            offsets, off = [], 0
            for n in shapes:
                offsets.append(off)
                off += T * n * 4
            pg = PeerGroup(world, rank, max(off, 256))
            acts = tb.Acts(wt, K, max(8, T))
            acts.quantize(torch.from_numpy(X).cuda(), T)
            for rep in range(3):   # several epochs through the same flags
                if live:
                    pg.mul_mat_gather(ws, acts, [offsets[i] for i in live], [shapes[i] for i in live], [spans[i][0] for i in live])
                pg.barrier()
            tb.synchronize()
            assert not pg.timed_out(), "peer barrier timed out"
            for n, W, offb in zip(shapes, Ws, offsets):
                got = pg.read(offb, T * n).reshape(T, n)
                ref = o.mul_mat(W, wt, X)
                good = H.scaled_err(got, ref) <= H.SCALED_TOL and not np.isnan(got).any()
                ok_all &= good
                if not good:
                    print(f"rank {rank}: MISMATCH {H.TYPE_NAME[wt]} n={n} T={T} err={H.scaled_err(got, ref):.3e}", flush=True)
            dist.barrier()
    t = torch.tensor([1 if ok_all else 0], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("P2P_GATHER_OK" if int(t.item()) == 1 else "P2P_GATHER_FAILED", flush=True)
    sys.stdout.flush()
    os._exit(0 if int(t.item()) == 1 else 1)


if __name__ == "__main__":
    main()
