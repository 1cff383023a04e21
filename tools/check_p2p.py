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


def check_plan(torch, dist, tb, H, o, B, PeerGroup, shard_rows, rank, world):
    """The decode plan on row-sharded weights (tb200_plan_create_peers): a chain x0 -> [W0 sharded] -> y0 -> [W1a, W1b sharded] -> y1a, y1b
    -> [W2 replicated, private output] -> y2 (from y1a) -> [W3 sharded] -> y3 (from y1b), every sharded output gathered into every
    rank's symmetric buffer by the kernel itself. Each stage is checked against the oracle fed with the GPU's previous stage."""
    ok = True
    for wt in (H.Q4_0, H.Q4_K):
        K0, N0, N1, N2, N3 = 512, 1024, 1280, 300, 128256 // 8
        rng = np.random.default_rng(3)
        x0 = (rng.standard_normal(K0) * 0.5).astype(np.float32)
        mk = lambda n, k, seed: H.make_weights(wt, n, k, seed=seed, realistic_q4k=False)   # noqa: E731
        W0, W1a, W1b, W2, W3 = mk(N0, K0, 1), mk(N1, N0, 2), mk(N1, N0, 3), mk(N2, N1, 4), mk(N3, N1, 5)
        offs, off = {}, 0
        for name, n in (("y0", N0), ("y1a", N1), ("y1b", N1), ("y3", N3)):
            offs[name] = off
            off += n * 4
        pg = PeerGroup(world, rank, off)

        def shard(W, n, k):
            lo, hi = shard_rows(n, rank, world)
            rb = B.row_bytes(wt, k)
            return tb.Weights(W[lo * rb:hi * rb], wt, hi - lo, k), lo, hi
        (w0, l0, h0), (w1a, l1, h1), (w1b, _, _), (w3, l3, h3) = shard(W0, N0, K0), shard(W1a, N1, N0), shard(W1b, N1, N0), shard(W3, N3, N1)
        assert min(h0 - l0, h1 - l1, h3 - l3) > 0
        w2 = tb.Weights(W2, wt, N2, N1)
        x0d = torch.from_numpy(x0).cuda()
        y2 = torch.full((N2,), float("nan"), device="cuda")
        sym = lambda name, lo: pg.out + offs[name] + lo * 4   # noqa: E731  (device address of this rank's slice)
        plan = tb.Plan([{"w": [w0], "out": [sym("y0", l0)], "x": x0d},
                        {"w": [w1a, w1b], "out": [sym("y1a", l1), sym("y1b", l1)], "x": pg.out + offs["y0"]},
                        {"w": [w2], "out": [y2], "x": pg.out + offs["y1a"]},
                        {"w": [w3], "out": [sym("y3", l3)], "x": pg.out + offs["y1b"]}], peers=pg)
        # a second plan on the same peer group (the first two groups only): its launches share the peers' flag words with the first
        plan_b = tb.Plan([{"w": [w0], "out": [sym("y0", l0)], "x": x0d},
                          {"w": [w1a, w1b], "out": [sym("y1a", l1), sym("y1b", l1)], "x": pg.out + offs["y0"]}], peers=pg)
        for rep in range(4):
            plan.run()
            if rep == 1:
                plan.run(0, 2); plan.run(2, 2)   # the same as two launches (the second starts from gathered data of the first)
            if rep == 2:
                plan_b.run()
        tb.synchronize()
        assert plan.status() == 0, f"plan wait timed out (status {plan.status()})"
        y0 = pg.read(offs["y0"], N0); y1a = pg.read(offs["y1a"], N1); y1b = pg.read(offs["y1b"], N1); y3 = pg.read(offs["y3"], N3)
        checks = (("y0", y0, o.mul_mat(W0, wt, x0[None])[0]), ("y1a", y1a, o.mul_mat(W1a, wt, y0[None])[0]), ("y1b", y1b, o.mul_mat(W1b, wt, y0[None])[0]),
                  ("y2", y2.cpu().numpy(), o.mul_mat(W2, wt, y1a[None])[0]), ("y3", y3, o.mul_mat(W3, wt, y1b[None])[0]))
        for name, got, ref in checks:
            good = not np.isnan(got).any() and H.scaled_err(got, ref) <= H.SCALED_TOL
            ok &= good
            if not good:
                print(f"rank {rank}: PLAN MISMATCH {H.TYPE_NAME[wt]} {name} err={H.scaled_err(got, ref):.3e} nan={int(np.isnan(got).sum())}", flush=True)
        dist.barrier()
        plan.free(); plan_b.free()
    if rank == 0:
        print("P2P_PLAN_OK" if ok else "P2P_PLAN_FAILED", flush=True)
    return ok


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
    ok_all &= check_plan(torch, dist, tb, H, o, B, PeerGroup, shard_rows, rank, world)
    t = torch.tensor([1 if ok_all else 0], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("P2P_GATHER_OK" if int(t.item()) == 1 else "P2P_GATHER_FAILED", flush=True)
    sys.stdout.flush()
    os._exit(0 if int(t.item()) == 1 else 1)


if __name__ == "__main__":
    main()
