"""CPU, world_size 2 and 3 over gloo: the host-side row sharding + all-gather used by the multi-GPU path reproduces
the unsharded result exactly (each rank's slice computed by the oracle)."""
import os
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent


@pytest.mark.parametrize("world,N", [(2, 1000), (3, 520), (2, 128)])
def test_row_shard_all_gather_matches_unsharded(world, N):
    import json
    import subprocess
    port = 29500 + (os.getpid() + N) % 2000
    procs = [subprocess.Popen([sys.executable, str(ROOT / "tests" / "_shard_worker.py"), str(r), str(world), str(port), str(N), "256", "3", "2"],
                              stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True) for r in range(world)]
    res = []
    for p in procs:
        out, err = p.communicate(timeout=180)
        assert p.returncode == 0, err[-2000:]
        res.append(json.loads(out.strip().splitlines()[-1]))
    assert all(r["equal"] for r in res), res
    spans = sorted(tuple(r["span"]) for r in res)
    assert spans[0][0] == 0 and spans[-1][1] == N and all(a[1] == b[0] for a, b in zip(spans, spans[1:]))


def test_shard_rows_properties():
    from llamafile_b200.shard import max_shard, shard_rows
    for N in (128256, 4096, 1024, 14336, 100, 128, 28672):
        for world in (1, 2, 4, 8):
            spans = [shard_rows(N, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == N
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            assert all(lo % 128 == 0 for lo, _ in spans)
            assert max_shard(N, world) - min(h - l for l, h in spans) <= 128 or N < 128 * world
