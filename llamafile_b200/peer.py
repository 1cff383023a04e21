"""Host side of the fused all-gather (tb200_peer.cu): one process per GPU; every rank allocates a symmetric output buffer
and a flag buffer with tb200_malloc, the CUDA IPC handles are exchanged with torch.distributed (plumbing only), and the
decode kernels then store their row slices straight into every rank's buffer over NVLink."""
from __future__ import annotations

import ctypes as C

import numpy as np

from .lib import TB200Error, _check, lib
from .shard import shard_rows


class PeerGroup:
    def __init__(self, world: int, rank: int, out_bytes: int, group=None):
        import torch.distributed as dist
        L = lib()
        self.world, self.rank, self.out_bytes = world, rank, out_bytes
        self.out = L.tb200_malloc(out_bytes)
        self.flags = L.tb200_malloc(256)
        if not self.out or not self.flags:
            raise TB200Error("peer buffers: " + L.tb200_last_error().decode(errors="replace"))
        _check(L.tb200_memset(self.out, 0, out_bytes), "memset")
        _check(L.tb200_memset(self.flags, 0, 256), "memset")
        h_out, h_flag = (C.c_ubyte * 64)(), (C.c_ubyte * 64)()
        _check(L.tb200_ipc_get_handle(self.out, h_out), "ipc_get_handle")
        _check(L.tb200_ipc_get_handle(self.flags, h_flag), "ipc_get_handle")
        mine = (bytes(h_out), bytes(h_flag))
        everyone = [None] * world
        dist.all_gather_object(everyone, mine, group=group)
        self._opened = []
        outs, flags = (C.c_void_p * world)(), (C.c_void_p * world)()
        for p, (ho, hf) in enumerate(everyone):
            if p == rank:
                outs[p], flags[p] = self.out, self.flags
                continue
            po = L.tb200_ipc_open_handle((C.c_ubyte * 64).from_buffer_copy(ho))
            pf = L.tb200_ipc_open_handle((C.c_ubyte * 64).from_buffer_copy(hf))
            if not po or not pf:
                raise TB200Error("ipc_open_handle: " + L.tb200_last_error().decode(errors="replace"))
            self._opened += [po, pf]
            outs[p], flags[p] = po, pf
        self.handle = L.tb200_peers_create(world, rank, outs, flags)
        if not self.handle:
            raise TB200Error("peers_create: " + L.tb200_last_error().decode(errors="replace"))
        dist.barrier(group=group)   # every rank has zeroed and mapped its buffers before anyone stores

    def mul_mat_gather(self, ws, acts, out_offsets, full_rows, row_los):
        """ws: this rank's row-sharded weights; matrix m's slice starts at row row_los[m] of a full output of full_rows[m]
        rows stored at byte offset out_offsets[m] of the symmetric buffer."""
        n = len(ws)
        W = (C.c_void_p * n)(*[w.handle for w in ws])
        Off = (C.c_size_t * n)(*out_offsets)
        Ld = (C.c_long * n)(*full_rows)
        Lo = (C.c_long * n)(*row_los)
        _check(lib().tb200_mul_mat_acts_multi_gather(self.handle, n, W, acts.handle, Off, Ld, Lo), "mul_mat_acts_multi_gather")

    def barrier(self):
        _check(lib().tb200_peers_barrier(self.handle), "peers_barrier")

    def timed_out(self) -> bool:
        return lib().tb200_peers_timed_out(self.handle) != 0

    def read(self, offset: int, count: int) -> np.ndarray:
        """fp32 values of the local copy of the symmetric buffer (verification)."""
        out = np.empty(count, np.float32)
        _check(lib().tb200_memcpy_d2h(out.ctypes.data, self.out + offset, count * 4), "memcpy_d2h")
        return out


__all__ = ["PeerGroup", "shard_rows"]
