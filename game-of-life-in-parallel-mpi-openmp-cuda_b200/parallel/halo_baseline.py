"""BASELINE transport (not the product): k-deep halo exchange with torch.distributed point-to-point
ops - ncclSend/ncclRecv on GPUs, gloo on CPUs - plus an all_reduce for the exit flags.

This is what "our kernel + NCCL" looks like, i.e. the structure of the reference's MPI programs
(16 requests + Startall/Waitall + Allreduce per generation, src/game_mpi_collective.c:280-344,
70-109) moved to GPUs: compute an epoch, pack the border strips into contiguous messages, exchange
them in two phases (rows, then columns over the full padded height so the corners ride along),
unpack into the ghost zones.  The fused path in the strip kernel (peer stores + flags, no packing,
no host involvement) has to beat it; `scripts/mgpu_check.py --baseline` measures both.

Two block stores share the exchange code:
  * `PackedBlockCPU`  - uint64 words + CPU oracle, used by the gloo tests of the protocol logic;
  * `EngineBlockGPU`  - a single-domain engine per rank; its padded device buffer is viewed as a
    tensor and the ghost zones the kernel filled with the block's own wrap are overwritten.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist

from .topology import Topology


class _Exchange:
    """Two-phase halo exchange on a 2-D tensor `t` with interior [r0:r1, c0:c1] and ghost thickness
    (ky rows, kx columns)."""

    def __init__(self, topo: Topology, gid: int, group=None):
        self.topo, self.gid, self.group = topo, gid, group
        nb = topo.neighbors(gid)
        self.north, self.south, self.west, self.east = nb[0], nb[1], nb[2], nb[3]
        self.messages = 0
        self.bytes = 0

    def _pair(self, t, send_first, send_second, recv_first, recv_second, peers):
        """peers = (dst of send_first, dst of send_second, src of recv_first, src of recv_second).
        Sends and receives between a pair are matched in issue order (NCCL has no tags), so the
        order here is chosen to stay consistent when both neighbours are the same rank."""
        d1, d2, s1, s2 = peers
        bufs, ops, copies = [], [], []
        for (view, dst) in ((send_first, d1), (send_second, d2)):
            if dst == self.gid:
                continue
            b = view.contiguous()
            bufs.append(b)
            ops.append(dist.P2POp(dist.isend, b, dst, group=self.group))
            self.messages += 1
            self.bytes += b.numel() * b.element_size()
        for (view, src) in ((recv_first, s1), (recv_second, s2)):
            if src == self.gid:
                continue
            b = torch.empty(view.shape, dtype=view.dtype, device=view.device)
            ops.append(dist.P2POp(dist.irecv, b, src, group=self.group))
            copies.append((view, b))
        if d1 == self.gid:  # a 1-wide process dimension: the neighbour is this rank, wrap locally
            recv_first.copy_(send_first)     # e.g. top ghost rows <- my own bottom rows
            recv_second.copy_(send_second)
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        for view, b in copies:
            view.copy_(b)

    def run(self, t, r0, r1, c0, c1, ky, kx, rows_prefilled=False):
        if not (rows_prefilled and self.north == self.gid):
            # rows: my bottom rows go south (first), my top rows go north
            self._pair(t, t[r1 - ky:r1, c0:c1], t[r0:r0 + ky, c0:c1],
                       t[r0 - ky:r0, c0:c1], t[r1:r1 + ky, c0:c1],
                       (self.south, self.north, self.north, self.south))
        ra, rb = r0 - ky, r1 + ky
        if not (rows_prefilled and self.west == self.gid):
            # columns over the padded height: my right columns go east (first), my left go west
            self._pair(t, t[ra:rb, c1 - kx:c1], t[ra:rb, c0:c0 + kx],
                       t[ra:rb, c0 - kx:c0], t[ra:rb, c1:c1 + kx],
                       (self.east, self.west, self.west, self.east))


class PackedBlockCPU:
    """One rank's block as uint64 words with a (k rows, 1 word) ghost frame; compute = CPU oracle.
    Block widths must be multiples of 64 (the engine handles ragged widths, the baseline does not)."""

    def __init__(self, topo: Topology, gid: int, k: int, group=None):
        from ..ops import cpu
        self.cpu = cpu
        self.topo, self.gid, self.k = topo, gid, k
        self.x0, self.y0, self.w, self.h = topo.extent(gid)
        if self.w % 64 or self.x0 % 64:
            raise ValueError("baseline blocks must be 64-cell aligned")
        if k > 64 or k > self.h:
            raise ValueError("halo depth too large for the block")
        self.nq = self.w // 64
        self.t = torch.zeros((self.h + 2 * k, self.nq + 2), dtype=torch.int64)
        self.ex = _Exchange(topo, gid, group)

    def load_global(self, packed: np.ndarray):
        q0 = self.x0 // 64
        blk = packed[self.y0:self.y0 + self.h, q0:q0 + self.nq].astype(np.uint64).view(np.int64)
        self.t[self.k:self.k + self.h, 1:1 + self.nq] = torch.from_numpy(blk.copy())

    def interior(self) -> np.ndarray:
        return self.t[self.k:self.k + self.h, 1:1 + self.nq].numpy().view(np.uint64).copy()

    def epoch(self):
        """exchange k-deep halos, then k local generations (garbage from the frame's outer edge
        creeps in one cell per generation and never reaches the interior)."""
        k = self.k
        self.ex.run(self.t, k, k + self.h, 1, 1 + self.nq, k, 1)
        arr = np.ascontiguousarray(self.t.numpy().view(np.uint64))
        out = self.cpu.step(arr, (self.nq + 2) * 64, k)
        self.t = torch.from_numpy(out.view(np.int64).copy())

    def flags_all(self, group=None):
        f = torch.tensor([int(self.interior().any())], dtype=torch.int32)
        dist.all_reduce(f, op=dist.ReduceOp.MAX, group=group)
        return bool(f.item())


class EngineBlockGPU:
    """One rank's block in a single-domain engine; NCCL moves the ghost zones between epochs."""

    def __init__(self, topo: Topology, gid: int, k: int, device: int, group=None):
        from ..models.life import GameOfLife
        from .. import _native as N
        self.N = N
        self.topo, self.gid, self.k = topo, gid, k
        self.x0, self.y0, self.w, self.h = topo.extent(gid)
        if self.w % 64 or self.x0 % 64:
            raise ValueError("baseline blocks must be 64-cell aligned")
        self.life = GameOfLife(self.w, self.h, device=device, k_max=k, similarity_frequency=0)
        self.ex = _Exchange(topo, gid, group)
        self.group = group

    def load_global(self, packed: np.ndarray):
        q0 = self.x0 // 64
        nq = self.w // 64
        self.life.load_packed(np.ascontiguousarray(packed[self.y0:self.y0 + self.h, q0:q0 + nq]))
        self._exchange()

    def randomize(self, seed: int):
        from ..ops import dense
        g = dense.random_grid(self.w, self.h, seed, device=self.life.device, x0=self.x0, y0=self.y0)
        self.life.load_packed(g)
        self._exchange()

    def _exchange(self):
        t, lay = self.life.padded_tensor()
        g, gw = int(lay.ghost_rows), self.N.GHOST_WORDS32
        # ghost rows / columns the kernel wrote are the block's own torus wrap: right when this rank
        # is its own neighbour in that direction, overwritten here otherwise
        self.ex.run(t, g, g + self.h, gw, gw + int(lay.nw32), self.k, gw, rows_prefilled=True)

    def epoch(self) -> float:
        ms = self.life.step(self.k)
        self._exchange()
        return ms

    def interior(self) -> np.ndarray:
        return self.life.packed_numpy()

    def flags_all(self):
        from ..ops import dense
        alive, _ = dense.flags(self.life.packed().contiguous(), self.w)
        f = torch.tensor([int(alive)], dtype=torch.int32, device=self.life.device)
        dist.all_reduce(f, op=dist.ReduceOp.MAX, group=self.group)   # ncclAllReduce, like empty_all
        return bool(f.item())
