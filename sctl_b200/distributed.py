"""Multi-GPU driver: target sharding + source replication over NCCL (one process per GPU, torch.distributed).

Mirrors the per-rank structure of the reference's ParticleFMM::EvalDirect
(/root/reference/include/sctl/fmm-wrapper.txx:490-562):

  reference (MPI)                                          here (NCCL over NVLink/NVSwitch)
  -------------------------------------------------------  ----------------------------------------------------
  partition(): rank r owns block [N*r/np, N*(r+1)/np)      partition() below, same formula (:504-508)
  of targets AND of every source array (:504-529)
  np ring rounds: Isend/Irecv (Xs,Xn,F) of rank r-i,       every rank's source block is broadcast on a side stream
  blocking Wait, then Eval accumulates into U (:537-558)   (block k+1 in flight while block k is evaluated); Eval of
                                                           block k accumulates into U exactly like a ring round
  U zeroed first (:501-502), results stay with the         U overwritten by the first block, accumulated by the rest;
  target owner                                             results stay sharded (no reduction: targets are disjoint)

There is no collective on the data path other than this source replication. `evaluator` is the per-block
Eval call (GenericKernel.Eval on CUDA tensors by default); the CPU gloo tests inject the oracle there to
exercise the sharding / exchange logic with world_size 2 without a GPU.
"""
from __future__ import annotations

from typing import Callable, Optional

import torch
import torch.distributed as dist


def partition(n: int, rank: int, size: int) -> tuple[int, int]:
    """[begin, end) of rank's block — the formula of fmm-wrapper.txx:507."""
    return n * rank // size, n * (rank + 1) // size


class ShardedDirectSum:
    """Direct summation with targets and sources sharded over the ranks of a process group."""

    def __init__(self, kernel, group=None, evaluator: Optional[Callable] = None, overlap: bool = True):
        self.kernel = kernel
        self.group = group
        self.evaluator = evaluator or (lambda U, Xt, Xs, Xn, F, digits, overwrite: kernel.Eval(U, Xt, Xs, Xn, F, digits=digits, overwrite=overwrite))
        self.overlap = overlap
        self._comm_stream = None
        self._bufs = {}

    # -- helpers -------------------------------------------------------------------------------------
    def _world(self):
        if not dist.is_available() or not dist.is_initialized():
            return 0, 1
        return dist.get_rank(self.group), dist.get_world_size(self.group)

    def _buf(self, key, n, like):
        b = self._bufs.get(key)
        if b is None or b.numel() < n or b.dtype != like.dtype or b.device != like.device:
            b = torch.empty(max(n, 1), dtype=like.dtype, device=like.device)
            self._bufs[key] = b
        return b[:n]

    def source_counts(self, ns_local: int, device) -> list[int]:
        """number of sources held by every rank (tiny all_gather; pass `counts` to __call__ to skip it)"""
        rank, size = self._world()
        if size == 1:
            return [ns_local]
        t = torch.tensor([ns_local], dtype=torch.int64, device=device)
        out = [torch.empty_like(t) for _ in range(size)]
        dist.all_gather(out, t, group=self.group)
        return [int(o.item()) for o in out]

    # -- the sharded evaluation ----------------------------------------------------------------------
    def __call__(self, U, Xt, Xs, Xn, F, digits: int = -1, counts: Optional[list[int]] = None):
        """U (local targets, overwritten) = sum over ALL ranks' sources. Xt/Xs/Xn/F are this rank's blocks."""
        rank, size = self._world()
        K0, Nd, K1 = self.kernel.SrcDim(), self.kernel.NormalDim(), self.kernel.TrgDim()
        nt = Xt.numel() // 3
        if U is None or U.numel() != nt * K1:
            U = torch.zeros(nt * K1, dtype=Xt.dtype, device=Xt.device)
        if size == 1:
            return self.evaluator(U, Xt, Xs, Xn if Nd else None, F, digits, True)
        if counts is None:
            counts = self.source_counts(Xs.numel() // 3, Xs.device)
        cuda = Xs.is_cuda
        if cuda and self.overlap and self._comm_stream is None:
            self._comm_stream = torch.cuda.Stream(device=Xs.device)

        # replicate: one broadcast per (rank block, array); block k+1 is in flight while block k is evaluated
        blocks, ready = [], []
        main = torch.cuda.current_stream(Xs.device) if cuda else None
        if cuda and self.overlap:
            self._comm_stream.wait_stream(main)  # inputs produced on the main stream are visible to the comm stream
        for r in range(size):
            n = counts[r]
            if r == rank:
                blocks.append((Xs, Xn if Nd else None, F))
                ready.append(None)
                srcs = (Xs, Xn if Nd else None, F)
            else:
                srcs = (self._buf(("x", r), 3 * n, Xs), self._buf(("n", r), Nd * n, Xs) if Nd else None, self._buf(("f", r), K0 * n, Xs))
                blocks.append(srcs)
            ctx = torch.cuda.stream(self._comm_stream) if (cuda and self.overlap) else _null()
            with ctx:
                for t in srcs:
                    if t is not None and t.numel() > 0:
                        dist.broadcast(t, src=dist.get_global_rank(self.group, r) if self.group is not None else r, group=self.group)
                if cuda and self.overlap and r != rank:
                    ev = torch.cuda.Event()
                    ev.record(self._comm_stream)
                    ready.append(ev)
                elif r != rank:
                    ready.append(None)

        # evaluate own block first (needs no communication), then the others in ring order (rank+1, rank+2, ...)
        first = True
        for i in range(size):
            r = (rank + i) % size
            if ready[r] is not None:
                main.wait_event(ready[r])
            bx, bn, bf = blocks[r]
            if counts[r] == 0:
                continue
            U = self.evaluator(U, Xt, bx, bn, bf, digits, first)
            first = False
        if first:  # no sources anywhere
            U.zero_()
        if cuda and self.overlap:
            self._comm_stream.wait_stream(main)  # receive buffers are reused by the next call
        return U


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False
