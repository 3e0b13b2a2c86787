"""Multi-GPU driver: target sharding + source replication over NCCL (one process per GPU, torch.distributed).

Mirrors the per-rank structure of the reference's ParticleFMM::EvalDirect
(/root/reference/include/sctl/fmm-wrapper.txx:490-562):

  reference (MPI)                                          here (NCCL over NVLink/NVSwitch)
  -------------------------------------------------------  ----------------------------------------------------
  partition(): rank r owns block [N*r/np, N*(r+1)/np)      partition() below, same formula (:504-508)
  of targets AND of every source array (:504-529)
  np ring rounds: Isend/Irecv (Xs,Xn,F) of rank r-i,       mode "gather" (default): ONE all-gather per source array
  blocking Wait, then Eval accumulates into U (:537-558)   (ncclAllGather; chained ncclBroadcast when the blocks are
                                                           ragged) into a contiguous replica, then ONE Eval over all
                                                           sources: 1 pack + 1 eval + 1 reduce launch per step
                                                           mode "ring": every rank's block broadcast on a side stream
                                                           while earlier blocks are evaluated, Eval accumulating block
                                                           after block like the reference's ring rounds
  U zeroed first (:501-502), results stay with the         U overwritten (first Eval), results stay sharded: targets
  target owner                                             are disjoint, so there is no reduction step

Why "gather" is the default: the exchange moves 8*(3+Nd+K0) bytes per source against ~22*Nt/G FP64 instructions per
source, i.e. microseconds against seconds, so there is nothing to hide; and overlapping has a measured cost on B200:
NCCL kernels that wait for a slower peer WHILE the eval kernel runs keep their CTAs resident and push eval CTAs off
those SMs. With a start skew of a few hundred ms between two ranks the "ring" mode locked into steps 1.49x slower than
"gather" (979 ms vs 656 ms at N = 1M, profiles/r01i_multi_gpu_skew.txt); in "gather" mode a rank that is ahead merely
waits in the collective with no compute resident, so skew is re-absorbed at every step instead of amplified.

`evaluator` is the Eval call (GenericKernel.Eval on CUDA tensors by default); the CPU gloo tests inject the oracle there
to exercise the sharding / exchange logic with world_size 2 without a GPU.
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

    def __init__(self, kernel, group=None, evaluator: Optional[Callable] = None, mode: str = "gather", overlap: Optional[bool] = None):
        if overlap is not None:  # backward-compatible spelling: overlap=True -> ring on a side stream
            mode = "ring" if overlap else "gather"
        if mode not in ("gather", "ring"):
            raise ValueError("mode must be 'gather' or 'ring'")
        self.kernel = kernel
        self.group = group
        self.evaluator = evaluator or (lambda U, Xt, Xs, Xn, F, digits, overwrite: kernel.Eval(U, Xt, Xs, Xn, F, digits=digits, overwrite=overwrite))
        self.mode = mode
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

    def _global(self, r):
        return dist.get_global_rank(self.group, r) if self.group is not None else r

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
        if self.mode == "gather":
            return self._call_gather(U, Xt, Xs, Xn if Nd else None, F, digits, counts, rank, size)
        return self._call_ring(U, Xt, Xs, Xn if Nd else None, F, digits, counts, rank, size)

    def _allgather(self, key, local, width, counts, rank, size):
        """contiguous replica of every rank's block of one AoS array (width values per source)"""
        total = sum(counts)
        out = self._buf(key, width * total, local)
        if total == 0:
            return out
        if len(set(counts)) == 1:
            dist.all_gather_into_tensor(out, local.reshape(-1), group=self.group)  # one ncclAllGather
        else:  # ragged blocks (N not divisible by the group size): in-place chain of broadcasts into the replica
            off = [0]
            for c in counts:
                off.append(off[-1] + width * c)
            out[off[rank]:off[rank + 1]].copy_(local.reshape(-1))
            for r in range(size):
                if counts[r]:
                    dist.broadcast(out[off[r]:off[r + 1]], src=self._global(r), group=self.group)
        return out

    def _call_gather(self, U, Xt, Xs, Xn, F, digits, counts, rank, size):
        K0, Nd = self.kernel.SrcDim(), self.kernel.NormalDim()
        gx = self._allgather("gx", Xs, 3, counts, rank, size)
        gn = self._allgather("gn", Xn, Nd, counts, rank, size) if Nd else None
        gf = self._allgather("gf", F, K0, counts, rank, size)
        if sum(counts) == 0:
            U.zero_()
            return U
        return self.evaluator(U, Xt, gx, gn, gf, digits, True)

    def _call_ring(self, U, Xt, Xs, Xn, F, digits, counts, rank, size):
        K0, Nd = self.kernel.SrcDim(), self.kernel.NormalDim()
        cuda = Xs.is_cuda
        if cuda and self._comm_stream is None:
            self._comm_stream = torch.cuda.Stream(device=Xs.device)

        # replicate: one broadcast per (rank block, array); block k+1 is in flight while block k is evaluated
        blocks, ready = [], []
        main = torch.cuda.current_stream(Xs.device) if cuda else None
        if cuda:
            self._comm_stream.wait_stream(main)  # inputs produced on the main stream are visible to the comm stream
        for r in range(size):
            n = counts[r]
            if r == rank:
                srcs = (Xs, Xn, F)
            else:
                srcs = (self._buf(("x", r), 3 * n, Xs), self._buf(("n", r), Nd * n, Xs) if Nd else None, self._buf(("f", r), K0 * n, Xs))
            blocks.append(srcs)
            ctx = torch.cuda.stream(self._comm_stream) if cuda else _null()
            with ctx:
                for t in srcs:
                    if t is not None and t.numel() > 0:
                        dist.broadcast(t, src=self._global(r), group=self.group)
                if cuda and r != rank:
                    ev = torch.cuda.Event()
                    ev.record(self._comm_stream)
                    ready.append(ev)
                else:
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
        if cuda:
            self._comm_stream.wait_stream(main)  # receive buffers are reused by the next call
        return U


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False
