"""Multi-GPU driver: target sharding + source replication over NCCL (one process per GPU).

On CUDA tensors both classes below are thin bindings of the C-ABI driver inside libsctl_b200.so (sctl_b200_comm_* /
sctl_b200_eval_sharded* / sctl_b200_fmm_set_comm, csrc/comm.cu): the NCCL calls, the gather buffers and the evaluation are the
library's; torch.distributed only carries the 128-byte NCCL id to the other ranks once (class LibComm). The pure-Python
exchange further down is the same host logic over torch.distributed; it runs when a CPU `evaluator` / `local` registry is
injected (the world-size-2 gloo tests) and for the diagnostic "ring" mode.

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


class LibComm:
    """A communicator of libsctl_b200.so (ncclCommInitRank inside the library) spanning the ranks of a torch.distributed group:
    rank 0 asks the library for the NCCL id, torch.distributed broadcasts its 128 bytes, every rank creates its handle on
    its current CUDA device."""

    _cache = {}

    def __init__(self, group=None, device=None):
        import ctypes

        from . import capi

        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.size = dist.get_world_size(group) if dist.is_initialized() else 1
        dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        ident = torch.zeros(128, dtype=torch.uint8)
        if self.size > 1:
            if self.rank == 0:
                buf = (ctypes.c_ubyte * 128)()
                capi.check(capi.lib().sctl_b200_comm_unique_id(buf, 128))
                ident = torch.tensor(list(buf), dtype=torch.uint8)
            wire = ident.to(dev) if dist.get_backend(group) == "nccl" else ident
            dist.broadcast(wire, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
            ident = wire.cpu()
        raw = (ctypes.c_ubyte * 128)(*ident.tolist())
        h = ctypes.c_void_p()
        with torch.cuda.device(dev):
            capi.check(capi.lib().sctl_b200_set_device(dev.index))
            capi.check(capi.lib().sctl_b200_comm_create(ctypes.byref(h), self.size, self.rank, raw))
        self.handle, self.device = h, dev

    @classmethod
    def get(cls, group=None, device=None) -> "LibComm":
        dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        key = (id(group) if group is not None else None, dev.index)
        if key not in cls._cache:
            cls._cache[key] = cls(group, dev)
        return cls._cache[key]

    def destroy(self):
        from . import capi

        h, self.handle = self.handle, None
        if h:
            capi.lib().sctl_b200_comm_destroy(h)
        for k, v in list(self._cache.items()):
            if v is self:
                del self._cache[k]


def partition(n: int, rank: int, size: int) -> tuple[int, int]:
    """[begin, end) of rank's block — the formula of fmm-wrapper.txx:507."""
    return n * rank // size, n * (rank + 1) // size


def allgather_blocks(local, width, counts, rank, size, group=None, out=None):
    """Contiguous replica of every rank's block of an AoS array with `width` values per particle: ONE all_gather_into_tensor
    when the blocks are equal (ncclAllGather), a chain of in-place broadcasts when they are ragged. `out` (optional) is
    reused when large enough."""
    total = width * sum(counts)
    if local.numel() != width * counts[rank]:
        raise ValueError(f"expected {width} values for each of this rank's {counts[rank]} particles, got {local.numel()}")
    if out is None or out.numel() < total or out.dtype != local.dtype or out.device != local.device:
        out = torch.empty(max(total, 1), dtype=local.dtype, device=local.device)
    out = out[:total]
    if total == 0:
        return out
    if len(set(counts)) == 1:
        dist.all_gather_into_tensor(out, local.reshape(-1), group=group)
        return out
    off = [0]
    for c in counts:
        off.append(off[-1] + width * c)
    out[off[rank]:off[rank + 1]].copy_(local.reshape(-1))
    for r in range(size):
        if counts[r]:
            dist.broadcast(out[off[r]:off[r + 1]], src=dist.get_global_rank(group, r) if group is not None else r, group=group)
    return out


class ShardedDirectSum:
    """Direct summation with targets and sources sharded over the ranks of a process group."""

    def __init__(self, kernel, group=None, evaluator: Optional[Callable] = None, mode: str = "gather", overlap: Optional[bool] = None):
        if overlap is not None:  # backward-compatible spelling: overlap=True -> ring on a side stream
            mode = "ring" if overlap else "gather"
        if mode not in ("gather", "ring"):
            raise ValueError("mode must be 'gather' or 'ring'")
        self.kernel = kernel
        self.group = group
        self.native = evaluator is None and mode == "gather"  # CUDA tensors go to the C-ABI driver (sctl_b200_eval_sharded_async)
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
        if self.native and Xt.is_cuda:
            return self._call_native(U, Xt, Xs, Xn if Nd else None, F, digits, counts, size)
        if counts is None:
            counts = self.source_counts(Xs.numel() // 3, Xs.device)
        if self.mode == "gather":
            return self._call_gather(U, Xt, Xs, Xn if Nd else None, F, digits, counts, rank, size)
        return self._call_ring(U, Xt, Xs, Xn if Nd else None, F, digits, counts, rank, size)

    def _call_native(self, U, Xt, Xs, Xn, F, digits, counts, size):
        """the C path: counts exchange (when not given), one ncclAllGather per source array and pack -> eval -> reduce, all
        enqueued by libsctl_b200.so on torch's current stream; no synchronisation when `counts` is given"""
        import ctypes

        from . import capi

        K0, Nd = self.kernel.SrcDim(), self.kernel.NormalDim()
        tdt = {torch.float32: capi.F32, torch.float64: capi.F64}.get(Xt.dtype)
        if tdt is None:
            raise ValueError("Real must be float32 or float64")
        ns = Xs.numel() // 3
        for a, n, what in ((Xt, Xt.numel(), "r_trg"), (Xs, 3 * ns, "r_src"), (F, K0 * ns, "v_src"), (U, U.numel(), "v_trg")) + (((Xn, Nd * ns, "n_src"),) if Nd else ()):
            if a.dtype != Xt.dtype or a.device != Xt.device or not a.is_contiguous() or a.numel() != n:
                raise ValueError(f"{what}: expected a contiguous {Xt.dtype} tensor of {n} values on {Xt.device}")
        comm = LibComm.get(self.group, Xt.device)
        cnt = None
        if counts is not None:
            if len(counts) != size or counts[comm.rank] != ns:
                raise ValueError("counts must list the number of sources of every rank")
            cnt = (ctypes.c_int64 * size)(*counts)
        p = lambda a: None if a is None or a.numel() == 0 else ctypes.c_void_p(a.data_ptr())
        with torch.cuda.device(Xt.device):
            stream = torch.cuda.current_stream().cuda_stream
            capi.check(capi.lib().sctl_b200_eval_sharded_async(comm.handle, self.kernel.KERNEL_ID, tdt, Xt.numel() // 3, ns, p(Xt), p(Xs), p(Xn), p(F), p(U),
                                                               digits, capi.OVERWRITE, cnt, ctypes.c_void_p(stream)))
        return U

    def eval_host(self, U, Xt, Xs, Xn, F, digits: int = -1, counts: Optional[list[int]] = None, device=None):
        """HOST numpy blocks through the C ABI (sctl_b200_eval_sharded, host pointers): this rank's blocks are uploaded by the
        library, replicated over NCCL, evaluated, and U (numpy, overwritten) is downloaded - synchronous, like sctl_b200_eval."""
        import ctypes

        import numpy as np

        from . import capi

        rank, size = self._world()
        K0, Nd, K1 = self.kernel.SrcDim(), self.kernel.NormalDim(), self.kernel.TrgDim()
        dt = np.dtype(Xt.dtype)
        tdt = {np.dtype(np.float32): capi.F32, np.dtype(np.float64): capi.F64}[dt]
        nt, ns = Xt.size // 3, Xs.size // 3
        for a, n, what in ((Xt, 3 * nt, "r_trg"), (Xs, 3 * ns, "r_src"), (F, K0 * ns, "v_src"), (U, K1 * nt, "v_trg")) + (((Xn, Nd * ns, "n_src"),) if Nd else ()):
            if a.dtype != dt or not a.flags.c_contiguous or a.size != n:
                raise ValueError(f"{what}: expected a contiguous {dt} array of {n} values")
        comm = LibComm.get(self.group, device)
        cnt = (ctypes.c_int64 * size)(*counts) if counts is not None else None
        p = lambda a: None if a is None or a.size == 0 else a.ctypes.data_as(ctypes.c_void_p)
        capi.check(capi.lib().sctl_b200_eval_sharded(comm.handle, self.kernel.KERNEL_ID, tdt, nt, ns, p(Xt), p(Xs), p(Xn) if Nd else None, p(F), p(U), digits,
                                                     capi.OVERWRITE, cnt))
        return U

    def _allgather(self, key, local, width, counts, rank, size):
        """contiguous replica of every rank's block of one AoS array (width values per source), in a reused buffer"""
        out = allgather_blocks(local, width, counts, rank, size, self.group, self._bufs.get(key))
        if out.numel():
            base = out._base if out._base is not None else out
            self._bufs[key] = base
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


class ShardedParticleFMM:
    """ParticleFMM<Real,3> registry + EvalDirect over the ranks of a process group (one process per GPU).

    The reference's EvalDirect with an MPI communicator (fmm-wrapper.txx:490-562): every rank hands in ITS particles; targets
    are re-partitioned evenly (PartitionN, :504-510), every source block travels round an Isend/Irecv ring on EVERY call
    (:537-558) and the potentials are sent back to the ranks that own the targets (:560-561). Here:

      * sources (coordinates, normals, densities, weights) are all-gathered over NCCL when they are SET, not when they
        are evaluated: a solver that sets coordinates once and a new density per iteration (boundary_integral.txx:777-778
        vs :1054) moves 8*K0 bytes per source per iteration, and every Eval is a purely local launch sequence on the
        device-resident registry of sctl_b200.fmm.ParticleFMM;
      * targets stay with their owner when the ranks hold (almost) equal numbers; when they do not (more than `imbalance`
        above the mean) they are re-partitioned by the reference's formula - all-gather the coordinates, evaluate block
        [Nt*r/np, Nt*(r+1)/np), all-gather the potentials and keep the owner's rows - so that no GPU waits for another.

    `local` is the per-rank registry (default: the CUDA one); the CPU gloo tests inject an oracle-backed stand-in.
    """

    def __init__(self, dtype=None, group=None, local=None, device=None, imbalance: float = 0.05):
        import numpy as np

        self.dtype = np.dtype(dtype or np.float64)
        self.group = group
        self.device = torch.device(device) if device is not None else (torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu"))
        # native: the registry of libsctl_b200.so does the exchange itself (sctl_b200_fmm_set_comm); every Set* / EvalDirect below
        # then just forwards this rank's particles. With an injected `local` (CPU tests) the Python exchange further down runs.
        self.native = local is None
        if local is None:
            from .fmm import ParticleFMM

            local = ParticleFMM(self.dtype)
            if self._world()[1] > 1:
                local.SetComm(LibComm.get(group, self.device))
        self.local = local
        self.imbalance = float(imbalance)
        self._src_counts = {}  # source name -> particles per rank
        self._src_dims = {}  # source name -> (SrcDim, NormalDim)
        self._trg = {}  # target name -> dict(counts, moved, lo, hi, dim)
        self._bufs = {}

    # -- helpers -----------------------------------------------------------------------------------------
    def _world(self):
        if not dist.is_available() or not dist.is_initialized():
            return 0, 1
        return dist.get_rank(self.group), dist.get_world_size(self.group)

    def _tensor(self, a):
        if a is None:
            return None
        tdt = torch.float64 if self.dtype.itemsize == 8 else torch.float32
        if not isinstance(a, torch.Tensor):
            import numpy as np

            a = torch.from_numpy(np.ascontiguousarray(a, dtype=self.dtype).reshape(-1))
        return a.to(device=self.device, dtype=tdt).contiguous().reshape(-1)

    def _to_local(self, t):
        """what the per-rank registry accepts: CUDA tensors as they are, CPU tensors as numpy"""
        if t is None:
            return None
        return t if t.is_cuda else t.numpy()

    def _counts(self, n_local):
        rank, size = self._world()
        if size == 1:
            return [n_local]
        t = torch.tensor([n_local], dtype=torch.int64, device=self.device)
        out = torch.empty(size, dtype=torch.int64, device=self.device)
        dist.all_gather_into_tensor(out, t, group=self.group)
        return [int(v) for v in out.tolist()]

    def _gather(self, local, width, counts):
        """replica of every rank's block of an AoS array with `width` values per particle (fresh tensor)"""
        rank, size = self._world()
        if size == 1:
            return local
        return allgather_blocks(local, width, counts, rank, size, self.group)

    # -- registry: identical on every rank (fmm-wrapper.txx:240-443) ---------------------------------------------
    def SetAccuracy(self, digits):
        self.local.SetAccuracy(digits)

    def SetKernels(self, ker_m2m, ker_m2l, ker_l2l):
        self.local.SetKernels(ker_m2m, ker_m2l, ker_l2l)

    def AddSrc(self, name, ker_s2m, ker_s2l):
        self.local.AddSrc(name, ker_s2m, ker_s2l)
        self._src_counts[name] = None
        self._src_dims[name] = (ker_s2m.SrcDim(), ker_s2m.NormalDim())

    def AddTrg(self, name, ker_m2t, ker_l2t):
        self.local.AddTrg(name, ker_m2t, ker_l2t)
        self._trg[name] = dict(counts=None, moved=False, dim=ker_l2t.TrgDim(), out_dim=ker_l2t.TrgDim())

    def SetKernelS2T(self, src_name, trg_name, ker_s2t):
        self.local.SetKernelS2T(src_name, trg_name, ker_s2t)

    def DeleteSrc(self, name):
        self.local.DeleteSrc(name)
        self._src_counts.pop(name, None)

    def DeleteTrg(self, name):
        self.local.DeleteTrg(name)
        self._trg.pop(name, None)

    # -- particle data: this rank's particles in, replicas / blocks into the per-rank registry ---------------------
    def SetSrcCoord(self, name, src_coord, src_normal=None):
        if self.native:
            return self.local.SetSrcCoord(name, src_coord, src_normal)
        X, Xn = self._tensor(src_coord), self._tensor(src_normal)
        if X.numel() % 3:
            raise ValueError("coordinate arrays must hold 3 values per point")
        counts = self._counts(X.numel() // 3)
        self._src_counts[name] = counts
        nd = self._src_dims[name][1]
        gX = self._gather(X, 3, counts)
        gN = self._gather(Xn, nd, counts) if (nd and Xn is not None) else None
        self.local.SetSrcCoord(name, self._to_local(gX), self._to_local(gN))

    def _src_counts_of(self, name):
        counts = self._src_counts.get(name)
        if counts is None:
            raise ValueError(f"SetSrcCoord('{name}') must come first: it fixes how many sources every rank holds")
        return counts

    def SetSrcDensity(self, name, src_density):
        if self.native:
            return self.local.SetSrcDensity(name, src_density)
        F = self._tensor(src_density)
        self.local.SetSrcDensity(name, self._to_local(self._gather(F, self._src_dims[name][0], self._src_counts_of(name))))

    def SetSrcWeights(self, name, wts):
        if self.native or wts is None:
            return self.local.SetSrcWeights(name, wts)
        self.local.SetSrcWeights(name, self._to_local(self._gather(self._tensor(wts), 1, self._src_counts_of(name))))

    def _trg_block(self, name, t, width):
        """this rank's share of a per-target array: its own rows, or block `rank` of the gathered array when re-partitioned"""
        T = self._trg[name]
        if not T["moved"]:
            return t
        g = self._gather(t, width, T["counts"])
        return g[width * T["lo"]:width * T["hi"]].contiguous()

    def SetTrgCoord(self, name, trg_coord):
        if self.native:
            return self.local.SetTrgCoord(name, trg_coord)
        X = self._tensor(trg_coord)
        if X.numel() % 3:
            raise ValueError("coordinate arrays must hold 3 values per point")
        rank, size = self._world()
        T = self._trg[name]
        T["counts"] = counts = self._counts(X.numel() // 3)
        total = sum(counts)
        T["moved"] = size > 1 and total > 0 and max(counts) > (1.0 + self.imbalance) * total / size + 1
        T["lo"], T["hi"] = partition(total, rank, size) if T["moved"] else (sum(counts[:rank]), sum(counts[:rank + 1]))
        self.local.SetTrgCoord(name, self._to_local(self._trg_block(name, X, 3)))

    def SetTrgNormal(self, name, trg_normal):
        if self.native:
            return self.local.SetTrgNormal(name, trg_normal)
        T = self._trg[name]
        if trg_normal is None:
            T["out_dim"] = T["dim"]
            return self.local.SetTrgNormal(name, None)
        if T["counts"] is None:
            raise ValueError(f"SetTrgCoord('{name}') must come first")
        T["out_dim"] = T["dim"] // 3
        self.local.SetTrgNormal(name, self._to_local(self._trg_block(name, self._tensor(trg_normal), 3)))

    # -- evaluation ----------------------------------------------------------------------------------------------
    def EvalDirect(self, trg_name, U=None):
        """potentials of THIS rank's targets from ALL ranks' sources (overwritten; fmm-wrapper.txx:500-502, :560-561)"""
        if self.native:
            if U is None:
                nt, k = self.local.TrgCount(trg_name)
                U = torch.zeros(nt * k, dtype=torch.float64 if self.dtype.itemsize == 8 else torch.float32, device=self.device)
            return self.local.EvalDirect(trg_name, U)
        rank, size = self._world()
        T = self._trg[trg_name]
        if T["counts"] is None:
            raise ValueError(f"SetTrgCoord('{trg_name}') must come first")
        k = T["out_dim"]
        n_own = T["counts"][rank] * k
        tdt = torch.float64 if self.dtype.itemsize == 8 else torch.float32
        if not T["moved"]:
            if U is None or not isinstance(U, torch.Tensor) or U.numel() != n_own or U.device != self.device or U.dtype != tdt:
                U = torch.zeros(n_own, dtype=tdt, device=self.device)
            if U.is_cuda:
                self.local.EvalDirect(trg_name, U)
            else:
                U.copy_(torch.from_numpy(self.local.EvalDirect(trg_name)))
            return U
        blk = torch.zeros((T["hi"] - T["lo"]) * k, dtype=tdt, device=self.device)
        if blk.is_cuda:
            self.local.EvalDirect(trg_name, blk)
        else:
            blk.copy_(torch.from_numpy(self.local.EvalDirect(trg_name)))
        total = sum(T["counts"])
        g = self._gather(blk, k, [partition(total, r, size)[1] - partition(total, r, size)[0] for r in range(size)])
        own0 = sum(T["counts"][:rank]) * k
        out = g[own0:own0 + n_own]
        if U is not None and isinstance(U, torch.Tensor) and U.numel() == n_own and U.device == self.device and U.dtype == tdt:
            U.copy_(out)
            return U
        return out.clone()

    def Eval(self, trg_name, U=None):
        return self.EvalDirect(trg_name, U)
