"""CPU, world_size 2, gloo: the target-sharding / source-replication driver (sctl_b200/distributed.py). The per-block
evaluator is replaced by the CPU oracle so that only the host logic (partition, broadcast exchange, ring order,
overwrite-then-accumulate) is under test; the CUDA evaluator is covered by the -m gpu tests."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, kid, Nt, Ns, out_dir, mode):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import loader
    from sctl_b200 import KERNELS, capi, synth
    from sctl_b200.distributed import ShardedDirectSum, partition

    chk = loader.port()
    d = capi.kernel_dims(kid)
    t0, t1 = partition(Nt, rank, world)
    s0, s1 = partition(Ns, rank, world)
    Xt, Xs, Xn, F = synth.make_problem(d, Nt, Ns, np.float64, t0=t0, t1=t1, s0=s0, s1=s1, coincident=8)

    def evaluator(U, xt, xs, xn, f, digits, overwrite):
        u = U.numpy()
        if overwrite:
            u[:] = 0
        chk.eval(kid, xt.numpy(), xs.numpy(), None if xn is None else xn.numpy(), f.numpy(), v_trg=u)
        return U

    T = lambda a: None if a is None else torch.from_numpy(a)
    drv = ShardedDirectSum(KERNELS[kid](), evaluator=evaluator, mode=mode)
    U = torch.full((d["trg_dim"] * (t1 - t0),), 123.0, dtype=torch.float64)  # stale contents must be overwritten
    U = drv(U, T(Xt), T(Xs), T(Xn), T(F))
    U2 = drv(None, T(Xt), T(Xs), T(Xn), T(F), counts=[partition(Ns, r, world)[1] - partition(Ns, r, world)[0] for r in range(world)])
    assert torch.equal(U, U2)
    np.save(os.path.join(out_dir, f"u{rank}.npy"), U.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("kid,Nt,Ns,mode", [(3, 101, 57, "gather"), (4, 64, 34, "gather"), (0, 7, 1, "gather"), (4, 64, 33, "ring"), (3, 50, 58, "ring")])
def test_sharded_direct_sum_world2(tmp_path, port, kid, Nt, Ns, mode):
    """ragged blocks (chained broadcasts), even blocks (one all-gather), an empty block, and the ring mode"""
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), kid, Nt, Ns, str(tmp_path), mode), nprocs=world, join=True)
    from sctl_b200 import capi, synth

    d = capi.kernel_dims(kid)
    Xt, Xs, Xn, F = synth.make_problem(d, Nt, Ns, np.float64, coincident=8)
    want = port.eval(kid, Xt, Xs, Xn, F)
    got = np.concatenate([np.load(tmp_path / f"u{r}.npy") for r in range(world)])
    assert got.shape == want.shape
    assert np.linalg.norm(got - want) <= 1e-13 * np.linalg.norm(want)


def test_partition_formula():
    from sctl_b200.distributed import partition

    for n in (0, 1, 7, 4_000_000):
        for size in (1, 2, 3, 8):
            blocks = [partition(n, r, size) for r in range(size)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            assert all(blocks[i][1] == blocks[i + 1][0] for i in range(size - 1))
            assert max(b - a for a, b in blocks) - min(b - a for a, b in blocks) <= 1


class _OracleRegistry:
    """CPU stand-in for sctl_b200.fmm.ParticleFMM (same methods), evaluating with the oracle: lets the gloo test exercise
    the gather / re-partition logic of ShardedParticleFMM without a GPU."""

    def __init__(self, chk):
        self.chk, self.src, self.trg, self.s2t = chk, {}, {}, {}

    def SetAccuracy(self, digits):
        pass

    def SetKernels(self, *a):
        pass

    def AddSrc(self, name, k1, k2):
        self.src[name] = dict(X=None, Xn=None, F=None, W=None)

    def AddTrg(self, name, k1, k2):
        self.trg[name] = dict(X=None, N=None, dim=k2.TrgDim())

    def SetKernelS2T(self, s, t, ker):
        self.s2t[(s, t)] = ker.KERNEL_ID

    def DeleteSrc(self, name):
        del self.src[name]

    def DeleteTrg(self, name):
        del self.trg[name]

    def SetSrcCoord(self, name, X, Xn=None):
        self.src[name].update(X=np.array(X), Xn=None if Xn is None else np.array(Xn))

    def SetSrcDensity(self, name, F):
        self.src[name]["F"] = np.array(F)

    def SetSrcWeights(self, name, W):
        self.src[name]["W"] = None if W is None else np.array(W)

    def SetTrgCoord(self, name, X):
        self.trg[name]["X"] = np.array(X)

    def SetTrgNormal(self, name, N):
        self.trg[name]["N"] = None if N is None else np.array(N)

    def EvalDirect(self, name, U=None):
        T = self.trg[name]
        nt = T["X"].size // 3
        u = np.zeros(nt * T["dim"])
        for (s, t), kid in self.s2t.items():
            if t != name:
                continue
            S = self.src[s]
            F = S["F"] if S["W"] is None else (S["F"].reshape(S["W"].size, -1) * S["W"][:, None]).reshape(-1)
            if nt and S["X"].size:
                self.chk.eval(kid, T["X"], S["X"], S["Xn"], F, v_trg=u)
        if T["N"] is not None:
            u = np.einsum("tcd,td->tc", u.reshape(nt, -1, 3), T["N"].reshape(nt, 3)).reshape(-1)
        return u


def _fmm_worker(rank, world, port, trg_counts, src_counts, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import loader
    from sctl_b200 import KERNELS, capi, synth
    from sctl_b200.distributed import ShardedParticleFMM

    sl, dl = KERNELS[3](), KERNELS[4]()  # Stokes FxU, DxU
    Nt, Ns = sum(trg_counts), sum(src_counts)
    t0, s0 = sum(trg_counts[:rank]), sum(src_counts[:rank])
    t1, s1 = t0 + trg_counts[rank], s0 + src_counts[rank]
    d = capi.kernel_dims(4)
    Xt, Xs, Xn, F = synth.make_problem(d, Nt, Ns, np.float64, t0=t0, t1=t1, s0=s0, s1=s1, coincident=4)
    rng = np.random.default_rng(5)
    W = rng.random(Ns)[s0:s1]
    Nrm = rng.standard_normal(3 * Nt)[3 * t0:3 * t1]

    fmm = ShardedParticleFMM(np.float64, local=_OracleRegistry(loader.port()), device="cpu")
    fmm.SetKernels(sl, sl, sl)
    fmm.AddTrg("U", sl, sl)
    fmm.AddSrc("SL", sl, sl)
    fmm.AddSrc("DL", dl, dl)
    fmm.SetKernelS2T("SL", "U", sl)
    fmm.SetKernelS2T("DL", "U", dl)
    fmm.SetTrgCoord("U", Xt)
    fmm.SetSrcCoord("SL", Xs)
    fmm.SetSrcCoord("DL", Xs, Xn)
    fmm.SetSrcDensity("SL", F)
    fmm.SetSrcDensity("DL", 2 * F)
    u1 = fmm.Eval("U").numpy().copy()
    fmm.SetSrcWeights("SL", W)
    fmm.SetSrcDensity("DL", -F)  # density update only: coordinates are not gathered again
    fmm.SetTrgNormal("U", Nrm)
    u2 = fmm.EvalDirect("U", torch.full((trg_counts[rank],), 7.0, dtype=torch.float64)).numpy().copy()
    np.save(os.path.join(out_dir, f"a{rank}.npy"), u1)
    np.save(os.path.join(out_dir, f"b{rank}.npy"), u2)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("trg_counts,src_counts", [((40, 41), (30, 30)), ((75, 6), (13, 40)), ((0, 33), (21, 0))])
def test_sharded_particle_fmm_world2(tmp_path, port, trg_counts, src_counts):
    """even targets (stay with their owner), skewed targets (re-partitioned, potentials returned to the owner), empty blocks;
    two source types summed into one target, weights, density update, target-normal contraction"""
    world = 2
    mp.spawn(_fmm_worker, args=(world, _free_port(), trg_counts, src_counts, str(tmp_path)), nprocs=world, join=True)
    from sctl_b200 import capi, synth

    Nt, Ns = sum(trg_counts), sum(src_counts)
    Xt, Xs, Xn, F = synth.make_problem(capi.kernel_dims(4), Nt, Ns, np.float64, coincident=4)
    rng = np.random.default_rng(5)
    W = rng.random(Ns)
    Nrm = rng.standard_normal(3 * Nt)
    want1 = port.eval(3, Xt, Xs, None, F) + port.eval(4, Xt, Xs, Xn, 2 * F)
    full2 = port.eval(3, Xt, Xs, None, (F.reshape(Ns, 3) * W[:, None]).reshape(-1)) + port.eval(4, Xt, Xs, Xn, -F)
    want2 = np.einsum("td,td->t", full2.reshape(Nt, 3), Nrm.reshape(Nt, 3))
    got1 = np.concatenate([np.load(tmp_path / f"a{r}.npy") for r in range(world)])
    got2 = np.concatenate([np.load(tmp_path / f"b{r}.npy") for r in range(world)])
    assert got1.shape == want1.shape and got2.shape == want2.shape
    assert np.linalg.norm(got1 - want1) <= 1e-13 * np.linalg.norm(want1)
    assert np.linalg.norm(got2 - want2) <= 1e-12 * np.linalg.norm(want2)
