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
