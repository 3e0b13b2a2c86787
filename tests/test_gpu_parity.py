"""GPU parity tests proper: the CUDA path, called THROUGH THE C ABI (ctypes -> libsctl_b200.so), against the CPU
checker (compiled reference if oracle/_ref is present, else the C port) on identical seeded inputs."""
import numpy as np
import pytest

from sctl_b200 import KERNELS, capi, synth
from util import TOL, rel_l2

pytestmark = [pytest.mark.gpu, pytest.mark.usefixtures("eval_path")]

DTYPES = [np.float64, np.float32]


def _problem(kid, Nt, Ns, dt, seed=0, coincident=64):
    return synth.make_problem(capi.kernel_dims(kid), Nt, Ns, dt, seed=seed, coincident=coincident)


@pytest.mark.parametrize("dt", DTYPES)
@pytest.mark.parametrize("kid", range(8))
@pytest.mark.parametrize("Nt,Ns", [(1, 1), (7, 300), (300, 7), (1000, 1003), (2051, 5000), (4099, 2500), (5003, 700), (15001, 300)])
def test_eval_matches_checker(checker, kid, dt, Nt, Ns):
    Xt, Xs, Xn, F = _problem(kid, Nt, Ns, dt)
    want = checker.eval(kid, Xt, Xs, Xn, F)
    got = KERNELS[kid]().Eval(None, Xt, Xs, Xn, F)
    assert got.dtype == dt and got.shape == want.shape
    assert np.all(np.isfinite(got))
    assert rel_l2(got, want) <= TOL[np.dtype(dt)]


@pytest.mark.parametrize("dt", DTYPES)
@pytest.mark.parametrize("kid", [0, 3, 4])
def test_eval_split_sources_and_accumulate(checker, kid, dt, monkeypatch):
    """many source chunks (partial buffer + deterministic reduce) and += semantics (generic-kernel.txx:98-101)"""
    Xt, Xs, Xn, F = _problem(kid, 3000, 20000, dt)
    K1 = capi.kernel_dims(kid)["trg_dim"]
    U0 = synth.uniform(9, 0, 3000 * K1, dtype=dt)
    want = checker.eval(kid, Xt, Xs, Xn, F, v_trg=U0.copy())
    monkeypatch.setenv("SCTL_B200_FORCE_CHUNKS", "7")
    got = KERNELS[kid]().Eval(U0.copy(), Xt, Xs, Xn, F)
    got2 = KERNELS[kid]().Eval(U0.copy(), Xt, Xs, Xn, F)
    assert rel_l2(got, want) <= TOL[np.dtype(dt)]
    assert np.array_equal(got, got2), "fixed-order reduce must be run-to-run deterministic"
    monkeypatch.setenv("SCTL_B200_FORCE_CHUNKS", "1")
    got1 = KERNELS[kid]().Eval(U0.copy(), Xt, Xs, Xn, F)
    assert rel_l2(got1, want) <= TOL[np.dtype(dt)]
    # overwrite flag ignores previous contents (EvalDirect semantics, fmm-wrapper.txx:501-502)
    got_o = KERNELS[kid]().Eval(U0.copy(), Xt, Xs, Xn, F, overwrite=True)
    assert rel_l2(got_o, checker.eval(kid, Xt, Xs, Xn, F)) <= TOL[np.dtype(dt)]


@pytest.mark.parametrize("kid", range(8))
def test_coincident_points_contribute_zero(kid):
    """r2 == 0 -> exactly 0 (kernel_functions.hpp:28): all targets == all sources == one point"""
    d = capi.kernel_dims(kid)
    for dt in DTYPES:
        Xs = np.tile(np.array([0.25, -0.125, 0.5], dtype=dt), 300)
        Xn = np.ones(300 * d["normal_dim"], dtype=dt) if d["normal_dim"] else None
        F = np.ones(300 * d["src_dim"], dtype=dt)
        U = KERNELS[kid]().Eval(None, Xs[: 3 * 40], Xs, Xn, F)
        assert np.all(U == 0)


@pytest.mark.parametrize("dt", DTYPES)
@pytest.mark.parametrize("kid", range(8))
@pytest.mark.parametrize("digits", [3, 4, 5, 6, 8, 10, 11, 12])
def test_eval_digits(checker, kid, dt, digits):
    """lower digits only relax the rsqrt refinement; the result must still carry `digits` digits - in rel-L2 AND for the worst
    single target (the rinv^5 kernels amplify the rsqrt error 5x). 4|5, 5|6 (f32) and 10|11 are the boundaries of the
    digits -> refinement table (newton_level, capi.cu)."""
    Xt, Xs, Xn, F = _problem(kid, 500, 2000, dt)
    want = checker.eval(kid, Xt, Xs, Xn, F)
    got = KERNELS[kid]().Eval(None, Xt, Xs, Xn, F, digits=digits)
    tol = max(10.0 ** (-digits), TOL[np.dtype(dt)])
    assert rel_l2(got, want) <= tol
    assert float(np.max(np.abs(got - want)) / np.max(np.abs(want))) <= tol


@pytest.mark.parametrize("dt", DTYPES)
@pytest.mark.parametrize("kid", range(8))
@pytest.mark.parametrize("Nt,Ns", [(1, 37), (37, 1), (130, 67), (33, 200), (516, 40), (1031, 33)])
def test_kernel_matrix_matches_checker(checker, kid, dt, Nt, Ns):
    Xt, Xs, Xn, _ = _problem(kid, Nt, Ns, dt, coincident=8)
    want = checker.kernel_matrix(kid, Xt, Xs, Xn)
    got = KERNELS[kid]().KernelMatrix(Xt, Xs, Xn)
    assert got.shape == want.shape
    assert rel_l2(got, want) <= TOL[np.dtype(dt)]
    # U = F^T M consistency with Eval (generic-kernel.txx:85-89 vs :284)
    d = capi.kernel_dims(kid)
    F = synth.uniform(4, 0, Ns * d["src_dim"], dtype=dt)
    U = KERNELS[kid]().Eval(None, Xt, Xs, Xn, F)
    assert rel_l2(U, F.astype(np.float64) @ got.astype(np.float64)) <= 20 * TOL[np.dtype(dt)]


def test_device_pointer_entry(checker):
    import torch

    kid = 3
    Xt, Xs, Xn, F = _problem(kid, 5000, 6000, np.float64)
    want = checker.eval(kid, Xt, Xs, Xn, F)
    dev = torch.device("cuda:0")
    t = lambda a: torch.from_numpy(a).to(dev)
    U = KERNELS[kid]().Eval(None, t(Xt), t(Xs), None, t(F))
    torch.cuda.synchronize()
    assert rel_l2(U.cpu().numpy(), want) <= 1e-12
    U2 = KERNELS[kid]().Eval(U.clone(), t(Xt), t(Xs), None, t(F))  # accumulates
    torch.cuda.synchronize()
    assert rel_l2(U2.cpu().numpy(), 2 * want) <= 1e-12


def test_launch_counter_moves():
    n0 = capi.launch_count()
    Xt, Xs, Xn, F = _problem(0, 10, 10, np.float64)
    KERNELS[0]().Eval(None, Xt, Xs, Xn, F)
    assert capi.launch_count() >= n0 + 1  # the small path is ONE launch, the tiled path two or three


@pytest.mark.parametrize("kid,dt", [(3, np.float64), (1, np.float32)])
def test_eval_all_gpus_flag(checker, kid, dt):
    """SCTL_B200_ALL_GPUS: in-process target sharding over every visible device with the partition formula of
    fmm-wrapper.txx:507 (falls back to one device when only one is visible); accumulate semantics are kept per shard."""
    Xt, Xs, Xn, F = _problem(kid, 7001, 3000, dt)
    K1 = capi.kernel_dims(kid)["trg_dim"]
    U0 = synth.uniform(9, 0, 7001 * K1, dtype=dt)
    want = checker.eval(kid, Xt, Xs, Xn, F, v_trg=U0.copy())
    got = KERNELS[kid]().Eval(U0.copy(), Xt, Xs, Xn, F, all_gpus=True)
    assert rel_l2(got, want) <= TOL[np.dtype(dt)]


def test_init_finalize(checker):
    """sctl_b200_init selects the devices SCTL_B200_ALL_GPUS shards over; finalize releases the thread's streams and staging
    blocks; every entry point re-initialises lazily afterwards."""
    import ctypes

    L = capi.lib()
    n = capi.device_count()
    ids = (ctypes.c_int * 1)(0)
    capi.check(L.sctl_b200_init(1, ids))
    Xt, Xs, Xn, F = _problem(3, 5000, 9000, np.float64)
    want = checker.eval(3, Xt, Xs, Xn, F)
    got = KERNELS[3]().Eval(None, Xt, Xs, Xn, F, all_gpus=True)  # one selected device
    assert rel_l2(got, want) <= TOL[np.dtype(np.float64)]
    capi.check(L.sctl_b200_finalize())
    got = KERNELS[3]().Eval(None, Xt, Xs, Xn, F)  # lazily re-initialised
    assert rel_l2(got, want) <= TOL[np.dtype(np.float64)]
    bad = (ctypes.c_int * 1)(n)
    assert L.sctl_b200_init(1, bad) != 0 and b"out of range" in L.sctl_b200_last_error()
    twice = (ctypes.c_int * 2)(0, 0)
    assert L.sctl_b200_init(2, twice) != 0
    capi.check(L.sctl_b200_init(0, None))  # back to "every visible device"
    got = KERNELS[3]().Eval(None, Xt, Xs, Xn, F, all_gpus=True)
    assert rel_l2(got, want) <= TOL[np.dtype(np.float64)]
    capi.check(L.sctl_b200_finalize())
