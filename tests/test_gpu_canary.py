"""Out-of-bounds canaries (compute-sanitizer is closed on this pool): outputs live in the middle of larger device tensors whose
guard bands must be untouched after every call, at ragged sizes that exercise the tails of every tile shape."""
import ctypes

import numpy as np
import pytest

from sctl_b200 import KERNELS, capi, synth

pytestmark = [pytest.mark.gpu, pytest.mark.usefixtures("eval_path")]
GUARD = 4096


def _guarded(n, dt, dev):
    import torch

    tdt = {np.dtype(np.float32): torch.float32, np.dtype(np.float64): torch.float64}[np.dtype(dt)]
    buf = torch.full((n + 2 * GUARD,), 777.0, dtype=tdt, device=dev)
    return buf, buf[GUARD:GUARD + n]


def _check(buf, n):
    assert bool((buf[:GUARD] == 777.0).all()) and bool((buf[GUARD + n:] == 777.0).all()), "guard band overwritten"


@pytest.mark.parametrize("dt", [np.float64, np.float32])
@pytest.mark.parametrize("kid", range(8))
def test_outputs_stay_in_bounds(kid, dt):
    import torch

    dev = torch.device("cuda:0")
    d = capi.kernel_dims(kid)
    ker = KERNELS[kid]()
    t = lambda a: None if a is None else torch.from_numpy(a).to(dev)
    for Nt, Ns in ((1, 1), (767, 255), (769, 257), (2049, 513), (5, 3000)):
        Xt, Xs, Xn, F = synth.make_problem(d, Nt, Ns, dt, coincident=4)
        n = Nt * d["trg_dim"]
        buf, U = _guarded(n, dt, dev)
        U.zero_()
        ker.Eval(U, t(Xt), t(Xs), t(Xn), t(F))
        torch.cuda.synchronize()
        _check(buf, n)
        assert bool(torch.isfinite(U).all())
    Nt, Ns = 131, 67
    Xt, Xs, Xn, _ = synth.make_problem(d, Nt, Ns, dt, coincident=4)
    n = Ns * d["src_dim"] * Nt * d["trg_dim"]
    buf, M = _guarded(n, dt, dev)
    ker.KernelMatrix(t(Xt), t(Xs), t(Xn), out=M.view(Ns * d["src_dim"], Nt * d["trg_dim"]))
    torch.cuda.synchronize()
    _check(buf, n)
    trg_dsp, src_dsp = np.array([0, 131], np.int64), np.array([0, 67], np.int64)
    buf2, Mb = _guarded(n, dt, dev)
    ker.KernelMatrixBatch(trg_dsp, src_dsp, t(Xt), t(Xs), t(Xn), out=Mb)
    torch.cuda.synchronize()
    _check(buf2, n)
    assert torch.equal(Mb, M)
