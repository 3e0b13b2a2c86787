"""Thread-safety of the C ABI: the reference calls KernelMatrix from inside an OpenMP parallel region
(boundary_integral.txx:949-986) and Eval is const / re-entrant, so concurrent host threads must get the same answers as a
serial run (each thread owns its stream and workspace inside the library)."""
import threading

import numpy as np
import pytest

from sctl_b200 import KERNELS, capi, synth
from util import rel_l2

pytestmark = [pytest.mark.gpu, pytest.mark.usefixtures("eval_path")]


def test_concurrent_eval_and_kernel_matrix_from_host_threads():
    jobs = []
    for i, kid in enumerate([0, 3, 4, 2, 5, 1, 6, 7]):
        d = capi.kernel_dims(kid)
        dt = np.float64 if i % 2 == 0 else np.float32
        Xt, Xs, Xn, F = synth.make_problem(d, 900 + 37 * i, 1500 + 11 * i, dt, seed=i, coincident=16)
        jobs.append((kid, dt, Xt, Xs, Xn, F))
    serial = [(KERNELS[k]().Eval(None, Xt, Xs, Xn, F), KERNELS[k]().KernelMatrix(Xt[:300], Xs[:150], None if Xn is None else Xn[:150]))
              for k, dt, Xt, Xs, Xn, F in jobs]
    out, errs = [None] * len(jobs), []

    def work(i):
        try:
            k, dt, Xt, Xs, Xn, F = jobs[i]
            res = []
            for _ in range(5):
                res.append((KERNELS[k]().Eval(None, Xt, Xs, Xn, F), KERNELS[k]().KernelMatrix(Xt[:300], Xs[:150], None if Xn is None else Xn[:150])))
            out[i] = res
        except Exception as e:  # noqa: BLE001
            errs.append(e)

    th = [threading.Thread(target=work, args=(i,)) for i in range(len(jobs))]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errs, errs
    for i, res in enumerate(out):
        for U, M in res:
            assert np.array_equal(U, serial[i][0]), "Eval must be deterministic and unaffected by concurrent callers"
            assert np.array_equal(M, serial[i][1])
