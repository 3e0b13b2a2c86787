"""Generates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref/libsctl_ref_*.so, i.e. /root/reference's
own GenericKernel::Eval / KernelMatrix / ParticleFMM::EvalDirect / approx_rsqrt compiled by oracle/Makefile).

Run in the authoring container (needs /root/reference to have built oracle/_ref):  python tests/tools/make_golden.py
Inputs are NOT stored: they are regenerated from sctl_b200.synth with the recorded (Nt, Ns, seed, coincident).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import loader  # noqa: E402
from sctl_b200 import synth  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
CASES = [(1, 1), (5, 9), (33, 70), (131, 257)]  # (Nt, Ns): below / across the reference's VecLen fast path (generic-kernel.txx:104)
COINCIDENT = 4


def main():
    ref = loader.reference()
    assert ref is not None, "build oracle/_ref first (make -C oracle ref)"
    os.makedirs(OUT, exist_ok=True)
    for kid in range(8):
        d = ref.kernel_dims(kid)
        out = {"dims": np.array([d["coord_dim"], d["normal_dim"], d["src_dim"], d["trg_dim"], d["flops"]]), "scale": np.array(d["scale"]),
               "name": np.array(d["name"]), "cases": np.array(CASES), "coincident": np.array(COINCIDENT)}
        for dt, tag in ((np.float64, "f64"), (np.float32, "f32")):
            for ci, (Nt, Ns) in enumerate(CASES):
                Xt, Xs, Xn, F = synth.make_problem(d, Nt, Ns, dt, seed=100 + kid, coincident=COINCIDENT)
                out[f"eval_{tag}_{ci}"] = ref.eval(kid, Xt, Xs, Xn, F, omp=True)
                U0 = synth.uniform(9, 0, Nt * d["trg_dim"], seed=kid, dtype=dt)
                out[f"evalacc_{tag}_{ci}"] = ref.eval(kid, Xt, Xs, Xn, F, v_trg=U0.copy(), omp=False)  # += semantics, serial path
                out[f"direct_{tag}_{ci}"] = ref.eval_direct(kid, Xt, Xs, Xn, F)
                if Nt * Ns <= 33 * 70:
                    out[f"kmat_{tag}_{ci}"] = ref.kernel_matrix(kid, Xt, Xs, Xn)
        np.savez_compressed(os.path.join(OUT, f"kernel{kid}_{d['name']}.npz"), **out)
        print("wrote", d["name"])
    # approx_rsqrt<digits>(x, x > 0): vec-test.hpp:384-418 pins digits 4 and 7; we also record full precision
    x64 = np.concatenate([[0.0], np.exp(synth.uniform(7, 0, 255) * 40.0)])
    x32 = x64.astype(np.float32)
    np.savez_compressed(os.path.join(OUT, "rsqrt.npz"), x64=x64, x32=x32, **{f"y64_d{dg}": ref.rsqrt(x64, dg) for dg in (4, 7, -1)},
                        **{f"y32_d{dg}": ref.rsqrt(x32, dg) for dg in (4, 7, -1)})
    print("veclen f64/f32:", ref.veclen(np.float64), ref.veclen(np.float32), "lib:", ref.path)


if __name__ == "__main__":
    main()
