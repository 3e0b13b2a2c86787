import glob
import os

import numpy as np

from sctl_b200 import synth

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TAGS = {"f64": np.float64, "f32": np.float32}


def kernel_files():
    fs = sorted(glob.glob(os.path.join(GOLDEN, "kernel*_*.npz")))
    assert len(fs) == 8, "golden fixtures missing: run tests/tools/make_golden.py in the authoring container"
    return fs


def load(kid):
    g = np.load(kernel_files()[kid])
    d = dict(coord_dim=int(g["dims"][0]), normal_dim=int(g["dims"][1]), src_dim=int(g["dims"][2]), trg_dim=int(g["dims"][3]),
             flops=int(g["dims"][4]), scale=float(g["scale"]), name=str(g["name"]))
    return g, d


def inputs(g, d, kid, ci, dt):
    Nt, Ns = (int(v) for v in g["cases"][ci])
    return synth.make_problem(d, Nt, Ns, dt, seed=100 + kid, coincident=int(g["coincident"]))
