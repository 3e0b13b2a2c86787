import numpy as np

from sctl_b200 import synth


def test_counter_based_slices_are_consistent():
    a = synth.uniform(2, 0, 1000)
    b = np.concatenate([synth.uniform(2, 0, 400), synth.uniform(2, 400, 600)])
    assert np.array_equal(a, b)
    assert a.min() >= -0.5 and a.max() < 0.5 and abs(a.mean()) < 0.05
    assert not np.array_equal(a, synth.uniform(3, 0, 1000))


def test_problem_shards_match_global():
    d = dict(normal_dim=3, src_dim=3)
    Xt, Xs, Xn, F = synth.make_problem(d, 3000, 2000, np.float64)
    xt, xs, xn, f = synth.make_problem(d, 3000, 2000, np.float64, t0=500, t1=1500, s0=100, s1=900)
    assert np.array_equal(xt, Xt[1500:4500]) and np.array_equal(xs, Xs[300:2700])
    assert np.array_equal(xn, Xn[300:2700]) and np.array_equal(f, F[300:2700])
    assert np.array_equal(Xt[: 3 * 1024], Xs[: 3 * 1024])  # coincident block
