"""Batched small-block KernelMatrix (sctl_b200_kernel_matrix_batch) against per-block KernelMatrix of the CPU checker, with
the fused epilogues of the reference's call site (BoundaryIntegralOp::SetupNear, boundary_integral.txx:946-1008):
source-weight row scaling (:990-997) and target-normal contraction (:968-984)."""
import numpy as np
import pytest

from sctl_b200 import KERNELS, capi, synth
from util import TOL, rel_l2

pytestmark = pytest.mark.gpu


def _batch(rng, nb, max_t, max_s):
    nt = rng.integers(0, max_t, nb)
    ns = rng.integers(0, max_s, nb)
    nt[0], ns[0] = max_t, max_s  # one block larger than a CTA tile in both directions
    if nb > 3:
        nt[1] = 0  # empty blocks
        ns[2] = 0
    return np.concatenate([[0], np.cumsum(nt)]).astype(np.int64), np.concatenate([[0], np.cumsum(ns)]).astype(np.int64)


def _want(checker, kid, dt, trg_dsp, src_dsp, Xt, Xs, Xn, wts, Nt_n):
    d = capi.kernel_dims(kid)
    K0, K1, Nd = d["src_dim"], d["trg_dim"], d["normal_dim"]
    blocks = []
    for b in range(trg_dsp.size - 1):
        t0, t1, s0, s1 = trg_dsp[b], trg_dsp[b + 1], src_dsp[b], src_dsp[b + 1]
        if t1 == t0 or s1 == s0:
            blocks.append(np.zeros(0, dt))
            continue
        M = checker.kernel_matrix(kid, Xt[3 * t0:3 * t1].copy(), Xs[3 * s0:3 * s1].copy(), None if Xn is None else Xn[Nd * s0:Nd * s1].copy()).astype(np.float64)
        if wts is not None:
            M = M * np.repeat(wts[s0:s1].astype(np.float64), K0)[:, None]
        if Nt_n is not None:  # M[., (t*K1'+k)*3+l] * n[t,l]
            n = Nt_n[3 * t0:3 * t1].astype(np.float64).reshape(-1, 1, 3)
            M = (M.reshape(M.shape[0], t1 - t0, K1 // 3, 3) * n[None]).sum(-1).reshape(M.shape[0], -1)
        blocks.append(M.reshape(-1))
    return blocks


@pytest.mark.parametrize("dt", [np.float64, np.float32])
@pytest.mark.parametrize("kid", range(8))
def test_batch_matches_per_block_kernel_matrix(checker, kid, dt):
    rng = np.random.default_rng(kid)
    trg_dsp, src_dsp = _batch(rng, 37, 150, 70)
    d = capi.kernel_dims(kid)
    Xt = synth.uniform(1, 0, 3 * trg_dsp[-1], dtype=dt)
    Xs = synth.uniform(2, 0, 3 * src_dsp[-1], dtype=dt)
    Xn = synth.uniform(3, 0, 3 * src_dsp[-1], dtype=dt) if d["normal_dim"] else None
    wts = synth.uniform(5, 0, src_dsp[-1], dtype=dt) + dt(1.0)
    ker = KERNELS[kid]()
    for use_w in (False, True):
        M, M_dsp = ker.KernelMatrixBatch(trg_dsp, src_dsp, Xt, Xs, Xn, src_wts=wts if use_w else None)
        want = _want(checker, kid, dt, trg_dsp, src_dsp, Xt, Xs, Xn, wts if use_w else None, None)
        assert M.size == sum(w.size for w in want)
        for b, w in enumerate(want):
            if w.size:
                assert rel_l2(M[M_dsp[b]:M_dsp[b] + w.size], w) <= TOL[np.dtype(dt)], (b, use_w)


@pytest.mark.parametrize("kid", [2, 5])  # TrgDim 3 and 9
def test_batch_target_normal_contraction(checker, kid):
    dt = np.float64
    rng = np.random.default_rng(100 + kid)
    trg_dsp, src_dsp = _batch(rng, 11, 140, 40)
    Xt = synth.uniform(1, 0, 3 * trg_dsp[-1], dtype=dt)
    Xs = synth.uniform(2, 0, 3 * src_dsp[-1], dtype=dt)
    Nt_n = synth.uniform(7, 0, 3 * trg_dsp[-1], dtype=dt)
    wts = synth.uniform(5, 0, src_dsp[-1], dtype=dt) + 1.0
    M, M_dsp = KERNELS[kid]().KernelMatrixBatch(trg_dsp, src_dsp, Xt, Xs, None, src_wts=wts, n_trg=Nt_n)
    want = _want(checker, kid, dt, trg_dsp, src_dsp, Xt, Xs, None, wts, Nt_n)
    for b, w in enumerate(want):
        if w.size:
            assert rel_l2(M[M_dsp[b]:M_dsp[b] + w.size], w) <= 1e-12, b
    with pytest.raises(ValueError):  # Laplace FxU (TrgDim 1) cannot be contracted with a target normal
        KERNELS[0]().KernelMatrixBatch(trg_dsp, src_dsp, Xt, Xs, None, n_trg=Nt_n)


def test_batch_device_pointers_and_launch_count(checker):
    import torch

    kid, dt = 3, np.float64
    rng = np.random.default_rng(5)
    trg_dsp, src_dsp = _batch(rng, 200, 60, 30)
    Xt = synth.uniform(1, 0, 3 * trg_dsp[-1], dtype=dt)
    Xs = synth.uniform(2, 0, 3 * src_dsp[-1], dtype=dt)
    dev = torch.device("cuda:0")
    n0 = capi.launch_count()
    M, M_dsp = KERNELS[kid]().KernelMatrixBatch(trg_dsp, src_dsp, torch.from_numpy(Xt).to(dev), torch.from_numpy(Xs).to(dev))
    assert capi.launch_count() - n0 == 1, "200 blocks must be ONE launch"
    want = _want(checker, kid, dt, trg_dsp, src_dsp, Xt, Xs, None, None, None)
    Mh = M.cpu().numpy()
    for b, w in enumerate(want):
        if w.size:
            assert rel_l2(Mh[M_dsp[b]:M_dsp[b] + w.size], w) <= 1e-12


def test_batch_gaps_in_M_keep_the_callers_values(checker):
    """host pointers: "entries of M not covered by any block are left untouched" (include/sctl_b200.h) - blocks placed with
    gaps between them in a canary-filled M; and the gap-free layout (blocks tile M exactly, upload of M skipped) gives the same blocks"""
    kid, dt = 4, np.float64
    rng = np.random.default_rng(9)
    trg_dsp, src_dsp = _batch(rng, 9, 50, 20)
    d = capi.kernel_dims(kid)
    Xt = synth.uniform(1, 0, 3 * trg_dsp[-1], dtype=dt)
    Xs = synth.uniform(2, 0, 3 * src_dsp[-1], dtype=dt)
    Xn = synth.uniform(3, 0, 3 * src_dsp[-1], dtype=dt)
    ker = KERNELS[kid]()
    tight, tight_dsp = ker.KernelMatrixBatch(trg_dsp, src_dsp, Xt, Xs, Xn)
    sizes = np.diff(src_dsp) * d["src_dim"] * np.diff(trg_dsp) * d["trg_dim"]
    gap = 17
    M_dsp = (np.concatenate([[0], np.cumsum(sizes + gap)[:-1]]) + 5).astype(np.int64)
    M = np.full(int((M_dsp + sizes).max()) + 11, -7.25, dtype=dt)
    out, _ = ker.KernelMatrixBatch(trg_dsp, src_dsp, Xt, Xs, Xn, out=M, M_dsp=M_dsp)
    covered = np.zeros(M.size, bool)
    for b in range(sizes.size):
        covered[M_dsp[b]:M_dsp[b] + sizes[b]] = True
        assert np.array_equal(out[M_dsp[b]:M_dsp[b] + sizes[b]], tight[tight_dsp[b]:tight_dsp[b] + sizes[b]]), b
    assert np.all(out[~covered] == -7.25) and (~covered).sum() >= 5 + 11
