"""sctl_b200_eval_noalloc (include/sctl_b200.h, sctl_b200/csrc/noalloc.cu): the allocation-free variant of the device-pointer
potential evaluation. It has no counterpart in the reference, so the checker here
is the EXISTING entry point sctl_b200_eval_async on the same device buffers - both through the C ABI, device memory from the ABI's
own sctl_b200_buf_* calls - within the tolerances the parity tests use (rel-L2 1e-12 f64 / 1e-5 f32), plus the CPU oracle.

The file sorts last on purpose: it must pass alone in a fresh process and after everything else.
"""
import ctypes

import numpy as np
import pytest

from sctl_b200 import capi, synth

TOL = {np.dtype(np.float64): 1e-12, np.dtype(np.float32): 1e-5}
DT = {np.dtype(np.float32): capi.F32, np.dtype(np.float64): capi.F64}


class DevBuf:
    """device copy of a host array through sctl_b200_buf_* (this is the only allocation in the test, and it is the caller's)"""

    def __init__(self, a):
        self.a = np.ascontiguousarray(a)
        self.p = ctypes.c_void_p()
        capi.check(capi.lib().sctl_b200_buf_create(ctypes.byref(self.p), DT[self.a.dtype], max(self.a.size, 1)))
        self.put(self.a)

    def put(self, a):
        a = np.ascontiguousarray(a, dtype=self.a.dtype)
        assert a.size == self.a.size
        if a.size:
            capi.check(capi.lib().sctl_b200_buf_upload(self.p, a.ctypes.data_as(ctypes.c_void_p), DT[a.dtype], a.size))

    def get(self):
        out = np.empty_like(self.a)
        if out.size:
            capi.check(capi.lib().sctl_b200_buf_download(self.p, out.ctypes.data_as(ctypes.c_void_p), DT[out.dtype], out.size))
        return out

    def free(self):
        if self.p:
            capi.check(capi.lib().sctl_b200_buf_destroy(self.p))
            self.p = ctypes.c_void_p()


def rel_l2(got, want):
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    den = np.linalg.norm(want)
    return float(np.linalg.norm(got - want) / den) if den > 0 else float(np.linalg.norm(got - want))


def run_both(kid, dt, nt, ns, digits=-1, accumulate=False, checker=None, coincident=16, seed_shift=0):
    """one problem through the new entry point and through sctl_b200_eval_async; returns (no-alloc result, async result)"""
    L = capi.lib()
    dt = np.dtype(dt)
    d = capi.kernel_dims(kid)
    Xt, Xs, Xn, F = synth.make_problem(d, nt, ns, dt, coincident=min(coincident, nt, ns))
    if seed_shift:
        F = (F * (1 + 0.125 * seed_shift)).astype(dt)
    U0 = (0.25 * np.cos(np.arange(nt * d["trg_dim"]))).astype(dt)  # what v_trg holds before an accumulating call
    bufs = [DevBuf(Xt), DevBuf(Xs), DevBuf(Xn) if Xn is not None else None, DevBuf(F), DevBuf(U0), DevBuf(U0)]
    rt, rs, rn, vs, u_new, u_old = bufs
    flags = 0 if accumulate else capi.OVERWRITE
    pn = rn.p if rn is not None else None
    try:
        l0 = capi.launch_count()
        capi.check(L.sctl_b200_eval_noalloc(kid, DT[dt], nt, ns, rt.p, rs.p, pn, vs.p, u_new.p, digits, flags, None))
        launched = capi.launch_count() - l0
        capi.check(L.sctl_b200_eval_async(kid, DT[dt], nt, ns, rt.p, rs.p, pn, vs.p, u_old.p, digits, flags, None))
        got, want = u_new.get(), u_old.get()  # the downloads synchronise the device
    finally:
        for b in bufs:
            if b is not None:
                b.free()
    assert launched >= 2, "pack + eval_kernel (+ reduce) must have been launched by the new entry point"
    assert np.all(np.isfinite(got))
    err = rel_l2(got, want)
    assert err <= TOL[dt], (d["name"], dt, nt, ns, digits, accumulate, err)
    if checker is not None and digits == -1 and nt * ns <= 80_000_000:
        ref = checker.eval(kid, Xt, Xs, Xn, F)
        if accumulate:
            ref = ref + U0.astype(np.float64).reshape(ref.shape)
        assert rel_l2(got.reshape(ref.shape), ref) <= TOL[dt], (d["name"], dt, nt, ns)
    return got, want


def _need_gpu():
    if capi.device_count() < 1:
        pytest.fail("no CUDA device: the gpu-marked tests need one (sctl_b200 has no CPU fallback)")
    capi.check(capi.lib().sctl_b200_set_device(0))


@pytest.mark.gpu
def test_two_calls_in_a_row_match_the_existing_entry_point(checker):
    """Two DIFFERENT problems back to back on the shared static block: the second (smaller, another kernel and dtype) runs over
    whatever the first left there; then the first again, which must reproduce its own earlier result bit for bit."""
    _need_gpu()
    a1, _ = run_both(capi_id("Stokes3D-FxU"), np.float64, 6007, 9001, checker=checker)
    run_both(capi_id("Laplace3D-FxdU"), np.float32, 1501, 2049, checker=checker)
    a2, _ = run_both(capi_id("Stokes3D-FxU"), np.float64, 6007, 9001, checker=checker)
    assert np.array_equal(a1, a2), "same inputs, same plan: the result may not depend on what the block held before"


def capi_id(name):
    for k in range(8):
        if capi.kernel_dims(k)["name"] == name:
            return k
    raise KeyError(name)


@pytest.mark.gpu
@pytest.mark.parametrize("dt", [np.float64, np.float32], ids=["f64", "f32"])
@pytest.mark.parametrize("kid", range(8))
def test_every_kernel_accumulating_and_overwriting(kid, dt, checker):
    _need_gpu()
    run_both(kid, dt, 3001, 2600, accumulate=False, checker=checker)
    run_both(kid, dt, 3001, 2600, accumulate=True, checker=checker, seed_shift=1)


@pytest.mark.gpu
@pytest.mark.parametrize("nt,ns", [(1, 1), (5, 300), (700, 33), (257, 256), (129, 257), (40000, 20000)])
def test_ragged_and_small_shapes_take_the_tiled_path(nt, ns, checker):
    """the existing entry point routes small shapes to its one-launch kernel; the new one always runs pack -> eval -> reduce"""
    _need_gpu()
    for kid, dt in ((3, np.float64), (1, np.float32)):
        run_both(kid, dt, nt, ns, checker=checker, coincident=4)


@pytest.mark.gpu
@pytest.mark.parametrize("digits", [3, 8, 12])
def test_digits_levels(digits):
    _need_gpu()
    run_both(3, np.float64, 4000, 9000, digits=digits)
    run_both(0, np.float32, 4000, 9000, digits=min(digits, 7))


@pytest.mark.gpu
def test_largest_accepted_source_set_and_many_targets(checker):
    """Ns at the documented limit for the widest record (Stokes3D-DxU f64: 80 bytes -> exactly SCTL_B200_NOALLOC_RECORD_BYTES)
    with few targets, and a target set large enough that the partial sums do not fit at the planner's chunk count."""
    _need_gpu()
    run_both(4, np.float64, 300, capi.NOALLOC_MAX_SRC)
    run_both(3, np.float64, 400_000, 30_000)  # 9.6 MB of partial sums per chunk: at most 3 chunks fit beside the records


@pytest.mark.gpu
def test_sizes_beyond_the_static_block_are_rejected_and_nothing_is_launched():
    _need_gpu()
    L = capi.lib()
    kid, dt = 4, np.dtype(np.float64)
    ns = capi.NOALLOC_MAX_SRC + 1
    z = DevBuf(np.zeros(ns * 3, dt))
    u = DevBuf(np.full(30, 7.0, dt))
    try:
        l0 = capi.launch_count()
        rc = L.sctl_b200_eval_noalloc(kid, capi.F64, 10, ns, u.p, z.p, z.p, z.p, u.p, -1, capi.OVERWRITE, None)
        assert rc == capi.ERR_ARG
        msg = L.sctl_b200_last_error().decode()
        assert "sctl_b200_eval_noalloc" in msg and str(capi.NOALLOC_MAX_SRC) in msg and "sctl_b200_eval_async" in msg, msg
        assert capi.launch_count() == l0
        assert np.all(u.get() == 7.0)
        # Stokes FxU f64 has 48-byte records: 524288 sources fit, one more does not
        assert L.sctl_b200_eval_noalloc(3, capi.F64, 0, 524288 + 1, None, None, None, None, None, -1, capi.OVERWRITE, None) == 0  # Nt = 0: nothing to do
        rc = L.sctl_b200_eval_noalloc(3, capi.F64, 10, 524288 + 1, u.p, z.p, None, z.p, u.p, -1, capi.OVERWRITE, None)
        assert rc == capi.ERR_ARG and "524288" in L.sctl_b200_last_error().decode()
        # flags other than OVERWRITE / DEVICE_PTRS are refused
        assert L.sctl_b200_eval_noalloc(3, capi.F64, 10, 10, u.p, z.p, None, z.p, u.p, -1, capi.ALL_GPUS, None) == capi.ERR_ARG
        assert L.sctl_b200_eval_noalloc(9, capi.F64, 10, 10, u.p, z.p, None, z.p, u.p, -1, 0, None) == capi.ERR_KERNEL
        assert L.sctl_b200_eval_noalloc(3, capi.F64, 10, 10, None, z.p, None, z.p, u.p, -1, 0, None) == capi.ERR_ARG
    finally:
        z.free()
        u.free()


def test_argument_checks_need_no_gpu():
    """the checks that come before any CUDA call"""
    L = capi.lib()
    assert L.sctl_b200_eval_noalloc(9, capi.F64, 1, 1, None, None, None, None, None, -1, 0, None) == capi.ERR_KERNEL
    assert L.sctl_b200_eval_noalloc(3, 7, 1, 1, None, None, None, None, None, -1, 0, None) == capi.ERR_DTYPE
    assert L.sctl_b200_eval_noalloc(3, capi.F64, -1, 1, None, None, None, None, None, -1, 0, None) == capi.ERR_ARG
    assert L.sctl_b200_eval_noalloc(3, capi.F64, 1, 1, None, None, None, None, None, -1, 0, None) == capi.ERR_ARG  # null pointers
    assert capi.NOALLOC_MAX_SRC == capi.NOALLOC_RECORD_BYTES // 80 // 256 * 256
    # the binding's copies of the limits are the header's
    import os
    import re

    hdr = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "sctl_b200.h")).read()
    defs = {k: v for k, v in re.findall(r"#define\s+(SCTL_B200_NOALLOC_\w+)\s+\(?([^)\n]+)\)?", hdr)}
    val = lambda t: eval(t.replace("u", ""))  # "32u << 20" / "314368"
    assert val(defs["SCTL_B200_NOALLOC_SCRATCH_BYTES"]) == capi.NOALLOC_SCRATCH_BYTES
    assert val(defs["SCTL_B200_NOALLOC_RECORD_BYTES"]) == capi.NOALLOC_RECORD_BYTES
    assert val(defs["SCTL_B200_NOALLOC_MAX_SRC"]) == capi.NOALLOC_MAX_SRC
