"""The reference-side binding (include/sctl_b200/generic-kernel-b200.hpp) exercised by the reference's own C++ call
sites: tests/cpp/shim_test is compiled against the UNMODIFIED reference headers where /root/reference exists (the
authoring container, by __graft_entry__.build()) and travels to the GPU box as a built file."""
import os
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
BINS = [os.path.join(HERE, "cpp", "shim_test"), os.path.join(HERE, "cpp", "shim_test_memdebug")]


def _need(path):
    if not os.path.exists(path):
        pytest.skip(f"{os.path.basename(path)} not built (needs /root/reference; run __graft_entry__.build())")


@pytest.mark.parametrize("path", BINS, ids=["plain", "memdebug"])
def test_shim_fails_loudly_without_gpu(path):
    """no CPU fallback: without a usable device the reference's Eval must abort through SCTL_ERROR (common.hpp:53-70)"""
    _need(path)
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present; the loud-failure path is checked on the CPU box")
    p = subprocess.run([path, "--no-gpu"], capture_output=True, text=True, timeout=120)
    assert p.returncode != 0, p.stdout
    assert "no CPU fallback" in (p.stderr + p.stdout)
    assert "UNEXPECTED" not in p.stdout


@pytest.mark.gpu
def test_registry_wrapper_aborts_on_a_short_normal_vector():
    """sctl_b200::ParticleFMM::SetSrcCoord with 12 normal values for 10 sources must abort like the reference's assert
    (fmm-wrapper.txx:535) - the C ABI has no length argument for the normals"""
    _need(BINS[0])
    p = subprocess.run([BINS[0], "--bad-normal"], capture_output=True, text=True, timeout=120)
    assert p.returncode != 0, p.stdout
    assert "UNEXPECTED" not in p.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("path", BINS, ids=["plain", "memdebug"])
def test_shim_matches_reference_cpu_templates(path):
    """all 8 kernels x f32/f64: Eval, accumulate, type-erased static Eval, KernelMatrix shapes, ParticleFMM::EvalDirect"""
    _need(path)
    p = subprocess.run([path], capture_output=True, text=True, timeout=900)
    print(p.stdout[-6000:])
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-2000:]
    assert "SHIM TEST PASSED" in p.stdout


@pytest.mark.gpu
def test_nccl_sharding_driver_matches_reference_eval_direct():
    """tests/cpp/sharded_test: sctl_b200_comm_* / sctl_b200_eval_sharded / sctl_b200::ParticleFMM + SetComm with one rank per
    visible GPU (ranks = threads of one process, ncclCommInitRank) against the reference's single-rank
    ParticleFMM::EvalDirect (fmm-wrapper.txx:490-562) of the global problem; also SCTL_B200_ALL_GPUS. On a 1-GPU box the
    communicator has one rank; scripts/gpu_multi.sh runs the same binary on 2/4/8 GPUs (profiles/r02_sharded_test_*.txt)."""
    path = os.path.join(HERE, "cpp", "sharded_test")
    _need(path)
    p = subprocess.run([path], capture_output=True, text=True, timeout=900)
    print(p.stdout[-6000:])
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-2000:]
    assert "SHARDED TEST PASSED" in p.stdout


BIO_BINS = [os.path.join(HERE, "cpp", "bio_test"), os.path.join(HERE, "cpp", "bio_test_patched")]


@pytest.mark.gpu
@pytest.mark.parametrize("path,args", [(BIO_BINS[0], ["8", "2"]), (BIO_BINS[1], ["8", "2"]), (BIO_BINS[0], ["--near-on-device", "5", "1"])],
                         ids=["unmodified-reference+shim", "patched-reference", "shim+tiny-call-stress"])
def test_boundary_integral_op_runs_on_the_library(path, args):
    """The reference's BoundaryIntegralOp::Setup / ComputePotential (boundary_integral.txx:608-614: SetupNear's KernelMatrix calls
    inside its own omp parallel for :949-986, ComputeFarField -> fmm.Eval :1016-1077) with a test-only ElementListBase
    (tests/cpp/sphere_patch_list.hpp): GPU arm vs the reference's CPU templates, rel-L2 <= 1e-11, for Laplace FxU, Stokes DxU
    (on- and off-surface targets), Laplace FxdU and Stokes FxT with the target-normal contraction, Stokes FxU f32. `patched` =
    the same test built against a copy of boundary_integral.{hpp,txx} with tests/cpp/boundary_integral_b200.patch applied
    (batched SetupNear, device-resident registry with fused weights / normals). `tiny-call-stress` also routes the element
    list's own one-target near-quadrature KernelMatrix calls through the library (~1e5 concurrent calls from the reference's omp loop)."""
    _need(path)
    p = subprocess.run([path] + args, capture_output=True, text=True, timeout=1500)
    print(p.stdout[-6000:])
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-2000:]
    assert "BIO TEST PASSED" in p.stdout


def test_boundary_integral_op_reference_arm_runs_without_gpu():
    """the test-only element list drives the UNMODIFIED reference operator on the CPU (sanity: exterior single-layer potential of
    the unit sphere to 1e-8); keeps the BIO test infrastructure honest where it is authored"""
    _need(BIO_BINS[0])
    p = subprocess.run([BIO_BINS[0], "--cpu-only", "4", "1", "Laplace3D_FxU"], capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-2000:]
    assert "BIO TEST PASSED" in p.stdout and "sanity" in p.stdout
