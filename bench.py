#!/usr/bin/env python
"""bench.py — headline benchmark of the direct particle-kernel summation path.

Metric (BASELINE.json): pair interactions/s, Stokes3D_FxU FP64, Ns = Nt = 4M (config C3), targets sharded over
N GPUs (strong scaling: total work fixed), sources sharded and replicated over NCCL each step.

  python bench.py [--gpus N] [--steps K] [--warmup W]            our arm (CUDA, sm_100a)
  python bench.py --impl reference [...]                         the reference's own CPU Eval on the host cores
  torchrun ... bench.py --gpus N ...                             N > 1, one rank per GPU

A "step" is one full evaluation U = sum over all 4M sources for all 4M targets. `value` is timed with inputs
resident in HBM; `e2e` is the same metric through the public API with pinned HOST buffers, H2D/D2H inside the
timed region. Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

KERNEL_ID = 3  # Stokes3D_FxU
TRAFFIC_4M_BYTES = 1231038976 + 755320576  # ncu, one launch of the 4M x 4M eval_kernel (profiles/r02w_traffic_4m.csv; round 1: 1.33 + 0.76 GB)
FLOPS_REF = 23  # reference FLOPS() for Stokes3D_FxU (kernel_functions.hpp:79-81): the algorithmic flops per pair
FP64_INSTR_PER_PAIR = 22  # FP64-pipe instructions our kernel executes per pair (checked in SASS)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--n", type=int, default=int(os.environ.get("SCTL_B200_BENCH_N", 4_000_000)), help="Ns = Nt (default 4M = config C3)")
    ap.add_argument("--e2e-steps", type=int, default=3, help="end-to-end samples; the MEDIAN is reported")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of one reference sample")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------------------
# clocks: sample nvidia-smi during the timed region (B200_PROFILING.md recipe)
# ------------------------------------------------------------------------------------------------------------
class ClockSampler:
    """Samples SM clock, power and throttle reasons of one GPU DURING the timed region, in-process through NVML (the
    library nvidia-smi itself reads) from a background thread. A looping `nvidia-smi -lms 200` subprocess was measured to
    slow the 2-GPU run by up to 50 % (profiles/r01g_*: 3.9 s vs 2.63 s per step), so it is only the fallback when pynvml
    is missing."""
    PERIOD_S = float(os.environ.get("SCTL_B200_SAMPLER_PERIOD", "0.5"))
    MODE = os.environ.get("SCTL_B200_SAMPLER_MODE", "all")  # diagnostic: which NVML queries to issue

    def __init__(self, gpu_index: int):
        import threading

        self.samples = []  # (sm_mhz, sm_max_mhz, power_w, reasons bitmask)
        self._stop = threading.Event()
        self._thread = None
        self._proc = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self._nv = pynvml
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[gpu_index]) if vis and all(x.strip().isdigit() for x in vis.split(",")) else gpu_index
            self._h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self._max = pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM)
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()
        except Exception:
            self._nv = None
            self._start_smi(gpu_index)

    def _loop(self):
        nv = self._nv
        while not self._stop.is_set():
            try:
                m = self.MODE
                self.samples.append((float(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM)) if m in ("all", "clock") else 0.0, float(self._max),
                                     nv.nvmlDeviceGetPowerUsage(self._h) / 1000.0 if m in ("all", "power") else 0.0,
                                     int(nv.nvmlDeviceGetCurrentClocksEventReasons(self._h)) if m in ("all", "reasons") else 0))
            except Exception:
                pass
            self._stop.wait(self.PERIOD_S)

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def _start_smi(self, gpu_index):
        self._f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self._proc = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "1000"],
                                          stdout=self._f, stderr=subprocess.DEVNULL)
        except OSError:
            self._proc = None

    def stop(self) -> dict:
        out = {"sm_mhz": None, "sm_max_mhz": None, "power_w_max": None, "reasons": [], "samples": 0, "how": None}
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        sm, mx, pw, reasons = [], [], [], set()
        if self._thread is not None:
            self._stop.set()
            self._thread.join(timeout=5)
            nv = self._nv
            bits = {"hw_slowdown": nv.nvmlClocksEventReasonHwSlowdown, "hw_thermal_slowdown": nv.nvmlClocksEventReasonHwThermalSlowdown,
                    "sw_thermal_slowdown": nv.nvmlClocksEventReasonSwThermalSlowdown, "sw_power_cap": nv.nvmlClocksEventReasonSwPowerCap}
            for a, b, c, r in self.samples:
                sm.append(a); mx.append(b); pw.append(c)
                for nm in names:
                    if r & bits[nm]:
                        reasons.add(nm)
            out["how"] = f"NVML in-process thread, every {self.PERIOD_S} s"
        elif self._proc is not None:
            self._proc.terminate()
            try:
                self._proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self._proc.kill()
            self._f.flush()
            self._f.seek(0)
            for line in self._f:
                c = [x.strip() for x in line.split(",")]
                if len(c) < 7:
                    continue
                try:
                    sm.append(float(c[0])); mx.append(float(c[1])); pw.append(float(c[2]))
                except ValueError:
                    continue
                for nm, v in zip(names, c[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            self._f.close()
            try:
                os.unlink(self._f.name)
            except OSError:
                pass
            out["how"] = "nvidia-smi -lms 1000 subprocess"
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(mx), power_w_max=max(pw), reasons=sorted(reasons), samples=len(sm))
        return out


# ------------------------------------------------------------------------------------------------------------
# CPU reference arm / baseline: the reference's own Eval<double,true> (oracle/_ref) on all host threads
# ------------------------------------------------------------------------------------------------------------
HOST_THREADS = None  # (count, how): taken in main() BEFORE an OpenMP runtime is loaded (binding narrows the main thread's mask to one core)


def host_threads() -> tuple[int, str, int]:
    """Threads for the CPU arm: every PHYSICAL core this process may run on, whatever OMP_NUM_THREADS
    says - torch.distributed.run exports OMP_NUM_THREADS=1 to its workers, which is not a statement about the host.
    SCTL_B200_CPU_THREADS overrides."""
    forced = os.environ.get("SCTL_B200_CPU_THREADS")
    if forced and forced.isdigit() and int(forced) > 0:
        return int(forced), "SCTL_B200_CPU_THREADS", int(forced)
    try:
        allowed = os.sched_getaffinity(0)
    except AttributeError:
        allowed = set(range(os.cpu_count() or 1))
    cores = set()
    try:
        cpu = phys = core = None
        with open("/proc/cpuinfo") as f:
            for line in f.read().split("\n") + [""]:
                if not line.strip():
                    if cpu is not None and cpu in allowed:
                        cores.add((phys, core if core is not None else cpu))
                    cpu = phys = core = None
                    continue
                k, _, v = line.partition(":")
                k, v = k.strip(), v.strip()
                if k == "processor":
                    cpu = int(v)
                elif k == "physical id":
                    phys = int(v)
                elif k == "core id":
                    core = int(v)
    except (OSError, ValueError):
        cores = set()
    if cores:
        return len(cores), f"physical cores among the {len(allowed)} logical CPUs of the affinity mask (/proc/cpuinfo)", len(allowed)
    return max(1, len(allowed)), "logical CPUs of the affinity mask", max(1, len(allowed))


def cpu_sample_runner(n: int, seconds: float):
    """Returns (run(), pairs_per_call, info): run() evaluates a bounded sample of the workload — the first T targets
    against ALL n sources — with the reference's OpenMP+SIMD Eval (or the C port when oracle/_ref is absent). Never touches
    the product library: the kernel dims come from the checker itself."""
    from oracle import loader
    from sctl_b200 import synth  # input generator only (numpy); loads no native code

    chk = loader.best()
    d = chk.kernel_dims(KERNEL_ID)
    cores, how, logical = HOST_THREADS or host_threads()
    _, Xs, _, F = synth.make_problem(d, 1, n, np.float64, coincident=0)

    def calibrate(threads):
        """pairs/s on a small slab (64 targets per thread x all sources) with `threads` threads"""
        if chk.kind == "reference":
            chk.set_threads(threads)  # omp_set_num_threads: overrides a torchrun-injected OMP_NUM_THREADS=1
        Tc = 64 * threads
        Xt = synth.uniform(synth.STREAM_TRG, 0, 3 * Tc)
        chk.eval(KERNEL_ID, Xt[: 3 * 16], Xs[: 3 * 1000], None, F[: 3 * 1000])  # start the thread pool
        chk.eval(KERNEL_ID, Xt, Xs, None, F)  # warm (page faults, thread pool)
        t0 = time.perf_counter()
        chk.eval(KERNEL_ID, Xt, Xs, None, F)
        return Tc * n / (time.perf_counter() - t0)

    if chk.kind == "reference":
        # "all the host threads it can use": one per physical core or one per logical CPU, whichever
        # the reference's Eval runs faster with on this host
        trial = {t: calibrate(t) for t in sorted({cores, logical})}
        cores = max(trial, key=trial.get)
        rate = trial[cores]
        chk.set_threads(cores)
        assert chk.max_threads() == cores
        how += "; calibrated " + ", ".join(f"{t} threads: {r:.3g} pairs/s" for t, r in trial.items())
    else:
        os.environ["OMP_NUM_THREADS"] = str(cores)
        rate = calibrate(cores)
    Tc = 64 * cores
    T = int(max(Tc, min(n, rate * seconds / n)))
    T = (T // (8 * cores)) * (8 * cores) or Tc
    Xt = synth.uniform(synth.STREAM_TRG, 0, 3 * T)
    U = np.zeros(3 * T)

    def run():
        U[:] = 0
        chk.eval(KERNEL_ID, Xt, Xs, None, F, v_trg=U)
        return U

    info = {"kind": chk.kind, "cores": cores, "threads_from": how, "omp_proc_bind": os.environ.get("OMP_PROC_BIND"), "omp_places": os.environ.get("OMP_PLACES"),
            "sample": f"first {T} of {n} targets x all {n} sources, Stokes3D_FxU f64, "
            f"{'reference Eval<double,true> OpenMP+Vec (' + os.path.basename(chk.path) + ')' if chk.kind == 'reference' else 'C port (oracle_port.c), OpenMP'}"}
    return run, T * n, info


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # rank 0 alone runs the CPU arm
    run, pairs, info = cpu_sample_runner(args.n, args.cpu_seconds)
    for _ in range(args.warmup):
        run()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        run()
    dt = (time.perf_counter() - t0) / args.steps
    v = pairs / dt
    line = {"impl": "reference", "metric": "pair_interactions_per_s", "value": v, "unit": "pairs/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": workload_config(args),
            "cpu_baseline": dict(info, value=v, unit="pairs/s"),
            "e2e": {"value": v, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def workload_config(args) -> dict:
    return {"workload": f"Stokes3D_FxU direct summation, Ns=Nt={args.n}, FP64 (BASELINE.json configs[2])", "kernel": "Stokes3D_FxU",
            "n_sources": args.n, "n_targets": args.n, "parallelism": f"targets sharded x{args.gpus}, sources sharded + NCCL all-gather",
            "l2": f"no flush; source records streamed per step = {args.n * 48 / 1e6:.0f} MB vs 126 MB L2"
            + (" (inputs larger than L2)" if args.n * 48 > 126e6 else " (fits L2: compute-bound kernel, HBM traffic is not the limiter)")}


# ------------------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------------------
def run_b200_arm(args):
    import torch
    import torch.distributed as dist

    from sctl_b200 import Stokes3D_FxU, capi, synth
    from sctl_b200.distributed import ShardedDirectSum, partition

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if args.gpus > 1 and world == 1:
        raise SystemExit("for --gpus N > 1 launch with torch.distributed.run (one rank per GPU)")
    if capi.device_count() < 1:
        raise SystemExit("no CUDA device: the sctl_b200 arm has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    capi.check(capi.lib().sctl_b200_set_device(local_rank))
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def allsum(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    n = args.n
    ker = Stokes3D_FxU()
    d = capi.kernel_dims(KERNEL_ID)
    t0, t1 = partition(n, rank, world)
    s0, s1 = partition(n, rank, world)
    counts = [partition(n, r, world)[1] - partition(n, r, world)[0] for r in range(world)]
    Xt_np, Xs_np, _, F_np = synth.make_problem(d, n, n, np.float64, t0=t0, t1=t1, s0=s0, s1=s1)
    pin = lambda a: torch.from_numpy(a).pin_memory()
    Xt_h, Xs_h, F_h = pin(Xt_np), pin(Xs_np), pin(F_np)
    U_h = torch.zeros(3 * (t1 - t0), dtype=torch.float64).pin_memory()
    Xt, Xs, F = Xt_h.to(dev), Xs_h.to(dev), F_h.to(dev)
    U = torch.zeros(3 * (t1 - t0), dtype=torch.float64, device=dev)
    sharded = ShardedDirectSum(ker)

    def step():
        return sharded(U, Xt, Xs, None, F, counts=counts)

    # ---- value: inputs resident in HBM -------------------------------------------------------------------
    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local_rank) if (rank == 0 and not os.environ.get("SCTL_B200_BENCH_NO_SAMPLER")) else None
    l0 = capi.launch_count()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    evs[0].record()
    for i in range(args.steps):
        step()
        evs[i + 1].record()
    barrier()
    ms_steps = [evs[i].elapsed_time(evs[i + 1]) for i in range(args.steps)]
    if os.environ.get("SCTL_B200_BENCH_VERBOSE"):
        print(f"[rank {rank}] per-step ms: {ms_steps}", file=sys.stderr, flush=True)
    ms_step = allmax(evs[0].elapsed_time(evs[-1]) / args.steps)
    clocks = sampler.stop() if sampler else {}
    launches = int(allsum(capi.launch_count() - l0))
    value = float(n) * float(n) / (ms_step * 1e-3)

    # ---- parity spot check on the timed result (first 64 local targets of rank 0 vs the CPU oracle) --------
    parity = None
    if rank == 0:
        from oracle import loader

        chk = loader.best()
        T = min(64, t1 - t0)
        _, Xs_all, _, F_all = synth.make_problem(d, 1, n, np.float64, coincident=0)
        want = chk.eval(KERNEL_ID, Xt_np[: 3 * T].copy(), Xs_all, None, F_all)
        got = U[: 3 * T].cpu().numpy()
        parity = {"rel_l2": float(np.linalg.norm(got - want) / np.linalg.norm(want)), "targets": T, "checker": chk.kind, "tol": 1e-12}
        del Xs_all, F_all

    # ---- e2e: public API with pinned HOST buffers, H2D + D2H inside the timed region -----------------------
    h2d = (Xt_h.numel() + Xs_h.numel() + F_h.numel()) * 8
    d2h = U_h.numel() * 8
    eval_ms_kernel = None
    if world == 1:
        h2d += U_h.numel() * 8  # Eval accumulates into v_trg (generic-kernel.txx:98-101): v_trg is uploaded too
        Xt_v, Xs_v, F_v, U_v = Xt_h.numpy(), Xs_h.numpy(), F_h.numpy(), U_h.numpy()

        def e2e_step():
            U_v[:] = 0
            ker.Eval(U_v, Xt_v, Xs_v, None, F_v)  # synchronous C-ABI host-pointer call (what the C++ shim calls)
    else:
        Xt_v, Xs_v, F_v, U_v = Xt_h.numpy(), Xs_h.numpy(), F_h.numpy(), U_h.numpy()

        def e2e_step():
            # synchronous C-ABI call with HOST buffers on every rank: upload of the rank's blocks, NCCL replication, eval, download
            sharded.eval_host(U_v, Xt_v, Xs_v, None, F_v, counts=counts, device=dev)

    e2e_step()  # warm-up (first host-path call creates the library stream / pool)
    barrier()
    l1 = capi.launch_count()
    e2e_samples = []
    for _ in range(max(1, args.e2e_steps)):
        barrier()
        w0 = time.perf_counter()
        e2e_step()
        torch.cuda.synchronize()
        e2e_samples.append(allmax((time.perf_counter() - w0) * 1e3))  # max over ranks of each step, then the median of the steps
    e2e_ms = statistics.median(e2e_samples)
    barrier()
    launches_e2e = int(allsum(capi.launch_count() - l1)) // max(1, args.e2e_steps)
    if world == 1:
        eval_ms_kernel = capi.last_timing()["eval_ms"]
    e2e_value = float(n) * float(n) / (e2e_ms * 1e-3)

    # ---- roofline: FP64 CUDA-core pipe (compute bound; HBM traffic is ~0.1 B/pair) -------------------------
    peak_tf = capi.measure_peak(capi.F64, 1.0) if rank == 0 else 0.0
    barrier()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    kern_ms = eval_ms_kernel if eval_ms_kernel else ms_step
    pairs_per_launch = float(n) * float(t1 - t0)
    ach = pairs_per_launch * FLOPS_REF / (kern_ms * 1e-3) * 1e-12
    roofline = {"bound": "fp64_cuda_core", "achieved": ach, "peak": peak_tf, "unit": "TFLOP/s", "frac": ach / peak_tf if peak_tf else None,
                "traffic": TRAFFIC_4M_BYTES if (args.n == 4_000_000 and world == 1) else None,
                "traffic_source": "profiles/r02w_traffic_4m.csv: ncu dram__bytes_read.sum + dram__bytes_write.sum of ONE eval_kernel launch at 4M x 4M "
                                  "(1.23 GB read, 0.76 GB written, over 20.94 s: 0.1 GB/s against 6.5 TB/s; the kernel is FP64-pipe-bound)",
                "kernel": "eval_kernel<KStokesFxU,double,R=6,NEWTON=2>", "kernel_ms": kern_ms,
                "peak_source": "measured live: register-resident DFMA chain (sctl_b200_measure_peak), 2 flop per FMA; MEASURED_PEAKS.json has no FP64 figure",
                "flops_per_pair_algorithmic": FLOPS_REF, "fp64_instr_per_pair_executed": FP64_INSTR_PER_PAIR,
                "fp64_pipe_frac": pairs_per_launch * FP64_INSTR_PER_PAIR * 2 / (kern_ms * 1e-3) * 1e-12 / peak_tf if peak_tf else None}
    line = {"metric": "pair_interactions_per_s", "value": value, "unit": "pairs/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args), "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "pairs/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms,
                    "steps": args.e2e_steps, "ms_samples": e2e_samples, "statistic": "median of the per-step max over ranks", "gpu_launches_per_step": launches_e2e,
                    "api": "sctl_b200_eval (host pointers, C ABI)" if world == 1 else "sctl_b200_eval_sharded (host pointers, C ABI: H2D of the rank's blocks -> ncclAllGather -> eval -> D2H)"},
            "gpu_launches": launches, "roofline": roofline, "parity": parity}
    if not args.no_cpu_baseline:  # rank 0 only (the other ranks have left), at every N: the CPU path timed beside the GPU number
        run, pairs, info = cpu_sample_runner(n, args.cpu_seconds)
        run()
        c0 = time.perf_counter()
        run()
        line["cpu_baseline"] = dict(info, value=pairs / (time.perf_counter() - c0), unit="pairs/s")
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    # the CPU arm runs with OMP_PROC_BIND=close OMP_PLACES=cores. The OpenMP runtime reads these when it
    # is loaded, so they are set before anything imports it (an explicit setting of the caller wins).
    global HOST_THREADS
    HOST_THREADS = host_threads()
    # Only on the rank that runs CPU work: binding pins the initial thread to the first place, and N ranks must not share one core.
    if int(os.environ.get("RANK", "0")) == 0:
        os.environ.setdefault("OMP_PROC_BIND", "close")
        os.environ.setdefault("OMP_PLACES", "cores")
    # stdout carries exactly ONE JSON line: everything libraries print while we run (e.g. "NCCL version ..." from the first
    # communicator) is sent to stderr, and print() below writes to the saved descriptor.
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = real_stdout
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_b200_arm(args)
    real_stdout.flush()


if __name__ == "__main__":
    main()
