#!/usr/bin/env python
"""bench.py -- OCP QP solves/s of the batched computeDirection hot path on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl reference]

A "step" is one pass of the hot path (one `computeDirection` per problem) over one batch
of synthetic derivatives.  Default workload = BASELINE.json configs[1]: UR5-shaped SQP
(nx=12, nu=6, T=100), B=4096 problems per GPU (weak scaling, no data-path collective; the
only collective is the all-reduce of the 5-number batch summary).

Printed JSON (one line, rank 0):
  value     QP solves/s, device-resident inputs, K back-to-back launches, CUDA events,
            barrier + synchronize on both sides, max over ranks
  e2e       same metric through the C ABI with HOST (pinned) buffers: H2D of the step's
            derivatives and D2H of dx/du/lag_mul + scalars inside the timed region
  roofline  algorithmic bytes (SURVEY 8d bytes_QP x B) / mean kernel time vs measured HBM
  cpu_baseline  the CPU oracle (a port of the reference's Eigen code, see oracle/) on a
            bounded sample of the same workload on this box's host cores
`--impl reference` times that CPU implementation alone (the reference itself cannot be
built here: Eigen/Boost/crocoddyl are absent).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

from mim_solvers_b200 import _capi  # noqa: E402
from mim_solvers_b200.problems import CONFIGS, OCPBatch, make_config, tile_batch  # noqa: E402
from mim_solvers_b200.sharding import allreduce_max, allreduce_summary  # noqa: E402

METRIC = "OCP QP solves/sec (batched FP64 Riccati+ADMM)"
UNIT = "QP/s"

# workload -> (solver, per-GPU batch, CSQP protocol)
WORKLOADS = {
    "C2_ur5_sqp": dict(solver="sqp", B=4096),
    "C1_ur5_sqp": dict(solver="sqp", B=4096),
    "C1_ur5_csqp": dict(solver="csqp", B=4096),
    "C3_solo12_csqp": dict(solver="csqp", B=2048),
    "C3b_solo12_bench": dict(solver="csqp", B=2048),
    "C4_lq_nx64": dict(solver="sqp", B=1024),
    "C4_clq_nx64": dict(solver="csqp", B=256),
    "C4_lq_nx32": dict(solver="sqp", B=4096),
    "C4_clq_nx32": dict(solver="csqp", B=1024),
    "C4_lq_nx16": dict(solver="sqp", B=4096),
    "C4_clq_nx16": dict(solver="csqp", B=4096),
    "C4_clq_nx8": dict(solver="csqp", B=4096),
    "C4_lq_nx8": dict(solver="sqp", B=4096),
    "C5_talos_csqp": dict(solver="csqp", B=512),
    "C5_talos_sqp": dict(solver="sqp", B=512),
}


QP_ITERS = 50  # --qp-iters (experiments only; the reported protocol is 50)


def csqp_bench_params():
    """benchmarks/solo12.cpp:213-216: eps_abs = eps_rel = 0, max_qp_iters = 50, no
    equality-QP initial guess -> exactly 50 ADMM iterations per QP."""
    return _capi.default_csqp_params(eps_abs=0.0, eps_rel=0.0, max_qp_iters=QP_ITERS,
                                     equality_qp_initial_guess=0)


def make_workload(name: str, B: int, first: int, distinct: int) -> OCPBatch:
    """B problems; the first `distinct` are drawn independently (seed 5489 + index), the rest
    repeat them (timing of these kernels does not depend on the data values)."""
    d = min(B, distinct)
    base = make_config(name, d, first=first)
    return base if d == B else tile_batch(base, B)


def base_config(name: str, B: int, default_tol: bool = False) -> dict:
    """The `config` object both arms print (identical keys and values for the same workload)."""
    w = WORKLOADS[name]
    cfg = dict(CONFIGS[name])
    nb = 8 * B * (cfg["T"] * (2 * cfg["nx"] ** 2 + 2 * cfg["nx"] * cfg["nu"] + cfg["nu"] ** 2) + cfg["nx"] ** 2)
    return {"workload": name, "solver": w["solver"], **cfg, "B_per_gpu": B,
            "l2": f"inputs >= {nb / 1e6:.0f} MB per step > 126 MB L2 (no flush needed)" if nb > 2 * 126e6
                  else "inputs smaller than 2x L2",
            "csqp_protocol": (("default tolerances (eps 1e-4, max 1000), cold start each step" if default_tol
                               else f"eps_abs=eps_rel=0, {QP_ITERS} ADMM iterations, no eq. guess")
                              if w["solver"] == "csqp" else None)}


# ------------------------------------------------------------------ clocks sampler
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device: int):
        self.device, self.rows, self._stop, self._t = device, [], threading.Event(), None

    def _run(self):
        try:  # NVML in-process: ~1 ms per sample, so even a 20 ms timed region gets several
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.device)
            mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            bits = [("hw_slowdown", nv.nvmlClocksThrottleReasonHwSlowdown),
                    ("hw_thermal_slowdown", nv.nvmlClocksThrottleReasonHwThermalSlowdown),
                    ("sw_thermal_slowdown", nv.nvmlClocksThrottleReasonSwThermalSlowdown),
                    ("sw_power_cap", nv.nvmlClocksThrottleReasonSwPowerCap)]
            while not self._stop.is_set():
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                self.rows.append([str(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)), str(mx),
                                  str(nv.nvmlDeviceGetPowerUsage(h) / 1000.0)]
                                 + ["Active" if r & b else "Not Active" for _, b in bits])
                self._stop.wait(0.002)
            return
        except Exception:
            pass
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.device), f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True,
                                     timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.1)

    def __enter__(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=6)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        sm = sorted(float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": float(self.rows[0][1]),
                "power_w_max": max(float(r[2]) for r in self.rows), "samples": len(self.rows), "reasons": reasons}


# ------------------------------------------------------------------ CPU arm (oracle)
def _oracle_native():
    """Build the oracle with -O3 -march=native ON THIS BOX (the portable build travels
    with the snapshot; a native one must not) and return the module bound to it."""
    from oracle import oracle
    out = ROOT / "oracle" / "_build" / "libmim_oracle_native.so"
    src = ROOT / "oracle" / "mim_oracle.c"
    try:
        if not out.exists() or out.stat().st_mtime < src.stat().st_mtime:
            out.parent.mkdir(exist_ok=True)
            subprocess.run(["gcc", "-O3", "-march=native", "-std=c11", "-fPIC", "-fopenmp", "-shared", "-o",
                            str(out), str(src), "-lm"], check=True, capture_output=True)
        oracle._LIB_PATH = out
        oracle._lib = None
    except Exception:
        oracle.build()
    return oracle


def cpu_run(workload: str, solver: str, sample_B: int, steps: int, warmup: int, nthreads: int = 0):
    """Times the oracle's computeDirection over `sample_B` problems, all host threads."""
    oracle = _oracle_native()
    batch = make_workload(workload, sample_B, 0, distinct=min(sample_B, 64))
    if nthreads <= 0:  # all host cores, whatever OMP_NUM_THREADS says (torchrun sets it to 1)
        nthreads = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    o = oracle.Oracle(batch, nthreads=nthreads)
    cores = nthreads
    if solver == "csqp":
        p = csqp_bench_params()
        o.csqp_calc_norms()
        o.csqp_reset_rho(p)
        call = lambda: o.csqp_direction(p, with_kkt=True)
    else:
        p = _capi.default_sqp_params()
        call = lambda: o.sqp_direction(p, with_kkt=True)
    for _ in range(warmup):
        call()
    t0 = time.perf_counter()
    for _ in range(steps):
        call()
    dt = (time.perf_counter() - t0) / steps
    return dict(value=sample_B / dt, ms_per_step=dt * 1e3, cores=cores, sample_B=sample_B)


def pick_cpu_sample(workload: str, solver: str, target_s: float = 1.5) -> int:
    """Sample size so that one CPU step is about `target_s` seconds."""
    probe = cpu_run(workload, solver, 8, 1, 1)
    per_qp_all_cores = 1.0 / probe["value"]
    n = int(max(8, min(4096, target_s / per_qp_all_cores)))
    return max(8, (n // 8) * 8)


def reference_arm(args, rank, world):
    if rank != 0:
        return
    w = WORKLOADS[args.workload]
    steps, warmup = max(1, args.steps), max(1, min(args.warmup, 3))
    # bounded sample: the whole --steps K run ends within ~2 minutes whatever K is
    sample = pick_cpu_sample(args.workload, w["solver"], target_s=min(1.5, 100.0 / (steps + warmup)))
    r = cpu_run(args.workload, w["solver"], sample, steps, warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": base_config(args.workload, args.batch or w["B"]),
        "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample_B_per_step": sample,
                         "sample": f"{sample} problems of {args.workload} per step, oracle -O3 -march=native, "
                                   f"OpenMP over problems"},
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "reference C++ unbuildable here (Eigen/Boost/crocoddyl absent): CPU oracle port timed instead",
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------ GPU arm
def fp64_peak(local_rank: int):
    """FP64 roofline denominator: measured HERE by mimgpu_measure_peaks (register-resident DFMA loop);
    profiles/peaks_r01.json only if the measurement fails.  MEASURED_PEAKS.json carries no FP64 figure."""
    import ctypes as C
    try:
        lib = _capi.load_library()
        a, b, c = C.c_double(), C.c_double(), C.c_double()
        if lib.mimgpu_measure_peaks(local_rank, C.byref(a), C.byref(b), None) == 0 and a.value > 1.0:
            return max(a.value, b.value), "mimgpu_measure_peaks in this run (DFMA %.1f / DMMA %.1f TFLOP/s)" % (a.value, b.value)
    except Exception:
        pass
    f = ROOT / "profiles" / "peaks_r01.json"
    if f.exists():
        d = json.loads(f.read_text())
        return float(max(d["dfma_tflops"], d["dmma_tflops"])), "profiles/peaks_r01.json"
    return 37.0, "fallback 37 TFLOP/s"


def traffic_entry(workload: str, B: int):
    """dram bytes of one launch of the dominant kernel from the committed ncu --set full capture."""
    for name in ("traffic_r02.json", "traffic_r01.json"):
        tf = ROOT / "profiles" / name
        if tf.exists():
            tr = json.loads(tf.read_text()).get(workload)
            if tr and tr.get("B") == B:
                return tr
            if tr and tr.get("B"):   # one CTA per problem, no sharing between problems: bytes scale with B
                k = B / tr["B"]
                return {**tr, "dram_bytes_per_launch": tr["dram_bytes_per_launch"] * k,
                        "source": tr["source"] + f", scaled x{k:.3g} to B={B}"}
    return None


def run_workload(args, name, rank, world, local_rank, *, steps, warmup, e2e_steps, with_cpu, peaks):
    """Device-resident arm + end-to-end arm (+ CPU arm on rank 0 at N = 1) of one workload."""
    import torch
    import torch.distributed as dist
    from mim_solvers_b200.engine import BatchEngine

    dev = torch.device("cuda", local_rank)
    w = WORKLOADS[name]
    solver, B = w["solver"], (args.batch or w["B"])
    cfg = dict(CONFIGS[name])
    batch = make_workload(name, B, first=rank * B, distinct=args.distinct)
    stream = torch.cuda.current_stream(dev)
    shape = batch.shape()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # --- device-resident arm: inputs live in HBM before the timed region, in the layout a device-side
    # producer writes (knot records for the shapes that have them: no re-layout pass; else field-major)
    eng = BatchEngine(shape=shape, device=local_rank, stream=stream.cuda_stream)
    layout = "field-major"
    dev_in = None
    if solver == "sqp" and not args.field_major:
        try:
            eng.record_layout()
            rec = torch.from_numpy(eng.pack_records(batch)).to(dev)
            cost = torch.from_numpy(np.ascontiguousarray(batch.arrays["cost"])).to(dev)
            eng.attach_records(rec, cost)
            dev_in, layout = (rec, cost), "knot-records"
        except _capi.MimGpuError:
            pass
    if dev_in is None:
        dev_in = {k: torch.from_numpy(v).to(dev) for k, v in batch.arrays.items()}
        eng.attach_device(dev_in)
    input_bytes = (int(dev_in[0].numel() + dev_in[1].numel()) * 8 if layout == "knot-records" else batch.nbytes())
    if solver == "csqp":
        p = csqp_bench_params() if not args.default_tol else _capi.default_csqp_params()
        eng.csqp_calc_norms()
        eng.csqp_reset_rho(p)
        if args.default_tol:   # every step starts from the same cold solver state
            step = lambda: (eng.csqp_reset_rho(p), eng.set_field("y", np.zeros(shape.field_shape("y"))),
                            eng.csqp_direction(p, with_kkt=True))
        else:
            step = lambda: eng.csqp_direction(p, with_kkt=True)
    else:
        p = _capi.default_sqp_params()
        step = lambda: eng.sqp_direction(p, with_kkt=True)

    for _ in range(warmup):
        step()
    barrier()
    l0 = eng.launch_count
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    with ClockSampler(local_rank) as clk:
        barrier()
        evs[0].record(stream)
        for i in range(steps):
            step()
            evs[i + 1].record(stream)
        barrier()
    launches = eng.launch_count - l0
    per = [evs[i].elapsed_time(evs[i + 1]) for i in range(steps)]
    total_ms = evs[0].elapsed_time(evs[-1])
    # the only collectives -- max time over ranks, batch-wide summary
    summ = allreduce_summary(eng.summary(1e-6), device=dev)
    total_ms = allreduce_max(total_ms, device=dev)
    ms_per_step = total_ms / steps
    value = world * B / (ms_per_step * 1e-3)
    qp_iters_hist = None
    if solver == "csqp":
        it = eng.qp_iters
        u, c = np.unique(it, return_counts=True)
        qp_iters_hist = {int(a): int(b) for a, b in zip(u, c)}

    # --- roofline of the dominant kernel (one launch per step)
    hbm_peak, hbm_src, f64_peak, f64_src = peaks
    kernel_ms = float(np.mean(per))
    bytes_qp = batch.algorithmic_bytes_per_qp()
    n_it = float(np.mean(eng.qp_iters)) if solver == "csqp" else 0.0
    flops_qp = batch.flops_csqp_qp(int(round(min(n_it, p.max_qp_iters))), bool(p.equality_qp_initial_guess)) \
        if solver == "csqp" else batch.flops_sqp_qp()
    ach_gbs = bytes_qp * B / (kernel_ms * 1e-3) / 1e9
    tflops = flops_qp * B / (kernel_ms * 1e-3) / 1e12
    roof = {"bound": "hbm", "achieved": ach_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": ach_gbs / hbm_peak,
            "traffic": None, "peak_source": hbm_src, "kernel_ms": kernel_ms,
            "algorithmic_bytes_per_qp": bytes_qp, "flops_per_qp": flops_qp, "fp64_tflops_achieved": tflops}
    tr = traffic_entry(name, B)
    if tr:
        roof.update(traffic=tr["dram_bytes_per_launch"], traffic_source=tr["source"],
                    algorithmic_bytes_per_launch=bytes_qp * B)
    if flops_qp / bytes_qp > f64_peak * 1e12 / (hbm_peak * 1e9):
        # above the machine balance the FP64 pipe is the binding roof of the algorithmic work
        roof.update(bound="fp64", achieved=tflops, peak=f64_peak, unit="TFLOP/s", frac=tflops / f64_peak,
                    peak_source=f64_src)
    if solver == "csqp":
        # what the ADMM sweeps actually stream: the iteration working set, read once per sweep
        T, nx, nu, ng = batch.T, batch.nx, batch.nu, batch.ng
        ws = 8 * T * (nx * nx + 2 * nx * nu + nu * nu + ng * (nx + nu) + 6 * nx + 4 * nu + 6 * ng)
        st_gbs = ws * n_it * B / (kernel_ms * 1e-3) / 1e9
        roof["hbm_stream"] = {"iteration_working_set_bytes": ws, "mean_qp_iters": n_it,
                              "achieved_gbs_one_read_per_iteration": st_gbs, "frac_of_hbm_peak": st_gbs / hbm_peak,
                              "note": "the sequential sweeps need hundreds of problems in flight, whose working sets "
                                      "(B_in_flight x ws) exceed L2: the loop streams from HBM (DESIGN.md 4.3)"}

    # --- end-to-end arm: host (pinned) buffers through the C ABI, copies inside the timed region
    e2e = None
    if e2e_steps > 0:
        del dev_in
        eng.close()
        torch.cuda.empty_cache()
        eng2 = BatchEngine(shape=shape, device=local_rank, stream=stream.cuda_stream)

        def pin(a):
            t = torch.empty(a.shape, dtype=torch.float64).pin_memory()
            t.numpy()[...] = a
            return t
        use_rec = layout == "knot-records"
        if use_rec:
            prec, pcost = pin(eng2.pack_records(batch)), pin(batch.arrays["cost"])
            h2d = prec.numel() * 8 + pcost.numel() * 8
        else:
            pinned = OCPBatch(batch.B, batch.T, batch.nx, batch.nu, batch.ng, batch.ngT)
            keep = {k: pin(v) for k, v in batch.arrays.items()}
            pinned.arrays = {k: t.numpy() for k, t in keep.items()}
            h2d = sum(v.nbytes for v in pinned.arrays.values())
        outs = {n: torch.empty(tuple(shape.field_shape(n)), dtype=torch.float64).pin_memory()
                for n in (("dx_tilde", "du_tilde", "lag_mul", "y", "z") if solver == "csqp"
                          else ("dx", "du", "lag_mul"))}
        kkt = torch.empty(B, dtype=torch.float64).pin_memory()
        d2h = sum(t.numel() * 8 for t in outs.values()) + B * 8

        def e2e_step():
            if solver == "sqp" and use_rec:   # upload + direction + download in overlapped slices, one C-ABI call
                eng2.sqp_direction_host_records(prec.numpy(), pcost.numpy(), p, dx=outs["dx"].numpy(),
                                                du=outs["du"].numpy(), lag_mul=outs["lag_mul"].numpy(),
                                                kkt=kkt.numpy(), n_chunks=args.e2e_chunks)
                return
            if solver == "sqp":
                eng2.sqp_direction_host(pinned, p, dx=outs["dx"].numpy(), du=outs["du"].numpy(),
                                        lag_mul=outs["lag_mul"].numpy(), kkt=kkt.numpy(), n_chunks=args.e2e_chunks)
                return
            on = lambda k: outs[k].numpy() if outs[k].numel() else None
            eng2.csqp_direction_host(pinned, p, with_kkt=True, calc_norms=True, reset_rho=True,
                                     dx_tilde=on("dx_tilde"), du_tilde=on("du_tilde"), lag_mul=on("lag_mul"),
                                     y=on("y"), z=on("z"), kkt=kkt.numpy(), n_chunks=args.e2e_chunks)

        for _ in range(2):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step()
        barrier()
        dt = allreduce_max((time.perf_counter() - t0) / e2e_steps, device=dev)
        e2e = {"value": world * B / dt, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h), "ms_per_step": dt * 1e3, "steps": e2e_steps,
               "h2d_gbs_per_rank": h2d / dt / 1e9,
               "call": ("mimgpu_sqp_direction_host_records" if use_rec else "mimgpu_sqp_direction_host")
                       + " (%d overlapped slices)" % args.e2e_chunks if solver == "sqp"
                       else "mimgpu_csqp_direction_host (%d overlapped slices)" % args.e2e_chunks,
               "max_kkt": float(kkt.max())}
        eng2.close()
        torch.cuda.empty_cache()
    else:
        eng.close()

    cpu = None
    if with_cpu and rank == 0 and world == 1:
        sample = pick_cpu_sample(name, solver, target_s=1.0)
        r = cpu_run(name, solver, sample, 5, 1)
        s1 = max(8, sample // max(1, r["cores"]) // 8 * 8)
        r1 = cpu_run(name, solver, s1, 3, 1, nthreads=1)
        cpu = {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port",
               "sample": f"{sample} problems of {name} x 5 passes, CPU oracle (-O3 -march=native, "
                         f"OpenMP over problems)",
               "single_thread": {"value": r1["value"], "cores": 1, "sample": f"{s1} problems x 3 passes"}}
    return {
        "value": value, "unit": UNIT, "ms_per_step": ms_per_step, "steps": steps, "warmup": warmup,
        "config": base_config(name, B, args.default_tol),
        "config_detail": {"input_layout": layout, "distinct_problems_per_gpu": min(B, args.distinct),
                          "input_bytes_per_step": input_bytes},
        "roofline": roof, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches),
        "qp_iters_histogram": qp_iters_hist, "clocks": clk.summary(),
        "batch_summary": {"max_kkt": float(summ["max_kkt"]), "sum_qp_iters": int(summ["sum_qp_iters"]),
                          "n_converged": int(summ["n_converged"]), "n_failed": int(summ["n_failed"])},
    }


def gpu_arm(args, rank, world, local_rank):
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    bind_rank_to_numa(local_rank)
    peaks_file = ROOT / "MEASURED_PEAKS.json"
    if peaks_file.exists():
        hbm_peak, hbm_src = float(json.loads(peaks_file.read_text())["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    else:
        hbm_peak, hbm_src = 6650.0, "fallback (B200_PROFILING.md)"
    f64_peak, f64_src = fp64_peak(local_rank)
    peaks = (hbm_peak, hbm_src, f64_peak, f64_src)
    main = run_workload(args, args.workload, rank, world, local_rank, steps=args.steps, warmup=args.warmup,
                        e2e_steps=0 if args.no_e2e else max(2, min(args.steps, args.e2e_steps)),
                        with_cpu=not args.no_cpu, peaks=peaks)
    extra = {}
    for name in ([] if args.no_extra else [w for w in args.extra.split(",") if w and w != args.workload]):
        r = run_workload(args, name, rank, world, local_rank, steps=args.extra_steps, warmup=3,
                         e2e_steps=0 if args.no_e2e else 2, with_cpu=not args.no_cpu, peaks=peaks)
        extra[name] = {k: r[k] for k in ("value", "unit", "ms_per_step", "steps", "config", "config_detail", "roofline", "cpu_baseline",
                                          "e2e", "gpu_launches", "qp_iters_histogram", "batch_summary")}
    if rank != 0:
        return
    line = {
        "metric": METRIC, "value": main["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": main["ms_per_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": main["config"],
        "config_detail": main["config_detail"],
        "roofline": main["roofline"], "cpu_baseline": main["cpu_baseline"], "e2e": main["e2e"],
        "gpu_launches": main["gpu_launches"], "clocks": main["clocks"], "batch_summary": main["batch_summary"],
        "workloads": extra,
    }
    if main["qp_iters_histogram"]:
        line["qp_iters_histogram"] = main["qp_iters_histogram"]
    print(json.dumps(line), flush=True)


def bind_rank_to_numa(local_rank: int):
    """Pin this rank's threads (and therefore its first-touch pinned allocations and copy threads) to
    the CPUs nearest to its GPU, when the box exposes more than one NUMA node (nvidia-smi topo)."""
    try:
        import pynvml as nv
        nv.nvmlInit()
        h = nv.nvmlDeviceGetHandleByIndex(local_rank)
        mask = nv.nvmlDeviceGetCpuAffinity(h, max(1, ((os.cpu_count() or 64) + 63) // 64))
        cpus = [64 * i + b for i, m in enumerate(mask) for b in range(64) if (m >> b) & 1]
        world = int(os.environ.get("LOCAL_WORLD_SIZE", os.environ.get("WORLD_SIZE", "1")))
        if cpus and world > 1:   # split the GPU's CPU set among the ranks that share it
            share = [c for i, c in enumerate(sorted(cpus)) if i % world == local_rank % world] or cpus
            os.sched_setaffinity(0, share)
        elif cpus:
            os.sched_setaffinity(0, cpus)
    except Exception:
        pass


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=600)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="C2_ur5_sqp", choices=sorted(WORKLOADS))
    ap.add_argument("--batch", type=int, default=0, help="problems per GPU (default: the workload's)")
    ap.add_argument("--distinct", type=int, default=4096, help="independently drawn problems per GPU")
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--e2e-chunks", type=int, default=8, help="slices of the overlapped host path (SQP)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--extra", default="C3_solo12_csqp,C5_talos_csqp,C4_clq_nx64,C1_ur5_csqp",
                    help="comma-separated workloads reported in the `workloads` sub-dict of the line")
    ap.add_argument("--extra-steps", type=int, default=5)
    ap.add_argument("--no-extra", action="store_true")
    ap.add_argument("--field-major", action="store_true",
                    help="device-resident SQP inputs in the field-major layout (re-laid-out inside every step)")
    ap.add_argument("--default-tol", action="store_true",
                    help="CSQP at the default tolerances (eps 1e-4, max 1000 iterations) instead of the 50-iteration protocol")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--qp-iters", type=int, default=50, help="ADMM iterations per CSQP QP (protocol: 50)")
    args = ap.parse_args()
    global QP_ITERS
    QP_ITERS = args.qp_iters
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        reference_arm(args, rank, world)
        return
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    try:
        gpu_arm(args, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


def _json_only_stdout():
    """Libraries (NCCL's version banner) write to file descriptor 1 directly: point fd 1 at stderr for
    the whole run and give Python's sys.stdout a private copy of the real stdout, so that the one JSON
    line is the only thing the driver reads there."""
    sys.stdout.flush()
    real = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real, "w", buffering=1)


if __name__ == "__main__":
    _json_only_stdout()
    main()
