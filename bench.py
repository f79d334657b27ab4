#!/usr/bin/env python
"""bench.py -- GP log-likelihood + gradient throughput (points/s, FP64) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c3|c5]

Workload (BASELINE.json configs[1], "c2"): RotationTerm + Matern32Term (J = 6), one series of N = 1,000,000
points per GPU, synthetic light curve.  One step = one evaluation of the log-likelihood
and of its gradient with respect to the kernel coefficients:

    assemble a/U/V -> factor (S saved) -> solve_lower (F saved) -> logdet/chi2
                   -> solve_lower_rev -> factor_rev -> adjoint of the assembly

`value`  : points/s with t, y, diag and the coefficients already resident in HBM.
`e2e`    : the same step through the public API with HOST buffers: pinned H2D of t, y, diag and the
           coefficients, D2H of the log-likelihood and the coefficient gradients, every step.
`roofline`: the slowest operator of the step (all kernels of that C-ABI call), algorithmic bytes of
           SURVEY.md section 8(d) over its CUDA-event duration, against the measured HBM copy peak.
`cpu_baseline` / `--impl reference`: the reference's own C++ core (oracle/_ref, built from /root/reference)
           timed on this box's host cores; a single series is inherently single-threaded in the reference.

With N > 1 ranks (torchrun) every rank owns an independent series (weak scaling, no data-path collective).
"""
import argparse
import json
import math
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# stdout carries exactly one JSON line: keep NCCL's version banner (printed on stdout at NCCL_DEBUG=VERSION) off it
if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
    os.environ["NCCL_DEBUG"] = "WARN"

import numpy as np  # noqa: E402

from celerite2_b200 import workloads as wl  # noqa: E402

METRIC = "gp_loglike_grad_points_per_s"
UNIT = "points/s"


def bytes_per_point(J, K=1):
    """Algorithmic bytes per point of each operator (SURVEY.md section 8a / BASELINE.md section 2)."""
    return {
        "assemble": (3 + 2 * J) * 8,
        "factor": (3 + 3 * J + J * J) * 8,
        "solve_lower": (1 + 2 * J + 2 * K + J * K) * 8,
        "norm_logdet": 4 * 8,
        "solve_lower_rev": (2 + 4 * J + 3 * K + J * K) * 8,
        "factor_rev": (5 + 5 * J + J * J) * 8,
        "assemble_rev": (3 + 2 * J) * 8,
    }


OP_OF_ENTRY = {
    "c2b_get_celerite_matrices": "assemble",
    "c2b_factor": "factor",
    "c2b_solve_lower": "solve_lower",
    "c2b_norm_logdet": "norm_logdet",
    "c2b_solve_lower_rev": "solve_lower_rev",
    "c2b_factor_rev": "factor_rev",
    "c2b_get_celerite_matrices_rev": "assemble_rev",
    "c2b_add_inplace": "add",
}


def make_workload(name, rank):
    if name == "c2":
        cfg = wl.config_c2()
    elif name == "c2small":
        cfg = wl.config_c2(N=100_000)
    else:
        raise SystemExit("unknown workload %r" % name)
    t, y, diag = wl.light_curve(cfg["N"], cfg["seed"] + 1000 * rank)
    return cfg, t, y, diag


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ---- reference arm / CPU baseline ---------------------------------------------------------------------------
def reference_step(lib, co, t, y, diag):
    """One log-likelihood + gradient evaluation with the reference's C++ core (single thread)."""
    c, a, U, V = lib.get_celerite_matrices(*co, t, diag)
    flag, d, W, S = lib.factor(t, c, a, U, V)
    Y = y[:, None]
    Z, F = lib.sweep("solve_lower", t, c, U, W, Y)
    z = Z[:, 0]
    ll = -0.5 * (np.sum(z * z / d) + np.sum(np.log(d)) + len(t) * math.log(2 * math.pi))
    bz = (-z / d)[:, None]
    bd = 0.5 * z * z / (d * d) - 0.5 / d
    bt, bc, bU, bW, bY = lib.sweep_rev("solve_lower", t, c, U, W, Y, Z, F, bz)
    bt2, bc2, ba, bU2, bV = lib.factor_rev(t, c, a, U, V, d, W, S, bd, bW)
    return ll, bc + bc2


def cpu_reference_lib():
    import oracle

    lib = oracle.ref_fast()
    kind = "reference"
    build = "oracle/_ref/libc2ref_fast.so (reference C++ core, -O3 -march=x86-64-v3, fixed-size J dispatch)"
    if lib is None:
        lib = oracle.ref()
        build = "oracle/_ref/libc2ref.so (reference C++ core, -O2, stock J<=4 dispatch)"
    if lib is None:
        lib = oracle.port()
        kind = "port"
        build = "oracle/liboracle.so (plain-C restatement)"
    return lib, kind, build


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    lib, kind, build = cpu_reference_lib()
    cfg, t, y, diag = make_workload(args.workload, 0)
    N = cfg["N"]
    for _ in range(args.warmup):
        reference_step(lib, cfg["coeffs"], t, y, diag)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        reference_step(lib, cfg["coeffs"], t, y, diag)
    dt = (time.perf_counter() - t0) / args.steps
    val = N / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": workload_config(cfg, args),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": 1, "kind": kind,
                         "sample": "every step is the full workload (1 series, N=%d); %s; the reference has no threading "
                                   "and a single series cannot be split across cores" % (N, build)},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def workload_config(cfg, args):
    J = wl.width(cfg["coeffs"])
    return {
        "workload": "%s: RotationTerm+Matern32Term (J=%d) log-likelihood + gradient, N=%d, one series per GPU"
                    % (cfg["name"], J, cfg["N"]),
        "J": J, "N": cfg["N"], "series_per_gpu": 1,
        "l2": "inputs + saved workspaces of one step (about %.1f GB) exceed the 126 MB L2; no explicit flush"
              % (sum(bytes_per_point(J).values()) * cfg["N"] / 1e9),
        "multi_gpu": "independent series per rank, no data-path collective",
        "launch": "the timed step replays the evaluation from two CUDA graphs (gp.GraphedLogLikelihood); eager_ms_per_step is the "
                  "same step launched kernel by kernel",
    }


# ---- our arm ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in out.strip().splitlines():
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def run_ours(args):
    import torch
    import torch.distributed as dist

    from celerite2_b200 import _lib, gp, ops

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device visible; there is no CPU fallback for the product path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    cfg, t, y, diag = make_workload(args.workload, rank)
    N = cfg["N"]
    J = wl.width(cfg["coeffs"])
    co_h = [torch.from_numpy(np.ascontiguousarray(v)).pin_memory() for v in cfg["coeffs"]]
    t_h, y_h, d_h = [torch.from_numpy(v).pin_memory() for v in (t, y, diag)]
    co_d = [v.to(dev) for v in co_h]
    t_d, y_d, d_d = t_h.to(dev), y_h.to(dev), d_h.to(dev)

    def step_eager():
        return gp.log_likelihood_and_grad(co_d, t_d, y_d, d_d)

    # the same evaluation replayed from CUDA graphs (inputs resident in HBM): this is the timed device-resident step
    graphed = gp.GraphedLogLikelihood(co_d, t_d, y_d, d_d)

    def step_resident():
        graphed.replay_front()
        return graphed.replay_back()

    # end-to-end path: the host-buffer entry point (pinned host arrays in, pinned host result out)
    host_api = gp.HostLogLikelihood(N, [v.shape for v in co_h], dev)
    n_out = host_api.n_out

    def step_e2e():
        out, _ = host_api(co_h, t_h, y_h, d_h)
        return out

    # the same through the streaming form (uploads of evaluation i+1 overlap the kernels of evaluation i); reported
    # next to the synchronous number, which stays the headline
    stream_api = gp.HostLogLikelihoodStream(N, [v.shape for v in co_h], dev)

    def run_streamed(steps):
        stream_api.submit(co_h, t_h, y_h, d_h)
        for _ in range(steps - 1):
            stream_api.submit(co_h, t_h, y_h, d_h)
            stream_api.collect()
        return stream_api.collect()[0]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        w0 = time.perf_counter()
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        wall = (time.perf_counter() - w0) * 1e3
        ms = e0.elapsed_time(e1)
        barrier()
        if world > 1:
            tt = torch.tensor([ms, wall], dtype=torch.float64, device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            ms, wall = tt.tolist()
        return ms / steps, wall / steps

    for _ in range(max(args.warmup, 3)):
        r = step_resident()
    torch.cuda.synchronize()
    status = int(r["status"].max().item())
    if status != 0:
        raise SystemExit("factorisation failed at row %d" % status)

    sampler = ClockSampler(local) if rank == 0 else None
    ms_step, wall_step = timed(step_resident, args.steps)
    for _ in range(3):
        step_eager()
    ms_eager, _ = timed(step_eager, args.steps)
    # second pass with the library's per-kernel CUDA events switched on (roofline timings, launch count)
    _lib.lib.c2b_profile_enable(1)
    ms_prof, _ = timed(step_eager, args.steps)
    report = _lib.lib.c2b_profile_report().decode()
    _lib.lib.c2b_profile_enable(0)
    for _ in range(2):
        step_e2e()
    ms_e2e, wall_e2e = timed(step_e2e, args.steps)
    run_streamed(3)
    ms_e2e_stream, _ = timed(lambda: run_streamed(args.steps), 1)
    ms_e2e_stream /= args.steps
    clocks = sampler.stop() if sampler else None

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- per-operator durations from the profile pass --------------------------------------------------------
    per_op, per_kernel, launches = {}, {}, 0
    for ln in report.strip().splitlines():
        key, cnt, tot = ln.rsplit(" ", 2)
        entry, kern = key.split("/")
        op = OP_OF_ENTRY.get(entry, entry)
        per_op[op] = per_op.get(op, 0.0) + float(tot) / args.steps
        per_kernel[op + "/" + kern] = round(float(tot) / args.steps, 5)
        launches += int(cnt)
    bpp = bytes_per_point(J)
    peak, peak_src = load_peaks()
    dom = max((o for o in per_op if o in bpp), key=lambda o: per_op[o])
    dom_bytes = bpp[dom] * N
    achieved = dom_bytes / (per_op[dom] * 1e-3) / 1e9
    step_bytes = sum(bpp.values()) * N
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "r01_traffic.json")
    if cfg["name"] == "c2" and os.path.exists(tpath):
        with open(tpath) as f:
            traffic = json.load(f).get(dom)  # DRAM bytes per call from the committed ncu --set full captures
    roofline = {
        "bound": "hbm", "kernel": dom + " (all kernels of the C-ABI call: summary + scan + apply)",
        "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "frac_of_nominal_8000_gbs": achieved / 8000.0,
        "traffic": traffic,
        "peak_source": peak_src, "algorithmic_bytes_per_launch": dom_bytes, "ms_per_launch": per_op[dom],
        "whole_step": {"algorithmic_bytes": step_bytes, "achieved_gbs": step_bytes / (ms_step * 1e-3) / 1e9,
                       "frac": step_bytes / (ms_step * 1e-3) / 1e9 / peak},
        "per_op_ms": {k: round(v, 5) for k, v in per_op.items()},
        "per_op_frac": {k: round(bpp[k] * N / (v * 1e-3) / 1e9 / peak, 4) for k, v in per_op.items() if k in bpp},
        "per_kernel_ms": per_kernel,
        "timing": "CUDA events around every kernel, second pass of %d steps (%.3f ms/step with events on)" % (args.steps, ms_prof),
    }

    # ---- CPU baseline on this box's host cores (rank 0, N = 1 only) -------------------------------------------
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        lib, kind, build = cpu_reference_lib()
        reps = 2
        best = float("inf")
        ll_ref = None
        for _ in range(reps):
            t0 = time.perf_counter()
            ll_ref, _ = reference_step(lib, cfg["coeffs"], t, y, diag)
            best = min(best, time.perf_counter() - t0)
        ll_gpu = float(step_resident()["loglike"].item())
        cpu = {"value": N / best, "unit": UNIT, "cores": 1, "kind": kind,
               "sample": "full workload (1 series, N=%d), best of %d passes; %s; single-threaded because the reference "
                         "cannot split one series across cores" % (N, reps, build),
               "loglike_rel_diff_vs_gpu": abs(ll_gpu - ll_ref) / abs(ll_ref)}

    h2d = sum(v.numel() for v in (t_h, y_h, d_h)) * 8 + sum(v.numel() for v in co_h) * 8
    line = {
        "metric": METRIC, "value": world * N / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "config": workload_config(cfg, args),
        "clocks": clocks,
        "e2e": {"value": world * N / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": n_out * 8,
                "ms_per_step": ms_e2e, "wall_ms_per_step": wall_e2e,
                "streamed": {"value": world * N / (ms_e2e_stream * 1e-3), "ms_per_step": ms_e2e_stream,
                             "note": "HostLogLikelihoodStream, two slots: the same copies every step, uploads of evaluation i+1 overlap the kernels of evaluation i"}},
        "gpu_launches": launches,
        "eager_ms_per_step": ms_eager,
        "wall_ms_per_step": wall_step,
        "roofline": roofline,
        "cpu_baseline": cpu,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        if args.steps > 10:
            pass  # each step is about a second of CPU work; K steps stay within a few minutes up to K ~ 100
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
