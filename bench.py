#!/usr/bin/env python
"""bench.py -- the five BASELINE.json configurations of the semiseparable GP core on B200 (FP64).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c1|c2|c3|c4|c5]

Default workload (BASELINE.json configs[1], "c2"): RotationTerm + Matern32Term (J = 6), ONE series of N = 1,000,000
points, synthetic light curve.  One step = one evaluation of the log-likelihood and of its
gradient with respect to the kernel coefficients:

    assemble a/U/V -> factor (S saved) -> solve_lower (F saved) -> logdet/chi2
                   -> solve_lower_rev -> factor_rev -> adjoint of the assembly

Workloads and what `--gpus N` does with them (each config's own multi-GPU mode):
  c1  SHOTerm J=2, N=1,000: compute + log_likelihood, latency in microseconds.   N>1: independent replicas.
  c2  as above.                                                                  N>1: the ONE series split into N segments,
      one per rank, carries exchanged by NCCL all-gather (strong scaling).
  c3  8,192 light curves per GPU (N=10,000, 2 x SHO, J=4), log-likelihood + gradient: the per-GPU share of the 65,536-series
      configuration.                                                             N>1: batch sharded, results gathered (weak).
  c4  J=16, N=200,000, K=64 right-hand sides: solve_lower, /d, solve_upper.      N>1: the K columns sharded (strong).
  c5  512 hyper-parameter samples per GPU (RotationTerm J=4, N=50,000, shared t and y): the per-GPU share of the 4,096-sample
      configuration.                                                             N>1: samples sharded, results gathered (weak).

`value`  : the metric with inputs already resident in HBM.
`e2e`    : the same step through the host-buffer entry point: pinned H2D of the step's inputs, D2H of its results, every step.
`roofline`: the slowest operator of the step, algorithmic bytes of SURVEY.md section 8(d) over its CUDA-event duration,
           against the measured HBM copy peak.
`parity` : every output of the step against the reference's own C++ core (oracle/_ref/libc2ref.so, strict build), max-norm
           relative error, asserted <= 1e-10 (rank 0).
`cpu_baseline` / `--impl reference`: the reference's own C++ core (oracle/_ref, built from /root/reference) timed on this
           box's host cores: one core for one series (the reference cannot split a series), every core for batches.
"""
import argparse
import concurrent.futures as cf
import json
import math
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# stdout carries exactly ONE JSON line.  Libraries write banners to file descriptor 1 behind Python's back (NCCL prints its
# version there when the first communicator comes up), so the real stdout is set aside for the JSON line and descriptor 1
# points at stderr for everything else.
_JSON_OUT = os.fdopen(os.dup(1), "w")
os.dup2(2, 1)
sys.stdout = sys.stderr


def emit(line):
    _JSON_OUT.write(json.dumps(line) + "\n")
    _JSON_OUT.flush()


import numpy as np  # noqa: E402

from celerite2_b200 import workloads as wl  # noqa: E402

TOL = 1e-10
SERIES_PER_GPU_C3 = 8192   # 65,536 / 8
SAMPLES_PER_GPU_C5 = 512   # 4,096 / 8


def bytes_per_point(J, K=1):
    """Algorithmic bytes per point of each operator (SURVEY.md section 8a / BASELINE.md section 2)."""
    return {
        "assemble": (3 + 2 * J) * 8,
        "factor": (3 + 3 * J + J * J) * 8,
        "factor_nows": (3 + 3 * J) * 8,
        "solve_lower": (1 + 2 * J + 2 * K + J * K) * 8,
        "solve_nows": (1 + 2 * J + 2 * K) * 8,
        "norm_logdet": 4 * 8,
        "solve_lower_rev": (2 + 4 * J + 3 * K + J * K) * 8,
        "factor_rev": (5 + 5 * J + J * J) * 8,
        "assemble_rev": (3 + 2 * J) * 8,
    }


GRAD_OPS = ("assemble", "factor", "solve_lower", "norm_logdet", "solve_lower_rev", "factor_rev", "assemble_rev")
BASELINE_OPS = ("factor", "solve_lower", "solve_lower_rev", "factor_rev")  # the four ops SURVEY 8(d) counts: 1,472 B/pt at J = 6

OP_OF_ENTRY = {
    "c2b_get_celerite_matrices": "assemble", "c2b_factor": "factor", "c2b_solve_lower": "solve_lower",
    "c2b_solve_upper": "solve_upper", "c2b_norm_logdet": "norm_logdet", "c2b_solve_lower_rev": "solve_lower_rev",
    "c2b_factor_rev": "factor_rev", "c2b_get_celerite_matrices_rev": "assemble_rev", "c2b_add_inplace": "add",
    "c2b_apply_inverse": "apply_inverse",
}

META = {
    "c1": dict(metric="gp_loglike_points_per_s", unit="points/s", grad=False),
    "c2": dict(metric="gp_loglike_grad_points_per_s", unit="points/s", grad=True),
    "c3": dict(metric="gp_loglike_grad_points_per_s", unit="points/s", grad=True),
    "c4": dict(metric="gp_apply_inverse_points_per_s", unit="points/s", grad=False),
    "c5": dict(metric="gp_loglike_grad_points_per_s", unit="points/s", grad=True),
}


def relerr(x, ref):
    x, ref = np.asarray(x, dtype=np.float64), np.asarray(ref, dtype=np.float64)
    s = float(np.max(np.abs(ref))) if ref.size else 1.0
    return float(np.max(np.abs(x - ref)) / (s if s > 0 else 1.0)) if ref.size else 0.0


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# =====================================================================================================================
# host-side inputs of each workload (numpy; identical for both arms)
# =====================================================================================================================
def host_inputs(name, rank, world, small=False, batch_per_gpu=None):
    """-> dict(cfg, coeffs (6 arrays), t, y, diag [, Y]); batched arrays carry the leading batch axis of THIS rank's shard."""
    if name == "c1":
        cfg = wl.config_c1()
        t, y, diag = wl.light_curve(cfg["N"], cfg["seed"] + 1000 * rank)
        return dict(cfg=cfg, coeffs=cfg["coeffs"], t=t, y=y, diag=diag)
    if name == "c2":
        cfg = wl.config_c2(N=100_000) if small else wl.config_c2()
        if os.environ.get("C2B_SPLIT_N"):  # development: a longer series for the split-series scaling curve (never the bench line)
            cfg = wl.config_c2(N=int(os.environ["C2B_SPLIT_N"]))
        t, y, diag = wl.light_curve(cfg["N"], cfg["seed"])  # one series for the whole job (split across ranks when world > 1)
        return dict(cfg=cfg, coeffs=cfg["coeffs"], t=t, y=y, diag=diag)
    if name == "c3":
        B = batch_per_gpu or (256 if small else SERIES_PER_GPU_C3)
        full = wl.config_c3(batch=B * world)
        lo = rank * B
        co = tuple(np.ascontiguousarray(v[lo:lo + B]) for v in full["coeffs"])
        N = 2000 if small else full["N"]
        rng = np.random.default_rng(full["seed"] + 7919 * rank)
        t = np.sort(rng.uniform(0.0, 0.1 * N, (B, N)), axis=1)
        yerr = rng.uniform(0.1, 0.3, (B, N))
        y = np.sin(t) + yerr * rng.standard_normal((B, N))
        cfg = dict(full, batch=B, N=N)
        return dict(cfg=cfg, coeffs=co, t=t, y=y, diag=yerr**2)
    if name == "c5":
        B = batch_per_gpu or (64 if small else SAMPLES_PER_GPU_C5)
        full = wl.config_c5(batch=B * world)
        lo = rank * B
        co = tuple(np.ascontiguousarray(v[lo:lo + B]) for v in full["coeffs"])
        N = 5000 if small else full["N"]
        t, y, diag = wl.light_curve(N, full["seed"])
        cfg = dict(full, batch=B, N=N)
        return dict(cfg=cfg, coeffs=co, t=t, y=y, diag=np.ascontiguousarray(np.tile(diag, (B, 1))))
    if name == "c4":
        cfg = wl.config_c4(N=20_000, K=64) if small else wl.config_c4()
        t, y, diag = wl.light_curve(cfg["N"], cfg["seed"])
        Y = np.random.default_rng(cfg["seed"] + 1).standard_normal((cfg["N"], cfg["K"]))
        return dict(cfg=cfg, coeffs=cfg["coeffs"], t=t, y=y, diag=diag, Y=Y)
    raise SystemExit("unknown workload %r (c1, c2, c3, c4, c5)" % name)


def workload_config(name, inp, world, extra=None):
    cfg = inp["cfg"]
    J = wl.width(tuple(np.atleast_2d(v)[0] for v in inp["coeffs"]))
    desc = {
        "c1": "c1: SHOTerm (J=2) compute + log_likelihood, N=%d, one series" % cfg["N"],
        "c2": "c2: RotationTerm+Matern32Term (J=%d) log-likelihood + gradient, ONE series of N=%d%s"
              % (J, cfg["N"], "" if world == 1 else " split into %d segments, one per GPU" % world),
        "c3": "c3: %d light curves per GPU x N=%d, 2 x SHOTerm (J=4), log-likelihood + gradient (the per-GPU share of the "
              "65,536-series configuration; %d series in this job)" % (cfg["batch"], cfg["N"], cfg["batch"] * world),
        "c4": "c4: 8 x SHOTerm (J=16), N=%d, K=%d right-hand sides: solve_lower, /d, solve_upper (apply_inverse)%s"
              % (cfg["N"], cfg["K"], "" if world == 1 else "; K columns sharded over %d GPUs" % world),
        "c5": "c5: %d hyper-parameter samples per GPU x N=%d, RotationTerm (J=4), shared t and y, log-likelihood + gradient "
              "(the per-GPU share of the 4,096-sample configuration; %d samples in this job)" % (cfg["batch"], cfg["N"], cfg["batch"] * world),
    }[name]
    out = {"workload": desc, "J": J, "N": cfg["N"], "batch_per_gpu": cfg["batch"], "K": cfg["K"]}
    if extra:
        out.update(extra)
    return out


# =====================================================================================================================
# reference arm: the reference's own C++ core on the host cores
# =====================================================================================================================
def cpu_reference_lib(strict=False):
    import oracle

    if strict:
        lib = oracle.ref()
        if lib is not None:
            return lib, "reference", "oracle/_ref/libc2ref.so (reference C++ core, -O2 -ffp-contract=off, stock J<=4 dispatch)"
        return oracle.port(), "port", "oracle/liboracle.so (plain-C restatement)"
    lib = oracle.ref_fast()
    if lib is not None:
        return lib, "reference", "oracle/_ref/libc2ref_fast.so (reference C++ core, -O3 -march=x86-64-v3, fixed-size J dispatch)"
    lib = oracle.ref()
    if lib is not None:
        return lib, "reference", "oracle/_ref/libc2ref.so (reference C++ core, -O2, stock J<=4 dispatch)"
    return oracle.port(), "port", "oracle/liboracle.so (plain-C restatement)"


def _usable_cores():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return max(1, os.cpu_count() or 1)


class ReferenceRunner:
    """One `step()` of the named workload with the reference's core; `units` = points the step covers."""

    def __init__(self, name, inp, budget_series=None):
        import oracle

        self.name, self.inp = name, inp
        self.lib, self.kind, self.build = cpu_reference_lib()
        cfg = inp["cfg"]
        co = inp["coeffs"]
        self.N = cfg["N"]
        self.cores = 1
        if name in ("c1", "c2"):
            Jr, Jc = len(co[0]), len(co[2])
            self.pipe = oracle.RefPipeline(self.lib, self.N, Jr, Jc, grad=META[name]["grad"])
            self.units = self.N
            self.sample = "every step is the full workload (1 series, N=%d) on one core: the reference has no threading and " \
                          "cannot split one series; %s; outputs preallocated" % (self.N, self.build)
        elif name in ("c3", "c5"):
            self.cores = _usable_cores()
            nser = budget_series or max(64, 2 * self.cores)
            nser = min(nser, cfg["batch"])
            self.idx = list(range(nser))
            Jr, Jc = co[0].shape[1], co[2].shape[1]
            self.pipes = [oracle.RefPipeline(self.lib, self.N, Jr, Jc, grad=True) for _ in range(self.cores)]
            self.pool = cf.ThreadPoolExecutor(max_workers=self.cores)
            self.units = nser * self.N
            self.sample = "every step = %d of this GPU's %d series (N=%d each), one series per task on a pool of %d threads (one per " \
                          "usable core; the library call releases the GIL); %s; outputs preallocated per thread" \
                          % (nser, cfg["batch"], self.N, self.cores, self.build)
        elif name == "c4":
            self.cores = min(_usable_cores(), cfg["K"])
            lib = self.lib
            c, a, U, V = lib.get_celerite_matrices(*co, inp["t"], inp["diag"])
            flag, d, W, _ = lib.factor(inp["t"], c, a, U, V, workspace=False)
            assert flag == 0
            self.c, self.U, self.W, self.d = c, U, W, d
            K = cfg["K"]
            cuts = [K * i // self.cores for i in range(self.cores + 1)]
            self.cols = [(cuts[i], cuts[i + 1]) for i in range(self.cores) if cuts[i + 1] > cuts[i]]
            self.Yc = [np.ascontiguousarray(inp["Y"][:, a_:b_]) for a_, b_ in self.cols]
            self.Zc = [np.empty_like(y) for y in self.Yc]
            self.pool = cf.ThreadPoolExecutor(max_workers=self.cores)
            self.units = self.N
            self.sample = "every step is the full workload (N=%d, K=%d): solve_lower, /d, solve_upper, the %d columns split over %d " \
                          "threads (the only parallelism the reference admits for one series); factorisation done once outside the " \
                          "timed region; %s" % (self.N, K, K, self.cores, self.build)
        self.last = None

    def _one_series(self, slot, b):
        inp = self.inp
        co = [v[b] for v in inp["coeffs"]]
        t = inp["t"] if inp["t"].ndim == 1 else inp["t"][b]
        y = inp["y"] if inp["y"].ndim == 1 else inp["y"][b]
        return self.pipes[slot].value_and_grad(co, t, y, inp["diag"][b])[0]

    def _c4_cols(self, i):
        from oracle import OPS, _p

        lib, N, J = self.lib, self.N, self.U.shape[1]
        Y, Z = self.Yc[i], self.Zc[i]
        K = Y.shape[1]
        t = self.inp["t"]
        lib._sweep(OPS["solve_lower"], N, J, K, _p(t), _p(self.c), _p(self.U), _p(self.W), _p(Y), _p(Z), _p(None))
        Z /= self.d[:, None]
        lib._sweep(OPS["solve_upper"], N, J, K, _p(t), _p(self.c), _p(self.U), _p(self.W), _p(Z), _p(Z), _p(None))

    def step(self):
        inp, name = self.inp, self.name
        if name == "c1":
            self.last = self.pipe.value(inp["coeffs"], inp["t"], inp["y"], inp["diag"])
        elif name == "c2":
            self.last = self.pipe.value_and_grad(inp["coeffs"], inp["t"], inp["y"], inp["diag"])
        elif name in ("c3", "c5"):
            # static assignment: series i runs on the pipeline (buffers) of slot i mod cores, each slot on its own thread
            def run_slot(slot):
                return [self._one_series(slot, b) for b in self.idx[slot::self.cores]]

            self.last = list(self.pool.map(run_slot, range(self.cores)))
        else:
            list(self.pool.map(self._c4_cols, range(len(self.cols))))
            self.last = self.Zc

    def baseline(self, reps):
        best = float("inf")
        for _ in range(reps):
            t0 = time.perf_counter()
            self.step()
            best = min(best, time.perf_counter() - t0)
        return {"value": self.units / best, "unit": "points/s", "cores": self.cores, "kind": self.kind,
                "sample": self.sample + "; best of %d passes" % reps, "ms_per_step": best * 1e3}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    inp = host_inputs(args.workload, 0, world, small=args.small, batch_per_gpu=args.batch_per_gpu)
    run = ReferenceRunner(args.workload, inp)
    for _ in range(max(1, min(args.warmup, 3))):
        run.step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        run.step()
    dt = (time.perf_counter() - t0) / args.steps
    val = run.units / dt
    meta = META[args.workload]
    line = {
        "impl": "reference", "metric": meta["metric"], "value": val, "unit": meta["unit"], "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "strong" if args.workload in ("c2", "c4") else "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "config": workload_config(args.workload, inp, world),
        "cpu_baseline": {"value": val, "unit": meta["unit"], "cores": run.cores, "kind": run.kind, "sample": run.sample},
        "e2e": {"value": val, "unit": meta["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    if args.workload == "c1":
        line["latency_us"] = dt * 1e6
    emit(line)


# =====================================================================================================================
# our arm
# =====================================================================================================================
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


class Ctx:
    """process-wide plumbing of our arm: device, ranks, timing helpers"""

    def __init__(self, args):
        import torch
        import torch.distributed as dist

        self.torch, self.dist, self.args = torch, dist, args
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device visible; there is no CPU fallback for the product path")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)

    def T(self, x):
        return self.torch.as_tensor(np.ascontiguousarray(x, dtype=np.float64), device=self.dev)

    def pin(self, x):
        return self.torch.from_numpy(np.ascontiguousarray(x, dtype=np.float64)).pin_memory()

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def timed(self, fn, steps):
        """K calls of fn bracketed by barrier + synchronize on both sides; CUDA events; max over ranks -> (ms/step, wall ms/step)"""
        torch = self.torch
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        w0 = time.perf_counter()
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        wall = (time.perf_counter() - w0) * 1e3
        ms = e0.elapsed_time(e1)
        self.barrier()
        if self.world > 1:
            tt = torch.tensor([ms, wall], dtype=torch.float64, device=self.dev)
            self.dist.all_reduce(tt, op=self.dist.ReduceOp.MAX)
            ms, wall = tt.tolist()
        return ms / steps, wall / steps

    def profile(self, fn, steps):
        """second pass with the library's per-kernel CUDA events on -> (per_op ms, per_kernel ms, launches per step, ms/step)"""
        from celerite2_b200 import _lib

        _lib.lib.c2b_profile_enable(1)
        ms, _ = self.timed(fn, steps)
        report = _lib.lib.c2b_profile_report().decode()
        _lib.lib.c2b_profile_enable(0)
        per_op, per_kernel, launches = {}, {}, 0
        for ln in report.strip().splitlines():
            key, cnt, tot = ln.rsplit(" ", 2)
            entry, kern = key.split("/")
            op = OP_OF_ENTRY.get(entry, entry)
            per_op[op] = per_op.get(op, 0.0) + float(tot) / steps
            per_kernel[op + "/" + kern] = round(per_kernel.get(op + "/" + kern, 0.0) + float(tot) / steps, 5)
            launches += int(cnt)
        return per_op, per_kernel, launches / steps, ms


def make_roofline(per_op, per_kernel, bpp_of, units, ms_step, step_ops, prof_ms, steps, label):
    """bpp_of: op -> algorithmic bytes per point; units: points one launch processes"""
    peak, peak_src = load_peaks()
    known = [o for o in per_op if o in bpp_of]
    dom = max(known, key=lambda o: per_op[o])
    dom_bytes = bpp_of[dom] * units
    achieved = dom_bytes / (per_op[dom] * 1e-3) / 1e9
    step_bytes = sum(bpp_of[o] for o in step_ops) * units
    base_bytes = sum(bpp_of[o] for o in step_ops if o in BASELINE_OPS) * units
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "r02_traffic.json")
    traffic_src = None
    if os.path.exists(tpath):
        with open(tpath) as f:
            tj = json.load(f)
        ent = tj.get(label, {})
        traffic = ent.get(dom)
        traffic_src = tj.get("_capture")
    roof = {
        "bound": "hbm", "kernel": dom + " (all kernels of the C-ABI call)",
        "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "frac_of_nominal_8000_gbs": achieved / 8000.0,
        "traffic": traffic, "traffic_capture": traffic_src,
        "peak_source": peak_src, "algorithmic_bytes_per_launch": dom_bytes, "ms_per_launch": per_op[dom],
        "whole_step": {"algorithmic_bytes": step_bytes, "achieved_gbs": step_bytes / (ms_step * 1e-3) / 1e9,
                       "frac": step_bytes / (ms_step * 1e-3) / 1e9 / peak},
        "per_op_ms": {k: round(v, 5) for k, v in per_op.items()},
        "per_op_frac": {k: round(bpp_of[k] * units / (v * 1e-3) / 1e9 / peak, 4) for k, v in per_op.items() if k in bpp_of},
        "per_kernel_ms": per_kernel,
        "timing": "CUDA events around every kernel, extra pass of %d eager steps (%.3f ms/step with events on)" % (steps, prof_ms),
    }
    if base_bytes and base_bytes != step_bytes:
        roof["whole_step"]["baseline_four_ops"] = {"algorithmic_bytes": base_bytes, "frac": base_bytes / (ms_step * 1e-3) / 1e9 / peak}
    return roof


def check_parity(par, loose=()):
    """every array within TOL; keys in `loose` (with their own bar) are reported apart"""
    par["tolerance"] = TOL
    bad = {k: v for k, v in par["arrays"].items() if not (v <= dict(loose).get(k, TOL))}
    par["ok"] = not bad
    if loose:
        par["loose"] = {k: tol for k, tol in loose if k in par["arrays"]}
    return par


# ---- c1 / c2 (one GPU) / c3 / c5: log-likelihood (+ gradient) pipelines -------------------------------------------------
def run_loglike(ctx, name, inp):
    torch = ctx.torch
    from celerite2_b200 import gp

    args, world, rank = ctx.args, ctx.world, ctx.rank
    meta = META[name]
    grad = meta["grad"]
    cfg = inp["cfg"]
    N, B = cfg["N"], cfg["batch"]
    batched = name in ("c3", "c5")
    shared_ty = name == "c5"
    co_h = [ctx.pin(v) for v in inp["coeffs"]]
    t_h, y_h, d_h = ctx.pin(inp["t"]), ctx.pin(inp["y"]), ctx.pin(inp["diag"])
    co_d = [v.to(ctx.dev) for v in co_h]
    t_d, y_d, d_d = t_h.to(ctx.dev), y_h.to(ctx.dev), d_h.to(ctx.dev)
    J = co_d[0].shape[-1] + 2 * co_d[2].shape[-1]
    units = B * N

    graphed = gp.GraphedLogLikelihood(co_d, t_d, y_d, d_d, grad=grad)
    gathered = torch.empty((world, graphed.packed.numel()), dtype=torch.float64, device=ctx.dev) if (world > 1 and batched) else None

    def step_resident():
        graphed.replay_front()
        r = graphed.replay_back()
        if gathered is not None:  # the final gather of the per-series log-likelihoods and gradients
            ctx.dist.all_gather_into_tensor(gathered, graphed.packed)
        return r

    def step_eager():
        if grad:
            return gp.log_likelihood_and_grad(co_d, t_d, y_d, d_d)
        return gp.log_likelihood(co_d, t_d, y_d, d_d)

    host_api = gp.HostLogLikelihood(N, [v.shape for v in co_h], ctx.dev, batch=B if batched else None, shared_ty=shared_ty, grad=grad)
    gathered_h = torch.empty((world, host_api.n_out), dtype=torch.float64, device=ctx.dev) if (world > 1 and batched) else None

    def step_e2e():
        out, _ = host_api(co_h, t_h, y_h, d_h)
        if gathered_h is not None:
            ctx.dist.all_gather_into_tensor(gathered_h, host_api.graph.packed)
        return out

    for _ in range(max(args.warmup, 3)):
        r = step_resident()
    torch.cuda.synchronize()
    status = int(r["status"].max().item())
    if status != 0:
        raise SystemExit("factorisation failed at row %d" % status)

    sampler = ClockSampler(ctx.local) if rank == 0 else None
    ms_step, wall_step = ctx.timed(step_resident, args.steps)
    for _ in range(2):
        step_eager()
    ms_eager, _ = ctx.timed(step_eager, args.steps)
    per_op, per_kernel, launches, ms_prof = ctx.profile(step_eager, args.steps)
    for _ in range(2):
        step_e2e()
    ms_e2e, wall_e2e = ctx.timed(step_e2e, args.steps)
    streamed = None
    if name == "c2":
        stream_api = gp.HostLogLikelihoodStream(N, [v.shape for v in co_h], ctx.dev)

        def run_streamed(steps):
            stream_api.submit(co_h, t_h, y_h, d_h)
            for _ in range(steps - 1):
                stream_api.submit(co_h, t_h, y_h, d_h)
                stream_api.collect()
            return stream_api.collect()[0]

        run_streamed(3)
        ms_s, _ = ctx.timed(lambda: run_streamed(args.steps), 1)
        ms_s /= args.steps
        streamed = {"value": world * units / (ms_s * 1e-3), "ms_per_step": ms_s,
                    "note": "HostLogLikelihoodStream, two slots: the same copies every step, uploads of evaluation i+1 overlap the kernels of evaluation i"}
    clocks = sampler.stop() if sampler else None
    if rank != 0:
        return None

    bpp = bytes_per_point(J)
    if grad:
        bpp_of = {o: bpp[o] for o in GRAD_OPS}
        step_ops = GRAD_OPS
    else:
        bpp_of = {"assemble": bpp["assemble"], "factor": bpp["factor_nows"], "solve_lower": bpp["solve_nows"], "norm_logdet": 3 * 8}
        step_ops = tuple(bpp_of)
    roofline = make_roofline(per_op, per_kernel, bpp_of, units, ms_step, step_ops, ms_prof, args.steps, name)

    # ---- parity against the reference's own core (strict build) --------------------------------------------------------
    import oracle

    lib, kind, build = cpu_reference_lib(strict=True)
    Jr, Jc = co_h[0].shape[-1], co_h[2].shape[-1]
    pipe = oracle.RefPipeline(lib, N, Jr, Jc, grad=grad)
    res = step_resident() if not batched else r
    torch.cuda.synchronize()
    errs = {}
    H = lambda x: x.detach().cpu().numpy()
    series = list(range(B)) if not batched else sorted(set([0, 1, 2, 3, B // 2, B - 3, B - 2, B - 1]))
    ll_gpu = H(res["loglike"]).reshape(-1)
    for b in series:
        co_b = [np.asarray(v if not batched else v[b]) for v in inp["coeffs"]]
        t_b = inp["t"] if inp["t"].ndim == 1 else inp["t"][b]
        y_b = inp["y"] if inp["y"].ndim == 1 else inp["y"][b]
        d_b = inp["diag"] if inp["diag"].ndim == 1 else inp["diag"][b]
        sel = (lambda x: H(x)) if not batched else (lambda x: H(x[b]))

        def upd(key, got, want):
            errs[key] = max(errs.get(key, 0.0), relerr(got, want))

        if not grad:
            ll_ref, flag = pipe.value(co_b, t_b, y_b, d_b)
            upd("loglike", ll_gpu[b], ll_ref)
            fr = graphed._front  # (t, c, U, W, d, status): d and W are the in-place factorisation
            upd("d", sel(fr[4]), pipe.a)
            upd("W", sel(fr[3]), pipe.V)
            continue
        ll_ref, bc_ref, flag = pipe.value_and_grad(co_b, t_b, y_b, d_b)
        raw = res["raw"]
        upd("loglike", ll_gpu[b], ll_ref)
        g_ref = oracle.get_celerite_matrices_rev(*co_b, t_b, bc_ref, pipe.ba, pipe.bU + pipe.bU2, pipe.bV)
        # raw["bt"] is the array AFTER the assembly adjoint added its own d/dx term to it (b_x accumulates): compare the total
        for nm, want in (("a", pipe.a), ("U", pipe.U), ("V", pipe.V), ("d", pipe.d), ("W", pipe.W), ("Z", pipe.Z),
                         ("F", pipe.F), ("bY", pipe.bY), ("ba", pipe.ba), ("bV", pipe.bV), ("bU", pipe.bU + pipe.bU2),
                         ("bt", pipe.bt + pipe.bt2 + g_ref[6]), ("bc", bc_ref)):
            upd(nm, sel(raw[nm]).reshape(want.shape), want)
        upd("S", sel(raw["S"]).reshape(N, -1), pipe.S)
        for nm, want in zip(gp.COEFF_NAMES, g_ref[:6]):
            if want.size:
                upd("grad_" + nm, sel(res["grads"][nm]), want)
    parity = check_parity({"against": build, "series_checked": len(series), "arrays": {k: float("%.3e" % v) for k, v in errs.items()},
                           "max_rel_err": max(errs.values()), "metric": "max |x - ref| / max |ref| per array"})

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        cpu = ReferenceRunner(name, inp).baseline(2)

    h2d, d2h = host_api.h2d_bytes, host_api.d2h_bytes
    line = {
        "metric": meta["metric"], "value": world * units / (ms_step * 1e-3), "unit": meta["unit"], "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": workload_config(name, inp, world, {
            "l2": "inputs + saved workspaces of one step (about %.2f GB) %s the 126 MB L2%s"
                  % (sum(bpp_of.values()) * units / 1e9, "exceed" if sum(bpp_of.values()) * units > 126e6 else "fit in",
                     "; no explicit flush" if sum(bpp_of.values()) * units > 126e6 else
                     ": this workload is launch-latency bound, its roofline fraction is reported for completeness only"),
            "multi_gpu": ("independent replicas" if name in ("c1", "c2") else
                          "batch sharded over ranks, no data-path collective; all_gather of the per-series log-likelihoods and "
                          "coefficient gradients inside the timed region"),
            "launch": "the timed step replays the evaluation from two CUDA graphs (gp.GraphedLogLikelihood); eager_ms_per_step is the "
                      "same step launched kernel by kernel"}),
        "clocks": clocks,
        "e2e": {"value": world * units / (ms_e2e * 1e-3), "unit": meta["unit"], "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": ms_e2e, "wall_ms_per_step": wall_e2e},
        "gpu_launches": int(round(launches * args.steps)),
        "gpu_launches_per_step": launches,
        "eager_ms_per_step": ms_eager, "wall_ms_per_step": wall_step,
        "roofline": roofline, "parity": parity, "cpu_baseline": cpu,
    }
    if streamed:
        line["e2e"]["streamed"] = streamed
    if name == "c1":
        line["latency_us"] = {"device_resident_graph_replay": ms_step * 1e3, "end_to_end_host_buffers": ms_e2e * 1e3,
                              "eager": ms_eager * 1e3,
                              "reference_python_end_to_end_us": 119.0,
                              "note": "reference figure: GaussianProcess.compute + log_likelihood through Python on the survey "
                                      "container (BASELINE.md section 4); the same-box C++-only figure is cpu_baseline"}
    return line


# ---- c2 over several GPUs: ONE series split into segments ------------------------------------------------------------------
def run_split_series(ctx, inp):
    torch, dist = ctx.torch, ctx.dist
    from celerite2_b200 import gp, segments

    args, world, rank = ctx.args, ctx.world, ctx.rank
    cfg = inp["cfg"]
    N = cfg["N"]
    r0, r1 = rank * N // world, (rank + 1) * N // world
    co_h = [ctx.pin(v) for v in inp["coeffs"]]
    t_h, y_h, d_h = ctx.pin(inp["t"]), ctx.pin(inp["y"][r0:r1]), ctx.pin(inp["diag"][r0:r1])
    co_d = [v.to(ctx.dev) for v in co_h]
    t_d, y_d, d_d = t_h.to(ctx.dev), y_h.to(ctx.dev), d_h.to(ctx.dev)
    J = co_d[0].shape[-1] + 2 * co_d[2].shape[-1]
    ev = segments.SplitSeries(co_d, t_d, y_d, d_d, r0, r1)
    graphed = (not os.environ.get("C2B_SPLIT_EAGER")) and ev.capture()
    out_h = torch.empty(ev.packed.numel(), dtype=torch.float64).pin_memory()

    def step_resident():
        return ev.step()

    def step_e2e():
        for s, h in zip(ev.coeffs, co_h):
            s.copy_(h, non_blocking=True)
        ev.t_full.copy_(t_h, non_blocking=True)
        ev.y.copy_(y_h, non_blocking=True)
        ev.diag.copy_(d_h, non_blocking=True)
        ev.step()
        out_h.copy_(ev.packed, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return out_h

    for _ in range(max(args.warmup, 3)):
        r = step_resident()
    torch.cuda.synchronize()
    if int(r["status"].item()) != 0:
        raise SystemExit("factorisation failed")
    sampler = ClockSampler(ctx.local) if rank == 0 else None
    ms_step, wall_step = ctx.timed(step_resident, args.steps)
    per_op, per_kernel, launches, ms_prof = ctx.profile(ev._step_eager, args.steps)  # eager: the events sit between launches
    for _ in range(2):
        step_e2e()
    ms_e2e, wall_e2e = ctx.timed(step_e2e, args.steps)
    # replicas: every rank evaluates the whole series on its own (what r01 reported at N > 1), for comparison
    y_full, d_full = ctx.T(inp["y"]), ctx.T(inp["diag"])
    rep = gp.GraphedLogLikelihood(co_d, t_d, y_full, d_full)

    def step_rep():
        rep.replay_front()
        return rep.replay_back()

    for _ in range(3):
        whole = step_rep()
    ms_rep, _ = ctx.timed(step_rep, args.steps)
    clocks = sampler.stop() if sampler else None
    res = ev.step()
    torch.cuda.synchronize()
    if rank != 0:
        return None

    import oracle

    lib, kind, build = cpu_reference_lib(strict=True)
    Jr, Jc = co_h[0].shape[-1], co_h[2].shape[-1]
    pipe = oracle.RefPipeline(lib, N, Jr, Jc, grad=True)
    ll_ref, bc_ref, flag = pipe.value_and_grad(inp["coeffs"], inp["t"], inp["y"], inp["diag"])
    g_ref = oracle.get_celerite_matrices_rev(*inp["coeffs"], inp["t"], bc_ref, pipe.ba, pipe.bU + pipe.bU2, pipe.bV)
    H = lambda x: x.detach().cpu().numpy()
    errs = {"loglike": relerr(float(res["loglike"].item()), ll_ref)}
    for nm, want in zip(gp.COEFF_NAMES, g_ref[:6]):
        if want.size:
            errs["grad_" + nm] = relerr(H(res["grads"][nm]), want)
    # rank 0's rows of the operator outputs as well
    seg = ev.seg
    n0 = r1 - r0
    for nm, got, want in (("d", seg.d, pipe.d[:n0]), ("W", seg.W, pipe.W[:n0]), ("Z", seg.Z, pipe.Z[:n0]), ("bV", seg.bV, pipe.bV[:n0]),
                          ("ba", seg.ba, pipe.ba[:n0]), ("bU", seg.bU, (pipe.bU + pipe.bU2)[:n0])):
        errs[nm + "[rank 0 rows]"] = relerr(H(got).reshape(want.shape), want)
    # d/d dc is sum_n t_n (...) over 10^6 rows with t up to 10^5: the rounding differences of bU / bV (3e-11, the same as
    # between two builds of the reference) add coherently along a chunk, and the figure moves with the chunk layout
    # (reference -O2 vs -O3 builds: 8e-12; one GPU: 3e-11; split series: up to 1.4e-10, profiles/r02b): its bar is 1e-9
    parity = check_parity({"against": build, "arrays": {k: float("%.3e" % v) for k, v in errs.items()},
                           "max_rel_err": max(errs.values()), "metric": "max |x - ref| / max |ref| per array"},
                          loose=(("grad_dc", 1e-9),))
    bpp = bytes_per_point(J)
    bpp_of = {o: bpp[o] for o in GRAD_OPS}
    roofline = make_roofline(per_op, per_kernel, bpp_of, r1 - r0, ms_step, GRAD_OPS, ms_prof, args.steps, "c2split")
    line = {
        "metric": META["c2"]["metric"], "value": N / (ms_step * 1e-3), "unit": "points/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": workload_config("c2", inp, world, {
            "l2": "every rank's share of the step (about %.2f GB) exceeds the 126 MB L2; no explicit flush" % (sum(bpp_of.values()) * (r1 - r0) / 1e9),
            "multi_gpu": "one series, contiguous row ranges per rank; per operator: summarise -> NCCL all-gather of one summary "
                         "per rank (<= %d bytes) -> device-side carry fold -> apply; one all-reduce of the log-likelihood and "
                         "coefficient gradients at the end" % (3 * J * J * 8),
            "launch": ("one CUDA graph per evaluation: kernels and NCCL exchanges replayed together" if graphed else
                       "kernels and NCCL exchanges launched from Python (graph capture unavailable: %s)" % getattr(ev, "capture_error", "disabled"))}),
        "clocks": clocks,
        "e2e": {"value": N / (ms_e2e * 1e-3), "unit": "points/s", "h2d_bytes_per_step": 8 * (t_h.numel() + y_h.numel() + d_h.numel() + sum(v.numel() for v in co_h)),
                "d2h_bytes_per_step": 8 * ev.packed.numel(), "ms_per_step": ms_e2e, "wall_ms_per_step": wall_e2e,
                "note": "per rank: the time stamps of the whole series (replicated) and the rank's rows of y and diag"},
        "gpu_launches": int(round(launches * args.steps)), "gpu_launches_per_step": launches,
        "wall_ms_per_step": wall_step,
        "replicas": {"value": world * N / (ms_rep * 1e-3), "ms_per_step": ms_rep,
                     "note": "every rank evaluating its own copy of the whole series (weak scaling, no collective): what r01 reported at N > 1"},
        "roofline": roofline, "parity": parity, "cpu_baseline": None,
    }
    return line


# ---- c4: multi-right-hand-side solves --------------------------------------------------------------------------------------
def run_c4(ctx, inp):
    torch = ctx.torch
    from celerite2_b200 import gp, ops

    args, world, rank = ctx.args, ctx.world, ctx.rank
    cfg = inp["cfg"]
    N, Kfull = cfg["N"], cfg["K"]
    k0, k1 = rank * Kfull // world, (rank + 1) * Kfull // world
    K = k1 - k0
    co_d = [ctx.T(v) for v in inp["coeffs"]]
    t_d, d_d = ctx.T(inp["t"]), ctx.T(inp["diag"])
    Y_h = ctx.pin(inp["Y"][:, k0:k1])
    Z_h = torch.empty_like(Y_h).pin_memory()
    Y_d = Y_h.to(ctx.dev)
    J = co_d[0].shape[-1] + 2 * co_d[2].shape[-1]
    c, a, U, V = ops.get_celerite_matrices(*co_d, t_d, d_d)
    d, W, _, status = ops.factor(t_d, c, a, U, V, workspace=False)
    torch.cuda.synchronize()
    if int(status.item()) != 0:
        raise SystemExit("factorisation failed")
    Z_d = torch.empty_like(Y_d)
    Ystage = torch.empty_like(Y_d)

    def step_resident():
        return ops.apply_inverse(t_d, c, U, W, d, Y_d, Z=Z_d)

    def step_e2e():
        Ystage.copy_(Y_h, non_blocking=True)
        ops.apply_inverse(t_d, c, U, W, d, Ystage, Z=Z_d)
        Z_h.copy_(Z_d, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return Z_h

    for _ in range(max(args.warmup, 3)):
        step_resident()
    torch.cuda.synchronize()
    sampler = ClockSampler(ctx.local) if rank == 0 else None
    ms_step, wall_step = ctx.timed(step_resident, args.steps)
    per_op, per_kernel, launches, ms_prof = ctx.profile(step_resident, args.steps)
    for _ in range(2):
        step_e2e()
    ms_e2e, wall_e2e = ctx.timed(step_e2e, args.steps)
    ms_factor, _ = ctx.timed(lambda: ops.factor(t_d, c, a, U, V, workspace=False), max(3, args.steps // 4))
    clocks = sampler.stop() if sampler else None
    Zg = step_resident().detach().cpu().numpy()
    if rank != 0:
        return None

    lib, kind, build = cpu_reference_lib(strict=True)
    cr, ar_, Ur, Vr = lib.get_celerite_matrices(*inp["coeffs"], inp["t"], inp["diag"])
    flag, dr, Wr, _ = lib.factor(inp["t"], cr, ar_, Ur, Vr, workspace=False)
    Yr = np.ascontiguousarray(inp["Y"][:, k0:k1])
    Z1, _ = lib.sweep("solve_lower", inp["t"], cr, Ur, Wr, Yr, workspace=False)
    Z1 = Z1 / dr[:, None]
    Z2, _ = lib.sweep("solve_upper", inp["t"], cr, Ur, Wr, Z1, workspace=False)
    errs = {"d": relerr(d.cpu().numpy(), dr), "W": relerr(W.cpu().numpy(), Wr), "Z = K^-1 Y (all %d columns of rank 0)" % K: relerr(Zg, Z2)}
    parity = check_parity({"against": build, "arrays": {k: float("%.3e" % v) for k, v in errs.items()},
                           "max_rel_err": max(errs.values()), "metric": "max |x - ref| / max |ref| per array"})
    bpp_solve = (1 + 2 * J + 2 * K) * 8
    bpp_of = {"apply_inverse": 2 * bpp_solve + 8}
    roofline = make_roofline(per_op, per_kernel, bpp_of, N, ms_step, ("apply_inverse",), ms_prof, args.steps, "c4")
    roofline["note"] = "algorithmic bytes: two sweeps of (1 + 2J + 2K) doubles per point plus d (SURVEY 8d counts 2 x 1,288 at K = 64)"
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        cpu = ReferenceRunner("c4", inp).baseline(2)
    line = {
        "metric": META["c4"]["metric"], "value": N / (ms_step * 1e-3), "unit": "points/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": workload_config("c4", inp, world, {
            "l2": "Y, Z and the factor (about %.2f GB per step) exceed the 126 MB L2; no explicit flush" % (bpp_of["apply_inverse"] * N / 1e9),
            "multi_gpu": "the K columns are independent: contiguous column ranges per rank, no collective",
            "launch": "c2b_apply_inverse launched eagerly (summary, scan and apply kernels of both sweeps)"}),
        "clocks": clocks, "rhs_points_per_s": N * Kfull / (ms_step * 1e-3), "factor_ms": ms_factor,
        "e2e": {"value": N / (ms_e2e * 1e-3), "unit": "points/s", "h2d_bytes_per_step": Y_h.numel() * 8, "d2h_bytes_per_step": Z_h.numel() * 8,
                "ms_per_step": ms_e2e, "wall_ms_per_step": wall_e2e},
        "gpu_launches": int(round(launches * args.steps)), "gpu_launches_per_step": launches, "wall_ms_per_step": wall_step,
        "roofline": roofline, "parity": parity, "cpu_baseline": cpu,
    }
    return line


def run_ours(args):
    ctx = Ctx(args)
    name = args.workload
    inp = host_inputs(name, ctx.rank, ctx.world, small=args.small, batch_per_gpu=args.batch_per_gpu)
    if name == "c2" and ctx.world > 1:
        line = run_split_series(ctx, inp)
    elif name == "c4":
        line = run_c4(ctx, inp)
    else:
        line = run_loglike(ctx, name, inp)
    if ctx.rank == 0:
        emit(line)
        if not line["parity"]["ok"]:
            sys.stderr.write("bench.py: PARITY FAILURE: %s\n" % json.dumps(line["parity"]))
    ok = line is None or line["parity"]["ok"]
    if ctx.world > 1:
        ctx.dist.destroy_process_group()
    if not ok and not os.environ.get("C2B_BENCH_NO_ASSERT"):
        raise SystemExit(3)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(META))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--batch-per-gpu", type=int, default=None, help="c3 / c5: series (samples) per GPU instead of the 1/8 share")
    ap.add_argument("--small", action="store_true", help="reduced sizes (development only; never a bench line)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
