"""Side measurements (not the bench line): the batched configurations c3 / c5 (one GPU's shard) and the c4 multi-RHS
solves, device-resident, CUDA events.  Results go to profiles/ by hand."""
import json, math, sys, os, time
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import numpy as np, torch
from celerite2_b200 import gp, ops, workloads as wl

dev = torch.device("cuda:0")
T = lambda x: torch.as_tensor(np.ascontiguousarray(x, dtype=np.float64), device=dev)

def timeit(fn, steps=5, warm=2):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps

out = {}
# ---- c3 shard: 8192 independent light curves, N = 10,000, J = 4 (2 x SHO), loglike + grad
B, N = 8192, 10000
cfg = wl.config_c3(batch=B, N=N)
co = [T(v) for v in cfg["coeffs"]]
g = torch.Generator(device=dev); g.manual_seed(3)
t = torch.sort(torch.rand((B, N), dtype=torch.float64, device=dev, generator=g) * 0.1 * N, dim=1).values
yerr = 0.1 + 0.2 * torch.rand((B, N), dtype=torch.float64, device=dev, generator=g)
y = torch.sin(t) + yerr * torch.randn((B, N), dtype=torch.float64, device=dev, generator=g)
diag = yerr * yerr
del yerr
r = gp.log_likelihood_and_grad(co, t, y, diag)
assert int(r["status"].max().item()) == 0 and bool(torch.isfinite(r["loglike"]).all())
ms = timeit(lambda: gp.log_likelihood_and_grad(co, t, y, diag), steps=3, warm=1)
out["c3_shard"] = dict(batch=B, N=N, J=4, ms_per_step=ms, points_per_s=B * N / ms * 1e3, frac_hbm=896 * B * N / (ms * 1e-3) / 6544e9)
print(out["c3_shard"], flush=True)
del r, t, y, diag, co; torch.cuda.empty_cache()
# ---- c5 shard: 512 hyper-parameter samples, shared t / y, N = 50,000, J = 4 (RotationTerm)
B, N = 512, 50000
cfg = wl.config_c5(batch=B, N=N)
co = [T(v) for v in cfg["coeffs"]]
t_, y_, d_ = wl.light_curve(N, 5)
t, y, diag = T(t_), T(y_), T(np.tile(d_, (B, 1)))
r = gp.log_likelihood_and_grad(co, t, y, diag)
assert int(r["status"].max().item()) == 0
ms = timeit(lambda: gp.log_likelihood_and_grad(co, t, y, diag), steps=5, warm=2)
out["c5_shard"] = dict(batch=B, N=N, J=4, ms_per_step=ms, points_per_s=B * N / ms * 1e3, frac_hbm=896 * B * N / (ms * 1e-3) / 6544e9)
print(out["c5_shard"], flush=True)
del r, t, y, diag, co; torch.cuda.empty_cache()
# ---- c4: J = 16, N = 200,000, K = 64: solve_lower, /d, solve_upper (generic multi-RHS path)
cfg = wl.config_c4()
N, K = cfg["N"], cfg["K"]
t_, y_, d_ = wl.light_curve(N, 4)
co = [T(v) for v in cfg["coeffs"]]
c, a, U, V = ops.get_celerite_matrices(*co, T(t_), T(d_))
tt = T(t_)
d, W, _, st = ops.factor(tt, c, a, U, V, workspace=False)
assert int(st.item()) == 0
Y = torch.randn((N, K), dtype=torch.float64, device=dev)
def solves():
    Z, _ = ops.solve_lower(tt, c, U, W, Y)
    Z = Z / d[:, None]
    Z, _ = ops.solve_upper(tt, c, U, W, Z, Z=Z)
    return Z
ms = timeit(solves, steps=5, warm=2)
msf = timeit(lambda: ops.factor(tt, c, a, U, V, workspace=False), steps=3, warm=1)
out["c4"] = dict(N=N, J=16, K=K, solves_ms=ms, points_per_s=N / ms * 1e3, frac_hbm=2576 * N / (ms * 1e-3) / 6544e9, factor_ms=msf)
print(out["c4"], flush=True)
json.dump(out, open(os.path.join(os.path.dirname(__file__), "..", "gpurun_out", "batched.json"), "w"), indent=1)
