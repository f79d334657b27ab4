"""ms per call of the K > 1 products (matmul_lower / matmul_upper / dot_tril) at the c4 shape: J = 16, N = 2e5, K = 64;
run once as is and once with C2B_NO_GROUP=1 (generic thread-per-column kernels) for the comparison."""
import os, sys, json
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import numpy as np, torch
from celerite2_b200 import ops, workloads as wl
dev = torch.device("cuda:0")
T = lambda x: torch.as_tensor(np.ascontiguousarray(x, dtype=np.float64), device=dev)
cfg = wl.config_c4()
N, K = cfg["N"], cfg["K"]
t_, y_, d_ = wl.light_curve(N, cfg["seed"])
t, diag = T(t_), T(d_)
co = [T(v) for v in cfg["coeffs"]]
c, a, U, V = ops.get_celerite_matrices(*co, t, diag)
d, W, S, st = ops.factor(t, c, a, U, V)
Y = torch.randn(N, K, dtype=torch.float64, device=dev)
def timeit(fn, steps=10, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps): fn()
    e1.record(); torch.cuda.synchronize()
    return round(e0.elapsed_time(e1) / steps, 4)
out = dict(mode="generic" if os.environ.get("C2B_NO_GROUP") else "tile kernels",
           matmul_lower=timeit(lambda: ops.matmul_lower(t, c, U, V, Y)), matmul_upper=timeit(lambda: ops.matmul_upper(t, c, U, V, Y)),
           dot_tril=timeit(lambda: ops.dot_tril(t, c, U, W, d, Y)), solve_lower=timeit(lambda: ops.solve_lower(t, c, U, W, Y)))
Z, F = ops.solve_lower(t, c, U, W, Y, workspace=True)
bZ = torch.randn_like(Z)
out["solve_lower_rev"] = timeit(lambda: ops.solve_lower_rev(t, c, U, W, Y, Z, F, bZ), steps=3, warm=1)
Zm, Fm = ops.matmul_lower(t, c, U, V, Y, workspace=True)
out["matmul_lower_rev"] = timeit(lambda: ops.matmul_lower_rev(t, c, U, V, Y, Zm, Fm, bZ), steps=3, warm=1)
print(json.dumps(out))
