"""Per-call times of every operator of the C ABI at the c2 size (N = 1e6, J = 6, K = 1), CUDA events."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import numpy as np, torch
from celerite2_b200 import ops, workloads as wl
dev = torch.device("cuda:0")
T = lambda x: torch.as_tensor(np.ascontiguousarray(x, dtype=np.float64), device=dev)
cfg = wl.config_c2(); N = int(os.environ.get("N", cfg["N"]))
t_, y_, d_ = wl.light_curve(N, 1)
co = [T(v) for v in cfg["coeffs"]]
t, y, diag = T(t_), T(y_), T(d_)
c, a, U, V = ops.get_celerite_matrices(*co, t, diag)
d, W, S, st = ops.factor(t, c, a, U, V)
Y = y[:, None].contiguous()
Z, F = ops.solve_lower(t, c, U, W, Y, workspace=True)
bZ = torch.randn_like(Z)
def timeit(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
res = {}
res["factor"] = timeit(lambda: ops.factor(t, c, a, U, V))
res["factor_rev"] = timeit(lambda: ops.factor_rev(t, c, a, U, V, d, W, S, y, W))
for op in ("solve_lower", "solve_upper", "matmul_lower", "matmul_upper"):
    B = W if op.startswith("solve") else V
    res[op] = timeit(lambda: getattr(ops, op)(t, c, U, B, Y, workspace=True))
    Zr, Fr = getattr(ops, op)(t, c, U, B, Y, workspace=True)
    res[op + "_rev"] = timeit(lambda: getattr(ops, op + "_rev")(t, c, U, B, Y, Zr, Fr, bZ))
t2 = T(np.sort(np.random.default_rng(0).uniform(t_[0], t_[-1], N // 2)))
for op in ("general_matmul_lower", "general_matmul_upper"):
    res[op] = timeit(lambda: getattr(ops, op)(t2, t, c, U[: N // 2].contiguous(), V, Y), n=5)
for k, v in res.items(): print("%-24s %.3f ms" % (k, v))
