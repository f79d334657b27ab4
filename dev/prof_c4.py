"""Per-kernel times of the c4 multi-RHS solves (J = 16, N = 2e5, K = 64) with the library's built-in profiling."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import numpy as np, torch
from celerite2_b200 import _lib, ops, workloads as wl
dev = torch.device("cuda:0")
T = lambda x: torch.as_tensor(np.ascontiguousarray(x, dtype=np.float64), device=dev)
cfg = wl.config_c4(); N, K = cfg["N"], cfg["K"]
t_, y_, d_ = wl.light_curve(N, 4)
co = [T(v) for v in cfg["coeffs"]]
tt = T(t_)
c, a, U, V = ops.get_celerite_matrices(*co, tt, T(d_))
d, W, _, st = ops.factor(tt, c, a, U, V, workspace=False)
Y = torch.randn((N, K), dtype=torch.float64, device=dev)
def solves():
    Z, _ = ops.solve_lower(tt, c, U, W, Y)
    Z = Z / d[:, None]
    Z, _ = ops.solve_upper(tt, c, U, W, Z)
    return Z
for _ in range(3): solves(); ops.factor(tt, c, a, U, V, workspace=False)
torch.cuda.synchronize()
_lib.lib.c2b_profile_enable(1)
for _ in range(5): solves(); ops.factor(tt, c, a, U, V, workspace=False)
torch.cuda.synchronize()
print(_lib.lib.c2b_profile_report().decode())
_lib.lib.c2b_profile_enable(0)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5): solves()
e1.record(); torch.cuda.synchronize()
print("solves ms", e0.elapsed_time(e1) / 5)
