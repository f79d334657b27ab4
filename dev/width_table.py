"""Per-width timing table: ms per operator at N = 1e6, K = 1 for every compiled width J = 1..16 and
one run-time width (J = 17) for comparison.  Kernels: J real terms / SHO terms as needed.  Output: gpurun_out/r02_width_table.json"""
import json, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import numpy as np, torch
from celerite2_b200 import ops, workloads as wl

dev = torch.device("cuda:0")
T = lambda x: torch.as_tensor(np.ascontiguousarray(x, dtype=np.float64), device=dev)
N = 1_000_000
t_, y_, d_ = wl.light_curve(N, 11)
t, y, diag = T(t_), T(y_), T(d_)

def kernel(J):
    terms = [wl.sho_coeffs(1.0 / (i + 1), 0.5 + 0.7 * i, 1.0 + 0.5 * i) for i in range(J // 2)]
    if J % 2:
        terms.append(wl.real_coeffs(0.7, 0.3))
    return wl.sum_coeffs(*terms)

def timeit(fn, steps=10, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps

out = {}
for J in tuple(range(1, 18)):
    co = [T(v) for v in kernel(J)]
    c, a, U, V = ops.get_celerite_matrices(*co, t, diag)
    d, W, S, st = ops.factor(t, c, a, U, V)
    assert int(st.item()) == 0, J
    Y = y.unsqueeze(-1)
    Z, F = ops.solve_lower(t, c, U, W, Y, workspace=True)
    bz = torch.randn_like(Z)
    bt, bc, bU, bW, bY = ops.solve_lower_rev(t, c, U, W, Y, Z, F, bz)
    bd = torch.randn_like(d)
    row = dict(
        factor=timeit(lambda: ops.factor(t, c, a, U, V)),
        solve_lower=timeit(lambda: ops.solve_lower(t, c, U, W, Y, workspace=True)),
        solve_upper=timeit(lambda: ops.solve_upper(t, c, U, W, Y, workspace=True)),
        matmul_lower=timeit(lambda: ops.matmul_lower(t, c, U, V, Y, workspace=True)),
        solve_lower_rev=timeit(lambda: ops.solve_lower_rev(t, c, U, W, Y, Z, F, bz)),
        factor_rev=timeit(lambda: ops.factor_rev(t, c, a, U, V, d, W, S, bd, bW)),
    )
    out["J=%d" % J] = {k: round(v, 4) for k, v in row.items()}
    print(J, out["J=%d" % J], flush=True)
    del c, a, U, V, d, W, S, Z, F, bt, bc, bU, bW, bY
    torch.cuda.empty_cache()
json.dump({"_what": "ms per operator call, N = 1e6, K = 1, one B200, CUDA events, 10 timed calls after 3 warm-up calls; J = 17 takes the run-time-width path", "rows": out},
          open(os.path.join(os.path.dirname(__file__), "..", "gpurun_out", "r02_width_table.json"), "w"), indent=1)
