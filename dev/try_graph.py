import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import numpy as np, torch
from celerite2_b200 import gp, workloads as wl
dev = torch.device("cuda:0")
T = lambda x: torch.as_tensor(np.ascontiguousarray(x, dtype=np.float64), device=dev)
cfg = wl.config_c2(); N = cfg["N"]
t_, y_, d_ = wl.light_curve(N, 1)
co = [T(v) for v in cfg["coeffs"]]; t, y, diag = T(t_), T(y_), T(d_)
def timeit(fn, n=50):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
step = lambda: gp.log_likelihood_and_grad(co, t, y, diag)
print("eager ms", timeit(step))
s = torch.cuda.Stream()
s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    for _ in range(3): step()
torch.cuda.current_stream().wait_stream(s)
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    r = step()
torch.cuda.synchronize()
print("graph ms", timeit(g.replay))
ref = step()
g.replay(); torch.cuda.synchronize()
print("loglike eager/graph", float(ref["loglike"]), float(r["loglike"]))
