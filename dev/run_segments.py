"""One c2 series (N = 1e6, J = 6) split over the ranks of a torchrun job: parity against the single-GPU pipeline and
step time (max over ranks, CUDA events).  gpurun --gpus 2 -- 'python -m torch.distributed.run --nnodes=1 --nproc-per-node 2
--master-addr 127.0.0.1 --master-port 29512 dev/run_segments.py'"""
import json, os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
    os.environ["NCCL_DEBUG"] = "WARN"
import numpy as np, torch, torch.distributed as dist
from celerite2_b200 import gp, segments, workloads as wl

world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
cfg = wl.config_c2()
N = int(os.environ.get("C2B_SPLIT_N", cfg["N"]))
t, y, diag = wl.light_curve(N, cfg["seed"])
T = lambda x: torch.as_tensor(np.ascontiguousarray(x, dtype=np.float64), device=dev)
co = [T(v) for v in cfg["coeffs"]]
r0, r1 = rank * N // world, (rank + 1) * N // world
tt, ys, ds = T(t), T(y[r0:r1]), T(diag[r0:r1])
step = lambda: segments.loglike_and_grad_distributed(co, tt, ys, ds, r0, r1)
for _ in range(3):
    r = step()
torch.cuda.synchronize(); dist.barrier()
K = 20
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(K):
    r = step()
e1.record(); torch.cuda.synchronize()
ms = torch.tensor([e0.elapsed_time(e1) / K], dtype=torch.float64, device=dev)
dist.all_reduce(ms, op=dist.ReduceOp.MAX)
if rank == 0:
    whole = gp.log_likelihood_and_grad(co, tt, T(y), T(diag))
    err = max(float((r["grads"][nm] - whole["grads"][nm]).abs().max() / whole["grads"][nm].abs().max()) for nm in gp.COEFF_NAMES if whole["grads"][nm].numel())
    print(json.dumps({"workload": "c2 kernel, one series of N = %d split over %d GPUs (strong scaling)" % (N, world), "n_gpus": world, "ms_per_step": float(ms), "points_per_s": N / float(ms) * 1e3,
                      "loglike_rel_diff": abs(float(r["loglike"]) - float(whole["loglike"])) / abs(float(whole["loglike"])), "grad_rel_diff": err,
                      "status": int(r["status"])}))
dist.destroy_process_group()
