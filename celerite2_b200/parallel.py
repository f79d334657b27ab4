"""Multi-GPU plumbing: one process per GPU, torch.distributed (NCCL on GPUs, gloo in the CPU tests).

Two ways the path shards:

1. Independent series / hyper-parameter samples / right-hand sides: contiguous ranges per rank, no data-path collective.
   `shard()` gives the range, `gather_rows()` collects the per-series results (a few doubles each) at the end.
2. One long series over several ranks: every rank owns a contiguous row range and needs the recursion state that
   enters its first row.  The product path does this on the device: `segments.SplitSeries`, the `c2b_*_segment` entry
   points and `c2b_segment_carry`.

`affine_prefix`, `diag_prefix`, `cong_prefix`, `lft_prefix` below are the same fold written in numpy.  They are TEST
ALGEBRA: `tests/test_multi_rank_gloo.py` uses them to check, on CPU with world size 2, that folding per-rank summaries
reproduces the sequential recursion; nothing on the product path calls them.
"""
import os

import numpy as np
import torch
import torch.distributed as dist


def shard(n_items, rank, world):
    """Contiguous, balanced range [lo, hi) of `n_items` units for `rank` of `world`."""
    base, rem = divmod(int(n_items), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def init_from_env(backend=None):
    """Initialise torch.distributed from RANK / WORLD_SIZE / MASTER_* (torchrun); no-op for a single process.
    -> (rank, world, local_rank)"""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        kw = {}
        if backend == "nccl":
            torch.cuda.set_device(local)
            kw["device_id"] = torch.device("cuda", local)
        dist.init_process_group(backend, **kw)
    return rank, world, local


def _world():
    return dist.get_world_size() if dist.is_initialized() else 1


def gather_rows(local, counts=None, group=None):
    """All-gather per-series results along dim 0.  `local`: (n_local, ...) tensor (CPU with gloo, CUDA with NCCL);
    counts: rows per rank when they differ.  -> (n_total, ...) tensor on every rank."""
    world = _world()
    if world == 1:
        return local
    if counts is None:
        out = [torch.empty_like(local) for _ in range(world)]
        dist.all_gather(out, local.contiguous(), group=group)
        return torch.cat(out, dim=0)
    mx = max(counts)
    pad = torch.zeros((mx,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    out = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(out, pad, group=group)
    return torch.cat([o[:c] for o, c in zip(out, counts)], dim=0)


def allgather_summary(summary, group=None):
    """All-gather one flat float64 summary per rank.  -> list of numpy arrays, index = rank."""
    world = _world()
    flat = torch.as_tensor(np.ascontiguousarray(summary, dtype=np.float64).reshape(-1))
    if world == 1:
        return [flat.numpy()]
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")
    flat = flat.to(dev)
    out = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(out, flat, group=group)
    return [o.cpu().numpy() for o in out]


# ---- folding rank summaries into the state entering a rank (host, O(world J^3)) ------------------------------------
def affine_prefix(summaries, rank, J, K):
    """solves (AffV, csrc/monoids.cuh): summary = [M (J*J row-major) | b (J*K, j + J*k)]; state G (J,K).
    `summaries` in PROCESSING order of the ranks; -> state entering `rank`."""
    G = np.zeros((J, K))
    for s in summaries[:rank]:
        M = s[: J * J].reshape(J, J)
        b = s[J * J:].reshape(K, J).T
        G = M @ G + b
    return G


def diag_prefix(summaries, rank, J, K):
    """matmuls (DiagV): summary = [m (J) | b (J*K)]"""
    G = np.zeros((J, K))
    for s in summaries[:rank]:
        G = s[:J, None] * G + s[J:].reshape(K, J).T
    return G


def cong_prefix(summaries, rank, J):
    """factor_rev (Cong): summary = [M | B]; state G -> M^T G M + B"""
    G = np.zeros((J, J))
    for s in summaries[:rank]:
        M = s[: J * J].reshape(J, J)
        B = s[J * J:].reshape(J, J)
        G = M.T @ G @ M + B
    return G


def lft_prefix(summaries, rank, J):
    """factor (Lft): summary = [A | C | Q]; state X = -Sigma -> A Z A^T + C with Z = (I + X Q)^-1 X"""
    X = np.zeros((J, J))
    I = np.eye(J)
    for s in summaries[:rank]:
        A = s[: J * J].reshape(J, J)
        C = s[J * J: 2 * J * J].reshape(J, J)
        Q = s[2 * J * J:].reshape(J, J)
        Z = np.linalg.solve(I + X @ Q, X)
        X = A @ Z @ A.T + C
        X = 0.5 * (X + X.T)
    return X
