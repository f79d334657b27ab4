"""CPU, world_size = 2, gloo: the host side of the multi-GPU paths (celerite2_b200/parallel.py).

  * batch sharding: ranges partition the batch; per-series results gathered in order;
  * one long series over two ranks: rank 1 starts from the state obtained by all-gathering the rank summaries and
    folding rank 0's -- for solve_lower (affine carry) and factor (linear-fractional carry) -- and reproduces the
    oracle's single-process result.
"""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    try:
        for p in (ROOT, os.path.join(ROOT, "tests")):
            if p not in sys.path:
                sys.path.insert(0, p)
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
        import chunk_algebra as ca
        import oracle
        from celerite2_b200 import parallel
        from celerite2_b200 import workloads as wl
        from helpers import relerr

        r, w, _ = parallel.init_from_env(backend="gloo")
        assert (r, w) == (rank, world)

        # ---- batch sharding ---------------------------------------------------------------------------------------
        B = 7
        lo, hi = parallel.shard(B, rank, world)
        spans = [parallel.shard(B, k, world) for k in range(world)]
        assert spans[0][0] == 0 and spans[-1][1] == B and all(spans[k][1] == spans[k + 1][0] for k in range(world - 1))
        local = torch.arange(lo, hi, dtype=torch.float64)[:, None] * torch.ones(1, 3, dtype=torch.float64)
        allr = parallel.gather_rows(local, counts=[b - a for a, b in spans])
        assert torch.equal(allr[:, 0], torch.arange(B, dtype=torch.float64))

        # ---- one long series over two ranks -------------------------------------------------------------------------
        P = oracle.port()
        N, J = 600, 6
        cfg = wl.config_c2(N)
        t, y, diag = wl.light_curve(N, 9)
        c, a, U, V = P.get_celerite_matrices(*cfg["coeffs"], t, diag)
        flag, d_ref, W_ref, S_ref = P.factor(t, c, a, U, V)
        Y = np.stack([y, np.cos(t)], axis=1)
        Z_ref, F_ref = P.sweep("solve_lower", t, c, U, W_ref, Y)
        n0, n1 = parallel.shard(N, rank, world)

        # factor: linear-fractional carry
        summ = parallel.allgather_summary(ca.factor_summary(t, c, a, U, V, n0, n1))
        X = parallel.lft_prefix(summ, rank, J)
        d, W, S, _ = ca.factor_chunk(t, c, a, U, V, n0, n1, -X)
        assert relerr(d, d_ref[n0:n1]) < 1e-10 and relerr(W, W_ref[n0:n1]) < 1e-10 and relerr(S, S_ref[n0:n1]) < 1e-10

        # solve_lower: affine carry (uses the exact W so that the two checks are independent)
        summ = parallel.allgather_summary(ca.solve_lower_summary(t, c, U, W_ref, Y, n0, n1))
        G = parallel.affine_prefix(summ, rank, J, Y.shape[1])
        Z, F, _ = ca.solve_lower_chunk(t, c, U, W_ref, Y, n0, n1, G)
        assert relerr(Z, Z_ref[n0:n1]) < 1e-10 and relerr(F, F_ref[n0:n1]) < 1e-10

        # the log-likelihood pieces reduce with one all_reduce of two doubles
        part = torch.tensor([np.sum(np.log(d)), np.sum(Z[:, 0] ** 2 / d)], dtype=torch.float64)
        dist.all_reduce(part)
        want = np.array([np.sum(np.log(d_ref)), np.sum(Z_ref[:, 0] ** 2 / d_ref)])
        assert relerr(part.numpy(), want) < 1e-12
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, "ok"))
    except Exception as e:  # pragma: no cover
        import traceback

        q.put((rank, "FAIL: %s\n%s" % (e, traceback.format_exc())))


@pytest.mark.timeout(300)
def test_two_rank_gloo_paths():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in results:
        assert msg == "ok", "rank %d: %s" % (rank, msg)


def test_shard_is_a_partition():
    from celerite2_b200 import parallel

    for n in (1, 5, 8, 65536, 4097):
        for world in (1, 2, 3, 8):
            spans = [parallel.shard(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
