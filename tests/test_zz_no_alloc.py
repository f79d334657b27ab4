"""GPU: c2b_solve_lower_noalloc, the solve_lower variant for callers that cannot allocate device memory.

A host that was handed a fixed set of device buffers (an XLA/JAX custom-call handler, a loop that gave the
whole GPU to its own pool, a callback that must not trigger the allocator's device-wide synchronisation)
cannot provide the scan scratch the ordinary entry point wants and must not let the library allocate it
either.  This variant keeps its scratch in an 8 MiB array built into the library.  The test drives it through
the C ABI twice in a row, with different shapes so that the second call finds the first call's leftovers in
the shared array, and checks both results against c2b_solve_lower (tolerance 1e-10 max-norm relative, the
bar used throughout tests/); it also checks that an oversized problem is refused with a clear message.

Collected last on purpose (file name), and written to pass both alone in a fresh process and after the rest.
"""
import ctypes

import numpy as np
import pytest

from helpers import ALL_KERNELS, relerr

pytestmark = pytest.mark.gpu

TOL = 1e-10


def _case(name, N, K, seed):
    import oracle
    from celerite2_b200 import workloads as wl

    P = oracle.port()
    t, y, diag = wl.light_curve(N, seed)
    c, a, U, V = P.get_celerite_matrices(*ALL_KERNELS[name], t, diag)
    flag, d, W, _ = P.factor(t, c, a, U, V, workspace=False)
    assert flag == 0
    Y = np.ascontiguousarray(np.stack([y * (k + 1) + np.cos(k * t) for k in range(K)], axis=1))
    return t, c, U, W, Y


def test_noalloc_matches_regular_entry_point_twice_in_a_row():
    import torch

    from celerite2_b200 import _lib

    lib = _lib.lib
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    T = lambda x: torch.as_tensor(np.ascontiguousarray(x, dtype=np.float64), device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    lib.c2b_testing_set_chunks_per_series(0)
    limit = lib.c2b_noalloc_scratch_limit()
    assert limit == 8 << 20

    primed = set()
    for name, N, K, seed in (("c2_J6", 20000, 1, 31), ("sum4", 30011, 2, 32), ("c2_J6", 20000, 1, 33)):
        t, c, U, W, Y = _case(name, N, K, seed)
        J = U.shape[1]
        need = lib.c2b_scratch_bytes(_lib.OP_IDS["solve_lower"], N, 0, J, K, 1)
        assert 0 < need <= limit, "the case must use scratch (chunked) and fit the built-in array"
        tt, cc, UU, WW, YY = map(T, (t, c, U, W, Y))
        # reference result: the ordinary entry point with a caller-owned scratch buffer
        scratch = torch.empty(need, dtype=torch.uint8, device=dev)
        Z_ref = torch.empty_like(YY)
        F_ref = torch.empty((N, J * K), dtype=torch.float64, device=dev)
        _lib.check(lib.c2b_solve_lower(N, J, K, 1, tt.data_ptr(), cc.data_ptr(), UU.data_ptr(), WW.data_ptr(), YY.data_ptr(),
                                       Z_ref.data_ptr(), F_ref.data_ptr(), _lib.SHARED_T | _lib.SHARED_C, scratch.data_ptr(), need,
                                       stream))
        Z = torch.full_like(YY, float("nan"))
        F = torch.full_like(F_ref, float("nan"))
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info(dev)[0]  # device-wide: anything the call allocated (and kept) would show here
        _lib.check(lib.c2b_solve_lower_noalloc(N, J, K, 1, tt.data_ptr(), cc.data_ptr(), UU.data_ptr(), WW.data_ptr(), YY.data_ptr(),
                                               Z.data_ptr(), F.data_ptr(), _lib.SHARED_T | _lib.SHARED_C, stream))
        torch.cuda.synchronize()
        if (J, K) in primed:  # the first launch of a kernel may load its code; a repeat of the same shape must not move the free memory
            assert torch.cuda.mem_get_info(dev)[0] == free_before
        primed.add((J, K))
        assert relerr(Z.cpu().numpy(), Z_ref.cpu().numpy()) < TOL
        assert relerr(F.cpu().numpy(), F_ref.cpu().numpy()) < TOL
        # and against the CPU oracle, so that a fault shared by both entry points cannot hide
        import oracle

        Z_o, F_o = oracle.port().sweep("solve_lower", t, c, U, W, Y)
        assert relerr(Z.cpu().numpy(), Z_o) < TOL
        assert relerr(F.cpu().numpy(), F_o) < TOL


def test_noalloc_rejects_oversized_problem():
    import torch

    from celerite2_b200 import _lib

    lib = _lib.lib
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    lib.c2b_testing_set_chunks_per_series(0)
    N, J, K, B = 4_000_000, 16, 1, 4
    assert lib.c2b_scratch_bytes(_lib.OP_IDS["solve_lower"], N, 0, J, K, B) > lib.c2b_noalloc_scratch_limit()
    x = torch.zeros(64, dtype=torch.float64, device=dev)  # never dereferenced: the size check comes first
    p = x.data_ptr()
    rc = lib.c2b_solve_lower_noalloc(N, J, K, B, p, p, p, p, p, p, None, 0, torch.cuda.current_stream().cuda_stream)
    assert rc == -2  # C2B_ERR_SCRATCH_TOO_SMALL
    msg = lib.c2b_last_error().decode()
    assert "built-in array" in msg and "c2b_solve_lower" in msg
