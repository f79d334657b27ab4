"""GPU: the reference's OWN Python front-end -- python/celerite2/{core,numpy,terms,testing}.py, unmodified -- running on top
of this repository's driver, i.e. `GaussianProcess.compute / log_likelihood / apply_inverse / dot_tril / predict / sample`
dispatching to the CUDA path through the names `celerite2.driver` and `celerite2.backprop`
(python/celerite2/driver.cpp:482-499, backprop.cpp:906-919; callers python/celerite2/numpy.py:35-121, terms.py:117-207).

oracle/build_pyref.py lays the reference's files out under oracle/_ref/pyref (GPU driver installed by
celerite2_b200.install_into) and oracle/_ref/pyref_cpu (the reference's own pybind11 modules compiled from its sources);
both trees are build products that travel with the snapshot, nothing of the reference is committed.  Each test runs in a
fresh interpreter per tree (two packages named `celerite2` cannot share a process):

  1. the reference's own test files python/test/test_driver.py and test_backprop.py pass on the GPU driver;
  2. a GaussianProcess session gives the same numbers on both drivers (rel 1e-10) and the same error behaviour
     (python/test/test_celerite2.py:178-237: unsorted input, wrong shapes, non-positive-definite matrix, `quiet`).
"""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GPU_TREE = os.path.join(ROOT, "oracle", "_ref", "pyref")
CPU_TREE = os.path.join(ROOT, "oracle", "_ref", "pyref_cpu")

needs_trees = pytest.mark.skipif(not os.path.isdir(os.path.join(GPU_TREE, "celerite2")),
                                 reason="oracle/_ref/pyref missing: run python oracle/build_pyref.py where /root/reference exists")


def _env(tree):
    env = dict(os.environ)
    env["PYTHONPATH"] = os.pathsep.join([tree, ROOT, env.get("PYTHONPATH", "")])
    return env


@needs_trees
@pytest.mark.parametrize("which", ["test_driver.py", "test_backprop.py"])
def test_reference_own_tests_pass_on_the_gpu_driver(which):
    r = subprocess.run([sys.executable, "-m", "pytest", "-q", "-x", "-p", "no:cacheprovider", os.path.join(GPU_TREE, "tests", which)],
                       env=_env(GPU_TREE), cwd=GPU_TREE, capture_output=True, text=True, timeout=1500)
    assert r.returncode == 0, r.stdout[-4000:] + r.stderr[-2000:]
    assert " passed" in r.stdout


_SESSION = r'''
import json, sys, warnings
import numpy as np
import celerite2
from celerite2 import terms, driver
out = {"driver_file": driver.__file__}
rng = np.random.default_rng(40582)
x = np.sort(rng.uniform(0, 10, 300))
t = np.sort(rng.uniform(-1, 12, 120))
diag = rng.uniform(0.1, 0.3, len(x))
y = np.sin(x) + 0.1 * rng.standard_normal(len(x))
Y = np.stack([y, np.cos(x), x ** 2 / 50.0], axis=1)
kernels = {
    "sho": terms.SHOTerm(S0=1.0, w0=0.5, Q=3.0),
    "rot+mat": terms.RotationTerm(sigma=1.5, period=3.45, Q0=1.3, dQ=1.05, f=0.5) + terms.Matern32Term(sigma=1.0, rho=2.0),
    "real+sho_over": terms.RealTerm(a=1.0, c=0.3) + terms.SHOTerm(S0=1.2, w0=0.3, Q=0.1),
    "product": terms.SHOTerm(S0=1.0, w0=0.5, Q=3.0) * terms.RealTerm(a=0.5, c=0.2),
}
for name, k in kernels.items():
    gp = celerite2.GaussianProcess(k, mean=0.1)
    gp.compute(x, diag=diag)
    r = {}
    r["loglike"] = float(gp.log_likelihood(y))
    r["apply_inverse"] = gp.apply_inverse(Y).tolist()
    r["apply_inverse_vec"] = gp.apply_inverse(y).tolist()
    r["dot_tril"] = gp.dot_tril(Y).tolist()
    mu, var = gp.predict(y, t, return_var=True)
    r["predict_mean"] = mu.tolist()
    r["predict_var"] = var.tolist()
    r["predict_self"] = gp.predict(y).tolist()
    r["kernel_dot"] = k.dot(x, diag, Y).tolist()
    np.random.seed(7)
    r["sample"] = gp.sample(size=2).tolist()
    gp.recompute()
    r["loglike_again"] = float(gp.log_likelihood(y))
    out[name] = r
# error behaviour (python/test/test_celerite2.py:178-237)
errs = {}
gp = celerite2.GaussianProcess(terms.SHOTerm(S0=1.0, w0=0.5, Q=3.0))
def raises(exc, fn):
    try:
        fn()
    except exc:
        return True
    except Exception as e:
        return "other: %s" % type(e).__name__
    return False
errs["recompute_before_compute"] = raises(RuntimeError, gp.recompute)
errs["loglike_before_compute"] = raises(RuntimeError, lambda: gp.log_likelihood(y))
errs["unsorted"] = raises(ValueError, lambda: gp.compute(x[::-1], diag=diag))
errs["not_1d"] = raises(ValueError, lambda: gp.compute(np.tile(x[:, None], (1, 5)), diag=diag))
errs["diag_and_yerr"] = raises(ValueError, lambda: gp.compute(x, diag=diag, yerr=np.sqrt(diag)))
errs["not_pd_raises_linalgerror"] = raises(driver.LinAlgError, lambda: gp.compute(x, diag=-10 * diag))
gp.compute(x, diag=-10 * diag, quiet=True)
errs["quiet_logdet_minus_inf"] = bool(np.isinf(gp._log_det) and gp._log_det < 0)
gp.compute(x, diag=diag)
errs["shape_mismatch"] = raises(ValueError, lambda: gp.log_likelihood(y[:-1]))
errs["rhs_2d_loglike"] = raises(ValueError, lambda: gp.log_likelihood(np.tile(y[:, None], (1, 5))))
errs["predict_t_2d"] = raises(ValueError, lambda: gp.predict(y, t=np.tile(t[:, None], (1, 5))))
out["errors"] = errs
json.dump(out, open(sys.argv[1], "w"))
'''


def _run_session(tree, tmp_path, tag):
    script = tmp_path / ("session_%s.py" % tag)
    script.write_text(_SESSION)
    res = tmp_path / ("session_%s.json" % tag)
    r = subprocess.run([sys.executable, str(script), str(res)], env=_env(tree), cwd=str(tmp_path), capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    return json.load(open(res))


@needs_trees
def test_gaussian_process_session_gpu_driver_vs_reference_driver(tmp_path):
    gpu = _run_session(GPU_TREE, tmp_path, "gpu")
    assert gpu["driver_file"].endswith("driver.py"), gpu["driver_file"]  # the shim written by install_into
    assert all(v is True for v in gpu["errors"].values()), gpu["errors"]
    if not os.path.isdir(os.path.join(CPU_TREE, "celerite2")):
        pytest.skip("oracle/_ref/pyref_cpu missing (the reference's pybind11 modules): only the GPU half ran")
    cpu = _run_session(CPU_TREE, tmp_path, "cpu")
    assert cpu["driver_file"].endswith(".so"), cpu["driver_file"]  # the reference's own compiled module
    assert cpu["errors"] == gpu["errors"]
    for name in gpu:
        if name in ("errors", "driver_file"):
            continue
        for key, want in cpu[name].items():
            got = np.asarray(gpu[name][key], dtype=np.float64)
            want = np.asarray(want, dtype=np.float64)
            scale = max(float(np.max(np.abs(want))), 1e-300)
            tol = 1e-10 if key != "predict_var" else 1e-9  # the variance is a difference of two O(1) numbers
            assert float(np.max(np.abs(got - want))) / scale < tol, (name, key, float(np.max(np.abs(got - want))) / scale)
