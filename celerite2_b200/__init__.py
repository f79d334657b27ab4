"""celerite2_b200: the celerite2 semiseparable Gaussian-process core as hand-written FP64 CUDA for B200.

`ops`      device-level operators on torch CUDA tensors (zero-copy into the C ABI)
`driver`   host mirror of the reference's `celerite2.driver` module (numpy in / numpy out)
`backprop` host mirror of the reference's `celerite2.backprop` module
`gp`       log-likelihood (+ gradient) pipelines built from the operators
`workloads` synthetic kernels / light curves of the benchmark configurations

Importing any compute submodule loads celerite2_b200/libcelerite2_b200.so and fails loudly if it is
missing (build it with `python -m celerite2_b200.build`); there is no CPU fallback.
"""
__version__ = "0.1.0"


_SHIM = '''"""`celerite2.%(name)s` served by celerite2_b200 (CUDA, sm_100a): written by celerite2_b200.install_into().

Stands where the reference ships its pybind11 module of the same name (python/celerite2/%(name)s.cpp); the functions
keep the reference's names, positional signatures, in-place conventions and exceptions.  No CPU fallback."""
from celerite2_b200.%(name)s import *  # noqa: F401,F403
from celerite2_b200.%(name)s import LinAlgError, __version__  # noqa: F401
'''


def install_into(package_dir):
    """Re-register the reference's extension-module names for the GPU: write `driver.py` and `backprop.py` into a
    `celerite2` package directory (the place python/celerite2/driver.cpp:482-499 and backprop.cpp:906-919 would be
    installed to), so that the unmodified front-end (`celerite2.GaussianProcess`, `celerite2.terms`, the PyTensor ops)
    dispatches to libcelerite2_b200.so.  Refuses to overwrite a compiled module of the same name."""
    import glob
    import os

    for name in ("driver", "backprop"):
        clash = glob.glob(os.path.join(package_dir, name + ".*.so")) + glob.glob(os.path.join(package_dir, name + ".pyd"))
        if clash:
            raise RuntimeError("celerite2_b200.install_into: %s holds a compiled %s module; remove it first" % (package_dir, name))
        with open(os.path.join(package_dir, name + ".py"), "w") as f:
            f.write(_SHIM % {"name": name})
