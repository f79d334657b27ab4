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
