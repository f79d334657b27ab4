#!/bin/bash
# compute-sanitizer over the golden-vector suite (run on the GPU box): memcheck and racecheck, logs under gpurun_out/
# (copied to profiles/ afterwards).  The lane-group kernels exchange values between lanes through shared memory with
# __syncwarp in between, which is what racecheck is for.
mkdir -p gpurun_out
SAN=/usr/local/cuda/bin/compute-sanitizer
SEL="tests/test_gpu_parity.py -k golden"
timeout 1500 $SAN --tool memcheck --error-exitcode 9 python -m pytest $SEL -m gpu -q -x > gpurun_out/r02_sanitizer_memcheck.log 2>&1; echo "memcheck rc=$?" >> gpurun_out/r02_sanitizer_memcheck.log
timeout 2400 $SAN --tool racecheck --error-exitcode 9 python -m pytest $SEL -m gpu -q -x > gpurun_out/r02_sanitizer_racecheck.log 2>&1; echo "racecheck rc=$?" >> gpurun_out/r02_sanitizer_racecheck.log
tail -4 gpurun_out/r02_sanitizer_memcheck.log gpurun_out/r02_sanitizer_racecheck.log
