#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --maxfail=15 > gpurun_out/r03e_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r03e_pytest.log
tail -12 gpurun_out/r03e_pytest.log
