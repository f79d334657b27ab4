#!/bin/bash
# one GPU: K > 1 products on the tile kernels against the generic kernels; full suite
mkdir -p gpurun_out
O=gpurun_out
timeout 300 python dev/bench_matmul_k.py > $O/r03a_matmul_tile.json 2> $O/r03a_matmul_tile.err
C2B_NO_GROUP=1 timeout 300 python dev/bench_matmul_k.py > $O/r03a_matmul_generic.json 2> $O/r03a_matmul_generic.err
cat $O/r03a_matmul_tile.json $O/r03a_matmul_generic.json; tail -3 $O/r03a_matmul_tile.err
timeout 1500 python -m pytest tests -m gpu -q --maxfail=15 > $O/r03a_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r03a_pytest.log
tail -6 $O/r03a_pytest.log
