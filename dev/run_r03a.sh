#!/bin/bash
# one GPU: K > 1 products and reverse sweeps on the tile kernels; full suite
mkdir -p gpurun_out
O=gpurun_out
T=${TAG:-r03c}
timeout 300 python dev/bench_matmul_k.py > $O/${T}_matmul_tile.json 2> $O/${T}_matmul_tile.err
cat $O/${T}_matmul_tile.json; tail -3 $O/${T}_matmul_tile.err
timeout 1500 python -m pytest tests -m gpu -q --maxfail=15 > $O/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> $O/${T}_pytest.log
tail -8 $O/${T}_pytest.log
