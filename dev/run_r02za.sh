#!/bin/bash
# one GPU: source-level profile of the fold kernels on c4
mkdir -p gpurun_out
O=gpurun_out
export C2B_BENCH_NO_ASSERT=1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:'fold_' --launch-skip 12 -c 3 -f -o $O/r02za python bench.py --workload c4 --steps 2 --warmup 3 --no-cpu-baseline > $O/r02za_ncu.log 2>&1
timeout 300 ncu -i $O/r02za.ncu-rep --page source --csv > $O/r02za_source.csv 2> $O/r02za_source.err
timeout 300 ncu -i $O/r02za.ncu-rep --page raw --csv > $O/r02za_raw.csv 2>> $O/r02za_source.err
ls -la $O/r02za*; rm -f $O/r02za.ncu-rep
