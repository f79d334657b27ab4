#!/bin/bash
# profile of the fold kernels (C2B_SCAN=fold): raw metric pages as CSV
mkdir -p gpurun_out /tmp/ncu
O=gpurun_out
export C2B_BENCH_NO_ASSERT=1 C2B_SCAN=fold
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"fold_" --launch-skip 18 -c 9 -o /tmp/ncu/fold python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/r02d_ncu_fold.log 2>&1
ncu -i /tmp/ncu/fold.ncu-rep --page raw --csv > $O/r02d_fold_raw.csv 2>/dev/null
ncu -i /tmp/ncu/fold.ncu-rep --page source --csv --kernel-name regex:fold_groups > $O/r02d_fold_groups_source.csv 2>/dev/null
head -c 3000000 $O/r02d_fold_groups_source.csv > $O/r02d_fold_groups_source_head.csv; rm -f $O/r02d_fold_groups_source.csv
du -sh $O
