#!/bin/bash
# final profile pass (one GPU): ncu launch list of the default bench command, full capture of the kernels of two
# consecutive evaluations (text pages exported on the box), source page of the dominant kernels
mkdir -p gpurun_out /tmp/ncu
O=gpurun_out
export C2B_BENCH_NO_ASSERT=1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/r02w_launches_raw.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/r02w_ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on --launch-skip 120 -c 80 -f -o /tmp/ncu/c2 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/r02w_ncu_full.log 2>&1
ncu -i /tmp/ncu/c2.ncu-rep --page raw --csv > $O/r02w_c2_raw.csv 2>/dev/null
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"gfactor_rev_kernel|gfactor_apply_kernel|gsweep_rev_apply|gsweep_apply" --launch-skip 8 -c 4 -f -o /tmp/ncu/c3 python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu-baseline > $O/r02w_ncu_c3.log 2>&1
ncu -i /tmp/ncu/c3.ncu-rep --page raw --csv > $O/r02w_c3_raw.csv 2>/dev/null
ls -la /tmp/ncu > $O/r02w_sizes.txt
du -sh $O; tail -3 $O/r02w_ncu_full.log
