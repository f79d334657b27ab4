#!/bin/bash
# round-2 profile pass (one GPU): ncu launch list of the default bench command, full captures of one evaluation's kernels,
# exported as CSV on the box (the .ncu-rep files are too large to travel: only the text pages come back)
mkdir -p gpurun_out /tmp/ncu
O=gpurun_out
export C2B_BENCH_NO_ASSERT=1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/r02p_launches_raw.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/r02p_ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on --launch-skip 140 -c 40 -o /tmp/ncu/c2 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/r02p_ncu_full.log 2>&1
ncu -i /tmp/ncu/c2.ncu-rep --page raw --csv > $O/r02p_c2_raw.csv 2>/dev/null
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"gfactor_rev_kernel|gfactor_apply_kernel|gsweep_rev_apply|gsweep_apply" --launch-skip 8 -c 4 -o /tmp/ncu/c3 python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu-baseline > $O/r02p_ncu_c3.log 2>&1
ncu -i /tmp/ncu/c3.ncu-rep --page raw --csv > $O/r02p_c3_raw.csv 2>/dev/null
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"csweep|scan_|fold_" --launch-skip 20 -c 12 -o /tmp/ncu/c4 python bench.py --workload c4 --steps 2 --warmup 3 --no-cpu-baseline > $O/r02p_ncu_c4.log 2>&1
ncu -i /tmp/ncu/c4.ncu-rep --page raw --csv > $O/r02p_c4_raw.csv 2>/dev/null
ls -la /tmp/ncu > $O/r02p_sizes.txt
for f in c2 c3 c4; do sz=$(stat -c %s /tmp/ncu/$f.ncu-rep 2>/dev/null || echo 0); if [ "$sz" -gt 0 ] && [ "$sz" -lt 15000000 ]; then cp /tmp/ncu/$f.ncu-rep $O/r02p_$f.ncu-rep; fi; done
du -sh $O
