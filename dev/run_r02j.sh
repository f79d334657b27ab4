#!/bin/bash
# one GPU: source-level ncu capture of the four heaviest c2 kernels (per-SASS-line instruction counts and stall samples)
mkdir -p gpurun_out
O=gpurun_out
export C2B_BENCH_NO_ASSERT=1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:'gfactor_apply_kernel|gfactor_rev_kernel|gfactor_summary_kernel|gsweep_summary_kernel' --launch-skip 16 -c 5 -f -o $O/r02j python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/r02j_ncu.log 2>&1
ls -la $O/r02j.ncu-rep
timeout 300 ncu -i $O/r02j.ncu-rep --page source --csv > $O/r02j_source.csv 2> $O/r02j_source.err
timeout 300 ncu -i $O/r02j.ncu-rep --page raw --csv > $O/r02j_raw.csv 2>> $O/r02j_source.err
ls -la $O/r02j*; head -c 600 $O/r02j_source.csv
sz=$(stat -c %s $O/r02j.ncu-rep); if [ "$sz" -gt 20000000 ]; then rm $O/r02j.ncu-rep; fi
