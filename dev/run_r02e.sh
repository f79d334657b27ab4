#!/bin/bash
# occupancy A/B: rebuild the library on the box with different resident-CTA targets per kernel family, bench c2 (and c5 / c3)
mkdir -p gpurun_out
O=gpurun_out
export C2B_BENCH_NO_ASSERT=1
B="timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu-baseline"
i=0
run() {  # run "<nvcc flags>" "<env>" workloads...
  flags="$1"; envs="$2"; shift 2
  i=$((i+1))
  C2B_NVCC_FLAGS="$flags" python -m celerite2_b200.build > $O/r02e_build_$i.log 2>&1 || { echo "build $i failed" >> $O/r02e_status.txt; return; }
  echo "$i: flags=[$flags] env=[$envs]" >> $O/r02e_status.txt
  cuobjdump -res-usage celerite2_b200/libcelerite2_b200.so 2>/dev/null | grep -A1 "gfactor_apply_kernelILi6ELb[01]\|gsweep_apply_kernelILi6ELb1ELb1\|gfactor_rev_kernelILi6ELb[01]\|gfactor_summary_kernelILi6\|gsweep_rev_apply_kernelILi6ELb1ELb1\|gsweep_summary_kernelILi6ELb1\|gsweep_rev_summary_kernelILi6ELb1" | grep -v "^--" > $O/r02e_res_$i.txt
  for w in "$@"; do
    env $envs $B --workload $w > $O/r02e_${i}_$w.json 2> $O/r02e_${i}_$w.err
  done
}
run "-DC2B_MINB_SWEEP=8" "" c2
run "-DC2B_MINB_SWEEP=10" "" c2
run "-DC2B_MINB_SWEEP=12" "" c2 c5 c3
run "-DC2B_MINB_SWEEP=10 -DC2B_MINB_FACTOR=8" "C2B_S_DIRECT=1" c2
run "-DC2B_MINB_SWEEP=12 -DC2B_MINB_FACTOR=10" "C2B_S_DIRECT=1" c2 c5
run "-DC2B_MINB_FREV=8 -DC2B_FREV_TILE=2" "" c2
run "-DC2B_MINB_SWEEP=12 -DC2B_MINB_FACTOR=10 -DC2B_MINB_FREV=8 -DC2B_FREV_TILE=2" "C2B_S_DIRECT=1" c2 c5 c3
run "-DC2B_MINB_SWEEP=12 -DC2B_MINB_FACTOR=8 -DC2B_MINB_FREV=10 -DC2B_FREV_TILE=2" "C2B_S_DIRECT=1" c2
run "-DC2B_MINB_SWEEP=12 -DC2B_MINB_FACTOR=10 -DC2B_MINB_FREV=8 -DC2B_FREV_TILE=2" "C2B_S_DIRECT=1 C2B_SCAN=prefix" c2
cat $O/r02e_status.txt
