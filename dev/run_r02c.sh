#!/bin/bash
# round-2 third GPU pass: the fold scan (v2: deep prefetch ring, transposed heads) under the full suite and the benches
mkdir -p gpurun_out
O=gpurun_out
export C2B_BENCH_NO_ASSERT=1
C2B_SCAN=fold timeout 1500 python -m pytest tests -m gpu -q --maxfail=15 --deselect tests/test_zz_no_alloc.py > $O/r02c_pytest_fold.log 2>&1; echo "pytest(fold) rc=$?" > $O/r02c_status.txt
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edges.py tests/test_gpu_round2.py -m gpu -q --maxfail=15 > $O/r02c_pytest_tree.log 2>&1; echo "pytest(tree subset) rc=$?" >> $O/r02c_status.txt
B="timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu-baseline"
for w in c2 c1 c4 c5; do
  $B --workload $w > $O/r02c_tree_$w.json 2> $O/r02c_tree_$w.err
  C2B_SCAN=fold $B --workload $w > $O/r02c_fold_$w.json 2> $O/r02c_fold_$w.err
done
for v in "C2B_SCAN=fold C2B_CHUNKS=512" "C2B_SCAN=fold C2B_CHUNKS=2048" "C2B_SCAN=fold C2B_CHUNKS=4096"; do
  tag=$(echo "$v" | tr ' =' '__')
  env $v $B --workload c4 > $O/r02c_ab_c4_$tag.json 2> $O/r02c_ab_c4_$tag.err
done
for v in "C2B_SCAN=fold C2B_TARGET_UNITS=9472" "C2B_SCAN=fold C2B_TARGET_UNITS=8288" "C2B_SCAN=fold C2B_S_DIRECT=1"; do
  tag=$(echo "$v" | tr ' =' '__')
  env $v $B --workload c2 > $O/r02c_ab_c2_$tag.json 2> $O/r02c_ab_c2_$tag.err
done
cat $O/r02c_status.txt
bash dev/run_r02p.sh > $O/r02c_prof.log 2>&1
