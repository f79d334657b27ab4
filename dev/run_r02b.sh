#!/bin/bash
# round-2 second GPU pass (one GPU): full suite, A/B switches, side configurations, ncu launch list + full capture, sanitizer
mkdir -p gpurun_out
O=gpurun_out
timeout 1800 python -m pytest tests -m gpu -q --maxfail=15 --durations=15 > $O/r02b_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r02b_pytest.log
C2B_SCAN=fold+prefix timeout 1500 python -m pytest tests -m gpu -q --maxfail=15 --deselect tests/test_zz_no_alloc.py > $O/r02b_pytest_fold.log 2>&1; echo "pytest(fold+prefix) rc=$?" >> $O/r02b_pytest.log
B="timeout 400 python bench.py --steps 10 --warmup 3"
export C2B_BENCH_NO_ASSERT=1
$B --workload c2 > $O/r02b_bench_c2.json 2> $O/r02b_bench_c2.err
for v in "C2B_SCAN=fold" "C2B_SCAN=fold+prefix" "C2B_SCAN=fold+prefix C2B_S_DIRECT=1" "C2B_SCAN=fold C2B_TARGET_UNITS=9472" "C2B_S_DIRECT=1" "C2B_SCAN=prefix" "C2B_S_DIRECT=1 C2B_SCAN=prefix" "C2B_SCAN=prefix C2B_TARGET_UNITS=9472" "C2B_SCAN=prefix C2B_TARGET_UNITS=4736" "C2B_TARGET_UNITS=4736" "C2B_TARGET_UNITS=5920"; do
  tag=$(echo "$v" | tr ' =' '__')
  env $v $B --workload c2 --no-cpu-baseline > $O/r02b_ab_c2_$tag.json 2> $O/r02b_ab_c2_$tag.err
done
for v in "C2B_MIN_CHUNK_ROWS=16" "C2B_MIN_CHUNK_ROWS=32" "C2B_MIN_CHUNK_ROWS=8" "C2B_MIN_CHUNK_ROWS=16 C2B_SCAN=fold" "C2B_SCAN=fold"; do
  tag=$(echo "$v" | tr ' =' '__')
  env $v $B --workload c1 --no-cpu-baseline > $O/r02b_ab_c1_$tag.json 2> $O/r02b_ab_c1_$tag.err
done
$B --workload c1 > $O/r02b_bench_c1.json 2> $O/r02b_bench_c1.err
$B --workload c4 > $O/r02b_bench_c4.json 2> $O/r02b_bench_c4.err
for v in "C2B_SCAN=fold" "C2B_SCAN=fold C2B_CHUNKS=2048" "C2B_SCAN=fold C2B_CHUNKS=512" "C2B_CHUNKS=256" "C2B_CHUNKS=512" "C2B_CHUNKS=2048" "C2B_CHUNKS=512 C2B_SCAN=prefix" "C2B_SCAN=prefix"; do
  tag=$(echo "$v" | tr ' =' '__')
  env $v $B --workload c4 --no-cpu-baseline > $O/r02b_ab_c4_$tag.json 2> $O/r02b_ab_c4_$tag.err
done
$B --workload c5 > $O/r02b_bench_c5.json 2> $O/r02b_bench_c5.err
C2B_SCAN=fold $B --workload c5 --no-cpu-baseline > $O/r02b_ab_c5_fold.json 2> $O/r02b_ab_c5_fold.err
C2B_S_DIRECT=1 $B --workload c5 --no-cpu-baseline > $O/r02b_ab_c5_sdirect.json 2> $O/r02b_ab_c5_sdirect.err
$B --workload c3 > $O/r02b_bench_c3.json 2> $O/r02b_bench_c3.err
C2B_S_DIRECT=1 $B --workload c3 --no-cpu-baseline > $O/r02b_ab_c3_sdirect.json 2> $O/r02b_ab_c3_sdirect.err
$B --workload c3 --batch-per-gpu 14208 --no-cpu-baseline > $O/r02b_ab_c3_wave.json 2> $O/r02b_ab_c3_wave.err
# reference arm lines (the driver's other arm)
for w in c2 c1 c4 c5 c3; do timeout 300 python bench.py --impl reference --workload $w --steps 5 --warmup 1 > $O/r02b_ref_$w.json 2> $O/r02b_ref_$w.err; done
bash dev/sanitize.sh > $O/r02b_sanitize.log 2>&1
tail -6 $O/r02b_pytest.log
