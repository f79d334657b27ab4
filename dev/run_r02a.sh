#!/bin/bash
# round-2 first GPU pass: full GPU test suite, then every workload's bench line
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r02a_gpus.txt 2>&1; nproc >> gpurun_out/r02a_gpus.txt
timeout 1500 python -m pytest tests -m gpu -q --maxfail=12 --deselect tests/test_zz_no_alloc.py > gpurun_out/r02a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02a_pytest.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q >> gpurun_out/r02a_pytest.log 2>&1
for w in c2 c1 c4 c5 c3; do
  C2B_BENCH_NO_ASSERT=1 timeout 600 python bench.py --workload $w --steps 10 --warmup 3 > gpurun_out/r02a_bench_$w.json 2> gpurun_out/r02a_bench_$w.err; echo "bench $w rc=$?" >> gpurun_out/r02a_pytest.log
done
C2B_SCAN=prefix C2B_BENCH_NO_ASSERT=1 timeout 600 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02a_bench_c2_prefix.json 2> gpurun_out/r02a_bench_c2_prefix.err
C2B_SCAN=prefix timeout 1200 python -m pytest tests -m gpu -q --maxfail=12 --deselect tests/test_zz_no_alloc.py > gpurun_out/r02a_pytest_prefix.log 2>&1; echo "pytest(prefix) rc=$?" >> gpurun_out/r02a_pytest.log
for v in "C2B_S_DIRECT=1" "C2B_S_DIRECT=1 C2B_SCAN=prefix" "C2B_SCAN=prefix C2B_TARGET_UNITS=9472" "C2B_SCAN=prefix C2B_TARGET_UNITS=4736" "C2B_SCAN=prefix C2B_TARGET_UNITS=14208"; do
  tag=$(echo "$v" | tr ' =' '__')
  env $v C2B_BENCH_NO_ASSERT=1 timeout 300 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02a_ab_$tag.json 2> gpurun_out/r02a_ab_$tag.err
done
tail -8 gpurun_out/r02a_pytest.log
