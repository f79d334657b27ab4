#!/bin/bash
# one GPU: per-width table, the J = 4 workloads after the factor_rev padding revert, full suite
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python dev/width_table.py > $O/r02g_width.log 2>&1
export C2B_BENCH_NO_ASSERT=1
for w in c5 c3 c2; do timeout 400 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline > $O/r02g_bench_$w.json 2> $O/r02g_bench_$w.err; done
for n in 125000 250000 500000; do for m in 64 32 16; do C2B_SPLIT_N=$n C2B_MIN_CHUNK_ROWS=$m timeout 300 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > $O/r02g_c2_N${n}_rows$m.json 2> $O/r02g_c2_N${n}_rows$m.err; done; done
for v in "C2B_CPT=1" "C2B_CPT=2" "C2B_CPT=2 C2B_CHUNKS=2048" "C2B_CPT=2 C2B_CHUNKS=1536" "C2B_CPT=1 C2B_CHUNKS=2048"; do
  tag=$(echo "$v" | tr ' =' '__')
  env $v timeout 300 python bench.py --workload c4 --steps 10 --warmup 3 --no-cpu-baseline > $O/r02g_c4_$tag.json 2> $O/r02g_c4_$tag.err
done
C2B_SCAN=prefix timeout 400 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > $O/r02g_bench_c2_prefix.json 2> $O/r02g_bench_c2_prefix.err
timeout 1800 python -m pytest tests -m gpu -q --maxfail=15 > $O/r02g_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r02g_pytest.log
tail -4 $O/r02g_pytest.log; tail -12 $O/r02g_width.log
