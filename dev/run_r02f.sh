#!/bin/bash
# round-2 check pass (one GPU): full GPU suite, every workload's bench line and reference-arm line with the library as committed
mkdir -p gpurun_out
O=gpurun_out
T=${1:-r02f}
timeout 1800 python -m pytest tests -m gpu -q --maxfail=15 > $O/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> $O/${T}_pytest.log
for w in c2 c1 c4 c5 c3; do
  timeout 500 python bench.py --workload $w --steps 20 --warmup 3 > $O/${T}_bench_$w.json 2> $O/${T}_bench_$w.err; echo "bench $w rc=$?" >> $O/${T}_pytest.log
  timeout 300 python bench.py --impl reference --workload $w --steps 5 --warmup 1 > $O/${T}_ref_$w.json 2> $O/${T}_ref_$w.err
done
python bench.py > $O/${T}_bench_default.json 2> $O/${T}_bench_default.err; echo "bench default rc=$?" >> $O/${T}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/${T}_smoke.log 2>&1; echo "smoke rc=$?" >> $O/${T}_pytest.log
tail -9 $O/${T}_pytest.log
