#!/bin/bash
# one GPU: all 16 compiled widths (width table + every-width test), full suite, c1..c5 with the new defaults
# (prefix scans for K = 1, bulk workspace rows, full-tile copies)
mkdir -p gpurun_out
O=gpurun_out
export C2B_BENCH_NO_ASSERT=1
timeout 1800 python -m pytest tests -m gpu -q --maxfail=15 > $O/r02l_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r02l_pytest.log
tail -5 $O/r02l_pytest.log
timeout 900 python dev/width_table.py > $O/r02l_width.log 2>&1; tail -18 $O/r02l_width.log
for w in c2 c1 c4 c5 c3; do timeout 400 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline > $O/r02l_bench_$w.json 2> $O/r02l_bench_$w.err; done
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/r02l_bench_*.json")):
    try: d=json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e: print(f,"ERR"); continue
    print(f.split("/")[-1], round(d["ms_per_step"],4), "e2e", d["e2e"].get("ms_per_step"), d.get("parity",{}).get("max_rel_err"))
PY
