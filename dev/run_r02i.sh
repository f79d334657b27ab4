#!/bin/bash
# one GPU: workspace rows through the TMA engine (C2B_S_BULK=1: bulk stores in factor, bulk loads in factor_rev)
mkdir -p gpurun_out
O=gpurun_out
export C2B_BENCH_NO_ASSERT=1
run() { tag=$1; shift; env "$@" timeout 400 python bench.py --workload ${W:-c2} --steps 10 --warmup 3 --no-cpu-baseline > $O/r02i_$tag.json 2> $O/r02i_$tag.err; }
run c2_base C2B_S_BULK=0
run c2_bulk C2B_S_BULK=1
W=c3 run c3_bulk C2B_S_BULK=1
W=c5 run c5_bulk C2B_S_BULK=1
C2B_S_BULK=1 timeout 1200 python -m pytest tests -m gpu -q --maxfail=15 > $O/r02i_pytest_bulk.log 2>&1; echo "pytest rc=$?" >> $O/r02i_pytest_bulk.log
tail -4 $O/r02i_pytest_bulk.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/r02i_*.json")):
    try: d=json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e: print(f,"ERR"); continue
    pk=d["roofline"].get("per_kernel_ms",{})
    print(f.split("r02i_")[1], round(d["ms_per_step"],4), d.get("parity",{}).get("max_rel_err"), {k:v for k,v in pk.items() if k.startswith("factor")})
PY
