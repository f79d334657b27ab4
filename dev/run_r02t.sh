#!/bin/bash
# one GPU: programmatic dependent launches: C2B_PDL = 0 (none), 1 (small kernels, default), 2 (every launch)
mkdir -p gpurun_out
O=gpurun_out
export C2B_BENCH_NO_ASSERT=1
run() { tag=$1; shift; env "$@" timeout 300 python bench.py --workload ${W:-c2} --steps 10 --warmup 3 --no-cpu-baseline > $O/r02u_$tag.json 2> $O/r02u_$tag.err; echo "$tag rc=$?"; }
for m in 0 1 2; do
  run c2_pdl$m C2B_PDL=$m
  W=c1 run c1_pdl$m C2B_PDL=$m
  W=c5 run c5_pdl$m C2B_PDL=$m
done
W=c3 run c3_pdl1 C2B_PDL=1
W=c4 run c4_pdl1 C2B_PDL=1
timeout 1500 python -m pytest tests -m gpu -q --maxfail=15 > $O/r02u_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r02u_pytest.log
tail -5 $O/r02u_pytest.log
C2B_PDL=2 timeout 1500 python -m pytest tests -m gpu -q --maxfail=15 > $O/r02u_pytest_pdl2.log 2>&1; echo "pytest rc=$?" >> $O/r02u_pytest_pdl2.log
tail -5 $O/r02u_pytest_pdl2.log
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/r02u_*.json")):
    try: d=json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e: print(f,"ERR"); continue
    print(f.split("/")[-1], round(d["ms_per_step"],4), "eager", d.get("eager_ms_per_step"), "e2e", d["e2e"].get("ms_per_step"), d.get("parity",{}).get("max_rel_err"))
PY
