#!/bin/bash
# one GPU: full-tile copy fast path + division fix; with and without the bulk workspace rows; prefix scans; full suite
mkdir -p gpurun_out
O=gpurun_out
T=${TAG:-r02k}
export C2B_BENCH_NO_ASSERT=1
run() { tag=$1; shift; env "$@" timeout 400 python bench.py --workload ${W:-c2} --steps 10 --warmup 3 --no-cpu-baseline > $O/${T}_$tag.json 2> $O/${T}_$tag.err; }
run c2 C2B_S_BULK=0
run c2_bulk C2B_S_BULK=1
run c2_bulk_prefix C2B_S_BULK=1 C2B_SCAN=prefix
W=c3 run c3 C2B_S_BULK=0
W=c3 run c3_bulk C2B_S_BULK=1
W=c5 run c5 C2B_S_BULK=0
W=c5 run c5_bulk C2B_S_BULK=1
W=c1 run c1 C2B_S_BULK=0
W=c4 run c4 C2B_S_BULK=0
timeout 1200 python -m pytest tests -m gpu -q --maxfail=15 > $O/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> $O/${T}_pytest.log
tail -4 $O/${T}_pytest.log
C2B_SCAN=prefix C2B_S_BULK=1 timeout 1200 python -m pytest tests -m gpu -q --maxfail=15 > $O/${T}_pytest_prefix_bulk.log 2>&1; echo "pytest rc=$?" >> $O/${T}_pytest_prefix_bulk.log
tail -4 $O/${T}_pytest_prefix_bulk.log
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/${T}_*.json")):
    try: d=json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e: print(f,"ERR"); continue
    pk=d.get("roofline",{}).get("per_kernel_ms",{})
    print(f.split("/")[-1], round(d["ms_per_step"],4), d.get("parity",{}).get("max_rel_err"), {k.split("/")[0][:6]+"/"+k.split("/")[1][:16]:v for k,v in pk.items() if "scan" not in k})
PY
