#!/bin/bash
# one GPU: sweep plans sized to the kernels' true occupancy, rev_finalize with four partial sums
mkdir -p gpurun_out
O=gpurun_out
T=${TAG:-r02r}
export C2B_BENCH_NO_ASSERT=1
for w in c2 c1 c5 c3; do timeout 400 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline > $O/${T}_bench_$w.json 2> $O/${T}_bench_$w.err; done
timeout 1800 python -m pytest tests -m gpu -q --maxfail=15 > $O/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> $O/${T}_pytest.log
tail -5 $O/${T}_pytest.log
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/${T}_bench_*.json")):
    try: d=json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e: print(f,"ERR"); continue
    pk=d.get("roofline",{}).get("per_kernel_ms",{})
    print(f.split("/")[-1], round(d["ms_per_step"],4), d.get("parity",{}).get("max_rel_err"), {k.split("/")[0][:9]+"/"+k.split("/")[1][:14]:v for k,v in pk.items() if k.startswith("solve") or "finalize" in k})
PY
