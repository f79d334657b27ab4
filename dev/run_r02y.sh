#!/bin/bash
# one GPU: c4 with the reworked fold (slices of J columns, unrolled dot products, head product and state side by side)
mkdir -p gpurun_out
O=gpurun_out
export C2B_BENCH_NO_ASSERT=1
for kb in 0 64 32 8; do
  C2B_FOLD_KB=$kb timeout 300 python bench.py --workload c4 --steps 10 --warmup 3 --no-cpu-baseline > $O/r02z_c4_kb$kb.json 2> $O/r02z_c4_kb$kb.err
done
timeout 1500 python -m pytest tests -m gpu -q --maxfail=15 > $O/r02z_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r02z_pytest.log
tail -4 $O/r02z_pytest.log
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/r02z_*.json")):
    try: d=json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e: print(f,"ERR"); continue
    pk=d.get("roofline",{}).get("per_kernel_ms",{})
    print(f.split("/")[-1], round(d["ms_per_step"],4), d.get("parity",{}).get("max_rel_err"), pk)
PY
