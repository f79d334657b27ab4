#!/bin/bash
mkdir -p gpurun_out
export C2B_BENCH_NO_ASSERT=1
for m in 2 4 6 8; do C2B_MIN_CHUNK_ROWS=$m timeout 200 python bench.py --workload c1 --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r03g_c1_rows$m.json 2> gpurun_out/r03g_c1_rows$m.err; done
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/r03g_*.json")):
    d=json.loads(open(f).read().strip().splitlines()[-1]); print(f.split("/")[-1], round(d["ms_per_step"],4), d["e2e"]["ms_per_step"], d["parity"]["max_rel_err"])
PY
