#!/bin/bash
# one GPU: c4, threads of the fold's top walk
mkdir -p gpurun_out
O=gpurun_out
export C2B_BENCH_NO_ASSERT=1
for tt in 0 128 64 512; do
  C2B_FOLD_TT=$tt timeout 300 python bench.py --workload c4 --steps 10 --warmup 3 --no-cpu-baseline > $O/r02zb_c4_tt$tt.json 2> $O/r02zb_c4_tt$tt.err
done
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/r02zb_*.json")):
    try: d=json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e: print(f,"ERR"); continue
    pk=d.get("roofline",{}).get("per_kernel_ms",{})
    print(f.split("/")[-1], round(d["ms_per_step"],4), d.get("parity",{}).get("max_rel_err"), {k.split("/")[1]:v for k,v in pk.items() if "fold" in k})
PY
