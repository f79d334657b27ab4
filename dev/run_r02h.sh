#!/bin/bash
# one GPU: multi-column factor apply kernel (C2B_MC=1) against the one-column-per-lane kernel, per-kernel times
mkdir -p gpurun_out
O=gpurun_out
export C2B_BENCH_NO_ASSERT=1
run() { tag=$1; shift; env "$@" timeout 400 python bench.py --workload ${W:-c2} --steps 10 --warmup 3 --no-cpu-baseline > $O/r02h_$tag.json 2> $O/r02h_$tag.err; }
run c2_base C2B_MC=0
for u in 7104 14208 18944 23680 28416; do run c2_mc_u$u C2B_MC=1 C2B_UNITS_FACTOR=$u; done
W=c3 run c3_base C2B_MC=0
W=c3 run c3_mc C2B_MC=1
for ch in 2 4; do W=c3 run c3_mc_ch$ch C2B_MC=1 C2B_CHUNKS=$ch; W=c3 run c3_base_ch$ch C2B_MC=0 C2B_CHUNKS=$ch; done
W=c5 run c5_base C2B_MC=0
W=c5 run c5_mc C2B_MC=1
W=c5 run c5_mc_u C2B_MC=1 C2B_UNITS_FACTOR=18944
C2B_MC=1 timeout 1200 python -m pytest tests -m gpu -q --maxfail=15 > $O/r02h_pytest_mc.log 2>&1; echo "pytest rc=$?" >> $O/r02h_pytest_mc.log
tail -4 $O/r02h_pytest_mc.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/r02h_*.json")):
    try: d=json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e: print(f,"ERR"); continue
    pk=d["roofline"].get("per_kernel_ms",{})
    print(f.split("r02h_")[1], round(d["ms_per_step"],4), d.get("parity",{}).get("max_rel_err"), {k.split("/")[1]:v for k,v in pk.items() if k.startswith("factor/")})
PY
