#!/bin/bash
# one GPU: warp-level fold walks for the one-right-hand-side AffV scans (default) against prefix blocks (C2B_SCAN=prefix)
mkdir -p gpurun_out
O=gpurun_out
export C2B_BENCH_NO_ASSERT=1
run() { tag=$1; shift; env "$@" timeout 400 python bench.py --workload ${W:-c2} --steps 10 --warmup 3 --no-cpu-baseline > $O/r02n_$tag.json 2> $O/r02n_$tag.err; }
run c2_auto C2B_X=0
run c2_prefix C2B_SCAN=prefix
W=c5 run c5_auto C2B_X=0
W=c1 run c1_auto C2B_X=0
for n in 125000 250000; do C2B_SPLIT_N=$n timeout 300 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > $O/r02n_c2_N$n.json 2> $O/r02n_c2_N$n.err; done
timeout 1800 python -m pytest tests -m gpu -q --maxfail=15 > $O/r02n_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r02n_pytest.log
tail -5 $O/r02n_pytest.log
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/r02n_*.json")):
    try: d=json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e: print(f,"ERR"); continue
    pk=d.get("roofline",{}).get("per_kernel_ms",{})
    print(f.split("/")[-1], round(d["ms_per_step"],4), d.get("parity",{}).get("max_rel_err"), {k.split("/")[0][:9]+"/"+k.split("/")[1][:14]:v for k,v in pk.items() if "scan" in k or "fold" in k})
PY
