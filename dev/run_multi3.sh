#!/bin/bash
# bash dev/run_multi3.sh N: c5 and c2 only (the two lines the last library changes moved)
N=${1:-4}
mkdir -p gpurun_out
O=gpurun_out
export C2B_BENCH_NO_ASSERT=1
P=29700
for w in c5 c2; do
  P=$((P+1))
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P bench.py --gpus $N --workload $w --steps 10 --warmup 3 --no-cpu-baseline > $O/r02q_${w}_g$N.json 2> $O/r02q_${w}_g$N.err
  echo "$w x$N rc=$?"
done
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/r02q_*_g$N.json")):
    try: d=json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e: print(f,"ERR"); continue
    print(f.split("/")[-1], round(d["ms_per_step"],4), d["value"], d.get("parity",{}).get("max_rel_err"))
PY
