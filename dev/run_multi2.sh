#!/bin/bash
# multi-GPU pass with the round's final defaults: bash dev/run_multi2.sh N  (inside gpurun --gpus N)
N=${1:-2}
mkdir -p gpurun_out
O=gpurun_out
export C2B_BENCH_NO_ASSERT=1
P=29600
for w in c2 c4 c5 c3; do
  P=$((P+1))
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P bench.py --gpus $N --workload $w --steps 10 --warmup 3 --no-cpu-baseline > $O/r02q_${w}_g$N.json 2> $O/r02q_${w}_g$N.err
  echo "$w x$N rc=$?" >> $O/r02q_status_$N.txt
done
P=$((P+1))
C2B_SPLIT_N=8000000 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P bench.py --gpus $N --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > $O/r02q_c2_8M_g$N.json 2> $O/r02q_c2_8M_g$N.err
cat $O/r02q_status_$N.txt
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/r02q_*_g$N.json")):
    try: d=json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e: print(f,"ERR"); continue
    print(f.split("/")[-1], round(d["ms_per_step"],4), d["value"], d.get("parity",{}).get("max_rel_err"))
PY
