#!/bin/bash
# multi-GPU pass: bash dev/run_multi.sh N  (inside gpurun --gpus N): every workload's own multi-GPU mode + the world-size-N tests
N=${1:-2}
mkdir -p gpurun_out
O=gpurun_out
export C2B_BENCH_NO_ASSERT=1
nvidia-smi topo -m > $O/r02m_topo_$N.txt 2>&1
P=29500
for w in c2 c4 c5 c3; do
  P=$((P+1))
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P bench.py --gpus $N --workload $w --steps 10 --warmup 3 > $O/r02m_${w}_g$N.json 2> $O/r02m_${w}_g$N.err
  echo "$w x$N rc=$?" >> $O/r02m_status_$N.txt
done
P=$((P+1))
C2B_SPLIT_EAGER=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P bench.py --gpus $N --workload c2 --steps 10 --warmup 3 > $O/r02m_c2_eager_g$N.json 2> $O/r02m_c2_eager_g$N.err
P=$((P+1))
C2B_SPLIT_N=8000000 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P bench.py --gpus $N --workload c2 --steps 10 --warmup 3 > $O/r02m_c2_8M_g$N.json 2> $O/r02m_c2_8M_g$N.err
cat $O/r02m_status_$N.txt
