#!/bin/bash
# build, run the GPU sweep script, print per-kernel times
cd /root/repo && python -m celerite2_b200.build 2>&1 | grep -v "^\[build\] /usr" | tail -3
/usr/local/graft/bin/gpurun --timeout 1200 -- 'bash dev/sweep.sh' 2>&1 | grep -v "^{" | tail -8
python - <<'PY'
import json
for T in (8192,16384):
    d=json.load(open('/root/repo/gpurun_out/b_%d.json'%T)); print(T, d['roofline']['per_op_ms']); print({k:v for k,v in d['roofline']['per_kernel_ms'].items()})
PY
