#!/bin/bash
mkdir -p gpurun_out
export C2B_BENCH_NO_ASSERT=1
timeout 900 python -m pytest tests -m gpu -q --maxfail=15 > gpurun_out/r03i_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r03i_pytest.log
tail -5 gpurun_out/r03i_pytest.log
timeout 200 python bench.py --workload c1 --steps 50 --warmup 5 > gpurun_out/r02x_bench_c1.json 2> gpurun_out/r02x_bench_c1.err
timeout 200 python bench.py --impl reference --workload c1 --steps 5 --warmup 1 > gpurun_out/r02x_ref_c1.json 2> gpurun_out/r02x_ref_c1.err
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02x_smoke.log 2>&1; tail -1 gpurun_out/r02x_smoke.log
python - <<PY
import json
d=json.loads(open("gpurun_out/r02x_bench_c1.json").read().strip().splitlines()[-1]); print("c1", d["ms_per_step"], d["e2e"]["ms_per_step"], d["parity"]["max_rel_err"])
PY
