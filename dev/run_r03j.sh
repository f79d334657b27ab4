#!/bin/bash
# last check of the round: GPU suite, default bench line, smoke, with the tree as committed
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --maxfail=15 > gpurun_out/r03j_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r03j_pytest.log
tail -4 gpurun_out/r03j_pytest.log
timeout 300 python bench.py > gpurun_out/r03j_bench_default.json 2> gpurun_out/r03j_bench_default.err; echo "bench rc=$?"
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python - <<PY
import json
d=json.loads(open("gpurun_out/r03j_bench_default.json").read().strip().splitlines()[-1]); print(d["ms_per_step"], d["value"], d["e2e"]["value"], d["parity"]["max_rel_err"], d["roofline"]["frac"])
PY
