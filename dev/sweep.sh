timeout 600 python -m pytest tests -x -q -m gpu 2>&1 | tail -5
for T in ${TS:-8192 16384 32768}; do
C2B_TARGET_UNITS=$T python bench.py --steps 5 --no-cpu-baseline > gpurun_out/b_$T.json 2>gpurun_out/b_$T.err
python - <<EOF
import json
try:
    d=json.load(open("gpurun_out/b_$T.json"))
    print($T, "ms/step", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["ms_per_step"],3), "wall", round(d["wall_ms_per_step"],3))
    print({k:v for k,v in d["roofline"]["per_kernel_ms"].items()})
except Exception as e:
    print("fail", e); print(open("gpurun_out/b_$T.err").read()[-1500:])
EOF
done
