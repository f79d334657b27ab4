"""DRAM traffic per C-ABI call from an ncu raw page (CSV) of one log-likelihood + gradient evaluation:
python dev/traffic_from_raw.py gpurun_out/r02w_c2_raw.csv "<capture note>" > profiles/r02_traffic.json
Kernels are attributed to the operator whose pass kernels surround them (scan / finalize kernels carry no operator name)."""
import csv, json, sys
csv.field_size_limit(10**9)
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[0]
I = {h: i for i, h in enumerate(hdr)}
OPS = [("assemble_rev", "assemble_rev"), ("assemble_kernel", "assemble"), ("gfactor_rev", "factor_rev"), ("gfactor_", "factor"),
       ("gsweep_rev", "solve_lower_rev"), ("gsweep_", "solve_lower"), ("norm_logdet", "norm_logdet")]
cur, started, out, seen = None, False, {}, []
for r in rows[2:]:
    name = r[I["Kernel Name"]]
    if not started:
        if "assemble_kernel" not in name:
            continue  # the capture window may open in the middle of an evaluation
        started = True
    for key, op in OPS:
        if key in name:
            cur = op
            break
    if cur is None or name.startswith("void at::") or "at::native" in name:
        continue
    if cur == "assemble" and "assemble" in seen and "assemble_kernel" in name:
        break  # second evaluation starts
    if cur not in seen:
        seen.append(cur)
    rd = float(r[I["dram__bytes_read.sum"]].replace(",", "")) if r[I["dram__bytes_read.sum"]] else 0.0
    wr = float(r[I["dram__bytes_write.sum"]].replace(",", "")) if r[I["dram__bytes_write.sum"]] else 0.0
    unit_r, unit_w = rows[1][I["dram__bytes_read.sum"]], rows[1][I["dram__bytes_write.sum"]]
    mul = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    out[cur] = out.get(cur, 0.0) + rd * mul.get(unit_r, 1) + wr * mul.get(unit_w, 1)
print(json.dumps({"_capture": sys.argv[2] if len(sys.argv) > 2 else "", "c2": {k: int(v) for k, v in out.items()}}, indent=1))
