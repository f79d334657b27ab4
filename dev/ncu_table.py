"""One line per kernel launch from an ncu raw page (CSV): python dev/ncu_table.py raw.csv [first_kernel_substring] [max_rows]"""
import csv, sys
csv.field_size_limit(10**9)
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
I = {h: i for i, h in enumerate(hdr)}
start = sys.argv[2] if len(sys.argv) > 2 else None
maxn = int(sys.argv[3]) if len(sys.argv) > 3 else 10**9
def val(r, k, scale=True):
    if k not in I or r[I[k]] == "":
        return float("nan")
    v = float(r[I[k]].replace(",", ""))
    if scale:
        u = units[I[k]]
        v *= {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "usecond": 1e-6, "msecond": 1e-3, "nsecond": 1e-9, "second": 1.0, "us": 1e-6, "ms": 1e-3, "ns": 1e-9, "s": 1.0}.get(u, 1.0)
    return v
stall = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
print("kernel | us | DRAM rd MB | DRAM wr MB | regs | grid | warps active % | issue active % | l1tex % | lts % | dram % | fp64 pipe % | top stalls (warps per issue)")
n, on = 0, start is None
for r in rows[2:]:
    name = r[I["Kernel Name"]]
    if not on:
        if start in name:
            on = True
        else:
            continue
    if name.startswith("void at::") or "at::native" in name:
        continue
    short = name.replace("void ", "").replace("c2b::", "").replace("(int)", "").replace("(bool)", "").split("(")[0]
    st = sorted(((val(r, h, False), h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", "")) for h in stall), reverse=True)[:3]
    print("%s | %.1f | %.1f | %.1f | %d | %d | %.1f | %.1f | %.1f | %.1f | %.1f | %.1f | %s" % (
        short[:48], val(r, "gpu__time_duration.sum") * 1e6, val(r, "dram__bytes_read.sum") / 1e6, val(r, "dram__bytes_write.sum") / 1e6,
        val(r, "launch__registers_per_thread", False), val(r, "launch__grid_size", False),
        val(r, "sm__warps_active.avg.pct_of_peak_sustained_active", False), val(r, "smsp__issue_active.avg.pct_of_peak_sustained_active", False),
        val(r, "l1tex__throughput.avg.pct_of_peak_sustained_active", False), val(r, "lts__throughput.avg.pct_of_peak_sustained_elapsed", False),
        val(r, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", False), val(r, "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", False),
        ", ".join("%s %.2f" % (nm, v) for v, nm in st)))
    n += 1
    if n >= maxn:
        break
