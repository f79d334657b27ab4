"""Static SASS opcode counts of the hot kernels: python dev/sass_summary.py > profiles/r02_sass_summary.txt
(cuobjdump -sass on the per-width objects of the last build; no GPU needed)."""
import collections, os, re, subprocess, sys
HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "celerite2_b200", "_build")
WANT = [("inst_factor_j6.o", "gfactor_apply_kernelILi6ELb0ELb1E", "gfactor_apply_kernel<6, bulk workspace stores>"),
        ("inst_factor_j6.o", "gfactor_apply_kernelILi6ELb1ELb0E", "gfactor_apply_kernel<6, direct workspace stores>"),
        ("inst_factor_j6.o", "gfactor_summary_kernelILi6E", "gfactor_summary_kernel<6>"),
        ("inst_factor_j6.o", "gfactor_rev_kernelILi6ELb1ELb1E", "gfactor_rev_kernel<6, apply, bulk workspace loads>"),
        ("inst_factor_j6.o", "gfactor_rev_kernelILi6ELb0ELb0E", "gfactor_rev_kernel<6, summary>"),
        ("inst_sweep_j6.o", "gsweep_apply_kernelILi6ELb1ELb1E", "gsweep_apply_kernel<6, solve, lower>"),
        ("inst_sweep_j6.o", "gsweep_summary_kernelILi6ELb1E", "gsweep_summary_kernel<6, lower>"),
        ("inst_sweep_j6.o", "gsweep_rev_apply_kernelILi6ELb1ELb1E", "gsweep_rev_apply_kernel<6, solve, lower>"),
        ("inst_factor_j6.o", "pscan_blockI3LftILi6E", "pscan_block<Lft<6>>"),
        ("inst_sweep_j16.o", "csweep_tile_kernelILi16ELb1ELb0ELb1E", "csweep_tile_kernel<16, lower, apply, solve>"),
        ("inst_sweep_j16.o", "csweep_tile_kernelILi16ELb1ELb0ELb0E", "csweep_tile_kernel<16, lower, apply, product>"),
        ("inst_sweep_j16.o", "csweep_rev_tile_kernelILi16ELb1ELb1E", "csweep_rev_tile_kernel<16, solve, lower>"),
        ("inst_sweep_j16.o", "fold_groups_kernelI4AffVILi16E", "fold_groups_kernel<AffV<16>>")]
cache = {}
def functions(obj):
    if obj not in cache:
        out = subprocess.run(["cuobjdump", "-sass", os.path.join(HERE, obj)], capture_output=True, text=True).stdout
        fn, cur = {}, None
        for line in out.splitlines():
            m = re.match(r"\s*Function : (\S+)", line)
            if m:
                cur = m.group(1)
                fn[cur] = []
                continue
            m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(.*?);", line)
            if m and cur:
                fn[cur].append(m.group(1))
        cache[obj] = fn
    return cache[obj]
print("# SASS of the hot kernels (cuobjdump -sass on the objects of the final build, sm_100a, nvcc 12.9): static opcode counts.")
print("# No HMMA / UTCMMA / tcgen05: the path has no contraction large enough for tensor cores (FP64, J <= 16 inside a dependent chain).")
print("# Tile movement is LDGSTS (cp.async, 8- and 16-byte); the J x J workspace rows of factor / factor_rev go through the TMA engine:")
print("# UBLKCP (cp.async.bulk, shared <-> global) with SYNCS (mbarrier) on the load side.  ACQBULK / griddepcontrol (programmatic")
print("# dependent launch) opens every kernel.  FP64 arithmetic = DFMA / DMUL / DADD; MUFU.RCP64H + DFMA chains are the divisions.")
print("# Dynamic counts per kernel: profiles/r02w_ncu_full_summary.txt, per SASS line: profiles/r02j_source_summary.txt.")
print()
for obj, key, label in WANT:
    fn = functions(obj)
    names = [n for n in fn if key in n]
    if not names:
        print("%s: not found (%s)" % (label, key))
        continue
    ins = fn[names[0]]
    ops = collections.Counter()
    for i in ins:
        t = i.split()
        op = t[1] if t[0].startswith("@") and len(t) > 1 else t[0]
        ops[op.split(".")[0]] += 1
    g = lambda *ks: sum(ops.get(k, 0) for k in ks)
    print("%s: %d instructions; FP64 arithmetic %d (DFMA %d DMUL %d DADD %d), LDGSTS %d, UBLKCP %d, SYNCS %d, LDS %d, STS %d, STG %d, LDG %d, SHFL %d, MUFU %d, BAR %d, ACQBULK %d, HMMA %d"
          % (label, len(ins), g("DFMA", "DMUL", "DADD"), ops["DFMA"], ops["DMUL"], ops["DADD"], ops["LDGSTS"], ops["UBLKCP"], ops["SYNCS"],
             ops["LDS"], ops["STS"], ops["STG"], ops["LDG"], ops["SHFL"], ops["MUFU"], ops["BAR"], ops["ACQBULK"], ops["HMMA"]))
    print("   top: " + ", ".join("%s %d" % kv for kv in ops.most_common(12)))
