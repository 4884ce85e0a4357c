"""Bin an ncu SASS source export by instruction index: samples %, dominant stalls, and marker opcodes per bin."""
import csv, gzip, io, sys, collections
path = sys.argv[1]; B = int(sys.argv[2]) if len(sys.argv) > 2 else 400
raw = gzip.open(path, "rt").read()
lines = raw.splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('"Address"'))
rows = list(csv.DictReader(io.StringIO("\n".join(lines[start:]))))
stalls = [k for k in rows[0] if k.startswith("stall_") and "Not Issued" not in k]
allsmp = sum(int(r["# Samples"] or 0) for r in rows)
for b0 in range(0, len(rows), B):
    rs = rows[b0:b0 + B]
    smp = sum(int(r["# Samples"] or 0) for r in rs)
    st = collections.Counter()
    ops = collections.Counter()
    ex = 0
    for r in rs:
        for k in stalls: st[k[6:]] += int(r[k] or 0)
        toks = r["Source"].split()
        op = toks[1] if toks[0].startswith("@") else toks[0]
        ops[op.split(".")[0]] += 1
        ex += int(r["Instructions Executed"] or 0)
    mark = {k: ops[k] for k in ("LDG", "STG", "LDS", "STS", "MUFU", "LDTM", "SYNCS", "BAR", "UTCHMMA", "LDL", "STL", "ATOMS", "F2FP") if ops[k]}
    top = ", ".join(f"{k} {v / max(1, sum(st.values())) * 100:.0f}%" for k, v in st.most_common(3))
    print(f"{b0:6d} smp {smp / allsmp * 100:5.1f}%  exec {ex / 1e6:8.1f}M  [{top}]  {mark}")
