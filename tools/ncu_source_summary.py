"""Summarise an `ncu --page source --csv` export (SASS view): stall reasons overall, top instructions by samples."""
import csv
import gzip
import io
import sys
import collections

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
raw = gzip.open(path, "rt").read() if path.endswith(".gz") else open(path).read()
lines = raw.splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('"Address"'))
rows = list(csv.DictReader(io.StringIO("\n".join(lines[start:]))))
stalls = [k for k in rows[0] if k.startswith("stall_") and "Not Issued" not in k]
tot = collections.Counter()
for r in rows:
    for k in stalls:
        tot[k] += int(r[k] or 0)
allsmp = sum(int(r["# Samples"] or 0) for r in rows)
print(f"{lines[0][:150]}\ninstructions {len(rows)}  samples {allsmp}")
print("stall reasons:", ", ".join(f"{k[6:]} {v / max(1, sum(tot.values())) * 100:.1f}%" for k, v in tot.most_common(8)))
byop = collections.Counter()
for r in rows:
    op = r["Source"].split()[0] if not r["Source"].strip().startswith("@") else r["Source"].split()[1]
    byop[op.split(".")[0]] += int(r["# Samples"] or 0)
print("samples by opcode:", ", ".join(f"{k} {v / allsmp * 100:.1f}%" for k, v in byop.most_common(14)))
exe = collections.Counter()
for r in rows:
    op = r["Source"].split()[0] if not r["Source"].strip().startswith("@") else r["Source"].split()[1]
    exe[op.split(".")[0]] += int(r["Instructions Executed"] or 0)
te = sum(exe.values())
print("warp-instructions by opcode:", ", ".join(f"{k} {v / te * 100:.1f}%" for k, v in exe.most_common(14)), f"(total {te})")
print(f"\ntop {top} instructions by samples (idx, samples%, main stall, SASS):")
order = sorted(range(len(rows)), key=lambda i: -int(rows[i]["# Samples"] or 0))[:top]
for i in sorted(order):
    r = rows[i]
    ms = max(stalls, key=lambda k: int(r[k] or 0))
    print(f"{i:6d} {int(r['# Samples']) / allsmp * 100:5.2f}%  {ms[6:]:12s} {r['Source'].strip()[:90]}")
