"""Per-kernel SASS opcode histogram of the built library (cuobjdump -sass): the mnemonics that prove the tcgen05 / TMEM /
bulk-copy path (UTCHMMA, UTCQMMA, LDTM, UBLKCP, UTCBAR, SYNCS) next to the FP32 pipe (FFMA, FMUL, FADD, MUFU) and
local-memory traffic (LDL, STL).   python tools/sass_histogram.py [lib.so] > profiles/rNN_sass_opcodes.md"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "meta_pde_b200", "csrc", "libmetapde_b200.so")
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
KEYS = ["UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UBLKCP", "UBLKPF", "UTMALDG", "SYNCS", "FFMA", "FMUL", "FADD", "MUFU", "F2FP",
        "LDG", "STG", "LDS", "STS", "LDL", "STL", "BAR", "ATOMS", "SHFL"]
cur, hist, order = None, {}, []
for line in txt.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        hist[cur] = collections.Counter()
        order.append(cur)
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and cur:
        hist[cur][m.group(1)] += 1
        hist[cur]["_total"] += 1
demangle = subprocess.run(["c++filt"], input="\n".join(order), capture_output=True, text=True).stdout.splitlines()
print(f"# SASS opcode histogram of `{os.path.relpath(lib, ROOT)}` (`cuobjdump -sass`, static instruction counts)\n")
print("| kernel | total | " + " | ".join(KEYS) + " |")
print("|---|---:|" + "---:|" * len(KEYS))
tot = collections.Counter()
for name, dm in zip(order, demangle):
    h = hist[name]
    short = re.sub(r"\(.*\)$", "", dm).replace("mpde::", "").replace("(int)", "").replace("void ", "")
    print(f"| `{short}` | {h['_total']} | " + " | ".join(str(h[k]) if h[k] else "" for k in KEYS) + " |")
    tot.update(h)
print(f"| **library** | {tot['_total']} | " + " | ".join(str(tot[k]) if tot[k] else "" for k in KEYS) + " |")
