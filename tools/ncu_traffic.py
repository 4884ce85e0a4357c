"""dram__bytes_read/write.sum, duration and pipe utilisation of one kernel from an `ncu --page raw --csv` export (.csv.gz).
usage: python tools/ncu_traffic.py NAME raw.csv.gz [NAME raw.csv.gz ...]  -> one JSON object on stdout"""
import csv
import gzip
import json
import sys

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
KEEP = {"dram__bytes_read.sum": "dram_read_bytes", "dram__bytes_write.sum": "dram_write_bytes",
        "gpu__time_duration.sum": "duration", "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active": "tensor_pipe_active_pct",
        "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_active_pct",
        "lts__t_sector_hit_rate.pct": "l2_hit_pct", "launch__registers_per_thread": "registers",
        "smsp__inst_executed.sum": "warp_instructions"}
out = {}
args = sys.argv[1:]
for name, path in zip(args[0::2], args[1::2]):
    rows = list(csv.reader(gzip.open(path, "rt")))
    rec = {"kernel": rows[2][rows[0].index("Kernel Name")] if "Kernel Name" in rows[0] else name}
    for h, u, v in zip(rows[0], rows[1], rows[2]):
        if h in KEEP:
            x = float(v.replace(",", ""))
            if u in UNIT:
                x *= UNIT[u]
            if h == "gpu__time_duration.sum":
                x *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u, 1.0)
                rec["duration_ms"] = x
                continue
            rec[KEEP[h]] = x
    if "dram_read_bytes" in rec:
        rec["traffic_bytes"] = rec["dram_read_bytes"] + rec.get("dram_write_bytes", 0.0)
    out[name] = rec
print(json.dumps(out, indent=1))
