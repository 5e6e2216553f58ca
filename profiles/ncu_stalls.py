"""Per-source-line stall reasons and instruction-cache counters of an .ncu-rep.
    python profiles/ncu_stalls.py gpurun_out/x.ncu-rep [top_n]"""
import csv
import subprocess
import sys


def I(x):
    try:
        return int(x)
    except ValueError:
        return 0


rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 20
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
for a, b, c in zip(rows[0], rows[1], rows[2]):
    if a in ("gpu__time_duration.sum", "sm__icc_request_hit_rate.pct", "sm__icc_requests.sum",
             "gcc__cache_requests_type_instruction.sum.pct_of_peak_sustained_elapsed",
             "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
             "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
             "lts__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread"):
        print(f"{a:80s} {c} {b}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda", "--launch-count", "1"],
                     capture_output=True, text=True).stdout
cur, hdr, agg = None, None, {}
keys = ["stall_no_inst", "stall_wait", "stall_long_sb", "stall_short_sb", "stall_barrier", "stall_mio",
        "stall_branch_resolving", "stall_selected", "stall_sleep", "Instructions Executed"]
for r in csv.reader(src.splitlines()):
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if r and r[0] == "Line No":
        hdr = {c: i for i, c in enumerate(r)}
        continue
    if hdr and cur and len(r) >= len(hdr) and r[0].isdigit():
        agg[(cur, int(r[0]))] = [I(r[hdr[k]]) for k in keys] + [r[1][:60]]
tot = [sum(v[i] for v in agg.values()) for i in range(len(keys))]
print("totals", dict(zip(keys, tot)))
print("--- top lines by stall samples (excluding selected): " + ", ".join(keys))
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1][:7]))[:top]:
    print(f"{k[0][:20]:20s} {k[1]:4d} {v[:-1]} {v[-1]}")
