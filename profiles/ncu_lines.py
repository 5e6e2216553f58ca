"""Summarise an .ncu-rep: key raw metrics + instructions / stall samples per source line.
    python profiles/ncu_lines.py gpurun_out/x.ncu-rep [top_n]"""
import csv
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
h = rows[0]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed.sum",
        "launch__grid_size", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "smsp__inst_executed.sum", "lts__t_bytes.sum", "l1tex__data_bank_conflicts_pipe_lsu.sum",
        "smsp__pcsamp_warps_issue_stalled"]
for i, c in enumerate(h):
    if any(c == w or c.startswith(w) for w in want):
        vals = [r[i] for r in rows[1:]]
        print(f"{c:75s} {vals}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda",
                      "--launch-count", "1"], capture_output=True, text=True).stdout
cur, out, tot, stot = None, [], 0, 0
for r in csv.reader(src.splitlines()):
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if len(r) > 10 and r[0] not in ("", "Line No"):
        try:
            n, s = int(r[7]), int(r[6])
        except ValueError:
            continue
        out.append((n, s, cur, r[0], r[1][:80]))
        tot += n
        stot += s
print("total warp-instructions", tot, "stall samples", stot)
print("--- by instructions")
for n, s, f, l, t in sorted(out, reverse=True)[:top]:
    print(f"{100*n/tot:5.1f}% inst {100*s/max(stot,1):5.1f}% stall  {f}:{l} {t}")
print("--- by stall samples")
for n, s, f, l, t in sorted(out, key=lambda x: -x[1])[:top]:
    print(f"{100*n/tot:5.1f}% inst {100*s/max(stot,1):5.1f}% stall  {f}:{l} {t}")
